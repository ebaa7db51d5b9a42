// ptb_scene.cuh - device-resident scene (immutable after ptb_scene_create) and its HBM layout.
#pragma once
#include "ptb_common.cuh"

namespace ptb {

// FP32 traversal record, 64 B per triangle (4 x float4, one 128 B line holds two triangles):
//   q0 = (n.x, n.y, n.z, d0)     unit plane normal as utils.py:112-114 builds it, d0 = n . v1
//   q1 = (mu.x, mu.y, mu.z, cu)  u(P) = mu . P + cu   barycentric weight of v2
//   q2 = (mv.x, mv.y, mv.z, cv)  v(P) = mv . P + cv   barycentric weight of v3
//   q3 = (eps, -, obj_id bits, -) eps = conservative FP32 error scale of u, v (see isect)
// in_triangle (utils.py:72-91) accepts exactly the points with u > 0, v > 0, u + v < 1.
struct TriRec { float4 q0, q1, q2, q3; };

// BVH2 node, 64 B: both children's boxes inline so one fetch decides both descents.
//   lo0.xyz hi0.xyz = child 0 box, lo1.xyz hi1.xyz = child 1 box,
//   c0 = child 0 (>= 0 inner node, < 0 : leaf, triangle slot = ~c).  In a Karras tree the children of a node are the
//   positions gamma and gamma + 1 (ptb_bvh_build.cuh), so child 1 is "index of child 0, plus one" and only its
//   leaf / inner flag needs storing: meta bit 31 = child 1 is a leaf.  The room that frees holds the PLANE ID of either
//   child, meta bits 0..7 / 8..15: k > 0 when every triangle below the child lies in supporting plane k, else 0.
//   A line that starts on a triangle of plane k meets every other triangle of that plane at t = 0, inside the
//   self-hit guard (main.py:47-48, 100: d^2 > 1e-5), so such subtrees are skipped without a box test (ptb_bvh_dyn.cuh).
//   n0 / n1 = lowest reference index below the child.
struct BvhNode {
    float lo0x, lo0y, lo0z, hi0x, hi0y, hi0z;
    float lo1x, lo1y, lo1z, hi1x, hi1y, hi1z;
    int c0, meta, n0, n1;
};
#define PTB_PLANE_IDS 255        // plane ids 1..255; 0 = none / mixed

// The TOP of a hierarchy - its first PTB_TOP_K inner nodes in breadth-first order - as a tile every block of the
// traversal kernels stages in SHARED memory (one TMA bulk copy at block start, ptb_bvh_dyn.cuh): a 64 B gather out of
// shared memory runs at 340-470 Gfetch/s against 200 out of L1 (ptb_gather_peak_smem / ptb_gather_peak), and every walk
// starts with these nodes.  64 B like BvhNode but with BOTH child links explicit (breadth-first numbering breaks the
// "child 1 = position of child 0 + 1" rule): d = (c0, c1, n0 | pid0 << 24, n1 | pid1 << 24).  A link < 0 is a leaf
// (~slot), 0 <= link < PTB_TOP_OFF another top node, link >= PTB_TOP_OFF the node link - PTB_TOP_OFF of the full array.
struct TopNode { float4 a, b, c; int4 d; };
#ifndef PTB_TOP_K
#define PTB_TOP_K 0              // nodes staged per block (0: off); 256 = 16 KB
#endif
#define PTB_TOP_OFF 4096         // first index of the "full array" range in a walk's node variable

struct DScene {
    // --- FP32 product path
    const TriRec *tri;        // [n_tri] in reference order (brute force) or BVH leaf order
    const int *tri_id;        // [n_tri] slot -> reference triangle index (identity for brute force)
    const float4 *tri_sn;     // [n_tri] by reference index: stored normal xyz (scene_reader.py:5-8), w = obj id bits
    const float *mat32;       // [n_obj+1][8] r g b ka kd ks kt n (row n_obj = light, zeros)
    const float4 *light_v32;  // [n_light_tri][3]
    // --- FP64 exact path (reference formulas on the parsed vertices)
    const double *verts64;    // [n_tri][9]
    const double *normals64;  // [n_tri][3]
    const double *mat64;      // [n_obj+1][8]
    const double *light_cdf;  // [n_light_tri+1] running sums of light_obj.areas (utils.py:31-35)
    const int *tri_obj;       // [n_tri]
    const unsigned *light_thr; // [n_light_tri] integer form of the light CDF: pick = largest i with thr[i] <= (x >> 8)
    const int *light_bucket;   // [2^k + 1] bucket b = m >> light_shift -> largest i with thr[i] <= b << light_shift: the
    int light_shift;           //   pick lies in [bucket[b], bucket[b+1]], a 1-2 step search instead of log2(n) dependent loads
    const int *obj_lobe_thr;  // [n_obj+1] diffuse lobe iff (x >> 8) < thr  (exact integer form of main.py:240-241)
    const int *obj_pow_int;   // [n_obj+1] exponent n if it is a small non-negative integer, else -1
    const int *obj_first;     // [n_obj+2] first reference triangle of every object (light = n_obj), then n_tri
    // --- acceleration structure (accel == PTB_ACCEL_LBVH)
    const BvhNode *nodes;     // [n_tri-1]; root = 0
    const BvhNode *nodes_occ; // BVH over the occlusion set (object triangles only)
    const TriRec *tri_occ;    // leaf-ordered records of the occlusion set
    const int *tri_id_occ;
    const unsigned char *tri_pid; // [n_tri] by reference index: supporting-plane id of the triangle (0: none); null when unused
    const TopNode *top, *top_occ; // [top_n], [top_occ_n] breadth-first top of either tree (null: not built)
    int top_n, top_occ_n;
    int n_tri, n_tri_obj, n_obj, n_light_tri;
    int accel;
    double eye[3], ortho[4], ambient, light_color[3];
    double light_total;       // sum(areas) (utils.py:30)
    float bb_lo[3], bb_inv[3]; // scene box origin and 256 / extent: Morton sort keys of path origins (ptb_sort.cuh)
    unsigned long long *counters;   // device ptb_counters (first 8 u64) + queue counters
};

// offsets into DScene::counters (u64 units)
enum { CTR_CLOSEST = 0, CTR_SHADOW = 1, CTR_SHADED = 2, CTR_SAMPLES = 3, CTR_RETRACED = 4, CTR_LAUNCHES = 5,
       CTR_NODES = 6, CTR_TRIS = 7, CTR_NODES_SH = 8, CTR_TRIS_SH = 9, CTR_COUNT = 10 };

}  // namespace ptb
