// ptb_bvh_build.cuh - GPU LBVH build (north_star (1); no reference counterpart):
//   1. per-triangle bounds + centroid, scene bounds            (k_tri_bounds)
//   2. 30-bit Morton codes inside each OBJECT's box, object rank above them (k_morton, k_rank_keys, k_key64)
//   3. LSD radix sort of (code, triangle) pairs, 8 bits/pass   (ptb_sort.cuh: k_ps_hist / k_ps_rowscan / k_ps_scatter)
//   4. Karras 2012 binary radix tree over the sorted codes     (k_karras)
//   5. bottom-up AABB refit, second arrival at a node proceeds (k_refit, atomic flags)
//   6. pack 64 B nodes holding BOTH children's boxes           (k_pack)
// Boxes are rounded outward and inflated by 2^-21 * scene scale (4 x the FMA slab error) so FP32 slab tests never cull a
// triangle whose reference test would accept.
#pragma once
#include "ptb_scene.cuh"
#include "ptb_sort.cuh"

namespace ptb {

struct Aabb { float lo[3], hi[3]; };

__device__ __forceinline__ unsigned f_order(float f) {
    unsigned b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float f_unorder(unsigned o) {
    unsigned b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
    return __uint_as_float(b);
}

// scene_box: 6 ordered uints (min xyz, max xyz), initialised to ~0 / 0
__global__ void k_bounds_init(unsigned *scene_box) {
    if (threadIdx.x < 3) scene_box[threadIdx.x] = 0xffffffffu;
    else if (threadIdx.x < 6) scene_box[threadIdx.x] = 0u;
}

// obj_box: 6 ordered uints per object (min xyz, max xyz), initialised like scene_box.  Triangles are stored object by
// object, so a warp nearly always holds one object: lanes of the same object are folded with match_any and the
// group leader issues the 6 atomics.
__global__ void k_obj_bounds_init(unsigned *obj_box, int n_objs) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n_objs * 6; i += gridDim.x * blockDim.x) obj_box[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
}

__global__ void __launch_bounds__(256)
k_tri_bounds(const double *__restrict__ verts, const int *__restrict__ tri_obj, int n, Aabb *__restrict__ box, float4 *__restrict__ centroid,
             unsigned *scene_box, unsigned *obj_box) {
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    const int n_round = (n + 31) & ~31;
    for (int i0 = blockIdx.x * blockDim.x + threadIdx.x; i0 < n_round; i0 += gridDim.x * blockDim.x) {
        const bool live = i0 < n;
        const int i = live ? i0 : n - 1;
        {   // per-object box
            const double *v = verts + 9 * (size_t)i;
            const int obj = live ? tri_obj[i] : -1 - (int)(threadIdx.x & 31);
            const unsigned peers = __match_any_sync(0xffffffffu, obj);
            float olo[3], ohi[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                olo[k] = __double2float_rd(fmin(v[k], fmin(v[3 + k], v[6 + k])));
                ohi[k] = __double2float_ru(fmax(v[k], fmax(v[3 + k], v[6 + k])));
            }
            const int leader = __ffs(peers) - 1;
            if (peers == 0xffffffffu) {          // the common case: whole warp, one object
#pragma unroll
                for (int k = 0; k < 3; ++k)
                    for (int o = 16; o > 0; o >>= 1) {
                        olo[k] = fminf(olo[k], __shfl_xor_sync(0xffffffffu, olo[k], o));
                        ohi[k] = fmaxf(ohi[k], __shfl_xor_sync(0xffffffffu, ohi[k], o));
                    }
                if ((int)(threadIdx.x & 31) == leader && live)
                    for (int k = 0; k < 3; ++k) { atomicMin(obj_box + 6 * obj + k, f_order(olo[k])); atomicMax(obj_box + 6 * obj + 3 + k, f_order(ohi[k])); }
            } else if (live) {
                for (int k = 0; k < 3; ++k) { atomicMin(obj_box + 6 * obj + k, f_order(olo[k])); atomicMax(obj_box + 6 * obj + 3 + k, f_order(ohi[k])); }
            }
        }
        if (!live) continue;
        const double *v = verts + 9 * (size_t)i;
        Aabb b;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double mn = fmin(v[k], fmin(v[3 + k], v[6 + k])), mx = fmax(v[k], fmax(v[3 + k], v[6 + k]));
            b.lo[k] = __double2float_rd(mn);
            b.hi[k] = __double2float_ru(mx);
            lo[k] = fminf(lo[k], b.lo[k]);
            hi[k] = fmaxf(hi[k], b.hi[k]);
        }
        box[i] = b;
        centroid[i] = make_float4((float)((v[0] + v[3] + v[6]) / 3.0), (float)((v[1] + v[4] + v[7]) / 3.0),
                                  (float)((v[2] + v[5] + v[8]) / 3.0), 0.f);
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(scene_box + k, f_order(lo[k]));
            atomicMax(scene_box + 3 + k, f_order(hi[k]));
        }
    }
}

__device__ __forceinline__ unsigned expand10(unsigned v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

// 30-bit Morton code of the centroid INSIDE THE BOX OF ITS OBJECT (obj_lo / obj_inv: origin and 1024 / extent per
// object): a flat wall spends its 10 + 10 bits on its own two dimensions instead of on the room's.
__global__ void __launch_bounds__(256)
k_morton(const float4 *__restrict__ centroid, const int *__restrict__ tri_obj, int n, const float *__restrict__ obj_lo,
         const float *__restrict__ obj_inv, unsigned *__restrict__ keys, unsigned *__restrict__ morton_orig, int *__restrict__ vals) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 c = centroid[i];
        const int o = tri_obj[i];
        const unsigned x = (unsigned)fminf(fmaxf((c.x - __ldg(obj_lo + 3 * o)) * __ldg(obj_inv + 3 * o), 0.f), 1023.f);
        const unsigned y = (unsigned)fminf(fmaxf((c.y - __ldg(obj_lo + 3 * o + 1)) * __ldg(obj_inv + 3 * o + 1), 0.f), 1023.f);
        const unsigned z = (unsigned)fminf(fmaxf((c.z - __ldg(obj_lo + 3 * o + 2)) * __ldg(obj_inv + 3 * o + 2), 0.f), 1023.f);
        const unsigned m = (expand10(x) << 2) | (expand10(y) << 1) | expand10(z);
        keys[i] = m; morton_orig[i] = m;
        vals[i] = i;
    }
}
// second sort key: the rank of the triangle's object (objects ordered along the Morton curve of their centres)
__global__ void __launch_bounds__(256)
k_rank_keys(const int *__restrict__ vals, const int *__restrict__ tri_obj, const int *__restrict__ rank, int n, unsigned *__restrict__ keys) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) keys[i] = (unsigned)__ldg(rank + tri_obj[vals[i]]);
}
// plane ids as the hash table numbered them (arrival order) -> ranks of the sorted table keys (deterministic)
__global__ void __launch_bounds__(256)
k_remap_pid(unsigned char *__restrict__ tri_pid, int n, const unsigned char *__restrict__ remap) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) tri_pid[i] = __ldg(remap + tri_pid[i]);
}
// 64-bit hierarchy key of the sorted sequence: object rank above the in-object Morton code, so the radix tree splits
// objects first (a top level over whole meshes) and each mesh gets its own spatial subtree.  (A plane id between the
// two - one subtree per face of a mesh - was measured: same frame rate, 15.9 against 15.6 shadow nodes per ray.)
__global__ void __launch_bounds__(256)
k_key64(const int *__restrict__ vals, const int *__restrict__ tri_obj, const int *__restrict__ rank, const unsigned *__restrict__ morton_orig,
        int n, unsigned long long *__restrict__ key64) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int t = vals[i];
        key64[i] = ((unsigned long long)(unsigned)__ldg(rank + tri_obj[t]) << 32) | morton_orig[t];
    }
}

// The sort of the (code, triangle) pairs is the stable LSD radix sort of ptb_sort.cuh (8 bits per pass, warp tiles,
// match_any ranking, per-digit row scans), the same kernels that re-order the live paths between bounces.

// ---------------------------------------------------------------- Karras hierarchy
// delta(i, j) = common-prefix length of (code_i, i) and (code_j, j); -1 outside [0, n)
__device__ __forceinline__ int lbvh_delta(const unsigned long long *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((unsigned)i ^ (unsigned)j);
    return __clzll((long long)(a ^ b));
}

// inner node i in [0, n-1); children encoded: >= 0 inner node, < 0 leaf with sorted position ~c
__global__ void __launch_bounds__(256)
k_karras(const unsigned long long *__restrict__ keys, int n, int2 *__restrict__ child, int *__restrict__ parent_inner, int *__restrict__ parent_leaf) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
        const int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
        const int dmin = lbvh_delta(keys, n, i, i - d);
        int lmax = 2;
        while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int t = lmax >> 1; t >= 1; t >>= 1)
            if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
        const int j = i + l * d;
        const int dnode = lbvh_delta(keys, n, i, j);
        int s = 0;
        int t = l;
        do {
            t = (t + 1) >> 1;
            if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        } while (t > 1);
        const int gamma = i + s * d + min(d, 0);
        const int lo = min(i, j), hi = max(i, j);
        const int c0 = (lo == gamma) ? ~gamma : gamma;
        const int c1 = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
        child[i] = make_int2(c0, c1);
        if (c0 < 0) parent_leaf[gamma] = i; else parent_inner[gamma] = i;
        if (c1 < 0) parent_leaf[gamma + 1] = i; else parent_inner[gamma + 1] = i;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) parent_inner[0] = -1;
}

__device__ __forceinline__ Aabb ld_box_cg(const Aabb *p) {
    Aabb r;
    const float *f = reinterpret_cast<const float *>(p);
#pragma unroll
    for (int k = 0; k < 3; ++k) { r.lo[k] = __ldcg(f + k); r.hi[k] = __ldcg(f + 3 + k); }
    return r;
}

// bottom-up refit: node_box[i] / node_minref[i] for inner nodes
__global__ void __launch_bounds__(256)
k_refit(int n, const int *__restrict__ sorted_tri, const Aabb *__restrict__ tri_box, const int2 *__restrict__ child,
        const int *__restrict__ parent_inner, const int *__restrict__ parent_leaf, Aabb *__restrict__ node_box,
        int *__restrict__ node_minref, unsigned *__restrict__ flags, const unsigned char *__restrict__ tri_pid,
        unsigned char *__restrict__ node_pid) {
    for (int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < n; leaf += gridDim.x * blockDim.x) {
        int node = parent_leaf[leaf];
        while (node >= 0) {
            if (atomicAdd(flags + node, 1u) == 0u) break;      // first arrival waits for the sibling
            __threadfence();
            const int2 c = child[node];
            Aabb a, b; int ra, rb;
            unsigned pa = 0, pb = 0;
            // children written by other threads: read through L2 (L1 may hold a stale neighbouring line)
            if (c.x < 0) { a = tri_box[sorted_tri[~c.x]]; ra = sorted_tri[~c.x]; if (tri_pid) pa = tri_pid[ra]; }
            else { a = ld_box_cg(node_box + c.x); ra = __ldcg(node_minref + c.x); if (tri_pid) pa = __ldcg(node_pid + c.x); }
            if (c.y < 0) { b = tri_box[sorted_tri[~c.y]]; rb = sorted_tri[~c.y]; if (tri_pid) pb = tri_pid[rb]; }
            else { b = ld_box_cg(node_box + c.y); rb = __ldcg(node_minref + c.y); if (tri_pid) pb = __ldcg(node_pid + c.y); }
            if (tri_pid) node_pid[node] = (unsigned char)(pa == pb ? pa : 0u);     // one plane below the node, or mixed
            Aabb r;
#pragma unroll
            for (int k = 0; k < 3; ++k) { r.lo[k] = fminf(a.lo[k], b.lo[k]); r.hi[k] = fmaxf(a.hi[k], b.hi[k]); }
            node_box[node] = r;
            node_minref[node] = min(ra, rb);
            __threadfence();
            node = parent_inner[node];
        }
    }
}

// final 64 B nodes; boxes inflated by `pad`
__global__ void __launch_bounds__(256)
k_pack(int n, const int *__restrict__ sorted_tri, const Aabb *__restrict__ tri_box, const int2 *__restrict__ child,
       const Aabb *__restrict__ node_box, const int *__restrict__ node_minref, float pad, BvhNode *__restrict__ nodes,
       const unsigned char *__restrict__ tri_pid, const unsigned char *__restrict__ node_pid) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
        const int2 c = child[i];
        const Aabb a = c.x < 0 ? tri_box[sorted_tri[~c.x]] : node_box[c.x];
        const Aabb b = c.y < 0 ? tri_box[sorted_tri[~c.y]] : node_box[c.y];
        BvhNode nd;
        nd.lo0x = a.lo[0] - pad; nd.lo0y = a.lo[1] - pad; nd.lo0z = a.lo[2] - pad;
        nd.hi0x = a.hi[0] + pad; nd.hi0y = a.hi[1] + pad; nd.hi0z = a.hi[2] + pad;
        nd.lo1x = b.lo[0] - pad; nd.lo1y = b.lo[1] - pad; nd.lo1z = b.lo[2] - pad;
        nd.hi1x = b.hi[0] + pad; nd.hi1y = b.hi[1] + pad; nd.hi1z = b.hi[2] + pad;
        nd.n0 = c.x < 0 ? sorted_tri[~c.x] : node_minref[c.x];     // lowest reference index below child 0 / 1
        nd.n1 = c.y < 0 ? sorted_tri[~c.y] : node_minref[c.y];
        unsigned p0 = 0, p1 = 0;
        if (tri_pid) { p0 = c.x < 0 ? tri_pid[nd.n0] : node_pid[c.x]; p1 = c.y < 0 ? tri_pid[nd.n1] : node_pid[c.y]; }
        nd.c0 = c.x;                                               // child 1 is position(child 0) + 1 (k_karras)
        nd.meta = (int)((c.y < 0 ? 0x80000000u : 0u) | (p1 << 8) | p0);
        nodes[i] = nd;
    }
}

// Scene extent, max |coordinate| and a tri_obj range check in one pass; every block writes 8 doubles
// (min xyz, max xyz, max abs, bad-object flag) that the host folds.
// Breadth-first top of a finished hierarchy (TopNode, ptb_scene.cuh): one warp walks the tree level by level, 32 queue
// entries at a time; an inner child gets the next top slot while slots last, else its link points into the full array.
__global__ void __launch_bounds__(32)
k_build_top(const BvhNode *__restrict__ nodes, int k_max, TopNode *__restrict__ top, int *__restrict__ top_count) {
    __shared__ int q[PTB_TOP_OFF];
    const int lane = threadIdx.x;
    if (lane == 0) q[0] = 0;
    __syncwarp();
    int head = 0, tail = 1;
    while (head < tail) {
        const int batch = min(32, tail - head);
        const bool on = lane < batch;
        BvhNode nd;
        int c0 = -1, c1 = -1;
        if (on) {
            nd = nodes[q[head + lane]];
            c0 = nd.c0;
            const int i1 = (c0 ^ (c0 >> 31)) + 1;
            c1 = i1 ^ (nd.meta >> 31);
        }
        const int in0 = (on && c0 >= 0) ? 1 : 0, in1 = (on && c1 >= 0) ? 1 : 0;
        int incl = in0 + in1;
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        const int total = __shfl_sync(0xffffffffu, incl, 31);
        const int base = tail + incl - (in0 + in1);
        if (on) {
            int l0 = c0, l1 = c1;
            if (in0) { if (base < k_max) { q[base] = c0; l0 = base; } else l0 = c0 + PTB_TOP_OFF; }
            if (in1) { const int s1 = base + in0; if (s1 < k_max) { q[s1] = c1; l1 = s1; } else l1 = c1 + PTB_TOP_OFF; }
            TopNode t;
            t.a = make_float4(nd.lo0x, nd.lo0y, nd.lo0z, nd.hi0x);
            t.b = make_float4(nd.hi0y, nd.hi0z, nd.lo1x, nd.lo1y);
            t.c = make_float4(nd.lo1z, nd.hi1x, nd.hi1y, nd.hi1z);
            t.d = make_int4(l0, l1, nd.n0 | ((nd.meta & 0xff) << 24), nd.n1 | (((nd.meta >> 8) & 0xff) << 24));
            top[head + lane] = t;
        }
        __syncwarp();
        tail = min(k_max, tail + total);
        head += batch;
    }
    if (lane == 0) *top_count = tail;
}

// Scene arrays that are already on the device (PTB_SCENE_ARRAYS_ON_DEVICE): the two small tables the host otherwise
// derives from its copy - the light's vertices in FP32 (utils.py:41-46 reads them per light sample) and the first
// triangle of every object (the lowest-occluder rule of the shadow walk) - come from these kernels instead.
__global__ void __launch_bounds__(256)
k_light_verts32(const double *__restrict__ verts, int n_tri_obj, int n_light_tri, float4 *__restrict__ out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < 3 * n_light_tri; i += gridDim.x * blockDim.x) {
        const double *p = verts + 9 * (size_t)n_tri_obj + 3 * (size_t)i;
        out[i] = make_float4((float)p[0], (float)p[1], (float)p[2], 0.f);
    }
}
__global__ void __launch_bounds__(256)
k_obj_first(const int *__restrict__ tri_obj, int n, int n_obj, int *__restrict__ first) {      // first[] preset to n
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int o = tri_obj[i];
        // triangles are listed object by object (ptb_scene_desc): only the first of a run needs the atomic
        if ((unsigned)o <= (unsigned)n_obj && (i == 0 || tri_obj[i - 1] != o)) atomicMin(first + o, i);
    }
}
__global__ void k_obj_first_fill(int *first, int n_obj, int n) {      // empty objects take their successor's first triangle
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        first[n_obj + 1] = n;
        for (int k = n_obj; k >= 0; --k) if (first[k] == n) first[k] = first[k + 1];
    }
}

__global__ void __launch_bounds__(256)
k_scene_extent(const double *__restrict__ verts, size_t n_pts, const int *__restrict__ tri_obj, int n_tri, int n_obj, double *__restrict__ out) {
    double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300}, mx = 0.0, bad = 0.0;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n_pts; i += (size_t)gridDim.x * blockDim.x) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double v = verts[3 * i + k];
            lo[k] = fmin(lo[k], v); hi[k] = fmax(hi[k], v); mx = fmax(mx, fabs(v));
        }
        if (i < (size_t)n_tri && (tri_obj[i] < 0 || tri_obj[i] > n_obj)) bad = 1.0;
    }
    __shared__ double s_red[8][8];
    double v[8] = {lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], mx, bad};
#pragma unroll
    for (int k = 0; k < 8; ++k)
        for (int o = 16; o > 0; o >>= 1) {
            const double w = __shfl_xor_sync(0xffffffffu, v[k], o);
            v[k] = k < 3 ? fmin(v[k], w) : fmax(v[k], w);
        }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 8; ++k) s_red[threadIdx.x >> 5][k] = v[k];
    __syncthreads();
    if (threadIdx.x < 8) {
        double r = s_red[0][threadIdx.x];
        for (int w = 1; w < 8; ++w) r = threadIdx.x < 3 ? fmin(r, s_red[w][threadIdx.x]) : fmax(r, s_red[w][threadIdx.x]);
        out[8 * (size_t)blockIdx.x + threadIdx.x] = r;
    }
}

// FP32 traversal records + shading normals from the uploaded FP64 vertices (large scenes: no host loop).
// Same operations in the same order as the host's make_trirec (ptb_api.cu), one IEEE rounding each, so a
// triangle gets the same record whichever side builds it.
__device__ __forceinline__ double hdot(const double *a, const double *b) {   // host dot_h: (a0 b0 + a1 b1) + a2 b2
    return xadd(xadd(xmul(a[0], b[0]), xmul(a[1], b[1])), xmul(a[2], b[2]));
}
__device__ __forceinline__ void hcross(const double *a, const double *b, double *c) {
    c[0] = xsub(xmul(a[1], b[2]), xmul(a[2], b[1])); c[1] = xsub(xmul(a[2], b[0]), xmul(a[0], b[2])); c[2] = xsub(xmul(a[0], b[1]), xmul(a[1], b[0]));
}
__global__ void __launch_bounds__(256)
k_make_records(const double *__restrict__ verts, const double *__restrict__ normals, const int *__restrict__ tri_obj, int n, double scale,
               TriRec *__restrict__ tri, float4 *__restrict__ tri_sn) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double *A = verts + 9 * (size_t)i, *B = A + 3, *C = A + 6;
        double ab[3], cb[3], nn[3], e1[3], e2[3], N[3], mu[3], mv[3], mw[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { ab[k] = xsub(A[k], B[k]); cb[k] = xsub(C[k], B[k]); e1[k] = xsub(B[k], A[k]); e2[k] = xsub(C[k], A[k]); }
        hcross(ab, cb, nn);
        const double nl = xsqrt(hdot(nn, nn));
#pragma unroll
        for (int k = 0; k < 3; ++k) nn[k] = xdiv(nn[k], nl);
        hcross(e1, e2, N);
        const double inv = xdiv(1.0, hdot(N, N));
        hcross(e2, N, mu);
        hcross(N, e1, mv);
#pragma unroll
        for (int k = 0; k < 3; ++k) { mu[k] = xmul(mu[k], inv); mv[k] = xmul(mv[k], inv); mw[k] = xadd(mu[k], mv[k]); }
        const double g = fmax(xsqrt(hdot(mu, mu)), fmax(xsqrt(hdot(mv, mv)), xsqrt(hdot(mw, mw))));
        const int obj = tri_obj[i];
        TriRec r;
        r.q0 = make_float4((float)nn[0], (float)nn[1], (float)nn[2], (float)hdot(nn, A));
        r.q1 = make_float4((float)mu[0], (float)mu[1], (float)mu[2], (float)(-hdot(A, mu)));
        r.q2 = make_float4((float)mv[0], (float)mv[1], (float)mv[2], (float)(-hdot(A, mv)));
        r.q3 = make_float4((float)xmul(xmul(xmul(32.0, 5.9604644775390625e-08), scale), g), 0.f, __int_as_float(obj), __int_as_float(i));
        tri[i] = r;
        const double *sn = normals + 3 * (size_t)i;
        tri_sn[i] = make_float4((float)sn[0], (float)sn[1], (float)sn[2], __int_as_float(obj));
    }
}

// Supporting-plane ids.  Two triangles get the same id when their unit normals (sign-normalised) agree to q_n per
// component and their plane offsets to q_d: with q_n * (scene diameter) and q_d far below sqrt(1e-5) * 1e-5 = 3.2e-8 a
// line that starts on one of them cannot meet the other one outside the reference's self-hit guard (it would need
// |t| > sqrt(1e-5) with |n.d| > 1e-5, main.py:47-48, 100, utils.py:117), so skipping it is exact, not a heuristic.
// table: PTB_PLANE_IDS + 1 keys (0 = empty) filled with atomicCAS; table[PTB_PLANE_IDS + 1] counts overflows - more
// planes than ids - and then the scene goes without (the caller checks): which planes got an id would depend on the
// race.  The NUMBERING depends on it too, but only equality of ids is ever used.
__device__ __forceinline__ unsigned long long pid_mix(unsigned long long h, long long v) {
    h ^= (unsigned long long)v + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
    h ^= h >> 30; h *= 0xbf58476d1ce4e5b9ull; h ^= h >> 27; h *= 0x94d049bb133111ebull; h ^= h >> 31;
    return h;
}
__global__ void __launch_bounds__(256)
k_plane_ids(const double *__restrict__ verts, int n, double q_n, double q_d, unsigned long long *table, unsigned char *__restrict__ tri_pid) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const double *A = verts + 9 * (size_t)i, *B = A + 3, *C = A + 6;
        double e1[3], e2[3], nn[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { e1[k] = xsub(B[k], A[k]); e2[k] = xsub(C[k], A[k]); }
        hcross(e1, e2, nn);
        const double nl = xsqrt(hdot(nn, nn));
        unsigned id = 0;
        if (nl > 0.0 && nl == nl) {
            int big = 0;
            if (fabs(nn[1]) > fabs(nn[big])) big = 1;
            if (fabs(nn[2]) > fabs(nn[big])) big = 2;
            const double sg = (nn[big] < 0.0 ? -1.0 : 1.0) / nl;
            const double d0 = hdot(nn, A) * sg;
            unsigned long long key = pid_mix(pid_mix(pid_mix(pid_mix(0x1234567ull, llrint(nn[0] * sg / q_n)), llrint(nn[1] * sg / q_n)),
                                                     llrint(nn[2] * sg / q_n)), llrint(d0 / q_d));
            if (key == 0ull) key = 1ull;
            unsigned slot = (unsigned)(key % PTB_PLANE_IDS);
            for (int probe = 0; probe < PTB_PLANE_IDS; ++probe) {
                unsigned long long cur = __ldcg(table + slot);
                if (cur == 0ull) cur = atomicCAS(table + slot, 0ull, key);
                if (cur == 0ull || cur == key) { id = slot + 1u; break; }
                slot = slot + 1u == PTB_PLANE_IDS ? 0u : slot + 1u;
            }
            if (id == 0u) atomicAdd(table + PTB_PLANE_IDS + 1, 1ull);
        }
        tri_pid[i] = (unsigned char)id;
    }
}

// leaf-ordered triangle records: slot s holds reference triangle sorted_tri[s]
__global__ void __launch_bounds__(256)
k_gather_tris(int n, const int *__restrict__ sorted_tri, const TriRec *__restrict__ src, TriRec *__restrict__ dst) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = src[sorted_tri[i]];
}

}  // namespace ptb
