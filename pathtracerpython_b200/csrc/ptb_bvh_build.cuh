// ptb_bvh_build.cuh - GPU LBVH build (north_star (1); no reference counterpart):
//   1. per-triangle bounds + centroid, scene bounds            (k_tri_bounds)
//   2. 30-bit Morton codes, 10 bits per axis                   (k_morton)
//   3. LSD radix sort of (code, triangle) pairs, 8 bits/pass   (k_rs_hist / k_rs_scan / k_rs_scatter)
//   4. Karras 2012 binary radix tree over the sorted codes     (k_karras)
//   5. bottom-up AABB refit, second arrival at a node proceeds (k_refit, atomic flags)
//   6. pack 64 B nodes holding BOTH children's boxes           (k_pack)
// Boxes are rounded outward and inflated by 2^-18 * scene scale so FP32 slab tests never cull a
// triangle whose reference test would accept.
#pragma once
#include "ptb_scene.cuh"

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

__global__ void __launch_bounds__(256)
k_tri_bounds(const double *__restrict__ verts, int n, Aabb *__restrict__ box, float4 *__restrict__ centroid, unsigned *scene_box) {
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
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

__global__ void __launch_bounds__(256)
k_morton(const float4 *__restrict__ centroid, int n, const unsigned *__restrict__ scene_box, unsigned *__restrict__ keys, int *__restrict__ vals) {
    const float lx = f_unorder(scene_box[0]), ly = f_unorder(scene_box[1]), lz = f_unorder(scene_box[2]);
    const float ex = fmaxf(f_unorder(scene_box[3]) - lx, 1e-30f), ey = fmaxf(f_unorder(scene_box[4]) - ly, 1e-30f),
                ez = fmaxf(f_unorder(scene_box[5]) - lz, 1e-30f);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 c = centroid[i];
        const unsigned x = (unsigned)fminf(fmaxf((c.x - lx) / ex * 1024.f, 0.f), 1023.f);
        const unsigned y = (unsigned)fminf(fmaxf((c.y - ly) / ey * 1024.f, 0.f), 1023.f);
        const unsigned z = (unsigned)fminf(fmaxf((c.z - lz) / ez * 1024.f, 0.f), 1023.f);
        keys[i] = (expand10(x) << 2) | (expand10(y) << 1) | expand10(z);
        vals[i] = i;
    }
}

// ---------------------------------------------------------------- radix sort (stable LSD, 8 bits per pass)
// Work is split into warp tiles of RS_TILE consecutive elements; a warp walks its tile 32 elements at a
// time, so order inside a tile is preserved and no block-level synchronisation is needed.
#define RS_TILE 2048
#define RS_WARPS 8

__global__ void __launch_bounds__(RS_WARPS * 32)
k_rs_hist(const unsigned *__restrict__ keys, int n, int shift, int n_tiles, unsigned *__restrict__ ghist /* [256][n_tiles] */) {
    __shared__ unsigned s_h[RS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * RS_WARPS + warp;
    for (int b = lane; b < 256; b += 32) s_h[warp][b] = 0;
    __syncwarp();
    if (tile < n_tiles) {
        const int beg = tile * RS_TILE, end = min(n, beg + RS_TILE);
        for (int i = beg + lane; i < end; i += 32) atomicAdd(&s_h[warp][(keys[i] >> shift) & 255u], 1u);
        __syncwarp();
        for (int b = lane; b < 256; b += 32) ghist[(size_t)b * n_tiles + tile] = s_h[warp][b];
    }
}

// exclusive scan of ghist (bin-major) in place; single block
__global__ void __launch_bounds__(1024) k_rs_scan(unsigned *ghist, size_t total) {
    __shared__ unsigned s_part[1024];
    const size_t per = (total + 1023) / 1024;
    const size_t beg = per * threadIdx.x, end = beg + per < total ? beg + per : total;
    unsigned sum = 0;
    for (size_t i = beg; i < end; ++i) sum += ghist[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    // Hillis-Steele inclusive scan of the 1024 partials
    for (int o = 1; o < 1024; o <<= 1) {
        unsigned v = threadIdx.x >= (unsigned)o ? s_part[threadIdx.x - o] : 0u;
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = threadIdx.x ? s_part[threadIdx.x - 1] : 0u;
    for (size_t i = beg; i < end; ++i) { const unsigned v = ghist[i]; ghist[i] = run; run += v; }
}

__global__ void __launch_bounds__(RS_WARPS * 32)
k_rs_scatter(const unsigned *__restrict__ keys, const int *__restrict__ vals, int n, int shift, int n_tiles,
             const unsigned *__restrict__ ghist, unsigned *__restrict__ keys_out, int *__restrict__ vals_out) {
    __shared__ unsigned s_off[RS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * RS_WARPS + warp;
    if (tile >= n_tiles) return;
    for (int b = lane; b < 256; b += 32) s_off[warp][b] = ghist[(size_t)b * n_tiles + tile];
    __syncwarp();
    const int beg = tile * RS_TILE, end = min(n, beg + RS_TILE);
    for (int base = beg; base < end; base += 32) {
        const int i = base + lane;
        const bool live = i < end;
        const unsigned k = live ? keys[i] : 0u;
        const int v = live ? vals[i] : 0;
        const unsigned d = live ? ((k >> shift) & 255u) : 256u + lane;   // dead lanes match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rank = __popc(peers & ((1u << lane) - 1u));
        unsigned pos = 0;
        if (live) pos = s_off[warp][d] + rank;
        __syncwarp();
        if (live && rank == 0) s_off[warp][d] += __popc(peers);
        __syncwarp();
        if (live) { keys_out[pos] = k; vals_out[pos] = v; }
    }
}

// ---------------------------------------------------------------- Karras hierarchy
// delta(i, j) = common-prefix length of (code_i, i) and (code_j, j); -1 outside [0, n)
__device__ __forceinline__ int lbvh_delta(const unsigned *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((unsigned)i ^ (unsigned)j);
    return __clz(a ^ b);
}

// inner node i in [0, n-1); children encoded: >= 0 inner node, < 0 leaf with sorted position ~c
__global__ void __launch_bounds__(256)
k_karras(const unsigned *__restrict__ keys, int n, int2 *__restrict__ child, int *__restrict__ parent_inner, int *__restrict__ parent_leaf) {
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
        int *__restrict__ node_minref, unsigned *__restrict__ flags) {
    for (int leaf = blockIdx.x * blockDim.x + threadIdx.x; leaf < n; leaf += gridDim.x * blockDim.x) {
        int node = parent_leaf[leaf];
        while (node >= 0) {
            if (atomicAdd(flags + node, 1u) == 0u) break;      // first arrival waits for the sibling
            __threadfence();
            const int2 c = child[node];
            Aabb a, b; int ra, rb;
            // children written by other threads: read through L2 (L1 may hold a stale neighbouring line)
            if (c.x < 0) { a = tri_box[sorted_tri[~c.x]]; ra = sorted_tri[~c.x]; }
            else { a = ld_box_cg(node_box + c.x); ra = __ldcg(node_minref + c.x); }
            if (c.y < 0) { b = tri_box[sorted_tri[~c.y]]; rb = sorted_tri[~c.y]; }
            else { b = ld_box_cg(node_box + c.y); rb = __ldcg(node_minref + c.y); }
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
       const Aabb *__restrict__ node_box, const int *__restrict__ node_minref, float pad, BvhNode *__restrict__ nodes) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n - 1; i += gridDim.x * blockDim.x) {
        const int2 c = child[i];
        const Aabb a = c.x < 0 ? tri_box[sorted_tri[~c.x]] : node_box[c.x];
        const Aabb b = c.y < 0 ? tri_box[sorted_tri[~c.y]] : node_box[c.y];
        BvhNode nd;
        nd.lo0x = a.lo[0] - pad; nd.lo0y = a.lo[1] - pad; nd.lo0z = a.lo[2] - pad;
        nd.hi0x = a.hi[0] + pad; nd.hi0y = a.hi[1] + pad; nd.hi0z = a.hi[2] + pad;
        nd.lo1x = b.lo[0] - pad; nd.lo1y = b.lo[1] - pad; nd.lo1z = b.lo[2] - pad;
        nd.hi1x = b.hi[0] + pad; nd.hi1y = b.hi[1] + pad; nd.hi1z = b.hi[2] + pad;
        nd.c0 = c.x; nd.c1 = c.y;
        nd.n0 = c.x < 0 ? sorted_tri[~c.x] : node_minref[c.x];     // lowest reference index below child 0 / 1
        nd.n1 = c.y < 0 ? sorted_tri[~c.y] : node_minref[c.y];
        nodes[i] = nd;
    }
}

// leaf-ordered triangle records: slot s holds reference triangle sorted_tri[s]
__global__ void __launch_bounds__(256)
k_gather_tris(int n, const int *__restrict__ sorted_tri, const TriRec *__restrict__ src, TriRec *__restrict__ dst) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) dst[i] = src[sorted_tri[i]];
}

}  // namespace ptb
