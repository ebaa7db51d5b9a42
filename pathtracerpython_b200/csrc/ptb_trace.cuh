// ptb_trace.cuh - brute-force (scene-in-shared-memory) traversal kernels for small scenes.
//
// closest-hit  = intersect_objects (main.py:83-122): every triangle of objects + light, LINE query,
//                key = squared distance > 1e-5, first minimum.
// shadow       = the occlusion loops of compute_shadow_rays (main.py:42-59): object triangles only,
//                1e-5 <= d2 < light_d2; the 3rd light sample also reports the lowest-index occluder's
//                object (the leaked loop variable `obj`, main.py:42, 70).
// All threads of a warp walk the same triangle at the same time (no divergence), triangle records
// are read from shared memory as broadcast 128-bit loads, grids are persistent (SM-count multiples)
// and read the live-path count from device memory so no host synchronisation sits between bounces.
#pragma once
#include "ptb_queue.cuh"

namespace ptb {

#define PTB_BRUTE_MAX_TRI 512     // 32 KB of TriRec in shared memory
#define PTB_BRUTE_AUTO_TRI 320    // PTB_ACCEL_AUTO: brute force (plane list) up to here, LBVH above
#define PTB_TRACE_THREADS 256

__device__ __forceinline__ void stage_tris(TriRec *s_tri, const TriRec *g_tri, int n) {
    const float4 *src = reinterpret_cast<const float4 *>(g_tri);
    float4 *dst = reinterpret_cast<float4 *>(s_tri);
    for (int i = threadIdx.x; i < n * 4; i += blockDim.x) dst[i] = __ldg(src + i);
    __syncthreads();
}

// ---------------------------------------------------------------- FP32 closest hit
template <bool TRACK>
__device__ __forceinline__ Hit32 closest_brute32(const TriRec *s_tri, int n_tri, f3 o, f3 d, float dt_scale, bool &uncertain) {
    Hit32 h; h.t = 0.f; h.abs_t = 3.0e38f; h.slot = -1;
    Unc32 uc; uc.init();
#pragma unroll 4
    for (int i = 0; i < n_tri; ++i) {
        float t, dt = 0.f; bool edge = false;
        const float eps = TRACK ? s_tri[i].q3.x : 0.f;
        const bool hit = tri_test32<TRACK>(s_tri[i].q0, s_tri[i].q1, s_tri[i].q2, eps, o, d, dt_scale, t, edge, dt);
        closest_update32<TRACK>(hit, t, edge, dt, i, h, uc);
    }
    uncertain = closest_uncertain32<TRACK>(h, uc);
    return h;
}

template <bool TRACK>
__global__ void __launch_bounds__(PTB_TRACE_THREADS)
k_closest_brute32(DScene sc, const float4 *__restrict__ qo, const float4 *__restrict__ qd, const unsigned *__restrict__ d_count,
                  float *__restrict__ hit_t, int *__restrict__ hit_tri, unsigned *unc_count, int *unc_list, float dt_scale) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TriRec *s_tri = reinterpret_cast<TriRec *>(smem_raw);
    stage_tris(s_tri, sc.tri, sc.n_tri);
    const unsigned count = *d_count;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float4 a = qo[i], b = qd[i];
        bool unc;
        const Hit32 h = closest_brute32<TRACK>(s_tri, sc.n_tri, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), dt_scale, unc);
        hit_t[i] = h.t;
        hit_tri[i] = h.slot;     // brute force: slot == reference triangle index
        if (TRACK && unc) unc_list[atomicAdd(unc_count, 1u)] = (int)i;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// ---------------------------------------------------------------- FP64 closest hit (exact)
__device__ __forceinline__ Hit64 closest_brute64(const DScene &sc, d3 pr, d3 dir_in) {
    const d3 vr = xdiv3(dir_in, xnorm(dir_in));      // utils.py:110
    Hit64 h; h.tri = -1; h.d2 = 0.0; h.P = mk3<double>(0, 0, 0);
    for (int i = 0; i < sc.n_tri; ++i) {
        d3 P;
        const bool hit = tri_test64(sc.verts64 + 9 * (size_t)i, pr, vr, P);
        closest_update64(hit, P, pr, i, h);
    }
    return h;
}

__global__ void __launch_bounds__(128)
k_closest_brute64(DScene sc, const double4 *__restrict__ qo, const double4 *__restrict__ qd, const unsigned *__restrict__ d_count,
                  double *__restrict__ hit_t, int *__restrict__ hit_tri) {
    const unsigned count = *d_count;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const double4 a = qo[i], b = qd[i];
        const d3 pr = mk3(a.x, a.y, a.z), dir = mk3(b.x, b.y, b.z);
        const Hit64 h = closest_brute64(sc, pr, dir);
        hit_tri[i] = h.tri;
        // k_shade64 recomputes P = pr + vr*t from the triangle with the same roundings, so only the
        // triangle id is needed downstream; the signed ray parameter is stored for callers.
        hit_t[i] = h.tri >= 0 ? xdot(xsub3(h.P, pr), xdiv3(dir, xnorm(dir))) : 0.0;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// FP64 re-trace of the FP32 primary rays whose discrete decisions were inside their error bound.
// The ray is rebuilt from the pixel index exactly as the reference builds it (utils.py:55-69).
__device__ __forceinline__ double linspace_at(double start, double stop, int num, int i) {
    if (num == 1) return start;
    if (i == num - 1) return stop;
    const double step = xdiv(xsub(stop, start), (double)(num - 1));
    return xadd(xmul((double)i, step), start);
}
__device__ __forceinline__ void primary_ray64(const DScene &sc, int W, int H, int r, d3 &o, d3 &d) {
    const int ix = r / H, iy = r - ix * H;
    const double x = linspace_at(sc.ortho[0], sc.ortho[2], W, ix);
    const double y = linspace_at(sc.ortho[1], sc.ortho[3], H, iy);
    o = mk3(sc.eye[0], sc.eye[1], sc.eye[2]);
    d = mk3(xsub(x, o.x), xsub(y, o.y), xsub(0.0, o.z));
}

}  // namespace ptb
