// ptb_planes.cuh - plane-grouped brute-force traversal (small scenes, FP32 product path).
//
// Meshes of walls and boxes are made of coplanar triangle groups (every Cornell quad is 2 triangles in
// one plane).  The ray/plane part of utils.py:116-120 (n.d, n.o, t, P) depends only on the plane, so it is
// evaluated once per group; only the inside test (utils.py:72-91) runs per triangle.  A warp vote skips
// a group when no lane can use it (plane farther than the lane's current best / outside the shadow
// range / sample already occluded).  For shadow rays the 3 light samples of a path share the origin, so
// n.o is evaluated once per plane and path.  Decisions are the same FP32 decisions as the per-triangle
// kernels up to rounding; bit-exact primary ids still come from the tracked kernel + FP64 re-trace.
#pragma once
#include "ptb_shadow.cuh"

namespace ptb {

struct PlaneRec {      // 96 B = 6 x float4: one supporting plane + up to two of its triangles
    float4 nd;         // unit normal, d0
    float4 a1, a2;     // triangle A barycentric gradients (TriRec q1, q2)
    float4 b1, b2;     // triangle B (or a never-inside dummy: u == -1)
    int ref_a, ref_b;  // reference triangle indices (ref_b = -1 for the dummy)
    int n_obj;         // how many of (A, B) belong to the occlusion set: 0, 1 (A only) or 2
    int kind;          // 1: A = (a1,a2,a3), B = (a1,a3,a1+a3-a2) form a parallelogram -> one test for both
};
struct PlaneScene {    // device pointer, small; staged into shared memory by every block
    const PlaneRec *planes;   // groups with n_obj > 0 first (n_occ of them), light-only groups last
    int n_planes, n_occ;
};

__host__ __device__ inline size_t plane_smem_bytes(int n_planes) { return (size_t)n_planes * sizeof(PlaneRec) + 16; }
// The plane list of a small scene (<= 48 KB) is staged into shared memory once per persistent block: ONE TMA bulk copy
// (cp.async.bulk.shared::cluster.global, SASS UBLKCP) issued by thread 0, its completion counted in bytes on an mbarrier
// that lives in the 16 spare bytes behind the records (plane_smem_bytes); every thread waits on phase 0 of the barrier.
// PTB_PLANES_TMA=0: the plain coalesced copy of round 1.
#ifndef PTB_PLANES_TMA
#define PTB_PLANES_TMA 1
#endif
__device__ __forceinline__ const PlaneRec *stage_planes(unsigned char *smem, const PlaneScene &ps) {
#if PTB_PLANES_TMA
    const unsigned bytes = (unsigned)ps.n_planes * (unsigned)sizeof(PlaneRec);          // a multiple of 32
    const unsigned dst = (unsigned)__cvta_generic_to_shared(smem), mb = dst + bytes;
    if (bytes) {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(ps.planes), "r"(bytes), "r"(mb) : "memory");
        }
        __syncthreads();                    // the barrier is initialised before anybody polls it
        unsigned done = 0;
        do {
            asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                         : "=r"(done) : "r"(mb) : "memory");
        } while (!done);
    }
    return reinterpret_cast<const PlaneRec *>(smem);
#else
    const float4 *a = reinterpret_cast<const float4 *>(ps.planes);
    float4 *da = reinterpret_cast<float4 *>(smem);
    for (int i = threadIdx.x; i < ps.n_planes * 6; i += blockDim.x) da[i] = __ldg(a + i);
    __syncthreads();
    return reinterpret_cast<const PlaneRec *>(smem);
#endif
}

// Parallelogram pair: with (u, v) the barycentrics of triangle A, P = a1 + (u+v)(a2-a1) + v(a4-a1), so P is in
// the quad iff 0 < v < 1 and 0 < u+v < 1; u > 0 selects A, u < 0 selects B, u == 0 is the shared diagonal
// (an edge of both triangles: miss, A-Q15).  6 FMA + 4 ALU instead of 2 x (6 FMA + 4 ALU).
__device__ __forceinline__ bool para_hit(const float4 q1, const float4 q2, float px, float py, float pz, float &u) {
    u = fmaf(q1.z, pz, fmaf(q1.y, py, fmaf(q1.x, px, q1.w)));
    const float v = fmaf(q2.z, pz, fmaf(q2.y, py, fmaf(q2.x, px, q2.w)));
    const float vh = v - 0.5f, sh = u + vh;
    return (fmaxf(fabsf(vh), fabsf(sh)) < 0.5f) && (u != 0.0f);
}

__device__ __forceinline__ float bary_min(const float4 q1, const float4 q2, float px, float py, float pz) {
    const float u = fmaf(q1.z, pz, fmaf(q1.y, py, fmaf(q1.x, px, q1.w)));
    const float v = fmaf(q2.z, pz, fmaf(q2.y, py, fmaf(q2.x, px, q2.w)));
    return fminf(fminf(u, v), 1.0f - u - v);
}


// ---------------------------------------------------------------- closest hit
// Warp-collective (votes): all 32 lanes must call it together; lanes with live == false never keep a group alive.
// PHYS (physical mode): a RAY query, only hits in front of the origin count (t > tmin).
// VOTE = false: no warp votes, straight-line predicated code the compiler can unroll and pipeline - for incoherent warps
// (every bounce after the first) some lane always needs every plane, so the votes only cost instructions and latency.
template <bool PHYS = false, bool VOTE = true>
__device__ __forceinline__ void closest_planes32(const PlaneRec *planes, int n_planes, const float4 a, const float4 b, bool live,
                                                 float &best_t, int &best_ref) {
    float best_a = live ? 3.0e38f : 0.f;
    best_t = 0.f;
    best_ref = -1;
#pragma unroll 2
    for (int g = 0; g < n_planes; ++g) {
        const float4 nd = planes[g].nd;
        const float dn = fmaf(nd.z, b.z, fmaf(nd.y, b.y, nd.x * b.x));
        const float on = nd.w - fmaf(nd.z, a.z, fmaf(nd.y, a.y, nd.x * a.x));
        const float t = on * rcp_approx(dn);
        const float at = fabsf(t);
        const bool ok = (fabsf(dn) > PTB_ZERO_F) && ((PHYS ? t : at) > PTB_TMIN_F) && (at < best_a);
        if (VOTE && !__any_sync(0xffffffffu, ok)) continue;
        const float px = fmaf(t, b.x, a.x), py = fmaf(t, b.y, a.y), pz = fmaf(t, b.z, a.z);
        if (planes[g].kind) {                   // warp-uniform
            float u;
            if (para_hit(planes[g].a1, planes[g].a2, px, py, pz, u) && ok) {
                best_a = at; best_t = t; best_ref = u > 0.0f ? planes[g].ref_a : planes[g].ref_b;
            }
        } else {
            const float ma = bary_min(planes[g].a1, planes[g].a2, px, py, pz);
            const float mb = bary_min(planes[g].b1, planes[g].b2, px, py, pz);
            // A precedes B in reference order; both lie in the same plane (same |t|), so A wins a tie
            const bool ha = ok && ma > 0.0f, hb = ok && mb > 0.0f;
            if (ha || hb) { best_a = at; best_t = t; best_ref = ha ? planes[g].ref_a : planes[g].ref_b; }
        }
    }
}

template <bool PHYS>
__global__ void __launch_bounds__(PTB_TRACE_THREADS)
k_closest_planes32(DScene sc, PlaneScene ps, const float4 *__restrict__ qo, const float4 *__restrict__ qd,
                   const unsigned *__restrict__ d_count, float *__restrict__ hit_t, int *__restrict__ hit_tri) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PlaneRec *planes = stage_planes(smem_raw, ps);
    const unsigned count = *d_count;
    const unsigned stride = gridDim.x * blockDim.x;
    // whole warps iterate together so the votes are legal
    const unsigned count_w = (count + 31u) & ~31u;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count_w; i += stride) {
        const bool live = i < count;
        const float4 a = live ? qo[i] : make_float4(0, 0, 0, 0), b = live ? qd[i] : make_float4(1, 0, 0, 0);
        float best_t; int best_ref;
        closest_planes32<PHYS>(planes, ps.n_planes, a, b, live, best_t, best_ref);
        if (live) {
            hit_t[i] = best_t;
            hit_tri[i] = best_ref;
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// ---------------------------------------------------------------- occlusion of the 3 light samples of one path
// Warp-collective.  pk.xyz = shaded point, d0..d2 = unit directions (w = squared distance to the light
// sample).  occ0 / occ1: sample occluded; leak_ref: lowest reference index among the occluders of sample
// 2, 0x7fffffff if it is unoccluded (-1 for dead lanes).
template <bool PHYS = false, bool VOTE = true>
__device__ __forceinline__ void shadow_planes32(const PlaneRec *planes, int n_occ, const float4 pk, const float4 d0, const float4 d1,
                                                const float4 d2, bool live, bool &occ0, bool &occ1, int &leak_ref) {
    occ0 = !live; occ1 = !live;            // dead lanes never keep a group alive
    leak_ref = live ? 0x7fffffff : -1;     // lowest reference index among occluders of sample 2
    for (int g = 0; g < n_occ; ++g) {
        const float4 nd = planes[g].nd;
        const float on = nd.w - fmaf(nd.z, pk.z, fmaf(nd.y, pk.y, nd.x * pk.x));
        const float dn0 = fmaf(nd.z, d0.z, fmaf(nd.y, d0.y, nd.x * d0.x));
        const float dn1 = fmaf(nd.z, d1.z, fmaf(nd.y, d1.y, nd.x * d1.x));
        const float dn2 = fmaf(nd.z, d2.z, fmaf(nd.y, d2.y, nd.x * d2.x));
        const float t0 = on * rcp_approx(dn0), t1 = on * rcp_approx(dn1), t2 = on * rcp_approx(dn2);
        const float s0 = t0 * t0, s1 = t1 * t1, s2 = t2 * t2;
        // main.py:47-51: 1e-5 <= d2 < light_squared_distance ; utils.py:117: |n.d| > 1e-5
        // PHYS: ray semantics, occluders behind the shaded point do not count
        const bool ok0 = !occ0 && (fabsf(dn0) > PTB_ZERO_F) && (s0 >= PTB_ZERO_F) && (s0 < d0.w) && (!PHYS || t0 > 0.f);
        const bool ok1 = !occ1 && (fabsf(dn1) > PTB_ZERO_F) && (s1 >= PTB_ZERO_F) && (s1 < d1.w) && (!PHYS || t1 > 0.f);
        const bool ok2 = (planes[g].ref_a < leak_ref) && (fabsf(dn2) > PTB_ZERO_F) && (s2 >= PTB_ZERO_F) && (s2 < d2.w) && (!PHYS || t2 > 0.f);
        if (VOTE && !__any_sync(0xffffffffu, ok0 || ok1 || ok2)) continue;
        const float p0x = fmaf(t0, d0.x, pk.x), p0y = fmaf(t0, d0.y, pk.y), p0z = fmaf(t0, d0.z, pk.z);
        const float p1x = fmaf(t1, d1.x, pk.x), p1y = fmaf(t1, d1.y, pk.y), p1z = fmaf(t1, d1.z, pk.z);
        const float p2x = fmaf(t2, d2.x, pk.x), p2y = fmaf(t2, d2.y, pk.y), p2z = fmaf(t2, d2.z, pk.z);
        const float4 a1 = planes[g].a1, a2 = planes[g].a2;
        if (planes[g].kind) {                  // warp-uniform: parallelogram pair, both triangles in the occlusion set
            float u0, u1, u2;
            const bool h0 = para_hit(a1, a2, p0x, p0y, p0z, u0);
            const bool h1 = para_hit(a1, a2, p1x, p1y, p1z, u1);
            const bool h2 = para_hit(a1, a2, p2x, p2y, p2z, u2);
            occ0 = occ0 || (ok0 && h0);
            occ1 = occ1 || (ok1 && h1);
            if (ok2 && h2) leak_ref = min(leak_ref, u2 > 0.0f ? planes[g].ref_a : planes[g].ref_b);
        } else {
            const bool ha0 = bary_min(a1, a2, p0x, p0y, p0z) > 0.0f;
            const bool ha1 = bary_min(a1, a2, p1x, p1y, p1z) > 0.0f;
            const bool ha2 = bary_min(a1, a2, p2x, p2y, p2z) > 0.0f;
            bool hb0 = false, hb1 = false, hb2 = false;
            if (planes[g].n_obj > 1) {             // warp-uniform
                const float4 b1 = planes[g].b1, b2 = planes[g].b2;
                hb0 = bary_min(b1, b2, p0x, p0y, p0z) > 0.0f;
                hb1 = bary_min(b1, b2, p1x, p1y, p1z) > 0.0f;
                hb2 = bary_min(b1, b2, p2x, p2y, p2z) > 0.0f;
            }
            occ0 = occ0 || (ok0 && (ha0 || hb0));
            occ1 = occ1 || (ok1 && (ha1 || hb1));
            if (ok2 && (ha2 || hb2)) leak_ref = min(leak_ref, ha2 ? planes[g].ref_a : planes[g].ref_b);
        }
    }
}

// ---------------------------------------------------------------- shadow rays + NEE accumulation
__global__ void __launch_bounds__(PTB_TRACE_THREADS)
k_shadow_planes32(DScene sc, PlaneScene ps, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d,
                  const int2 *__restrict__ sh_aux, const unsigned *__restrict__ d_count, float4 *__restrict__ L,
                  int cap, int bounce, RecPtrs rec) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PlaneRec *planes = stage_planes(smem_raw, ps);
    const unsigned count = *d_count;
    const unsigned stride = gridDim.x * blockDim.x;
    const unsigned count_w = (count + 31u) & ~31u;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float ambient = (float)sc.ambient;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count_w; j += stride) {
        const bool live = j < count;
        const unsigned jj = live ? j : 0u;
        const float4 pk = sh_p[jj];
        const float4 d0 = sh_d[jj], d1 = sh_d[(size_t)cap + jj], d2 = sh_d[2 * (size_t)cap + jj];
        bool occ0, occ1; int leak_ref;
        shadow_planes32(planes, ps.n_occ, pk, d0, d1, d2, live, occ0, occ1, leak_ref);
        if (!live) continue;
        const int2 aux = sh_aux[j];
        const float4 sn = __ldg(sc.tri_sn + aux.y);
        const int obj = __float_as_int(sn.w);
        const bool occ2 = leak_ref != 0x7fffffff;
        const int leak = occ2 ? __float_as_int(__ldg(sc.tri_sn + leak_ref).w) : sc.n_obj - 1;
        float dot_sum = 0.f;
        if (!occ0) dot_sum += fmaf(d0.z, sn.z, fmaf(d0.y, sn.y, d0.x * sn.x));
        if (!occ1) dot_sum += fmaf(d1.z, sn.z, fmaf(d1.y, sn.y, d1.x * sn.x));
        if (!occ2) dot_sum += fmaf(d2.z, sn.z, fmaf(d2.y, sn.y, d2.x * sn.x));
        const float dotv = dot_sum / 3.0f;
        const float *m = sc.mat32 + 8 * obj, *lm = sc.mat32 + 8 * leak;
        const float ka = __ldg(m + 3);
        const float c0 = __ldg(m + 0) * ka * ambient + lc0 * __ldg(lm + 0) * dotv;
        const float c1 = __ldg(m + 1) * ka * ambient + lc1 * __ldg(lm + 1) * dotv;
        const float c2 = __ldg(m + 2) * ka * ambient + lc2 * __ldg(lm + 2) * dotv;
        float4 l = L[aux.x];
        l.x = fmaf(c0, pk.w, l.x); l.y = fmaf(c1, pk.w, l.y); l.z = fmaf(c2, pk.w, l.z);
        L[aux.x] = l;
        if (rec.any) {
            const size_t ri = rec.index(aux.x, bounce);
            if (rec.occ) rec.occ[ri] = (int8_t)((occ0 ? 1 : 0) | (occ1 ? 2 : 0) | (occ2 ? 4 : 0));
            if (rec.leak) rec.leak[ri] = (int8_t)leak;
            if (rec.contrib) { rec.contrib[3 * ri] = c0 * pk.w; rec.contrib[3 * ri + 1] = c1 * pk.w; rec.contrib[3 * ri + 2] = c2 * pk.w; }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * count);
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)count);
    }
}

}  // namespace ptb
