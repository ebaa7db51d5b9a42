// ptb_shade.cuh - camera-ray generation, shading / bounce kernels with fused stream compaction.
//
// k_gen      : make_screen_pts + make_rays (utils.py:55-69, main.py:191) for a batch of paths.
// k_shade32  : per live path: miss / light hit / shaded (main.py:212-221); for shaded paths it
//              draws the 3 light samples (utils.py:21-46) and writes the shadow work item, then takes
//              the bounce step (main.py:236-268) and writes the next ray.  Survivors are compacted
//              with a warp ballot + block prefix sum + one atomic per block, so the next bounce's
//              queues are dense (north_star (2)).
// k_shade64  : the same in the reference's FP64 operation order (mirrors oracle/pt_oracle.c).
#pragma once
#include "ptb_shadow.cuh"

namespace ptb {

#ifndef PTB_SHADE_THREADS
#define PTB_SHADE_THREADS 256
#endif

struct BatchInfo {
    PixelMap pm;
    int npix;        // pixels in the call's set
    int S;           // samples in this batch
    int s_first;     // absolute index of the batch's first sample
    int bounce;
    uint32_t seed;
    const int *s_off;   // device word added to every sample index (ptb_trace_paths_at: a captured CUDA graph renders the NEXT
                        // samples at every replay without a new launch sequence), or null
    __device__ __forceinline__ uint32_t sample_index(int s_local) const {
        return (uint32_t)(s_first + s_local) + (s_off ? (uint32_t)__ldg(s_off) : 0u);
    }
};

// ---------------------------------------------------------------- camera rays
template <typename T>
__global__ void __launch_bounds__(256)
k_gen(DScene sc, BatchInfo bi, typename VecT<T>::v4 *__restrict__ qo, typename VecT<T>::v4 *__restrict__ qd,
      typename VecT<T>::v4 *__restrict__ L, unsigned *count, unsigned *unc_count, int n_counts) {
    const unsigned N = (unsigned)bi.npix * (unsigned)bi.S;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < N; j += gridDim.x * blockDim.x) {
        const int p = (int)(j % (unsigned)bi.npix);
        const int r = bi.pm.ray_index(p);
        d3 o, d;
        primary_ray64(sc, bi.pm.W, bi.pm.H, r, o, d);
        if (sizeof(T) == 4) d = xdiv3(d, xnorm(d));     // FP32 queues carry unit directions
        qo[j] = mk4((T)o.x, (T)o.y, (T)o.z, (T)1);
        qd[j] = mk4((T)d.x, (T)d.y, (T)d.z, id_pack((T)0, (int)j));
        L[j] = mk4((T)0, (T)0, (T)0, (T)0);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        count[0] = N;
        for (int i = 1; i < n_counts; ++i) count[i] = 0;
        unc_count[0] = 0;
        atomicAdd(sc.counters + CTR_SAMPLES, (unsigned long long)N);
    }
}

// block-wide exclusive scan of a predicate; returns this thread's output slot (valid when alive) and
// reserves `total` slots in *d_out_count with one atomic.
__device__ __forceinline__ unsigned compact_slot(bool alive, unsigned *d_out_count) {
    __shared__ unsigned s_warp[PTB_SHADE_THREADS / 32];
    __shared__ unsigned s_base;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned ballot = __ballot_sync(0xffffffffu, alive);
    if (lane == 0) s_warp[warp] = __popc(ballot);
    __syncthreads();
    unsigned before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < PTB_SHADE_THREADS / 32; ++w) {
        const unsigned c = s_warp[w];
        if (w < (int)warp) before += c;
        total += c;
    }
    if (threadIdx.x == 0) s_base = total ? atomicAdd(d_out_count, total) : 0u;
    __syncthreads();
    const unsigned slot = s_base + before + __popc(ballot & ((1u << lane) - 1u));
    __syncthreads();   // s_warp / s_base are reused by the next tile
    return slot;
}

// the same per WARP: ballot, one atomic by the leader, shuffle - no barrier, no shared memory; the order in which the warps
// of a launch get their runs of slots is not deterministic (callers: nothing downstream may depend on the queue order)
__device__ __forceinline__ unsigned compact_slot_warp(bool alive, unsigned *d_out_count) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned ballot = __ballot_sync(0xffffffffu, alive);
    unsigned base = 0;
    if (lane == 0 && ballot) base = atomicAdd(d_out_count, (unsigned)__popc(ballot));
    base = __shfl_sync(0xffffffffu, base, 0);
    return base + __popc(ballot & ((1u << lane) - 1u));
}
// light triangle pick: first i with cum[i] <= x < cum[i+1] (utils.py:28-39), FP64 so the discrete
// choice is identical to the reference's.
__device__ __forceinline__ int pick_light(const double *cdf, int n, double x) {
    int lo = 0, hi = n;               // invariant: cdf[lo] <= x (cdf[0] = 0), answer in [lo, hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(cdf + mid) <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// integer form of the same pick: thr[i] = smallest 24-bit m whose product total * m * 2^-24 reaches
// cum[i] (built on the host in FP64), so the choice is bit-identical and needs no FP64 on the device
__device__ __forceinline__ int pick_light_int(const DScene &sc, unsigned m) {
    const unsigned *thr = sc.light_thr;
    if (sc.n_light_tri == 2) return m >= __ldg(thr + 1) ? 1 : 0;
    const unsigned b = m >> sc.light_shift;
    int lo = __ldg(sc.light_bucket + b), hi = __ldg(sc.light_bucket + b + 1) + 1;      // thr[lo] <= m, answer in [lo, hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (__ldg(thr + mid) <= m) lo = mid; else hi = mid;
    }
    return lo;
}
// x^n for a small non-negative integer n (negative bases keep their sign like pow with an integral exponent)
__device__ __forceinline__ float pow_small_int(float x, int n) {
    float r = 1.0f;
    while (n) { if (n & 1) r *= x; x *= x; n >>= 1; }
    return r;
}

// The 3 light samples of one shaded point (main.py:30-40, utils.py:21-46): unit direction to the sampled
// light point in xyz, squared distance in w.  Philox dims 4s .. 4s+3.
// One light sample from one Philox quad (utils.py:21-46): area-weighted triangle pick, then the point
// (c0 v1 + c1 v2 + c2 v3) / (c0 + c1 + c2).  Also the device side of the known-answer entry ptb_kat.
__device__ __forceinline__ f3 light_point32(const DScene &sc, const uint4 w, int &li) {
    li = pick_light_int(sc, w.x >> 8);
    const float c0 = u01<float>(w.y), c1 = u01<float>(w.z), c2 = u01<float>(w.w);
    const float inv = rcp_approx((c0 + c1) + c2);
    const float wa = c0 * inv, wb = c1 * inv, wc = c2 * inv;
    const float4 v1 = __ldg(sc.light_v32 + 3 * li), v2 = __ldg(sc.light_v32 + 3 * li + 1), v3 = __ldg(sc.light_v32 + 3 * li + 2);
    return mk3(fmaf(wc, v3.x, fmaf(wb, v2.x, wa * v1.x)), fmaf(wc, v3.y, fmaf(wb, v2.y, wa * v1.y)),
               fmaf(wc, v3.z, fmaf(wb, v2.z, wa * v1.z)));
}

__device__ __forceinline__ void light_samples32(const DScene &sc, uint32_t seed, uint32_t r, uint32_t sample, uint32_t bounce,
                                                const f3 P, float4 (&out)[3]) {
#pragma unroll
    for (int s = 0; s < 3; ++s) {
        const uint4 w = rng4(seed, r, sample, bounce, (uint32_t)s);
        int li;
        const f3 lp = light_point32(sc, w, li);
        const f3 v = mk3(lp.x - P.x, lp.y - P.y, lp.z - P.z);
        const float d2 = dot(v, v);
        const float rn = rsqrtf(d2);
        out[s] = make_float4(v.x * rn, v.y * rn, v.z * rn, d2);
    }
}

// rotate (main.py:148-162) about (0,1,0) by the angle whose cosine / sine are cy / sy
__device__ __forceinline__ f3 rotate_y32(float cy, float sy, f3 v) {
    return mk3(fmaf(cy, v.x, sy * v.z), v.y, fmaf(-sy, v.x, cy * v.z));
}

// The bounce step (main.py:240-264): lobe pick, new direction wv, new accumulated_k.  Philox dims 12-14.
__device__ __forceinline__ void bounce32(const DScene &sc, uint32_t seed, uint32_t r, uint32_t sample, uint32_t bounce, int obj,
                                         const f3 n, const f3 d, const f3 P, float k, f3 &wv, float &knew, bool &diffuse) {
    const uint4 w = rng4(seed, r, sample, bounce, 3u);
    diffuse = (int)(w.x >> 8) < __ldg(sc.obj_lobe_thr + obj);
    const float cy = n.y, sy = sqrtf(fmaxf(0.f, fmaf(-cy, cy, 1.0f)));   // cos / sin of arccos(n.y)
    f3 v;
    if (diffuse) {
        const float u1 = u01<float>(w.y), u2 = u01<float>(w.z);
        const float sp = sqrtf(1.0f - u1), cp = sqrtf(u1);               // sin / cos of arccos(sqrt(u1))
        float st, ct;
        sincosf(6.28f * u2, &st, &ct);                                   // TAU = 6.28 (main.py:19)
        v = mk3(sp * ct, sp * st, cp);
    } else {
        const float dn2 = 2.0f * dot(n, d);
        v = mk3(fmaf(dn2, n.x, -d.x), fmaf(dn2, n.y, -d.y), fmaf(dn2, n.z, -d.z));
        const float rn = rsqrtf(dot(v, v));
        v = v * rn;
    }
    wv = rotate_y32(cy, sy, v);                                          // R_y(arccos(n.y)) (main.py:148-162)
    if (diffuse) {
        knew = k * (__ldg(sc.mat32 + 8 * obj + 4) * dot(wv, n));
    } else {
        f3 e = mk3((float)sc.eye[0] - P.x, (float)sc.eye[1] - P.y, (float)sc.eye[2] - P.z);
        e = e * rsqrtf(dot(e, e));
        const int pn = __ldg(sc.obj_pow_int + obj);
        const float ew = dot(e, wv);
        const float pw = pn >= 0 ? pow_small_int(ew, pn) : powf(ew, __ldg(sc.mat32 + 8 * obj + 7));
        knew = k * (__ldg(sc.mat32 + 8 * obj + 5) * pw);
    }
}

__global__ void __launch_bounds__(PTB_SHADE_THREADS)
k_shade32(DScene sc, BatchInfo bi, const float4 *__restrict__ qo, const float4 *__restrict__ qd,
          const float *__restrict__ hit_t, const int *__restrict__ hit_tri, const unsigned *__restrict__ d_count_in,
          unsigned *d_count_out, float4 *__restrict__ qo_out, float4 *__restrict__ qd_out,
          float4 *__restrict__ sh_p, float4 *__restrict__ sh_d, int2 *__restrict__ sh_aux, float4 *__restrict__ L,
          int cap, RecPtrs rec, unsigned *__restrict__ sort_keys, int share_primary, int *__restrict__ origin_pid = nullptr) {
    // share_primary (bounce 0 only): every sample of a pixel starts with the same primary ray (main.py:191,
    // A-Q14), so the ray and its closest hit are stored once per pixel slot and read by all S samples
    const unsigned count = share_primary ? (unsigned)bi.npix * (unsigned)bi.S : *d_count_in;
    const float inv_npix = 1.0f / (float)bi.npix;
    for (unsigned base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
        const unsigned i = base + threadIdx.x;
        bool alive = false;
        float4 a, b; float t = 0.f; int tri = -1, id = 0, obj = 0;
        float4 sn = make_float4(0, 0, 0, 0);
        if (i < count) {
            unsigned src = i;
            if (share_primary) { int q_, r_; fast_divmod((int)i, bi.npix, inv_npix, q_, r_); src = (unsigned)r_; }
            a = qo[src]; b = qd[src]; t = hit_t[src]; tri = hit_tri[src];
            id = share_primary ? (int)i : id_unpack(b.w);
            if (tri >= 0) {
                sn = __ldg(sc.tri_sn + tri);
                obj = __float_as_int(sn.w);
                if (obj == sc.n_obj) {                      // light hit: main.py:214-215, 265-266
                    float4 l = L[id];
                    const float c0 = (float)sc.light_color[0] * a.w, c1 = (float)sc.light_color[1] * a.w, c2 = (float)sc.light_color[2] * a.w;
                    l.x += c0; l.y += c1; l.z += c2;
                    L[id] = l;
                    if (rec.any) {
                        const size_t ri = rec.index(id, bi.bounce);
                        if (rec.contrib) { rec.contrib[3 * ri] = c0; rec.contrib[3 * ri + 1] = c1; rec.contrib[3 * ri + 2] = c2; }
                        if (rec.k_after) rec.k_after[ri] = a.w;
                    }
                } else {
                    alive = true;
                }
            }
            if (rec.any && rec.tri) rec.tri[rec.index(id, bi.bounce)] = tri;
        }
        // (per-warp slots were measured here too: shade kernel -4 %, shadow kernel +1 % through the changed order of equal sort
        // keys, frame unchanged - profiles/r02p_ab_full.log)
        const unsigned j = compact_slot(alive, d_count_out);
        if (!alive) continue;

        const f3 o = mk3(a.x, a.y, a.z), d = mk3(b.x, b.y, b.z), n = mk3(sn.x, sn.y, sn.z);
        const float k = a.w;
        const f3 P = mk3(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
        int s_local, p;
        fast_divmod(id, bi.npix, inv_npix, s_local, p);
        const uint32_t r = (uint32_t)bi.pm.ray_index(p), sample = bi.sample_index(s_local);

        // ---- 3 light samples (main.py:30-40)
        float4 ld[3];
        light_samples32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, P, ld);
#pragma unroll
        for (int s = 0; s < 3; ++s) sh_d[(size_t)s * cap + j] = ld[s];
        sh_p[j] = make_float4(P.x, P.y, P.z, k);
        sh_aux[j] = make_int2(id, tri);
        if (origin_pid) origin_pid[j] = (int)__ldg(sc.tri_pid + tri);     // plane the new rays start on (ptb_bvh_dyn.cuh)
        if (sort_keys) sort_keys[j] = ps_key(P.x, P.y, P.z, sc.bb_lo, sc.bb_inv);

        // ---- bounce (main.py:240-264)
        f3 wv; float knew; bool diffuse;
        bounce32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, obj, n, d, P, k, wv, knew, diffuse);
        qo_out[j] = make_float4(P.x, P.y, P.z, knew);
        qd_out[j] = make_float4(wv.x, wv.y, wv.z, __int_as_float(id));
        if (rec.any) {
            const size_t ri = rec.index(id, bi.bounce);
            if (rec.lobe) rec.lobe[ri] = diffuse ? 0 : 1;
            if (rec.k_after) rec.k_after[ri] = knew;
        }
    }
}

}  // namespace ptb
