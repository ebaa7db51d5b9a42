// ptb_shadow.cuh - shadow-ray (any-hit) kernels + NEE accumulation, brute-force flavour.
// One thread per shaded path walks its 3 light samples (main.py:30); the occlusion set is the object
// triangles only (main.py:42, A-Q4), the accepted range is 1e-5 <= d2 < light_d2 (main.py:47-51), the
// walk is in reference triangle order so the first occluder found is the one whose object the
// reference's loop variable keeps (main.py:42, 70; A-Q3).  The kernel then finishes compute_color
// (main.py:142-145): colour = ambient + light_color * rgb(leak) * mean(dir . n over unoccluded), and
// adds colour * k to the path's radiance (main.py:230-231).
#pragma once
#include "ptb_trace.cuh"

namespace ptb {

struct RecPtrs {   // device view of ptb_path_records + indexing
    int32_t *tri; float *contrib; float *k_after; int8_t *lobe; int8_t *occ; int8_t *leak;
    double *contrib64; double *k_after64;
    int npix, n_samples, bounces, s_off;   // s_off = batch first sample - sample_begin
    bool any;
    __device__ __forceinline__ size_t index(int path_id, int bounce) const {
        const int s_local = path_id / npix, p = path_id - s_local * npix;
        return ((size_t)p * n_samples + (size_t)(s_local + s_off)) * bounces + bounce;
    }
};

// first occluder in reference order, or -1
__device__ __forceinline__ int occluder_brute32(const TriRec *s_tri, int n_tri_obj, f3 o, f3 d, float d2max) {
    for (int i = 0; i < n_tri_obj; ++i) {
        float t, dt; bool edge;
        const bool hit = tri_test32<false>(s_tri[i].q0, s_tri[i].q1, s_tri[i].q2, 0.f, o, d, 0.f, t, edge, dt);
        const float t2 = t * t;
        if (hit && t2 >= PTB_ZERO_F && t2 < d2max) return i;
    }
    return -1;
}

__global__ void __launch_bounds__(PTB_TRACE_THREADS)
k_shadow_brute32(DScene sc, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d, const int2 *__restrict__ sh_aux,
                 const unsigned *__restrict__ d_count, float4 *__restrict__ L, int cap, int bounce, RecPtrs rec) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    TriRec *s_tri = reinterpret_cast<TriRec *>(smem_raw);
    stage_tris(s_tri, sc.tri, sc.n_tri_obj);
    const unsigned count = *d_count;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float ambient = (float)sc.ambient;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        const float4 pk = sh_p[j];
        const int2 aux = sh_aux[j];
        const f3 P = mk3(pk.x, pk.y, pk.z);
        const float4 sn = __ldg(sc.tri_sn + aux.y);
        const int obj = __float_as_int(sn.w);
        float dot_sum = 0.f;
        int occ = 0, leak = sc.n_obj - 1;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const float4 dv = sh_d[(size_t)s * cap + j];
            const f3 dir = mk3(dv.x, dv.y, dv.z);
            const int hit = occluder_brute32(s_tri, sc.n_tri_obj, P, dir, dv.w);
            if (hit >= 0) {
                occ |= 1 << s;
                if (s == 2) leak = __float_as_int(s_tri[hit].q3.z);
            } else {
                dot_sum += dot(dir, mk3(sn.x, sn.y, sn.z));
            }
        }
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
            if (rec.occ) rec.occ[ri] = (int8_t)occ;
            if (rec.leak) rec.leak[ri] = (int8_t)leak;
            if (rec.contrib) { rec.contrib[3 * ri] = c0 * pk.w; rec.contrib[3 * ri + 1] = c1 * pk.w; rec.contrib[3 * ri + 2] = c2 * pk.w; }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * count);
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)count);
    }
}

// FP64: main.py:23-73 op for op (mirrors compute_color in oracle/pt_oracle.c).
__global__ void __launch_bounds__(128)
k_shadow_brute64(DScene sc, const double4 *__restrict__ sh_p, const double4 *__restrict__ sh_d, const int2 *__restrict__ sh_aux,
                 const unsigned *__restrict__ d_count, double4 *__restrict__ L, int cap, int bounce, RecPtrs rec) {
    const unsigned count = *d_count;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        const double4 pk = sh_p[j];
        const int2 aux = sh_aux[j];
        const d3 P = mk3(pk.x, pk.y, pk.z);
        const d3 nrm = ld3(sc.normals64 + 3 * (size_t)aux.y);
        const int obj = sc.tri_obj[aux.y];
        double dot_sum = 0.0;
        int occ = 0, leak = sc.n_obj - 1;
        for (int s = 0; s < 3; ++s) {
            const double4 dv = sh_d[(size_t)s * cap + j];
            const d3 v = mk3(dv.x, dv.y, dv.z);
            const d3 vr = xdiv3(v, xnorm(v));                 // intersect() renormalises (utils.py:110)
            int hit = -1;
            for (int i = 0; i < sc.n_tri_obj; ++i) {
                d3 Q;
                if (!tri_test64(sc.verts64 + 9 * (size_t)i, P, vr, Q)) continue;
                const double d2 = xsqdist(Q, P);
                if (d2 < PTB_ZERO_D) continue;
                if (d2 < dv.w) { hit = i; break; }
            }
            if (hit >= 0) { occ |= 1 << s; if (s == 2) leak = sc.tri_obj[hit]; }
            else dot_sum = xadd(dot_sum, xdot(v, nrm));
        }
        const double dotv = xdiv(dot_sum, 3.0);
        const double *m = sc.mat64 + 8 * obj, *lm = sc.mat64 + 8 * leak;
        double c[3];
        for (int i = 0; i < 3; ++i)
            c[i] = xadd(xmul(xmul(m[i], m[3]), sc.ambient), xmul(xmul(sc.light_color[i], lm[i]), dotv));
        double4 l = L[aux.x];
        const double b0 = xmul(c[0], pk.w), b1 = xmul(c[1], pk.w), b2 = xmul(c[2], pk.w);
        l.x = xadd(l.x, b0); l.y = xadd(l.y, b1); l.z = xadd(l.z, b2);
        L[aux.x] = l;
        if (rec.any) {
            const size_t ri = rec.index(aux.x, bounce);
            if (rec.occ) rec.occ[ri] = (int8_t)occ;
            if (rec.leak) rec.leak[ri] = (int8_t)leak;
            if (rec.contrib64) { rec.contrib64[3 * ri] = b0; rec.contrib64[3 * ri + 1] = b1; rec.contrib64[3 * ri + 2] = b2; }
            if (rec.contrib) { rec.contrib[3 * ri] = (float)b0; rec.contrib[3 * ri + 1] = (float)b1; rec.contrib[3 * ri + 2] = (float)b2; }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * count);
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)count);
    }
}

}  // namespace ptb
