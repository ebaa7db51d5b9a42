// ptb_fused.cuh - one kernel per bounce for scenes whose geometry lives in shared memory (plane-grouped
// brute force, <= PTB_BRUTE_MAX_TRI triangles: configs C1-C3).
//
// With the separate shade / shadow / closest-hit kernels a shaded vertex moves ~330 B through HBM per bounce
// (shadow work item written and read back, ray written and read back, hit record, two L updates): at
// 4 G vertices per 1024-spp frame that is 1.3 TB and the shade kernel alone ran at 5 TB/s.  The geometry is
// on chip, so nothing forces those round trips: k_bounce_planes32 takes a queue of hit records and does, per
// path and in registers,
//     shade (miss / light / shaded split, 3 light samples)        main.py:212-221, utils.py:21-46
//     3 shadow rays against the occlusion planes + NEE colour     main.py:23-73, 142-145
//     bounce step                                                 main.py:236-268
//     closest hit of the NEW ray against all planes               main.py:83-122
//     light-hit / miss handling of that hit                       main.py:214-215
// and then compacts the paths that hit an object into the next queue (warp ballot + block prefix sum + one
// atomic per block).  Per vertex HBM sees: ray + hit in (40 B), L update (32 B, write-only at bounce 0), ray +
// hit out (40 B x survival).  Bounce 0 reads the ray and its FP64-verified hit once per PIXEL: all samples
// of a pixel share the primary ray (main.py:191, A-Q14), so camera rays and the tracked closest-hit kernel
// run on npix rays instead of npix x S.
//
// The arithmetic is the same device functions, in the same order, as the separate kernels (ptb_planes.cuh,
// ptb_shade.cuh), so both pipelines produce bit-identical radiance (tests/test_gpu_parity.py).
#pragma once
#include "ptb_planes.cuh"
#include "ptb_shade.cuh"

namespace ptb {

#ifndef PTB_FUSED_VOTE
#define PTB_FUSED_VOTE 0      // warp-vote plane skipping after the first bounce (incoherent warps: it never skips)
#endif
#ifndef PTB_FUSED_MIN_BLOCKS
#define PTB_FUSED_MIN_BLOCKS 4
#endif

// camera rays for the pixel set only (S samples share them); count[0] = npix drives the closest-hit kernel
__global__ void __launch_bounds__(256)
k_gen_pixels(DScene sc, BatchInfo bi, float4 *__restrict__ qo, float4 *__restrict__ qd, float4 *__restrict__ L, int zero_L,
             unsigned *count, unsigned *unc_count, int n_counts) {
    const unsigned N = (unsigned)bi.npix * (unsigned)bi.S;
    const unsigned n = zero_L ? N : (unsigned)bi.npix;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < n; j += gridDim.x * blockDim.x) {
        if (j < (unsigned)bi.npix) {
            const int r = bi.pm.ray_index((int)j);
            d3 o, d;
            primary_ray64(sc, bi.pm.W, bi.pm.H, r, o, d);
            d = xdiv3(d, xnorm(d));
            qo[j] = make_float4((float)o.x, (float)o.y, (float)o.z, 1.f);
            qd[j] = make_float4((float)d.x, (float)d.y, (float)d.z, __int_as_float((int)j));
        }
        if (zero_L) L[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        count[0] = (unsigned)bi.npix;
        for (int i = 1; i < n_counts; ++i) count[i] = 0;
        unc_count[0] = 0;
        atomicAdd(sc.counters + CTR_SAMPLES, (unsigned long long)N);
    }
}

// FIRST: bounce 0 - items are all npix x S paths, ray / hit read per pixel slot, L written (not updated).
// Otherwise items are the compacted paths whose previous ray hit an object.
template <bool FIRST>
__global__ void __launch_bounds__(PTB_SHADE_THREADS, PTB_FUSED_MIN_BLOCKS)
k_bounce_planes32(DScene sc, PlaneScene ps, BatchInfo bi, const float4 *__restrict__ qo, const float4 *__restrict__ qd,
                  const float *__restrict__ hit_t, const int *__restrict__ hit_tri, const unsigned *__restrict__ d_count_in,
                  unsigned *d_count_out, float4 *__restrict__ qo_out, float4 *__restrict__ qd_out, float *__restrict__ hit_t_out,
                  int *__restrict__ hit_tri_out, float4 *__restrict__ L, int last, RecPtrs rec) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PlaneRec *planes = stage_planes(smem_raw, ps);
    const unsigned count = FIRST ? (unsigned)bi.npix * (unsigned)bi.S : *d_count_in;
    const float inv_npix = 1.0f / (float)bi.npix;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float ambient = (float)sc.ambient;
    unsigned n_shaded = 0, n_traced = 0;
    for (unsigned base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
        const unsigned i = base + threadIdx.x;
        const bool live = i < count;
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = make_float4(1.f, 0.f, 0.f, 0.f), sn = make_float4(0.f, 1.f, 0.f, 0.f);
        float t = 0.f; int tri = -1, id = 0, obj = 0, s_local = 0, p = 0;
        bool shaded = false;
        float4 l = make_float4(0.f, 0.f, 0.f, 0.f);
        if (live) {
            if (FIRST) {
                id = (int)i;
                fast_divmod(id, bi.npix, inv_npix, s_local, p);
                a = qo[p]; b = qd[p]; t = hit_t[p]; tri = hit_tri[p];
            } else {
                a = qo[i]; b = qd[i]; t = hit_t[i]; tri = hit_tri[i];
                id = id_unpack(b.w);
                fast_divmod(id, bi.npix, inv_npix, s_local, p);
            }
            if (tri >= 0) {
                sn = __ldg(sc.tri_sn + tri);
                obj = __float_as_int(sn.w);
                shaded = obj != sc.n_obj;
            }
            if (FIRST) {
                if (tri >= 0 && !shaded) { l.x = lc0 * a.w; l.y = lc1 * a.w; l.z = lc2 * a.w; }   // light hit: main.py:214-215
                if (rec.any) {
                    const size_t ri = rec.index(id, 0);
                    if (rec.tri) rec.tri[ri] = tri;
                    if (tri >= 0 && !shaded) {
                        if (rec.contrib) { rec.contrib[3 * ri] = l.x; rec.contrib[3 * ri + 1] = l.y; rec.contrib[3 * ri + 2] = l.z; }
                        if (rec.k_after) rec.k_after[ri] = a.w;
                    }
                }
                if (!shaded) L[id] = l;
            }
        }
        bool alive = false;
        float4 o_new = a, d_new = b; float t_new = 0.f; int tri_new = -1;
        if (__any_sync(0xffffffffu, shaded)) {
            const f3 o = mk3(a.x, a.y, a.z), d = mk3(b.x, b.y, b.z), n = mk3(sn.x, sn.y, sn.z);
            const float k = a.w;
            const f3 P = mk3(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
            const uint32_t r = (uint32_t)bi.pm.ray_index(p), sample = bi.sample_index(s_local);
            // ---- next-event estimation: 3 light samples, occlusion, colour (main.py:23-73, 142-145)
            float4 ld[3];
            light_samples32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, P, ld);
            const float4 pk = make_float4(P.x, P.y, P.z, k);
            bool occ0, occ1; int leak_ref;
            shadow_planes32<false, FIRST || PTB_FUSED_VOTE>(planes, ps.n_occ, pk, ld[0], ld[1], ld[2], shaded, occ0, occ1, leak_ref);
            if (shaded) {
                ++n_shaded;
                const bool occ2 = leak_ref != 0x7fffffff;
                const int leak = occ2 ? __float_as_int(__ldg(sc.tri_sn + leak_ref).w) : sc.n_obj - 1;
                float dot_sum = 0.f;
                if (!occ0) dot_sum += fmaf(ld[0].z, sn.z, fmaf(ld[0].y, sn.y, ld[0].x * sn.x));
                if (!occ1) dot_sum += fmaf(ld[1].z, sn.z, fmaf(ld[1].y, sn.y, ld[1].x * sn.x));
                if (!occ2) dot_sum += fmaf(ld[2].z, sn.z, fmaf(ld[2].y, sn.y, ld[2].x * sn.x));
                const float dotv = dot_sum / 3.0f;
                const float *m = sc.mat32 + 8 * obj, *lm = sc.mat32 + 8 * leak;
                const float ka = __ldg(m + 3);
                const float c0 = __ldg(m + 0) * ka * ambient + lc0 * __ldg(lm + 0) * dotv;
                const float c1 = __ldg(m + 1) * ka * ambient + lc1 * __ldg(lm + 1) * dotv;
                const float c2 = __ldg(m + 2) * ka * ambient + lc2 * __ldg(lm + 2) * dotv;
                if (!FIRST) l = L[id];
                l.x = fmaf(c0, k, l.x); l.y = fmaf(c1, k, l.y); l.z = fmaf(c2, k, l.z);
                if (rec.any) {
                    const size_t ri = rec.index(id, bi.bounce);
                    if (rec.occ) rec.occ[ri] = (int8_t)((occ0 ? 1 : 0) | (occ1 ? 2 : 0) | (occ2 ? 4 : 0));
                    if (rec.leak) rec.leak[ri] = (int8_t)leak;
                    if (rec.contrib) { rec.contrib[3 * ri] = c0 * k; rec.contrib[3 * ri + 1] = c1 * k; rec.contrib[3 * ri + 2] = c2 * k; }
                }
            }
            // ---- bounce + closest hit of the new ray (main.py:236-268, 83-122); the ray made at the last bounce is
            // never traced by the reference (main.py:192), only its lobe / k are recorded
            if (!last || rec.any) {
                f3 wv = mk3(1.f, 0.f, 0.f); float knew = 0.f; bool diffuse = true;
                if (shaded) {
                    bounce32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, obj, n, d, P, k, wv, knew, diffuse);
                    if (rec.any) {
                        const size_t ri = rec.index(id, bi.bounce);
                        if (rec.lobe) rec.lobe[ri] = diffuse ? 0 : 1;
                        if (rec.k_after) rec.k_after[ri] = knew;
                    }
                }
                if (!last) {
                    o_new = make_float4(P.x, P.y, P.z, knew);
                    d_new = make_float4(wv.x, wv.y, wv.z, __int_as_float(id));
                    closest_planes32<false, PTB_FUSED_VOTE != 0>(planes, ps.n_planes, o_new, d_new, shaded, t_new, tri_new);
                    if (shaded) {
                        ++n_traced;
                        bool light_hit = false;
                        if (tri_new >= 0) {
                            light_hit = __float_as_int(__ldg(sc.tri_sn + tri_new).w) == sc.n_obj;
                            alive = !light_hit;
                        }
                        if (light_hit) { l.x += lc0 * knew; l.y += lc1 * knew; l.z += lc2 * knew; }   // main.py:214-215, 265-266
                        if (rec.any) {
                            const size_t ri = rec.index(id, bi.bounce + 1);
                            if (rec.tri) rec.tri[ri] = tri_new;
                            if (light_hit) {
                                if (rec.contrib) { rec.contrib[3 * ri] = lc0 * knew; rec.contrib[3 * ri + 1] = lc1 * knew; rec.contrib[3 * ri + 2] = lc2 * knew; }
                                if (rec.k_after) rec.k_after[ri] = knew;
                            }
                        }
                    }
                }
            }
            if (shaded) L[id] = l;
        }
        const unsigned j = compact_slot(alive, d_count_out);
        if (alive) {
            qo_out[j] = o_new; qd_out[j] = d_new;
            hit_t_out[j] = t_new; hit_tri_out[j] = tri_new;
        }
    }
    for (int s = 16; s > 0; s >>= 1) { n_shaded += __shfl_xor_sync(0xffffffffu, n_shaded, s); n_traced += __shfl_xor_sync(0xffffffffu, n_traced, s); }
    if ((threadIdx.x & 31) == 0 && (n_shaded | n_traced)) {
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)n_shaded);
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * n_shaded);
        atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)n_traced);
    }
}

}  // namespace ptb
