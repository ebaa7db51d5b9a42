// ptb_fused2.cuh - the fused per-bounce kernel of ptb_fused.cuh with PACKED FP32 arithmetic (sm_100a: fma.rn.f32x2 /
// mul.f32x2 / add.f32x2, SASS FFMA2).
//
// k_bounce_planes32 is issue-bound (ncu: 76 % of the issue slots, FMA pipe 39 %, ALU pipe 44 %; about 2 840 warp
// instructions per shaded vertex) and three quarters of those instructions are the two plane loops: 3 shadow rays
// against the occluder planes, then the new ray against all planes.  All FOUR rays leave the same point P, so
//   * the plane part that depends on the origin (n . P, utils.py:118) is evaluated once per plane for the four of them
//     - the two loops become one;
//   * the per-ray arithmetic (n . d, t, P + t d, the two barycentric plane equations, the parallelogram test's two
//     sums) runs on ray PAIRS - (shadow 0, shadow 1) and (shadow 2, new ray) - with one FFMA2 / FMUL2 / FADD2 per pair:
//     the plane coefficients enter as the broadcast scalar operand of FFMA2 (Rn.F32), no register holds a duplicate.
// Every lane of a packed operation rounds exactly like the scalar operation it replaces (same operands, same fused /
// unfused shape), so this kernel's radiance and records are BIT-IDENTICAL to k_bounce_planes32's and to the separate
// kernels' (tests/test_gpu_parity.py::test_fused_bounce_kernel_equals_separate_kernels runs all three).
// Measured issue rates behind the choice (scratch/ubench_fp32.cu, profiles/r02a_ubench_fp32_issue_rates.log): FFMA with
// three distinct register sources issues at 2.44 per clock and SM of the 4 an SM can issue (register bandwidth), FFMA2 at
// 1.32 - 8 % more FMAs per clock in 46 % fewer issue slots, and the slots are what this kernel is short of.
#pragma once
#include "ptb_fused.cuh"

namespace ptb {

__device__ __forceinline__ float2 bc2(float a) { return make_float2(a, a); }
__device__ __forceinline__ float2 fma2(float2 a, float2 b, float2 c) { return __ffma2_rn(a, b, c); }
__device__ __forceinline__ float2 mul2(float2 a, float2 b) { return __fmul2_rn(a, b); }
__device__ __forceinline__ float2 add2(float2 a, float2 b) { return __fadd2_rn(a, b); }

// plane equation q . P + q.w for a pair of points (the same chain as para_hit / bary_min, per lane)
__device__ __forceinline__ float2 plane2(const float4 q, float2 px, float2 py, float2 pz) {
    return fma2(bc2(q.z), pz, fma2(bc2(q.y), py, fma2(bc2(q.x), px, bc2(q.w))));
}
// parallelogram test of para_hit on a pair: inside iff max(|v - 1/2|, |u + v - 1/2|) < 1/2 and u != 0
__device__ __forceinline__ void para2(const float4 q1, const float4 q2, float2 px, float2 py, float2 pz, float2 &u, bool &h0, bool &h1) {
    u = plane2(q1, px, py, pz);
    const float2 v = plane2(q2, px, py, pz);
    const float2 vh = add2(v, bc2(-0.5f)), sh = add2(u, vh);
    h0 = (fmaxf(fabsf(vh.x), fabsf(sh.x)) < 0.5f) && (u.x != 0.0f);
    h1 = (fmaxf(fabsf(vh.y), fabsf(sh.y)) < 0.5f) && (u.y != 0.0f);
}

// The 3 light samples' occlusion (shadow_planes32) and the new ray's closest hit (closest_planes32) in ONE walk over the
// plane list.  pk.xyz = P (origin of all four lines), d0..d2 = unit directions to the light samples (w = squared
// distance), dc = direction of the new ray.  Warp-collective when VOTE.  CLOSEST = false: the last bounce, whose new
// ray is never traced (main.py:192); the second lane of pair B then idles.
template <bool VOTE, bool CLOSEST>
__device__ __forceinline__ void trace4_planes32(const PlaneRec *planes, int n_occ, int n_planes, const float4 pk, const float4 d0,
                                                const float4 d1, const float4 d2, const float4 dc, bool live, bool &occ0,
                                                bool &occ1, int &leak_ref, float &best_t, int &best_ref) {
    // "sample s is occluded" is carried in its distance limit: -1 once an occluder is found (and from the start for dead
    // lanes), so that s < lim is never true again - the same outcomes as shadow_planes32's `!occ && ...` without the
    // flag registers and their predicate conversions in every iteration
    float lim0 = live ? d0.w : -1.0f, lim1 = live ? d1.w : -1.0f;
    leak_ref = live ? 0x7fffffff : -1;
    float best_a = live ? 3.0e38f : 0.f;
    best_t = 0.f;
    best_ref = -1;
    const float2 dxA = make_float2(d0.x, d1.x), dyA = make_float2(d0.y, d1.y), dzA = make_float2(d0.z, d1.z);
    const float2 dxB = make_float2(d2.x, dc.x), dyB = make_float2(d2.y, dc.y), dzB = make_float2(d2.z, dc.z);
    const float2 Px = bc2(pk.x), Py = bc2(pk.y), Pz = bc2(pk.z);
    for (int g = 0; g < n_occ; ++g) {
        const float4 nd = planes[g].nd;
        const float on = nd.w - fmaf(nd.z, pk.z, fmaf(nd.y, pk.y, nd.x * pk.x));     // shared by the four lines
        const float2 dnA = fma2(bc2(nd.z), dzA, fma2(bc2(nd.y), dyA, mul2(bc2(nd.x), dxA)));
        const float2 dnB = fma2(bc2(nd.z), dzB, fma2(bc2(nd.y), dyB, mul2(bc2(nd.x), dxB)));
        const float2 tA = mul2(bc2(on), make_float2(rcp_approx(dnA.x), rcp_approx(dnA.y)));
        const float2 tB = mul2(bc2(on), make_float2(rcp_approx(dnB.x), rcp_approx(dnB.y)));
        const float2 sA = mul2(tA, tA);
        const float s2 = tB.x * tB.x;
        // main.py:47-51: 1e-5 <= d2 < light_squared_distance ; utils.py:117: |n.d| > 1e-5
        const bool ok0 = (fabsf(dnA.x) > PTB_ZERO_F) && (sA.x >= PTB_ZERO_F) && (sA.x < lim0);
        const bool ok1 = (fabsf(dnA.y) > PTB_ZERO_F) && (sA.y >= PTB_ZERO_F) && (sA.y < lim1);
        const bool ok2 = (planes[g].ref_a < leak_ref) && (fabsf(dnB.x) > PTB_ZERO_F) && (s2 >= PTB_ZERO_F) && (s2 < d2.w);
        const float atc = fabsf(tB.y);
        const bool okc = CLOSEST && (fabsf(dnB.y) > PTB_ZERO_F) && (atc > PTB_TMIN_F) && (atc < best_a);
        if (VOTE && !__any_sync(0xffffffffu, ok0 || ok1 || ok2 || okc)) continue;
        const float2 pxA = fma2(tA, dxA, Px), pyA = fma2(tA, dyA, Py), pzA = fma2(tA, dzA, Pz);
        const float2 pxB = fma2(tB, dxB, Px), pyB = fma2(tB, dyB, Py), pzB = fma2(tB, dzB, Pz);
        const float4 a1 = planes[g].a1, a2 = planes[g].a2;
        if (planes[g].kind) {                  // warp-uniform: parallelogram pair, both triangles in the occlusion set
            float2 uA, uB; bool h0, h1, h2, hc;
            para2(a1, a2, pxA, pyA, pzA, uA, h0, h1);
            para2(a1, a2, pxB, pyB, pzB, uB, h2, hc);
            if (ok0 && h0) lim0 = -1.0f;
            if (ok1 && h1) lim1 = -1.0f;
            if (ok2 && h2) leak_ref = min(leak_ref, uB.x > 0.0f ? planes[g].ref_a : planes[g].ref_b);
            if (CLOSEST && hc && okc) { best_a = atc; best_t = tB.y; best_ref = uB.y > 0.0f ? planes[g].ref_a : planes[g].ref_b; }
        } else {
            const bool ha0 = bary_min(a1, a2, pxA.x, pyA.x, pzA.x) > 0.0f;
            const bool ha1 = bary_min(a1, a2, pxA.y, pyA.y, pzA.y) > 0.0f;
            const bool ha2 = bary_min(a1, a2, pxB.x, pyB.x, pzB.x) > 0.0f;
            const float mac = CLOSEST ? bary_min(a1, a2, pxB.y, pyB.y, pzB.y) : -1.0f;
            bool hb0 = false, hb1 = false, hb2 = false;
            float mbc = -1.0f;
            if (planes[g].n_obj > 1) {             // warp-uniform (a single-triangle group's B is the never-inside dummy)
                const float4 b1 = planes[g].b1, b2 = planes[g].b2;
                hb0 = bary_min(b1, b2, pxA.x, pyA.x, pzA.x) > 0.0f;
                hb1 = bary_min(b1, b2, pxA.y, pyA.y, pzA.y) > 0.0f;
                hb2 = bary_min(b1, b2, pxB.x, pyB.x, pzB.x) > 0.0f;
                if (CLOSEST) mbc = bary_min(b1, b2, pxB.y, pyB.y, pzB.y);
            }
            if (ok0 && (ha0 || hb0)) lim0 = -1.0f;
            if (ok1 && (ha1 || hb1)) lim1 = -1.0f;
            if (ok2 && (ha2 || hb2)) leak_ref = min(leak_ref, ha2 ? planes[g].ref_a : planes[g].ref_b);
            if (CLOSEST) {
                // A precedes B in reference order; both lie in the same plane (same |t|), so A wins a tie
                const bool hac = okc && mac > 0.0f, hbc = okc && mbc > 0.0f;
                if (hac || hbc) { best_a = atc; best_t = tB.y; best_ref = hac ? planes[g].ref_a : planes[g].ref_b; }
            }
        }
    }
    occ0 = lim0 < 0.0f; occ1 = lim1 < 0.0f;
    if (CLOSEST) {
        // the light's planes: not occluders (main.py:42), the new ray alone (closest_planes32's loop body)
        for (int g = n_occ; g < n_planes; ++g) {
            const float4 nd = planes[g].nd;
            const float dn = fmaf(nd.z, dc.z, fmaf(nd.y, dc.y, nd.x * dc.x));
            const float on = nd.w - fmaf(nd.z, pk.z, fmaf(nd.y, pk.y, nd.x * pk.x));
            const float t = on * rcp_approx(dn);
            const float at = fabsf(t);
            const bool ok = (fabsf(dn) > PTB_ZERO_F) && (at > PTB_TMIN_F) && (at < best_a);
            if (VOTE && !__any_sync(0xffffffffu, ok)) continue;
            const float px = fmaf(t, dc.x, pk.x), py = fmaf(t, dc.y, pk.y), pz = fmaf(t, dc.z, pk.z);
            if (planes[g].kind) {
                float u;
                if (para_hit(planes[g].a1, planes[g].a2, px, py, pz, u) && ok) {
                    best_a = at; best_t = t; best_ref = u > 0.0f ? planes[g].ref_a : planes[g].ref_b;
                }
            } else {
                const float ma = bary_min(planes[g].a1, planes[g].a2, px, py, pz);
                const float mb = bary_min(planes[g].b1, planes[g].b2, px, py, pz);
                const bool ha = ok && ma > 0.0f, hb = ok && mb > 0.0f;
                if (ha || hb) { best_a = at; best_t = t; best_ref = ha ? planes[g].ref_a : planes[g].ref_b; }
            }
        }
    }
}

// Stall samples of the kernel (ncu source page, profiles/r02j_ncu_source_fused_x2.csv) put half of its time OUTSIDE the
// plane loop that holds two thirds of its instructions; the largest single item was the block-wide compaction - three
// barriers per tile, every warp of the block waiting out the slot counter's atomic (6 % of the samples).
// PTB_X2_WARP_COMPACT: slots are claimed per WARP - ballot, one atomic by the leader, shuffle - with no barrier: +3.2 %
// (2 930 -> 3 022 Msamples/s, profiles/r02k_ab_compact_prefetch.log).  The queue order then depends on the arrival of the
// warps; nothing downstream depends on it (a path's radiance lives in L[path id], records are indexed by path id), so
// radiance and records stay bit-identical and deterministic.  Measured with it and dropped: prefetching L[id] and the
// next tile's queue entries into L2 or L1 ahead of the plane loop (the two loads with the most long-scoreboard samples):
// -0.3 % / -0.2 .. -1.5 % (profiles/r02k_ab_prefetch_l1.log) - other warps were already covering those waits.
#ifndef PTB_X2_WARP_COMPACT
#define PTB_X2_WARP_COMPACT 1
#endif

#ifndef PTB_FUSED2_MIN_BLOCKS
#define PTB_FUSED2_MIN_BLOCKS 4
#endif

// Same contract as k_bounce_planes32 (ptb_fused.cuh).  CLOSEST = false is launched for the last bounce when no records
// are wanted (`last` && !rec.any): shading only.
template <bool FIRST, bool CLOSEST>
__global__ void __launch_bounds__(PTB_SHADE_THREADS, PTB_FUSED2_MIN_BLOCKS)
k_bounce_planes32x2(DScene sc, PlaneScene ps, BatchInfo bi, const float4 *__restrict__ qo, const float4 *__restrict__ qd,
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
            // ---- next-event estimation: 3 light samples (main.py:30-40, utils.py:21-46)
            float4 ld[3];
            light_samples32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, P, ld);
            // ---- the bounce step first (main.py:236-268): it needs nothing of the shadow rays' outcome, and with the new
            // direction known the four lines from P share one walk over the planes
            f3 wv = mk3(1.f, 0.f, 0.f); float knew = 0.f; bool diffuse = true;
            if ((CLOSEST || rec.any) && shaded) {
                bounce32(sc, bi.seed, r, sample, (uint32_t)bi.bounce, obj, n, d, P, k, wv, knew, diffuse);
                if (rec.any) {
                    const size_t ri = rec.index(id, bi.bounce);
                    if (rec.lobe) rec.lobe[ri] = diffuse ? 0 : 1;
                    if (rec.k_after) rec.k_after[ri] = knew;
                }
            }
            const float4 pk = make_float4(P.x, P.y, P.z, k);
            o_new = make_float4(P.x, P.y, P.z, knew);
            d_new = make_float4(wv.x, wv.y, wv.z, __int_as_float(id));
            bool occ0, occ1; int leak_ref;
            trace4_planes32<FIRST || PTB_FUSED_VOTE, CLOSEST>(planes, ps.n_occ, ps.n_planes, pk, ld[0], ld[1], ld[2], d_new, shaded,
                                                              occ0, occ1, leak_ref, t_new, tri_new);
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
                if (CLOSEST) {
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
                L[id] = l;
            }
        }
        if (CLOSEST) {
#if PTB_X2_WARP_COMPACT
            const unsigned j = compact_slot_warp(alive, d_count_out);
#else
            const unsigned j = compact_slot(alive, d_count_out);
#endif
            if (alive) {
                qo_out[j] = o_new; qd_out[j] = d_new;
                hit_t_out[j] = t_new; hit_tri_out[j] = tri_new;
            }
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
