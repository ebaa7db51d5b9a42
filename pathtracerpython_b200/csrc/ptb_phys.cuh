// ptb_phys.cuh - physical mode (PTB_FLAG_PHYSICAL; SURVEY.md 8a-n3 / 8f-f3, north_star (2)).
//
// The reference's estimator is kept bit for bit in reference mode, quirks included (line queries, leaked NEE
// colour, unclamped cosines, R_y frame, TAU = 6.28, fixed depth, `kt` unused).  Physical mode is the path
// tracer the scene description suggests, on the same wavefront pipeline and the same Philox stream layout:
//   * RAY queries: only hits in front of the origin, t > sqrt(1e-5) (traversal kernels, PHYS template);
//   * two-sided shading normal, true orthonormal basis about it, azimuth range 2 pi;
//   * materials: albedo (red, green, blue), kd / ks / kt used as lobe probabilities (scaled down when they sum to
//     more than 1 so energy is conserved), Phong exponent n, refraction index 1.5 for the transmissive lobe;
//       diffuse      cosine-weighted hemisphere, throughput *= albedo
//       specular     normalised Phong lobe about the mirror direction, throughput *= (n+2)/(n+1) cos(theta_o)
//       transmissive refraction (total internal reflection when it applies), throughput unchanged, "delta" flag set
//       remainder    absorption
//   * next-event estimation: 3 area-uniform light points per vertex with the geometry term
//     cos_s cos_l A / r^2, BRDF = albedo kd / pi + ks (n+2) / (2 pi) cos^n(alpha); emission is added on a light hit
//     only for camera rays and after a transmissive bounce (everything else is covered by NEE: no double counting);
//   * RGB throughput; Russian roulette from the third bounce on: survive with q = clamp(max(thr), 0.05, 0.95),
//     throughput /= q (Philox dim 15, the spare dimension of SURVEY.md 8c);
//   * the ambient term is dropped.
// Validated by its own tests (tests/test_gpu_physical.py): analytic direct lighting of a diffuse plane under the
// area light, white-furnace energy bound, Russian roulette unbiasedness, glass pane transmission, LBVH == planes.
#pragma once
#include "ptb_shade.cuh"

namespace ptb {

#define PTB_PHYS_IOR 1.5f
#define PTB_DELTA_BIT 0x80000000u

struct PhysMat { float r, g, b, kd, ks, kt, n; };
__device__ __forceinline__ PhysMat phys_mat(const DScene &sc, int obj) {
    const float *m = sc.mat32 + 8 * obj;
    PhysMat p;
    p.r = __ldg(m); p.g = __ldg(m + 1); p.b = __ldg(m + 2);
    p.kd = fmaxf(__ldg(m + 4), 0.f); p.ks = fmaxf(__ldg(m + 5), 0.f); p.kt = fmaxf(__ldg(m + 6), 0.f); p.n = fmaxf(__ldg(m + 7), 0.f);
    const float sum = p.kd + p.ks + p.kt;
    if (sum > 1.0f) { const float inv = 1.0f / sum; p.kd *= inv; p.ks *= inv; p.kt *= inv; }
    return p;
}

// branchless orthonormal basis (Duff et al. 2017): returns n's tangent t and bitangent b
__device__ __forceinline__ void onb(const f3 n, f3 &t, f3 &b) {
    const float sign = copysignf(1.0f, n.z);
    const float a = -1.0f / (sign + n.z);
    const float c = n.x * n.y * a;
    t = mk3(1.0f + sign * n.x * n.x * a, sign * c, -sign * n.x);
    b = mk3(c, sign + n.y * n.y * a, -n.y);
}

__global__ void __launch_bounds__(PTB_SHADE_THREADS)
k_shade_phys32(DScene sc, BatchInfo bi, const float4 *__restrict__ qo, const float4 *__restrict__ qd, const float2 *__restrict__ kgb,
               const float *__restrict__ hit_t, const int *__restrict__ hit_tri, const unsigned *__restrict__ d_count_in,
               unsigned *d_count_out, unsigned *d_count_sh, float4 *__restrict__ qo_out, float4 *__restrict__ qd_out,
               float2 *__restrict__ kgb_out, float4 *__restrict__ sh_p, float4 *__restrict__ sh_d, int2 *__restrict__ sh_aux,
               float2 *__restrict__ sh_w, int *__restrict__ sh_res, float4 *__restrict__ L, int cap,
               unsigned *__restrict__ sort_keys, int share_primary, int last, int no_rr) {
    const unsigned count = share_primary ? (unsigned)bi.npix * (unsigned)bi.S : *d_count_in;
    const float inv_npix = 1.0f / (float)bi.npix;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float area = (float)sc.light_total;
    unsigned n_shaded = 0;
    for (unsigned base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
        const unsigned i = base + threadIdx.x;
        bool surface = false;
        float4 a = make_float4(0, 0, 0, 0), b = make_float4(1, 0, 0, 0);
        float2 gb = make_float2(1.f, 1.f);
        float t = 0.f; int tri = -1, id = 0, obj = 0; bool delta = true;
        float4 sn = make_float4(0, 1, 0, 0);
        if (i < count) {
            unsigned src = i;
            if (share_primary) { int q_, r_; fast_divmod((int)i, bi.npix, inv_npix, q_, r_); src = (unsigned)r_; }
            a = qo[src]; b = qd[src]; t = hit_t[src]; tri = hit_tri[src];
            if (share_primary) { id = (int)i; delta = true; }
            else {
                const unsigned w = (unsigned)__float_as_int(b.w);
                id = (int)(w & ~PTB_DELTA_BIT); delta = (w & PTB_DELTA_BIT) != 0; gb = kgb[i];
            }
            if (tri >= 0) {
                sn = __ldg(sc.tri_sn + tri);
                obj = __float_as_int(sn.w);
                if (obj == sc.n_obj) {            // emitter: counted for camera rays and after a delta bounce only
                    if (delta) {
                        float4 l = L[id];
                        l.x = fmaf(lc0, a.w, l.x); l.y = fmaf(lc1, gb.x, l.y); l.z = fmaf(lc2, gb.y, l.z);
                        L[id] = l;
                    }
                } else {
                    surface = true;
                }
            }
        }
        // ---- material, frame, light samples
        const f3 o = mk3(a.x, a.y, a.z), d = mk3(b.x, b.y, b.z);
        const f3 P = mk3(fmaf(t, d.x, o.x), fmaf(t, d.y, o.y), fmaf(t, d.z, o.z));
        f3 n = mk3(sn.x, sn.y, sn.z);
        const bool entering = dot(n, d) < 0.f;            // stored normals point out of closed meshes (A-Q20)
        if (!entering) n = mk3(-n.x, -n.y, -n.z);         // shading normal faces the incoming ray
        PhysMat m; m.r = m.g = m.b = m.kd = m.ks = m.kt = m.n = 0.f;
        if (surface) m = phys_mat(sc, obj);
        const bool has_nee = surface && (m.kd + m.ks) > 0.f;
        int s_local, p;
        fast_divmod(id, bi.npix, inv_npix, s_local, p);
        const uint32_t r = (uint32_t)bi.pm.ray_index(p), sample = bi.sample_index(s_local);
        const float cosi = -dot(d, n);
        const f3 mirror = mk3(fmaf(2.0f * cosi, n.x, d.x), fmaf(2.0f * cosi, n.y, d.y), fmaf(2.0f * cosi, n.z, d.z));

        const unsigned j = compact_slot(has_nee, d_count_sh);
        if (has_nee) {
            ++n_shaded;
#pragma unroll
            for (int s = 0; s < 3; ++s) {
                const uint4 w = rng4(bi.seed, r, sample, (uint32_t)bi.bounce, (uint32_t)s);
                const int li = pick_light_int(sc, w.x >> 8);
                const float su = sqrtf(u01<float>(w.y)), u2 = u01<float>(w.z);
                const float wa = 1.0f - su, wb = su * (1.0f - u2), wc = su * u2;          // area-uniform barycentrics
                const float4 v1 = __ldg(sc.light_v32 + 3 * li), v2 = __ldg(sc.light_v32 + 3 * li + 1), v3 = __ldg(sc.light_v32 + 3 * li + 2);
                const f3 v = mk3(fmaf(wc, v3.x, fmaf(wb, v2.x, wa * v1.x)) - P.x,
                                 fmaf(wc, v3.y, fmaf(wb, v2.y, wa * v1.y)) - P.y,
                                 fmaf(wc, v3.z, fmaf(wb, v2.z, wa * v1.z)) - P.z);
                const float d2 = dot(v, v);
                const float rn = rsqrtf(d2);
                const f3 dir = v * rn;
                f3 nl = cross(mk3(v2.x - v1.x, v2.y - v1.y, v2.z - v1.z), mk3(v3.x - v1.x, v3.y - v1.y, v3.z - v1.z));
                nl = nl * rsqrtf(dot(nl, nl));
                const float cos_s = dot(n, dir), cos_l = fabsf(dot(nl, dir));
                float wd = 0.f, ws = 0.f;
                if (cos_s > 0.f && d2 > 1e-12f) {
                    const float G = cos_s * cos_l * area / (3.0f * d2);
                    wd = G * m.kd * 0.31830988618f;
                    const float ca = fmaxf(dot(mirror, dir), 0.f);
                    ws = G * m.ks * (m.n + 2.0f) * 0.15915494309f * __powf(ca, m.n);
                    if (!(ca > 0.f)) ws = 0.f;
                }
                sh_d[(size_t)s * cap + j] = make_float4(dir.x, dir.y, dir.z, d2);
                sh_w[(size_t)s * cap + j] = make_float2(wd, ws);
                sh_res[(size_t)s * cap + j] = (wd + ws) > 0.f ? 1 : 0;
            }
            sh_p[j] = make_float4(P.x, P.y, P.z, a.w);
            sh_w[(size_t)3 * cap + j] = gb;
            sh_aux[j] = make_int2(id, tri);
        }

        // ---- continuation: lobe pick, new direction, throughput, Russian roulette
        bool go_on = false;
        f3 wv = mk3(1.f, 0.f, 0.f);
        float tr = a.w, tg = gb.x, tb = gb.y;
        bool new_delta = false;
        if (surface && !last) {
            const uint4 w = rng4(bi.seed, r, sample, (uint32_t)bi.bounce, 3u);
            const float ul = u01<float>(w.x), u1 = u01<float>(w.y), u2 = u01<float>(w.z);
            float sphi, cphi;
            __sincosf(6.28318530718f * u2, &sphi, &cphi);
            if (ul < m.kd) {                                       // diffuse
                f3 tx, ty; onb(n, tx, ty);
                const float rr = sqrtf(u1), cz = sqrtf(fmaxf(0.f, 1.0f - u1));
                wv = mk3(rr * cphi * tx.x + rr * sphi * ty.x + cz * n.x, rr * cphi * tx.y + rr * sphi * ty.y + cz * n.y,
                         rr * cphi * tx.z + rr * sphi * ty.z + cz * n.z);
                tr *= m.r; tg *= m.g; tb *= m.b;
                go_on = true;
            } else if (ul < m.kd + m.ks) {                         // Phong lobe about the mirror direction
                f3 tx, ty; onb(mirror, tx, ty);
                const float ca = __powf(u1, 1.0f / (m.n + 1.0f)), sa = sqrtf(fmaxf(0.f, 1.0f - ca * ca));
                wv = mk3(sa * cphi * tx.x + sa * sphi * ty.x + ca * mirror.x, sa * cphi * tx.y + sa * sphi * ty.y + ca * mirror.y,
                         sa * cphi * tx.z + sa * sphi * ty.z + ca * mirror.z);
                const float co = dot(wv, n);
                const float f = (m.n + 2.0f) / (m.n + 1.0f) * co;
                tr *= f; tg *= f; tb *= f;
                go_on = co > 0.f;
            } else if (ul < m.kd + m.ks + m.kt) {                  // refraction, fixed index
                const float eta = entering ? (1.0f / PTB_PHYS_IOR) : PTB_PHYS_IOR;
                const float k2 = 1.0f - eta * eta * (1.0f - cosi * cosi);
                if (k2 < 0.f) wv = mirror;                         // total internal reflection
                else { const float c = eta * cosi - sqrtf(k2); wv = mk3(fmaf(eta, d.x, c * n.x), fmaf(eta, d.y, c * n.y), fmaf(eta, d.z, c * n.z)); }
                wv = wv * rsqrtf(dot(wv, wv));
                new_delta = true;
                go_on = true;
            }
            if (go_on && !no_rr && bi.bounce >= 2) {               // Russian roulette (Philox dim 15)
                const float q = fminf(fmaxf(fmaxf(tr, fmaxf(tg, tb)), 0.05f), 0.95f);
                if (u01<float>(w.w) >= q) go_on = false;
                else { const float iq = 1.0f / q; tr *= iq; tg *= iq; tb *= iq; }
            }
            go_on = go_on && (fmaxf(tr, fmaxf(tg, tb)) > 0.f);
        }
        const unsigned jr = compact_slot(go_on, d_count_out);
        if (go_on) {
            qo_out[jr] = make_float4(P.x, P.y, P.z, tr);
            qd_out[jr] = make_float4(wv.x, wv.y, wv.z, __int_as_float((int)((unsigned)id | (new_delta ? PTB_DELTA_BIT : 0u))));
            kgb_out[jr] = make_float2(tg, tb);
            if (sort_keys) sort_keys[jr] = ps_key(P.x, P.y, P.z, sc.bb_lo, sc.bb_inv);
        }
    }
    for (int s = 16; s > 0; s >>= 1) n_shaded += __shfl_xor_sync(0xffffffffu, n_shaded, s);
    if ((threadIdx.x & 31) == 0 && n_shaded) {
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)n_shaded);
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * n_shaded);
    }
}

// shadow rays of the small (plane-grouped) scenes, physical mode: cancels the weight of occluded samples
__global__ void __launch_bounds__(PTB_TRACE_THREADS)
k_shadow_planes_phys32(PlaneScene ps, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d,
                       const unsigned *__restrict__ d_count, int *__restrict__ sh_res, int cap) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const PlaneRec *planes = stage_planes(smem_raw, ps);
    const unsigned count = *d_count;
    const unsigned count_w = (count + 31u) & ~31u;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count_w; j += gridDim.x * blockDim.x) {
        const bool live = j < count;
        const unsigned jj = live ? j : 0u;
        const float4 pk = sh_p[jj];
        const float4 d0 = sh_d[jj], d1 = sh_d[(size_t)cap + jj], d2 = sh_d[2 * (size_t)cap + jj];
        bool occ0, occ1; int leak_ref;
        shadow_planes32<true>(planes, ps.n_occ, pk, d0, d1, d2, live, occ0, occ1, leak_ref);
        if (!live) continue;
        if (occ0) sh_res[j] = 0;
        if (occ1) sh_res[(size_t)cap + j] = 0;
        if (leak_ref != 0x7fffffff) sh_res[2 * (size_t)cap + j] = 0;
    }
}

// L += throughput * Le * (albedo * sum(visible diffuse weights) + sum(visible specular weights))
__global__ void __launch_bounds__(256)
k_nee_finish_phys32(DScene sc, const float4 *__restrict__ sh_p, const int2 *__restrict__ sh_aux, const float2 *__restrict__ sh_w,
                    const int *__restrict__ sh_res, const unsigned *__restrict__ d_count, float4 *__restrict__ L, int cap) {
    const unsigned count = *d_count;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        float wd = 0.f, ws = 0.f;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            if (sh_res[(size_t)s * cap + j] != 0) { const float2 w = sh_w[(size_t)s * cap + j]; wd += w.x; ws += w.y; }
        }
        if (wd + ws == 0.f) continue;
        const int2 aux = sh_aux[j];
        const int obj = __float_as_int(__ldg(sc.tri_sn + aux.y).w);
        const float *m = sc.mat32 + 8 * obj;
        const float tr = sh_p[j].w;
        const float2 gb = sh_w[(size_t)3 * cap + j];
        float4 l = L[aux.x];
        l.x = fmaf(tr * lc0, fmaf(__ldg(m + 0), wd, ws), l.x);
        l.y = fmaf(gb.x * lc1, fmaf(__ldg(m + 1), wd, ws), l.y);
        l.z = fmaf(gb.y * lc2, fmaf(__ldg(m + 2), wd, ws), l.z);
        L[aux.x] = l;
    }
}

}  // namespace ptb
