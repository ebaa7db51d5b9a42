// ptb_shade64.cuh - FP64 shading / bounce kernel in the reference's operation order, the per-batch
// reduction into the accumulation buffer and the make_image kernels.
#pragma once
#include "ptb_shade.cuh"

namespace ptb {

// main.py:148-162 with axis (0,1,0); structural zeros of the Euler-Rodrigues matrix dropped
// (they only add signed zeros), remaining operations in the reference's order.
__device__ __forceinline__ d3 rotate_y64(double angle, d3 v) {
    const double a = cos(xdiv(angle, 2.0)), s = sin(xdiv(angle, 2.0));
    const double c = -s;                                   // b = d = -0.0 * s
    const double aa = xmul(a, a), cc = xmul(c, c), ac = xmul(a, c);
    const double R00 = xsub(aa, cc), R02 = xmul(2.0, xsub(0.0, ac));
    const double R11 = xadd(aa, cc);
    const double R20 = xmul(2.0, ac), R22 = xsub(aa, cc);
    return mk3<double>(__fma_rn(R02, v.z, xmul(R00, v.x)), xmul(R11, v.y), __fma_rn(R22, v.z, xmul(R20, v.x)));
}

// One light sample in the reference's FP64 operation order (utils.py:21-46, 28-39); also behind ptb_kat
__device__ __forceinline__ d3 light_point64(const DScene &sc, const uint4 w, int &li) {
    li = pick_light(sc.light_cdf, sc.n_light_tri, xmul(sc.light_total, u01<double>(w.x)));
    const double c0 = u01<double>(w.y), c1 = u01<double>(w.z), c2 = u01<double>(w.w);
    const double sum = xadd(xadd(c0, c1), c2);
    const double wa = xdiv(c0, sum), wb = xdiv(c1, sum), wc = xdiv(c2, sum);
    const double *lt = sc.verts64 + 9 * (size_t)(sc.n_tri_obj + li);
    return mk3<double>(xadd(xadd(xmul(wa, lt[0]), xmul(wb, lt[3])), xmul(wc, lt[6])),
                       xadd(xadd(xmul(wa, lt[1]), xmul(wb, lt[4])), xmul(wc, lt[7])),
                       xadd(xadd(xmul(wa, lt[2]), xmul(wb, lt[5])), xmul(wc, lt[8])));
}

__global__ void __launch_bounds__(PTB_SHADE_THREADS)
k_shade64(DScene sc, BatchInfo bi, const double4 *__restrict__ qo, const double4 *__restrict__ qd,
          const int *__restrict__ hit_tri, const unsigned *__restrict__ d_count_in, unsigned *d_count_out,
          double4 *__restrict__ qo_out, double4 *__restrict__ qd_out, double4 *__restrict__ sh_p,
          double4 *__restrict__ sh_d, int2 *__restrict__ sh_aux, double4 *__restrict__ L, int cap, RecPtrs rec,
          int *__restrict__ origin_pid = nullptr) {
    const unsigned count = *d_count_in;
    const d3 eye = mk3(sc.eye[0], sc.eye[1], sc.eye[2]);
    for (unsigned base = blockIdx.x * blockDim.x; base < count; base += gridDim.x * blockDim.x) {
        const unsigned i = base + threadIdx.x;
        bool alive = false;
        double4 a, b; int tri = -1, id = 0, obj = 0;
        if (i < count) {
            a = qo[i]; b = qd[i]; tri = hit_tri[i];
            id = id_unpack(b.w);
            if (tri >= 0) {
                obj = sc.tri_obj[tri];
                if (obj == sc.n_obj) {
                    double4 l = L[id];
                    const double c0 = xmul(sc.light_color[0], a.w), c1 = xmul(sc.light_color[1], a.w), c2 = xmul(sc.light_color[2], a.w);
                    l.x = xadd(l.x, c0); l.y = xadd(l.y, c1); l.z = xadd(l.z, c2);
                    L[id] = l;
                    if (rec.any) {
                        const size_t ri = rec.index(id, bi.bounce);
                        if (rec.contrib64) { rec.contrib64[3 * ri] = c0; rec.contrib64[3 * ri + 1] = c1; rec.contrib64[3 * ri + 2] = c2; }
                        if (rec.contrib) { rec.contrib[3 * ri] = (float)c0; rec.contrib[3 * ri + 1] = (float)c1; rec.contrib[3 * ri + 2] = (float)c2; }
                        if (rec.k_after64) rec.k_after64[ri] = a.w;
                        if (rec.k_after) rec.k_after[ri] = (float)a.w;
                    }
                } else {
                    alive = true;
                }
            }
            if (rec.any && rec.tri) rec.tri[rec.index(id, bi.bounce)] = tri;
        }
        const unsigned j = compact_slot(alive, d_count_out);
        if (!alive) continue;

        const d3 pr = mk3(a.x, a.y, a.z), dir = mk3(b.x, b.y, b.z);
        double k = a.w;
        // hit point, recomputed with the closest-hit kernel's exact operations (utils.py:110-120)
        const d3 vr = xdiv3(dir, xnorm(dir));
        const double *tv = sc.verts64 + 9 * (size_t)tri;
        const d3 v1 = ld3(tv), v2 = ld3(tv + 3), v3 = ld3(tv + 6);
        d3 pn = xcross(xsub3(v1, v2), xsub3(v3, v2));
        pn = xdiv3(pn, xnorm(pn));
        const double t = xdiv(xsub(xdot(pn, v1), xdot(pn, pr)), xdot(pn, vr));
        const d3 P = mk3<double>(xadd(pr.x, xmul(vr.x, t)), xadd(pr.y, xmul(vr.y, t)), xadd(pr.z, xmul(vr.z, t)));
        const d3 nrm = ld3(sc.normals64 + 3 * (size_t)tri);
        const int s_local = id / bi.npix, p = id - s_local * bi.npix;
        const uint32_t r = (uint32_t)bi.pm.ray_index(p), sample = bi.sample_index(s_local);

        for (int s = 0; s < 3; ++s) {
            const uint4 w = rng4(bi.seed, r, sample, (uint32_t)bi.bounce, (uint32_t)s);
            int li;
            const d3 lp = light_point64(sc, w, li);
            d3 v = xsub3(lp, P);
            v = xdiv3(v, xnorm(v));
            sh_d[(size_t)s * cap + j] = make_double4(v.x, v.y, v.z, xsqdist(P, lp));
        }
        sh_p[j] = make_double4(P.x, P.y, P.z, k);
        sh_aux[j] = make_int2(id, tri);
        if (origin_pid) origin_pid[j] = (int)sc.tri_pid[tri];       // plane the new ray starts on (ptb_bvh_kernels.cuh)

        const uint4 w = rng4(bi.seed, r, sample, (uint32_t)bi.bounce, 3u);
        const double *m = sc.mat64 + 8 * obj;
        const double kd = m[4], ks = m[5], nexp = m[7];
        const double rnd = xmul(xadd(kd, ks), u01<double>(w.x));
        const double angle = acos(nrm.y);
        const bool diffuse = rnd <= kd;
        d3 wv;
        if (diffuse) {
            const double phi = acos(xsqrt(u01<double>(w.y)));
            const double theta = xmul(6.28, u01<double>(w.z));
            d3 v = mk3<double>(xmul(sin(phi), cos(theta)), xmul(sin(phi), sin(theta)), cos(phi));
            v = xdiv3(v, xnorm(v));
            wv = rotate_y64(angle, v);
            k = xmul(k, xmul(kd, xdot(wv, nrm)));
        } else {
            const double s2 = xmul(xdot(nrm, dir), 2.0);
            d3 v = mk3<double>(xsub(xmul(s2, nrm.x), dir.x), xsub(xmul(s2, nrm.y), dir.y), xsub(xmul(s2, nrm.z), dir.z));
            v = xdiv3(v, xnorm(v));
            d3 e = xsub3(eye, P);
            e = xdiv3(e, xnorm(e));
            wv = rotate_y64(angle, v);
            k = xmul(k, xmul(ks, pow(xdot(e, wv), nexp)));
        }
        qo_out[j] = make_double4(P.x, P.y, P.z, k);
        qd_out[j] = make_double4(wv.x, wv.y, wv.z, (double)id);
        if (rec.any) {
            const size_t ri = rec.index(id, bi.bounce);
            if (rec.lobe) rec.lobe[ri] = diffuse ? 0 : 1;
            if (rec.k_after64) rec.k_after64[ri] = k;
            if (rec.k_after) rec.k_after[ri] = (float)k;
        }
    }
}

// ---------------------------------------------------------------- batch reduction (main.py:269-271)
// accum[p] += sum over the batch's samples of L[s * npix + p], in sample order (deterministic).
template <typename T>
__global__ void __launch_bounds__(256)
k_reduce(const typename VecT<T>::v4 *__restrict__ L, float4 *__restrict__ accum, int npix, int S) {
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npix; p += gridDim.x * blockDim.x) {
        float4 acc = accum[p];
        T r = (T)acc.x, g = (T)acc.y, b = (T)acc.z;
        for (int s = 0; s < S; ++s) {
            const typename VecT<T>::v4 l = L[(size_t)s * npix + p];
            r += l.x; g += l.y; b += l.z;
        }
        accum[p] = make_float4((float)r, (float)g, (float)b, acc.w + (float)S);
    }
}

__global__ void k_init_records(RecPtrs rec, size_t n) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        if (rec.tri) rec.tri[i] = -2;
        if (rec.lobe) rec.lobe[i] = -1;
        if (rec.occ) rec.occ[i] = 0;
        if (rec.leak) rec.leak[i] = -1;
        if (rec.k_after) rec.k_after[i] = 0.f;
        if (rec.k_after64) rec.k_after64[i] = 0.0;
        if (rec.contrib) { rec.contrib[3 * i] = 0.f; rec.contrib[3 * i + 1] = 0.f; rec.contrib[3 * i + 2] = 0.f; }
        if (rec.contrib64) { rec.contrib64[3 * i] = 0.0; rec.contrib64[3 * i + 1] = 0.0; rec.contrib64[3 * i + 2] = 0.0; }
    }
}

}  // namespace ptb
