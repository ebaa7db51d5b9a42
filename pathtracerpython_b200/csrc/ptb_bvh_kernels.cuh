// ptb_bvh_kernels.cuh - LBVH flavours of the shadow / FP64 kernels (closest-hit FP32 is in
// ptb_bvh_trace.cuh).  Semantics are those of the brute-force kernels (ptb_trace.cuh, ptb_shadow.cuh);
// tests/test_gpu_bvh.py checks LBVH == brute force on the same scenes.
#pragma once
#include "ptb_bvh_trace.cuh"
#include "ptb_shadow.cuh"

namespace ptb {

// ---------------------------------------------------------------- FP32 shadow + NEE
template <bool COUNT>
__global__ void __launch_bounds__(PTB_BVH_THREADS)
k_shadow_bvh32(DScene sc, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d, const int2 *__restrict__ sh_aux,
               const unsigned *__restrict__ d_count, float4 *__restrict__ L, int cap, int bounce, RecPtrs rec) {
    const unsigned count = *d_count;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float ambient = (float)sc.ambient;
    unsigned nn = 0, nt = 0;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        const float4 pk = sh_p[j];
        const int2 aux = sh_aux[j];
        const f3 P = mk3(pk.x, pk.y, pk.z);
        const float4 sn = __ldg(sc.tri_sn + aux.y);
        const int obj = __float_as_int(sn.w);
        float dot_sum = 0.f;
        int occ = 0, leak = sc.n_obj - 1;
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const float4 dv = sh_d[(size_t)s * cap + j];
            const f3 dir = mk3(dv.x, dv.y, dv.z);
            const int hit = occluder_bvh32<COUNT>(sc, P, dir, dv.w, s == 2, nn, nt);
            if (hit >= 0) {
                occ |= 1 << s;
                if (s == 2) leak = __float_as_int(__ldg(sc.tri_sn + hit).w);
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
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(sc.counters + CTR_NODES_SH, (unsigned long long)nn); atomicAdd(sc.counters + CTR_TRIS_SH, (unsigned long long)nt); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(sc.counters + CTR_SHADOW, 3ull * count);
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)count);
    }
}

// ---------------------------------------------------------------- FP64 closest hit through the LBVH
// Exact reference tests at the leaves; boxes are float, padded, so pruning on them is conservative.
// opid > 0: supporting-plane id of the triangle the line starts on; children lying wholly in that plane are skipped
// (ptb_scene.cuh / ptb_bvh_dyn.cuh: the line meets them at t = 0, inside the d^2 > 1e-5 guard).  The FP64 mode runs the
// same rule as the FP32 product, so its record-for-record equality with the oracle (tests/test_gpu_bvh.py, up to 10^7
// triangles) is the evidence that the rule changes no outcome.
__device__ __forceinline__ Hit64 closest_bvh64(const DScene &sc, d3 pr, d3 dir_in, int opid = -1) {
    const d3 vr = xdiv3(dir_in, xnorm(dir_in));
    Hit64 h; h.tri = -1; h.d2 = 0.0; h.P = mk3<double>(0, 0, 0);
    double lim = 1.0e300;
    LineBox<double> lb; lb.set(pr.x, pr.y, pr.z, vr.x, vr.y, vr.z);
    int stack[PTB_BVH_STACK];
    int sp = 0, node = 0;
    while (true) {
        const NodeBoxes n = ld_node(sc.nodes, node);
        double tn0, tf0, tn1, tf1;
        slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
        slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
        const double k0 = near_abs(tn0, tf0), k1 = near_abs(tn1, tf1);
        bool go0 = (tn0 <= tf0) && (k0 <= lim), go1 = (tn1 <= tf1) && (k1 <= lim);
        if (opid > 0) { go0 = go0 && (n.pid0 != opid); go1 = go1 && (n.pid1 != opid); }
        const int c0 = n.d.x, c1 = n.d.y;
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const bool go = side ? go1 : go0;
            const int c = side ? c1 : c0;
            if (go && c < 0) {
                const int ref = side ? n.d.w : n.d.z;
                d3 P;
                const bool hit = tri_test64(sc.verts64 + 9 * (size_t)ref, pr, vr, P);
                closest_update64(hit, P, pr, ref, h);
                if (h.tri >= 0) lim = sqrt(h.d2) * (1.0 + 1e-9) + 1e-9;
            }
        }
        go0 = go0 && c0 >= 0; go1 = go1 && c1 >= 0;
        if (go0 && go1) {
            const bool first1 = k1 < k0;
            stack[sp++] = first1 ? c0 : c1;
            node = first1 ? c1 : c0;
        } else if (go0 || go1) node = go0 ? c0 : c1;
        else { if (sp == 0) break; node = stack[--sp]; }
    }
    return h;
}

__global__ void __launch_bounds__(PTB_BVH_THREADS)
k_closest_bvh64(DScene sc, const double4 *__restrict__ qo, const double4 *__restrict__ qd, const unsigned *__restrict__ d_count,
                double *__restrict__ hit_t, int *__restrict__ hit_tri, const int *__restrict__ origin_pid = nullptr) {
    const unsigned count = *d_count;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const double4 a = qo[i], b = qd[i];
        const d3 pr = mk3(a.x, a.y, a.z), dir = mk3(b.x, b.y, b.z);
        const Hit64 h = closest_bvh64(sc, pr, dir, origin_pid ? origin_pid[i] : -1);
        hit_tri[i] = h.tri;
        hit_t[i] = h.tri >= 0 ? xdot(xsub3(h.P, pr), xdiv3(dir, xnorm(dir))) : 0.0;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// lowest-reference-index occluder (or any, want_min = false) with the reference's FP64 tests
__device__ __forceinline__ int occluder_bvh64(const DScene &sc, d3 P0, d3 vr, double d2max, bool want_min, int opid = -1) {
    LineBox<double> lb; lb.set(P0.x, P0.y, P0.z, vr.x, vr.y, vr.z);
    const double lim = sqrt(d2max) * (1.0 + 1e-9) + 1e-9;
    int found = 0x7fffffff;
    int stack[PTB_BVH_STACK];
    int sp = 0, node = 0;
    while (true) {
        const NodeBoxes n = ld_node(sc.nodes, node);
        double tn0, tf0, tn1, tf1;
        slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
        slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
        bool go0 = (tn0 <= tf0) && (near_abs(tn0, tf0) <= lim) && (n.d.z < found) && (n.d.z < sc.n_tri_obj);
        bool go1 = (tn1 <= tf1) && (near_abs(tn1, tf1) <= lim) && (n.d.w < found) && (n.d.w < sc.n_tri_obj);
        if (opid > 0) { go0 = go0 && (n.pid0 != opid); go1 = go1 && (n.pid1 != opid); }
        const int c0 = n.d.x, c1 = n.d.y;
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const bool go = side ? go1 : go0;
            const int c = side ? c1 : c0;
            if (go && c < 0) {
                const int ref = side ? n.d.w : n.d.z;
                d3 Q;
                if (tri_test64(sc.verts64 + 9 * (size_t)ref, P0, vr, Q)) {
                    const double d2 = xsqdist(Q, P0);
                    if (!(d2 < PTB_ZERO_D) && d2 < d2max) {
                        if (!want_min) return ref;
                        found = min(found, ref);
                    }
                }
            }
        }
        go0 = go0 && c0 >= 0; go1 = go1 && c1 >= 0;
        if (go0 && go1) { stack[sp++] = c1; node = c0; }
        else if (go0 || go1) node = go0 ? c0 : c1;
        else { if (sp == 0) break; node = stack[--sp]; }
    }
    return found == 0x7fffffff ? -1 : found;
}

__global__ void __launch_bounds__(PTB_BVH_THREADS)
k_shadow_bvh64(DScene sc, const double4 *__restrict__ sh_p, const double4 *__restrict__ sh_d, const int2 *__restrict__ sh_aux,
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
        const int opid = sc.tri_pid ? (int)sc.tri_pid[aux.y] : -1;      // plane of the shaded triangle
#pragma unroll 1
        for (int s = 0; s < 3; ++s) {
            const double4 dv = sh_d[(size_t)s * cap + j];
            const d3 v = mk3(dv.x, dv.y, dv.z);
            const d3 vr = xdiv3(v, xnorm(v));
            const int hit = occluder_bvh64(sc, P, vr, dv.w, s == 2, opid);
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
