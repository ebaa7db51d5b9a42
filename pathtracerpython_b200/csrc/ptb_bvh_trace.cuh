// ptb_bvh_trace.cuh - stack-based LBVH traversal with the reference's LINE semantics (A-Q1):
// a node is relevant when the line's parameter interval inside its box, [tn, tf], meets [-tmax, +tmax]
// (closest hit: tmax = current best |t|; shadow: tmax = distance to the light sample).  Children are
// visited nearest-|t| first.  Leaves hold one triangle; the triangle tests are the same functions the
// brute-force kernels use, so both give identical decisions.  Ties in |t| go to the lowest reference
// index (main.py:115) whatever the visiting order.
#pragma once
#include "ptb_queue.cuh"

namespace ptb {

#define PTB_BVH_STACK 64
#define PTB_BVH_THREADS 128

template <typename T> struct LineBox {
    T ox, oy, oz, ix, iy, iz;   // origin, reciprocal direction
    T bx, by, bz;               // -origin * reciprocal direction (FP32 slab test in FMA form)
    __device__ __forceinline__ void set(T ox_, T oy_, T oz_, T dx, T dy, T dz) {
        ox = ox_; oy = oy_; oz = oz_;
        ix = (T)1 / dx; iy = (T)1 / dy; iz = (T)1 / dz;
        bx = -ox * ix; by = -oy * iy; bz = -oz * iz;
    }
};
// parameter interval of the line inside the box (no clamping to t >= 0); NaN-safe via fmin/fmax.
// t = plane * (1/d) - o * (1/d) as one FMA per plane: its error is <= 2^-23 * scale * |1/d|, 1/4 of what the box padding
// (2^-21 * scale in space, i.e. 2^-21 * scale * |1/d| in t) leaves, so no triangle the exact test accepts is culled.
__device__ __forceinline__ void slab(const LineBox<float> &r, float lx, float ly, float lz, float hx, float hy, float hz, float &tn, float &tf) {
    const float ax = fmaf(lx, r.ix, r.bx), bx = fmaf(hx, r.ix, r.bx);
    const float ay = fmaf(ly, r.iy, r.by), by = fmaf(hy, r.iy, r.by);
    const float az = fmaf(lz, r.iz, r.bz), bz = fmaf(hz, r.iz, r.bz);
    tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
    tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
}
__device__ __forceinline__ void slab(const LineBox<double> &r, float lx, float ly, float lz, float hx, float hy, float hz, double &tn, double &tf) {
    const double ax = ((double)lx - r.ox) * r.ix, bx = ((double)hx - r.ox) * r.ix;
    const double ay = ((double)ly - r.oy) * r.iy, by = ((double)hy - r.oy) * r.iy;
    const double az = ((double)lz - r.oz) * r.iz, bz = ((double)hz - r.oz) * r.iz;
    tn = fmax(fmax(fmin(ax, bx), fmin(ay, by)), fmin(az, bz));
    tf = fmin(fmin(fmax(ax, bx), fmax(ay, by)), fmax(az, bz));
}
// smallest |t| of [tn, tf] (tn <= tf): tn if the interval is in front, -tf if it is behind, else 0
__device__ __forceinline__ float near_abs(float tn, float tf) { return fmaxf(fmaxf(tn, -tf), 0.0f); }
__device__ __forceinline__ double near_abs(double tn, double tf) { return fmax(fmax(tn, -tf), 0.0); }
// largest |t| of [tn, tf]
__device__ __forceinline__ float far_abs(float tn, float tf) { return fmaxf(-tn, tf); }
__device__ __forceinline__ double far_abs(double tn, double tf) { return fmax(-tn, tf); }
// sqrt(1e-5) = 3.1623e-3 is the smallest |t| either query accepts (main.py:47-48, 100); boxes are padded and
// the FP32 ray parameter carries ~1e-5 of error, so prune only below 3.0e-3
#define PTB_TMIN_SAFE 3.0e-3f

// 256-bit read-only load (sm_100: LDG.E.256): the L1 tag stage serves scattered accesses one lane and one instruction
// at a time, so a 64 B node or triangle record costs two requests per lane this way instead of four LDG.128.
// p must be 32 B aligned (nodes and records are 64 B structs in their own cudaMalloc).
#ifndef PTB_LD256
#define PTB_LD256 1
#endif
__device__ __forceinline__ void ld256(const void *p, float4 &lo, float4 &hi) {
#if PTB_LD256
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(lo.x), "=f"(lo.y), "=f"(lo.z), "=f"(lo.w), "=f"(hi.x), "=f"(hi.y), "=f"(hi.z), "=f"(hi.w) : "l"(p));
#else
    lo = __ldg(reinterpret_cast<const float4 *>(p)); hi = __ldg(reinterpret_cast<const float4 *>(p) + 1);
#endif
}

struct NodeBoxes { float4 a, b, c; int4 d; int pid0, pid1; };
__device__ __forceinline__ NodeBoxes ld_node(const BvhNode *nodes, int i) {
    const float4 *p = reinterpret_cast<const float4 *>(nodes + i);
    NodeBoxes n;
    float4 w;
    ld256(p, n.a, n.b); ld256(p + 2, n.c, w);
    // child 1 = position of child 0 plus one, leaf flag in meta's sign (ptb_scene.cuh)
    const int c0 = __float_as_int(w.x), meta = __float_as_int(w.y);
    const int i1 = (c0 ^ (c0 >> 31)) + 1;
    n.d = make_int4(c0, i1 ^ (meta >> 31), __float_as_int(w.z), __float_as_int(w.w));
    n.pid0 = meta & 0xff; n.pid1 = (meta >> 8) & 0xff;      // dead code in the kernels that do not use them
    return n;
}
// BvhNode layout as float4s: a = lo0.xyz hi0.x ; b = hi0.yz lo1.xy ; c = lo1.z hi1.xyz ; d = c0 meta minref0 minref1
// (NodeBoxes.d is c0 c1 minref0 minref1)

// ---------------------------------------------------------------- FP32 closest hit
// TRACK: prune with best |t| + its error bound and collect uncertainty (ptb_isect.cuh) so uncertain
// primary rays can be re-traced in FP64.
template <bool TRACK, bool COUNT>
__device__ __forceinline__ Hit32 closest_bvh32(const DScene &sc, f3 o, f3 d, float dt_scale, bool &uncertain, int &best_ref,
                                               unsigned &n_nodes, unsigned &n_tris) {
    Hit32 h; h.t = 0.f; h.abs_t = 3.0e38f; h.slot = -1;
    best_ref = 0x7fffffff;
    Unc32 uc; uc.init();
    LineBox<float> lb; lb.set(o.x, o.y, o.z, d.x, d.y, d.z);
    int stack[PTB_BVH_STACK];
    int sp = 0, node = 0;
    while (true) {
        const NodeBoxes n = ld_node(sc.nodes, node);
        if (COUNT) n_nodes++;
        float tn0, tf0, tn1, tf1;
        slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
        slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
        const float lim = TRACK ? h.abs_t + uc.best_dt : h.abs_t;
        float k0 = near_abs(tn0, tf0), k1 = near_abs(tn1, tf1);
        bool go0 = (tn0 <= tf0) && (k0 <= lim), go1 = (tn1 <= tf1) && (k1 <= lim);
        int c0 = n.d.x, c1 = n.d.y;
        // leaves are tested on the spot
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const bool go = side ? go1 : go0;
            const int c = side ? c1 : c0;
            if (go && c < 0) {
                const int slot = ~c;
                const float4 *tp = reinterpret_cast<const float4 *>(sc.tri + slot);
                const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2), q3 = __ldg(tp + 3);
                if (COUNT) n_tris++;
                float t, dt = 0.f; bool edge = false;
                const bool hit = tri_test32<TRACK>(q0, q1, q2, q3.x, o, d, dt_scale, t, edge, dt);
                const int ref = __float_as_int(q3.w);
                const float a = fabsf(t);
                const bool ok = hit && (t * t > PTB_ZERO_F);
                if (ok && (a < h.abs_t || (a == h.abs_t && ref < best_ref))) {
                    if (TRACK) {
                        if (h.slot >= 0) uc.comp_lo = fminf(uc.comp_lo, h.abs_t - uc.best_dt);
                        uc.best_dt = dt; uc.best_edge = edge;
                    }
                    h.t = t; h.abs_t = a; h.slot = slot; best_ref = ref;
                } else if (TRACK && (hit || edge)) {
                    uc.comp_lo = fminf(uc.comp_lo, a - dt);
                }
            }
        }
        go0 = go0 && c0 >= 0; go1 = go1 && c1 >= 0;
        if (go0 && go1) {
            const bool first1 = k1 < k0;
            stack[sp++] = first1 ? c0 : c1;
            node = first1 ? c1 : c0;
        } else if (go0 || go1) {
            node = go0 ? c0 : c1;
        } else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
    uncertain = closest_uncertain32<TRACK>(h, uc);
    return h;
}

template <bool TRACK, bool COUNT>
__global__ void __launch_bounds__(PTB_BVH_THREADS)
k_closest_bvh32(DScene sc, const float4 *__restrict__ qo, const float4 *__restrict__ qd, const unsigned *__restrict__ d_count,
                float *__restrict__ hit_t, int *__restrict__ hit_tri, unsigned *unc_count, int *unc_list, float dt_scale) {
    const unsigned count = *d_count;
    unsigned nn = 0, nt = 0;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < count; i += gridDim.x * blockDim.x) {
        const float4 a = qo[i], b = qd[i];
        bool unc; int ref;
        const Hit32 h = closest_bvh32<TRACK, COUNT>(sc, mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), dt_scale, unc, ref, nn, nt);
        hit_t[i] = h.t;
        hit_tri[i] = h.slot >= 0 ? ref : -1;
        if (TRACK && unc) unc_list[atomicAdd(unc_count, 1u)] = (int)i;
    }
    if (COUNT) {
        for (int o = 16; o > 0; o >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, o); nt += __shfl_xor_sync(0xffffffffu, nt, o); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(sc.counters + CTR_NODES, (unsigned long long)nn); atomicAdd(sc.counters + CTR_TRIS, (unsigned long long)nt); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// ---------------------------------------------------------------- FP32 shadow query
// want_min: report the lowest reference index among occluders (light sample 2), else any occluder.
template <bool COUNT>
__device__ __forceinline__ int occluder_bvh32(const DScene &sc, f3 o, f3 d, float d2max, bool want_min, unsigned &n_nodes, unsigned &n_tris) {
    LineBox<float> lb; lb.set(o.x, o.y, o.z, d.x, d.y, d.z);
    const float lim = sqrtf(d2max) * 1.000001f;
    int found = 0x7fffffff;
    int stack[PTB_BVH_STACK];
    int sp = 0, node = 0;
    while (true) {
        const NodeBoxes n = ld_node(sc.nodes, node);
        if (COUNT) n_nodes++;
        float tn0, tf0, tn1, tf1;
        slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
        slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
        // only subtrees that still hold a lower reference index than the best occluder matter
        bool go0 = (tn0 <= tf0) && (near_abs(tn0, tf0) <= lim) && (n.d.z < found) && (n.d.z < sc.n_tri_obj);
        bool go1 = (tn1 <= tf1) && (near_abs(tn1, tf1) <= lim) && (n.d.w < found) && (n.d.w < sc.n_tri_obj);
        const int c0 = n.d.x, c1 = n.d.y;
#pragma unroll
        for (int side = 0; side < 2; ++side) {
            const bool go = side ? go1 : go0;
            const int c = side ? c1 : c0;
            if (go && c < 0) {
                const float4 *tp = reinterpret_cast<const float4 *>(sc.tri + ~c);
                const float4 q0 = __ldg(tp), q1 = __ldg(tp + 1), q2 = __ldg(tp + 2);
                if (COUNT) n_tris++;
                float t, dt; bool edge;
                const bool hit = tri_test32<false>(q0, q1, q2, 0.f, o, d, 0.f, t, edge, dt);
                const float t2 = t * t;
                if (hit && t2 >= PTB_ZERO_F && t2 < d2max) {
                    const int ref = side ? n.d.w : n.d.z;       // leaf: min-ref field is the triangle itself
                    if (!want_min) return ref;
                    found = min(found, ref);
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

}  // namespace ptb
