// ptb_isect.cuh - ray/triangle tests with the reference's semantics (utils.py:98-147, 72-91):
//   * the query is a LINE query: no t > 0 test, hits behind the origin count (A-Q1);
//   * |vr . n| <= 1e-5 is a miss (utils.py:117); inside test is strict, edge/vertex = miss (A-Q15).
// FP32 flavour: precomputed plane + barycentric-gradient record (TriRec), ~17 FMA + 1 reciprocal,
//   optionally tracking whether any discrete decision fell inside its FP32 error bound.
// FP64 flavour: the reference's formulas op for op on the parsed vertices (mirrors oracle/pt_oracle.c
//   line by line, one rounding per operation), used for the FP64 mode and for re-tracing uncertain
//   primary rays so that closest-hit triangle ids are bit-exact.
#pragma once
#include "ptb_scene.cuh"

namespace ptb {

#define PTB_ZERO_F 1e-5f
#define PTB_ZERO_D 1e-5
#define PTB_TMIN_F 3.1622776e-3f   // sqrt(1e-5): |t| threshold equivalent to squared_dist > ZERO

// ---------------------------------------------------------------- FP32
struct Hit32 {
    float t;       // signed ray parameter of the closest accepted hit
    float abs_t;   // |t| (ranking key; the reference ranks by squared distance)
    int slot;      // triangle slot (position in the TriRec array) or -1
};

// Uncertainty bookkeeping for one closest-hit query (primary rays only).
struct Unc32 {
    float best_dt;      // error bound of the current best |t|
    float comp_lo;      // smallest lower bound |t| - dt over all competitors of the best
    bool best_edge;     // the best hit's inside/parallel/self-hit decision was within its bound
    __device__ __forceinline__ void init() { best_dt = 0.f; comp_lo = 3.0e38f; best_edge = false; }
};

// One triangle against one line.  Returns true when FP32 arithmetic accepts the hit.
// TRACK: also classify the decision; `edge` = some decision within its error bound (and the point
// plausibly inside), `dt` = error bound of t.
template <bool TRACK>
__device__ __forceinline__ bool tri_test32(const float4 q0, const float4 q1, const float4 q2, const float eps_tri,
                                           const f3 o, const f3 d, const float dt_scale,
                                           float &t, bool &edge, float &dt) {
    const float dn = fmaf(q0.z, d.z, fmaf(q0.y, d.y, q0.x * d.x));
    const float on = q0.w - fmaf(q0.z, o.z, fmaf(q0.y, o.y, q0.x * o.x));
    const float inv = __frcp_rn(dn);
    t = on * inv;
    const float px = fmaf(t, d.x, o.x), py = fmaf(t, d.y, o.y), pz = fmaf(t, d.z, o.z);
    const float u = fmaf(q1.z, pz, fmaf(q1.y, py, fmaf(q1.x, px, q1.w)));
    const float v = fmaf(q2.z, pz, fmaf(q2.y, py, fmaf(q2.x, px, q2.w)));
    const float w = 1.0f - u - v;
    const float m = fminf(fminf(u, v), w);
    const float adn = fabsf(dn);
    const bool hit = (m > 0.0f) && (adn > PTB_ZERO_F);   // distance conditions are the caller's
    if (TRACK) {
        const float g = 1.0f + fabsf(inv);
        const float e = eps_tri * g;
        dt = dt_scale * g;
        const bool plausible = (m > -e) && (adn > 0.8e-5f);   // NaN-safe: false on NaN
        edge = plausible && ((m <= e) || (fabsf(adn - PTB_ZERO_F) < 2e-6f) ||
                             (fabsf(fabsf(t) - 3.1622776e-3f) <= dt));
    }
    return hit;
}

template <bool TRACK>
__device__ __forceinline__ void closest_update32(const bool hit, const float t, const bool edge, const float dt,
                                                 const int slot, Hit32 &h, Unc32 &uc) {
    const float a = fabsf(t);
    if (hit && (t * t > PTB_ZERO_F) && a < h.abs_t) {   // main.py:100 squared_dist > ZERO
        if (TRACK) {
            if (h.slot >= 0) uc.comp_lo = fminf(uc.comp_lo, h.abs_t - uc.best_dt);
            uc.best_dt = dt;
            uc.best_edge = edge;
        }
        h.t = t; h.abs_t = a; h.slot = slot;
    } else if (TRACK && (hit || edge)) {
        uc.comp_lo = fminf(uc.comp_lo, a - dt);
    }
}

template <bool TRACK>
__device__ __forceinline__ bool closest_uncertain32(const Hit32 &h, const Unc32 &uc) {
    if (!TRACK) return false;
    if (h.slot < 0) return uc.comp_lo < 3.0e38f;
    return uc.best_edge || (uc.comp_lo <= h.abs_t + uc.best_dt);
}

// ---------------------------------------------------------------- FP64 (reference formulas)
__device__ __forceinline__ d3 ld3(const double *p) { return mk3<double>(p[0], p[1], p[2]); }

// utils.py:72-91
__device__ __forceinline__ bool in_triangle64(const d3 p, const d3 v1, const d3 v2, const d3 v3) {
    d3 c1 = xcross(xsub3(v1, v2), xsub3(p, v2));
    d3 c2 = xcross(xsub3(v2, v3), xsub3(p, v3));
    d3 c3 = xcross(xsub3(v3, v1), xsub3(p, v1));
    c1 = xdiv3(c1, xnorm(c1));
    c2 = xdiv3(c2, xnorm(c2));
    c3 = xdiv3(c3, xnorm(c3));
    return (xdot(c1, c2) > 0.0) && (xdot(c1, c3) > 0.0);   // NaN -> false
}

// utils.py:98-147 with vr already normalised (utils.py:110).  Returns hit and P.
__device__ __forceinline__ bool tri_test64(const double *tri, const d3 pr, const d3 vr, d3 &P) {
    const d3 v1 = ld3(tri), v2 = ld3(tri + 3), v3 = ld3(tri + 6);
    d3 n = xcross(xsub3(v1, v2), xsub3(v3, v2));
    n = xdiv3(n, xnorm(n));
    const double dn = xdot(vr, n);
    if (!(fabs(dn) > PTB_ZERO_D)) return false;
    const double t = xdiv(xsub(xdot(n, v1), xdot(n, pr)), xdot(n, vr));
    P = mk3<double>(xadd(pr.x, xmul(vr.x, t)), xadd(pr.y, xmul(vr.y, t)), xadd(pr.z, xmul(vr.z, t)));
    return in_triangle64(P, v1, v2, v3);
}

struct Hit64 {
    double d2;   // squared distance (main.py:103), the reference's ranking key
    d3 P;
    int tri;     // reference triangle index or -1
};

// main.py:100-107, 115: accept d2 > 1e-5, keep the FIRST minimum in reference triangle order.
// Callers that visit triangles out of order (BVH) must pass `tri` = reference index so ties resolve
// to the lowest index.
__device__ __forceinline__ void closest_update64(const bool hit, const d3 P, const d3 pr, const int tri, Hit64 &h) {
    if (!hit) return;
    const double d2 = xsqdist(P, pr);
    if (!(d2 > PTB_ZERO_D)) return;
    if (h.tri < 0 || d2 < h.d2 || (d2 == h.d2 && tri < h.tri)) { h.d2 = d2; h.P = P; h.tri = tri; }
}

}  // namespace ptb
