// ptb_bvh_dyn.cuh - persistent LBVH traversal kernels with dynamic ray fetch.
//
// Secondary and shadow rays are incoherent and their traversal lengths have a heavy tail (a line that
// runs close to a tessellated wall enters hundreds of leaf boxes; the mean is ~50 nodes).  With one
// fixed ray per lane a warp runs as long as its longest ray: ncu showed 3.8 of 32 lanes active
// (profiles/r01c).  Here every warp is persistent: a lane that finishes its ray leaves the inner loop,
// and as soon as fewer than PTB_DYN_MIN lanes are still traversing the warp breaks out and the idle lanes
// claim the next rays from a global counter (one atomic per refill).  Results are written per ray when
// it terminates, so the processing order does not matter.
//
// Shadow rays are scheduled per (path, light sample) item - 3 independent items per shaded path - and
// only record the outcome (occluding object or -1); k_nee_finish32 then completes compute_color
// (main.py:142-145) per path.  For light sample 2 the query is "lowest OBJECT index among occluders"
// (the leaked loop variable, main.py:42, 70): once an occluder of object k is known, only subtrees that
// still contain a triangle of an earlier object (min reference index < first triangle of k) are entered.
#pragma once
#include "ptb_bvh_kernels.cuh"

namespace ptb {

#ifndef PTB_DYN_MIN
#define PTB_DYN_MIN 14     // refill when fewer lanes than this are still traversing (swept 8..26 on the 1 M-triangle scene)
#endif
#ifndef PTB_BVH_MIN_BLOCKS
#define PTB_BVH_MIN_BLOCKS 10
#endif
#ifndef PTB_SHADOW_ORDER
#define PTB_SHADOW_ORDER 2  // children of a shadow-ray node: 0 stored order, 1 nearer first, 2 farther first
#endif
#ifndef PTB_WW_MIN
#define PTB_WW_MIN 16      // leave the node loop for the triangle tests when fewer lanes than this are still walking (swept 4..24)
#endif
// straight-line (predicated) descend / push / pop step instead of a three-way branch: measured -6 % on the shadow
// kernel's time, +3 % on the closest-hit kernel's (whose step also carries the leaf tests), so only the former uses it
#ifndef PTB_FLAT_STEP_SHADOW
#define PTB_FLAT_STEP_SHADOW 1
#endif
#ifndef PTB_FLAT_STEP_CLOSEST
#define PTB_FLAT_STEP_CLOSEST 0
#endif
#ifndef PTB_SHADOW_GROUP
#define PTB_SHADOW_GROUP 32
#endif
// closest-hit kernel: one pending-leaf slot per lane.  A leaf child is parked in the slot instead of being tested on the
// spot; the triangle block runs for the whole warp when some lane needs its slot a second time (or finishes its ray with
// the slot occupied) - about every tenth node step with half the lanes taking part, instead of two steps out of three
// with one or two lanes (a lane meets a leaf in 7 % of its node steps, a 16-lane warp in 69 % of them).  Costs: the hit is
// known a few steps later, so the pruning bound shrinks later and a ray visits a few more nodes (counted: PTB_FLAG_COUNT_NODES).
#ifndef PTB_CLOSEST_DEFER
#define PTB_CLOSEST_DEFER 1
#endif
#define PTB_WW_DONE ((int)0x80000000)   // traversal state: nothing left (leaf codes are ~slot > INT_MIN)

// ---------------------------------------------------------------- the top of the tree in shared memory (PTB_TOP_K > 0)
#if PTB_TOP_K > 0
#define PTB_DYN_SMEM ((size_t)PTB_TOP_K * sizeof(TopNode) + 16)
// One TMA bulk copy (cp.async.bulk global -> shared, completion on an mbarrier) brings the breadth-first top of a tree
// into this block's shared memory; every thread of the block waits on the barrier's phase 0 before its first fetch.
__device__ __forceinline__ const TopNode *stage_top(const TopNode *__restrict__ top, int n_top) {
    extern __shared__ __align__(128) unsigned char s_top_raw[];
    const unsigned dst = (unsigned)__cvta_generic_to_shared(s_top_raw);
    const unsigned mb = dst + (unsigned)(PTB_TOP_K * sizeof(TopNode));
    const unsigned bytes = (unsigned)n_top * (unsigned)sizeof(TopNode);
    if (n_top <= 0) return reinterpret_cast<const TopNode *>(s_top_raw);     // no top (>= 2^24 triangles): walks start in the full array
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(top), "r"(bytes), "r"(mb) : "memory");
    }
    __syncthreads();                    // the barrier is initialised before anybody polls it
    unsigned done = 0;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(mb) : "memory");
    } while (!done);
    return reinterpret_cast<const TopNode *>(s_top_raw);
}
// a node of either range: top nodes (< PTB_TOP_OFF) out of shared memory, the rest with the two 256-bit loads of ld_node,
// their inner links shifted into the full-array range
__device__ __forceinline__ NodeBoxes ld_node_top(const BvhNode *__restrict__ nodes, const TopNode *s_top, int node) {
    NodeBoxes n;
    if (node < PTB_TOP_OFF) {
        const TopNode *t = s_top + node;
        n.a = t->a; n.b = t->b; n.c = t->c;
        const int4 d = t->d;
        n.d = make_int4(d.x, d.y, d.z & 0xffffff, d.w & 0xffffff);
        n.pid0 = (int)((unsigned)d.z >> 24); n.pid1 = (int)((unsigned)d.w >> 24);
    } else {
        n = ld_node(nodes, node - PTB_TOP_OFF);
        if (n.d.x >= 0) n.d.x += PTB_TOP_OFF;
        if (n.d.y >= 0) n.d.y += PTB_TOP_OFF;
    }
    return n;
}
#define PTB_LD_NODE(nodes, s_top, node) ld_node_top(nodes, s_top, node)
#define PTB_ROOT root_node
#else
#define PTB_DYN_SMEM ((size_t)0)
#define PTB_LD_NODE(nodes, s_top, node) ld_node(nodes, node)
#define PTB_ROOT 0
#endif

// warp-wide claim of new work items for the lanes with want == true; returns the claimed index or
// 0xffffffff when the pool [0, count) is exhausted.
__device__ __forceinline__ unsigned claim_work(bool want, unsigned *work, unsigned count) {
    const unsigned need = __ballot_sync(0xffffffffu, want);
    if (!need) return 0xffffffffu;
    const unsigned lane = threadIdx.x & 31;
    const int leader = __ffs(need) - 1;
    unsigned base = 0;
    if ((int)lane == leader) base = atomicAdd(work, (unsigned)__popc(need));
    base = __shfl_sync(0xffffffffu, base, leader);
    const unsigned idx = base + __popc(need & ((1u << lane) - 1u));
    return (want && idx < count) ? idx : 0xffffffffu;
}

// ---------------------------------------------------------------- closest hit
// PHYS (physical mode): ray semantics, only t > tmin counts; boxes behind the origin are skipped.
// PID: origin_pid[ray] = supporting-plane id of the triangle the ray starts on (0: none).  Children whose triangles all
// lie in that plane are skipped without a box test: the line meets them at t = 0, inside the self-hit guard
// (ptb_scene.cuh, k_plane_ids).  For rays leaving a wall this removes the walk down to the origin's own leaf.
template <bool TRACK, bool COUNT, bool PHYS = false, bool PID = false>
__global__ void __launch_bounds__(PTB_BVH_THREADS, PTB_BVH_MIN_BLOCKS)
k_closest_bvh32_dyn(DScene sc, const float4 *__restrict__ qo, const float4 *__restrict__ qd, const unsigned *__restrict__ d_count,
                    float *__restrict__ hit_t, int *__restrict__ hit_tri, unsigned *unc_count, int *unc_list, float dt_scale,
                    unsigned *work, const unsigned *__restrict__ order, const int *__restrict__ origin_pid = nullptr) {
    const unsigned count = *d_count;
#if PTB_TOP_K > 0
    const TopNode *s_top = stage_top(sc.top, sc.top_n);
    const int root_node = sc.top_n > 0 ? 0 : PTB_TOP_OFF;
#endif
    unsigned nn = 0, nt = 0;
    int stack[PTB_BVH_STACK];
    int sp = 0, node = PTB_ROOT, best_ref = 0x7fffffff, opid = -1, pend = 0;
    unsigned ray = 0xffffffffu;
    bool exhausted = false;                 // warp-uniform: the pool ran dry
    Hit32 h; h.t = 0.f; h.abs_t = 0.f; h.slot = -1;
    Unc32 uc; uc.init();
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(1.f, 0.f, 0.f);
    LineBox<float> lb; lb.set(0.f, 0.f, 0.f, 1.f, 1.f, 1.f);
    while (true) {
        if (!exhausted) {
            const bool want = ray == 0xffffffffu;
            const unsigned idx = claim_work(want, work, count);
            if (want && idx != 0xffffffffu) {
                ray = order ? order[idx] : idx;
                const float4 a = qo[ray], b = qd[ray];
                o = mk3(a.x, a.y, a.z); d = mk3(b.x, b.y, b.z);
                lb.set(o.x, o.y, o.z, d.x, d.y, d.z);
                h.t = 0.f; h.abs_t = 3.0e38f; h.slot = -1; best_ref = 0x7fffffff;
                uc.init();
                sp = 0; node = PTB_ROOT; pend = 0;
                if (PID) { const int pr = origin_pid[ray]; opid = pr ? pr : -1; }
            }
            exhausted = __any_sync(0xffffffffu, want && idx == 0xffffffffu);
        }
        if (!__any_sync(0xffffffffu, ray != 0xffffffffu)) break;
        while (ray != 0xffffffffu) {
            const NodeBoxes n = PTB_LD_NODE(sc.nodes, s_top, node);
            if (COUNT) nn++;
            float tn0, tf0, tn1, tf1;
            slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
            slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
            const int c0 = n.d.x, c1 = n.d.y, m0 = n.d.z, m1 = n.d.w;
            const float lim = TRACK ? h.abs_t + uc.best_dt : h.abs_t;
            const float k0 = PHYS ? fmaxf(tn0, 0.f) : near_abs(tn0, tf0), k1 = PHYS ? fmaxf(tn1, 0.f) : near_abs(tn1, tf1);
            bool go0 = (tn0 <= tf0) && (k0 <= lim), go1 = (tn1 <= tf1) && (k1 <= lim);
            if (PHYS) { go0 = go0 && (tf0 >= PTB_TMIN_SAFE); go1 = go1 && (tf1 >= PTB_TMIN_SAFE); }
            else if (!TRACK) {   // a box the line leaves before |t| reaches the self-hit guard holds no acceptable hit (main.py:100)
                go0 = go0 && (far_abs(tn0, tf0) >= PTB_TMIN_SAFE);
                go1 = go1 && (far_abs(tn1, tf1) >= PTB_TMIN_SAFE);
            }
            if (PID) { go0 = go0 && (n.pid0 != opid); go1 = go1 && (n.pid1 != opid); }
            if (PTB_CLOSEST_DEFER && !TRACK) {
                int l0 = (go0 && c0 < 0) ? c0 : 0, l1 = (go1 && c1 < 0) ? c1 : 0;      // incoming leaves (codes are < 0)
                if (pend == 0) { if (l0) { pend = l0; l0 = 0; } else { pend = l1; l1 = 0; } }
                // this step ends the ray when nothing is left to descend into or pop: its parked leaf must be tested first
                const bool ending = !((go0 && c0 >= 0) || (go1 && c1 >= 0)) && sp == 0;
                bool overflow = ((l0 | l1) != 0) || (ending && pend != 0);
                while (__any_sync(__activemask(), overflow)) {
                    if (pend) {
                        const int slot = ~pend;
                        const float4 *tp = reinterpret_cast<const float4 *>(sc.tri + slot);
                        float4 q0, q1, q2, q3;
                        ld256(tp, q0, q1); ld256(tp + 2, q2, q3);
                        if (COUNT) nt++;
                        float t, dt = 0.f; bool edge = false;
                        const bool hit = tri_test32<false>(q0, q1, q2, 0.f, o, d, dt_scale, t, edge, dt);
                        const int ref = __float_as_int(q3.w);
                        const float a = fabsf(t);
                        const bool ok = hit && (PHYS ? (t > PTB_TMIN_F) : (t * t > PTB_ZERO_F));
                        if (ok && (a < h.abs_t || (a == h.abs_t && ref < best_ref))) { h.t = t; h.abs_t = a; h.slot = slot; best_ref = ref; }
                        pend = 0;
                    }
                    if (l0) { pend = l0; l0 = 0; } else if (l1) { pend = l1; l1 = 0; }
                    overflow = ((l0 | l1) != 0) || (ending && pend != 0);
                }
            } else {
                // leaf children: one shared test block walked up to twice, so lanes with a single leaf child (either side)
                // run it together instead of the warp stepping through a copy per side
                const bool leaf0 = go0 && c0 < 0, leaf1 = go1 && c1 < 0;
                int lc = leaf0 ? c0 : c1, lref = leaf0 ? m0 : m1;
                int n_leaf = (leaf0 ? 1 : 0) + (leaf1 ? 1 : 0);
#pragma unroll 1
                while (n_leaf > 0) {
                    const int slot = ~lc;
                    const float4 *tp = reinterpret_cast<const float4 *>(sc.tri + slot);
                    float4 q0, q1, q2, q3;
                    ld256(tp, q0, q1);
                    if (TRACK) ld256(tp + 2, q2, q3); else q2 = __ldg(tp + 2);
                    const float eps = TRACK ? q3.x : 0.f;
                    if (COUNT) nt++;
                    float t, dt = 0.f; bool edge = false;
                    const bool hit = tri_test32<TRACK>(q0, q1, q2, eps, o, d, dt_scale, t, edge, dt);
                    const int ref = lref;                      // leaf: the min-ref field is the triangle itself
                    const float a = fabsf(t);
                    const bool ok = hit && (PHYS ? (t > PTB_TMIN_F) : (t * t > PTB_ZERO_F));
                    if (ok && (a < h.abs_t || (a == h.abs_t && ref < best_ref))) {
                        if (TRACK) {
                            if (h.slot >= 0) uc.comp_lo = fminf(uc.comp_lo, h.abs_t - uc.best_dt);
                            uc.best_dt = dt; uc.best_edge = edge;
                        }
                        h.t = t; h.abs_t = a; h.slot = slot; best_ref = ref;
                    } else if (TRACK && (hit || edge)) {
                        uc.comp_lo = fminf(uc.comp_lo, a - dt);
                    }
                    lc = c1; lref = m1;                        // a second leaf can only be child 1
                    --n_leaf;
                }
            }
            go0 = go0 && c0 >= 0; go1 = go1 && c1 >= 0;
#if PTB_FLAT_STEP_CLOSEST
            // one straight-line step for all lanes: descend into the nearer wanted child, push the other, or pop
            const bool both = go0 && go1, any = go0 || go1;
            const bool pick1 = both ? (k1 < k0) : go1;
            if (both) stack[sp] = pick1 ? c0 : c1;
            sp += both ? 1 : 0;
            const bool pop = !any && sp > 0;
            sp -= pop ? 1 : 0;
            int nxt = pick1 ? c1 : c0;
            if (pop) nxt = stack[sp];
            if (any || pop) node = nxt;
            else {                                              // ray finished
#else
            if (go0 && go1) {
                const bool first1 = k1 < k0;
                const int later = first1 ? c0 : c1;
                stack[sp++] = later;
                node = first1 ? c1 : c0;
            } else if (go0 || go1) {
                node = go0 ? c0 : c1;
            } else if (sp > 0) {
                node = stack[--sp];
            } else {                                            // ray finished
#endif
                hit_t[ray] = h.t;
                // without TRACK nothing else reads h.slot: 'a hit was found' is best_ref != INT_MAX, and the register h.slot would
                // occupy is one the kernel does not have (closest-hit kernel -4 %, profiles/r02n_ab_tos_prefetch.log, `base`)
                hit_tri[ray] = (TRACK ? h.slot >= 0 : best_ref != 0x7fffffff) ? best_ref : -1;
                if (TRACK && closest_uncertain32<TRACK>(h, uc)) unc_list[atomicAdd(unc_count, 1u)] = (int)ray;
                ray = 0xffffffffu;
            }
            if (!exhausted && __popc(__activemask()) < PTB_DYN_MIN) break;
        }
    }
    if (COUNT) {
        for (int s = 16; s > 0; s >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, s); nt += __shfl_xor_sync(0xffffffffu, nt, s); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(sc.counters + CTR_NODES, (unsigned long long)nn); atomicAdd(sc.counters + CTR_TRIS, (unsigned long long)nt); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(sc.counters + CTR_CLOSEST, (unsigned long long)count);
}

// ---------------------------------------------------------------- shadow rays: one item per (path, light sample)
// sh_res[s * cap + j] = occluding object index (lowest one for s == 2), or -1
// PHYS: plain any-hit ray query for every light sample (t in (tmin, distance)); the outcome array holds the
// sample's FLOAT weight written by the shade kernel and an occluded sample is zeroed (ptb_phys.cuh).
template <bool COUNT, bool PHYS = false, bool PID = false>
__global__ void __launch_bounds__(PTB_BVH_THREADS, PTB_BVH_MIN_BLOCKS)
k_shadow_bvh32_dyn(DScene sc, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d, const unsigned *__restrict__ d_count,
                   int *__restrict__ sh_res, int cap, unsigned *work, const unsigned *__restrict__ order,
                   const int *__restrict__ origin_pid = nullptr) {
    // occlusion tree: object triangles only (main.py:42); the full tree + min-ref pruning when it does not exist
    const BvhNode *__restrict__ nodes = sc.nodes_occ ? sc.nodes_occ : sc.nodes;
    const TriRec *__restrict__ tris = sc.nodes_occ ? sc.tri_occ : sc.tri;
    const unsigned paths = *d_count;
#if PTB_TOP_K > 0
    const TopNode *s_top = sc.nodes_occ ? stage_top(sc.top_occ, sc.top_occ_n) : stage_top(sc.top, sc.top_n);
    const int root_node = (sc.nodes_occ ? sc.top_occ_n : sc.top_n) > 0 ? 0 : PTB_TOP_OFF;
#endif
#if PTB_SHADOW_GROUP
    // work items in groups of PTB_SHADOW_GROUP paths x 3 samples: the lanes that finish sample s of a group of neighbouring
    // paths claim sample s + 1 of the same paths next, while the nodes those paths walk are still in this SM's L1
    const unsigned count = ((paths + PTB_SHADOW_GROUP - 1u) / PTB_SHADOW_GROUP) * (3u * PTB_SHADOW_GROUP);
#else
    const unsigned count = 3u * paths;
#endif
    unsigned nn = 0, nt = 0;
    int stack[PTB_BVH_STACK];
    int sp = 0, node = PTB_ROOT, found = 0x7fffffff, found_obj = -1, opid = -1;
    unsigned item = 0xffffffffu, out_slot = 0;
    bool exhausted = false, want_min = false;
    f3 o = mk3(0.f, 0.f, 0.f), d = mk3(1.f, 0.f, 0.f);
    float d2max = 0.f, lim = 0.f;
    LineBox<float> lb; lb.set(0.f, 0.f, 0.f, 1.f, 1.f, 1.f);
    while (true) {
        if (!exhausted) {
            const bool want = item == 0xffffffffu;
            const unsigned idx = claim_work(want, work, count);
#if PTB_SHADOW_GROUP
            const unsigned grp = idx / (3u * PTB_SHADOW_GROUP), rem = idx - grp * (3u * PTB_SHADOW_GROUP);
            const unsigned s = rem / PTB_SHADOW_GROUP, jj = grp * PTB_SHADOW_GROUP + (rem - s * PTB_SHADOW_GROUP);
            if (want && idx != 0xffffffffu && jj < paths) {
                item = idx;
#else
            if (want && idx != 0xffffffffu) {
                item = idx;
                const unsigned s = idx / paths, jj = idx - s * paths;    // sample-major: neighbouring lanes, neighbouring paths
#endif
                const unsigned j = order ? order[jj] : jj;
                out_slot = s * (unsigned)cap + j;
                const float4 pk = sh_p[j], dv = sh_d[out_slot];
                o = mk3(pk.x, pk.y, pk.z); d = mk3(dv.x, dv.y, dv.z);
                d2max = dv.w; lim = sqrtf(dv.w) * 1.000001f;
                lb.set(o.x, o.y, o.z, d.x, d.y, d.z);
                want_min = !PHYS && s == 2u;
                found = 0x7fffffff; found_obj = -1;
                sp = 0; node = PTB_ROOT;
                if (PID) { const int pr = origin_pid[j]; opid = pr ? pr : -1; }
            }
            exhausted = __any_sync(0xffffffffu, want && idx == 0xffffffffu);
        }
        if (!__any_sync(0xffffffffu, item != 0xffffffffu)) break;
        // while-while traversal: lanes walk inner nodes together; a lane that reaches a leaf parks on it (leaves travel
        // through the stack like inner nodes, encoded ~slot) and the triangle tests run when few lanes are still walking,
        // so the ~80-instruction test block executes with many lanes instead of one or two per node step.
        while (item != 0xffffffffu) {
            while (node >= 0) {
                const NodeBoxes n = PTB_LD_NODE(nodes, s_top, node);
                if (COUNT) nn++;
                float tn0, tf0, tn1, tf1;
                slab(lb, n.a.x, n.a.y, n.a.z, n.a.w, n.b.x, n.b.y, tn0, tf0);
                slab(lb, n.b.z, n.b.w, n.c.x, n.c.y, n.c.z, n.c.w, tn1, tf1);
                const int c0 = n.d.x, c1 = n.d.y, m0 = n.d.z, m1 = n.d.w;
                const int cut = min(found, sc.n_tri_obj);
                bool go0 = (tn0 <= tf0) && (tn0 <= lim) && (m0 < cut);
                bool go1 = (tn1 <= tf1) && (tn1 <= lim) && (m1 < cut);
                if (PHYS) { go0 = go0 && (tf0 >= PTB_TMIN_SAFE); go1 = go1 && (tf1 >= PTB_TMIN_SAFE); }
                else {
                    go0 = go0 && (tf0 >= -lim) && (fmaxf(-tn0, tf0) >= PTB_TMIN_SAFE);
                    go1 = go1 && (tf1 >= -lim) && (fmaxf(-tn1, tf1) >= PTB_TMIN_SAFE);
                }
                if (PID) { go0 = go0 && (n.pid0 != opid); go1 = go1 && (n.pid1 != opid); }
#if PTB_FLAT_STEP_SHADOW
                {
                    const bool both = go0 && go1, any = go0 || go1;
                    const float k0 = PHYS ? fmaxf(tn0, 0.f) : near_abs(tn0, tf0), k1 = PHYS ? fmaxf(tn1, 0.f) : near_abs(tn1, tf1);
                    const bool pick1 = both ? (k1 > k0) : go1;          // any-hit order: FARTHER child first (see below)
                    if (both) stack[sp] = pick1 ? c0 : c1;
                    sp += both ? 1 : 0;
                    const bool pop = !any && sp > 0;
                    sp -= pop ? 1 : 0;
                    int nxt = pick1 ? c1 : c0;
                    if (pop) nxt = stack[sp];
                    node = (any || pop) ? nxt : PTB_WW_DONE;
                }
#else
                if (go0 && go1) {
#if PTB_SHADOW_ORDER == 0
                    stack[sp++] = c1; node = c0;
#else
                    // any-hit order: FARTHER child first.  The boxes around the origin are the expensive ones (every
                    // ancestor of the shaded point is entered, tilted faces have fat boxes) and they rarely occlude: the
                    // point's own surface is rejected by the self-hit guard.  Occluders sit away from the origin.
                    const float k0 = PHYS ? fmaxf(tn0, 0.f) : near_abs(tn0, tf0), k1 = PHYS ? fmaxf(tn1, 0.f) : near_abs(tn1, tf1);
                    const bool first1 = PTB_SHADOW_ORDER == 2 ? (k1 > k0) : (k1 < k0);
                    const int later = first1 ? c0 : c1;
                    stack[sp++] = later; node = first1 ? c1 : c0;
#endif
                }
                else if (go0 || go1) node = go0 ? c0 : c1;
                else if (sp > 0) node = stack[--sp];
                else node = PTB_WW_DONE;
#endif
                if (__popc(__activemask()) < PTB_WW_MIN) break;
            }
            if (node < 0 && node != PTB_WW_DONE) {                   // parked on a leaf: ~slot
                const float4 *tp = reinterpret_cast<const float4 *>(tris + ~node);
                float4 q0, q1, q2, q3;
                ld256(tp, q0, q1); ld256(tp + 2, q2, q3);
                const int ref = __float_as_int(q3.w);
                if (COUNT) nt++;
                float t, dt; bool edge;
                const bool hit = tri_test32<false>(q0, q1, q2, 0.f, o, d, 0.f, t, edge, dt);
                const float t2 = t * t;
                bool finished = false;
                if (hit && t2 >= PTB_ZERO_F && t2 < d2max && (!PHYS || t > 0.f)) {
                    if (ref < found) {
                        found_obj = __ldg(sc.tri_obj + ref);
                        found = __ldg(sc.obj_first + found_obj);    // first triangle of that object
                    }
                    if (!want_min || found == 0) finished = true;
                }
                node = (finished || sp == 0) ? PTB_WW_DONE : stack[--sp];
            }
            if (node == PTB_WW_DONE) {
                if (!PHYS) sh_res[out_slot] = found_obj;
                else if (found_obj >= 0) sh_res[out_slot] = 0;      // 0 == 0.0f: the sample's weight is cancelled
                item = 0xffffffffu;
            }
            if (!exhausted && __popc(__activemask()) < PTB_DYN_MIN) break;
        }
    }
    if (COUNT) {
        for (int s = 16; s > 0; s >>= 1) { nn += __shfl_xor_sync(0xffffffffu, nn, s); nt += __shfl_xor_sync(0xffffffffu, nt, s); }
        if ((threadIdx.x & 31) == 0) { atomicAdd(sc.counters + CTR_NODES_SH, (unsigned long long)nn); atomicAdd(sc.counters + CTR_TRIS_SH, (unsigned long long)nt); }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(sc.counters + CTR_SHADOW, (unsigned long long)(3u * paths));
        atomicAdd(sc.counters + CTR_SHADED, (unsigned long long)paths);
    }
}

// ---------------------------------------------------------------- NEE finish (main.py:61-73, 142-145, 230-231)
__global__ void __launch_bounds__(256)
k_nee_finish32(DScene sc, const float4 *__restrict__ sh_p, const float4 *__restrict__ sh_d, const int2 *__restrict__ sh_aux,
               const int *__restrict__ sh_res, const unsigned *__restrict__ d_count, float4 *__restrict__ L, int cap, int bounce,
               RecPtrs rec) {
    const unsigned count = *d_count;
    const float lc0 = (float)sc.light_color[0], lc1 = (float)sc.light_color[1], lc2 = (float)sc.light_color[2];
    const float ambient = (float)sc.ambient;
    for (unsigned j = blockIdx.x * blockDim.x + threadIdx.x; j < count; j += gridDim.x * blockDim.x) {
        const float4 pk = sh_p[j];
        const int2 aux = sh_aux[j];
        const float4 sn = __ldg(sc.tri_sn + aux.y);
        const int obj = __float_as_int(sn.w);
        float dot_sum = 0.f;
        int occ = 0, leak = sc.n_obj - 1;
#pragma unroll
        for (int s = 0; s < 3; ++s) {
            const int r = sh_res[(size_t)s * cap + j];
            if (r >= 0) { occ |= 1 << s; if (s == 2) leak = r; }
            else {
                const float4 dv = sh_d[(size_t)s * cap + j];
                dot_sum += fmaf(dv.z, sn.z, fmaf(dv.y, sn.y, dv.x * sn.x));
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
}

}  // namespace ptb
