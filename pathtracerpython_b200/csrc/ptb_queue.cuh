// ptb_queue.cuh - wavefront state in HBM: SoA path queues, hit records, shadow work items.
//
// One batch = (pixels of the call) x (S samples).  Path id = s_local * npix + pixel_slot identifies the
// (pixel, sample) pair for the whole life of the path; queue position changes at every bounce because
// the shade kernel compacts survivors.  All per-path arrays are float4/double4 so every access is one
// (FP32) or two (FP64) 128-bit transactions, coalesced across the warp.
//
//   ray queue (2 copies, ping-pong)  qo[i] = (o.x, o.y, o.z, k)        k = accumulated_k (main.py:190)
//                                    qd[i] = (d.x, d.y, d.z, path id)
//   hit record                       hit_t[i], hit_tri[i]              written by closest-hit kernels
//   shadow work item (slot j of the NEXT queue: a path continues iff it is shaded, main.py:239)
//                                    sh_p[j]    = (P.x, P.y, P.z, k before the bounce)
//                                    sh_d[s][j] = (dir.x, dir.y, dir.z, light_squared_distance)  s = 0..2
//                                    sh_aux[j]  = (path id, hit triangle reference index)
//   per-path radiance                L[path id] = (r, g, b, -)         owned by one thread at a time
#pragma once
#include "ptb_isect.cuh"
#include "ptb_sort.cuh"

namespace ptb {

template <typename T> struct VecT;
template <> struct VecT<float> { typedef float4 v4; };
template <> struct VecT<double> { typedef double4 v4; };

__device__ __forceinline__ float4 mk4(float x, float y, float z, float w) { return make_float4(x, y, z, w); }
__device__ __forceinline__ double4 mk4(double x, double y, double z, double w) { return make_double4(x, y, z, w); }

__device__ __forceinline__ float id_pack(float, int id) { return __int_as_float(id); }
__device__ __forceinline__ double id_pack(double, int id) { return (double)id; }
__device__ __forceinline__ int id_unpack(float w) { return __float_as_int(w); }
__device__ __forceinline__ int id_unpack(double w) { return (int)w; }

template <typename T> struct Queues {
    typedef typename VecT<T>::v4 v4;
    v4 *qo[2], *qd[2];     // ping-pong ray queues
    T *hit_t;
    int *hit_tri;
    T *hit_t2;             // second hit record set: the fused per-bounce kernel writes hits of the NEXT queue
    int *hit_tri2;
    v4 *sh_p, *sh_d;       // sh_d is [3][cap]
    int2 *sh_aux;
    int *sh_res;           // [3][cap] LBVH shadow outcome per (light sample, path): occluding object or -1
    unsigned *work;        // dynamic-fetch counters of the persistent LBVH kernels
    v4 *L;
    unsigned *count;       // [bounces + 2] live-path count entering each bounce
    unsigned *unc_count;   // number of uncertain primary rays
    int *unc_list;         // their queue positions
    float2 *kgb[2];        // physical mode: green / blue throughput of the ray queues (red rides in qo.w)
    float2 *sh_w;          // physical mode: [4][cap] per light sample (diffuse, specular) weights, [3] = throughput g, b
    unsigned *count_sh;    // physical mode: [bounces + 2] shadow work items per bounce (own compaction)
    SortBufs sort;         // per-bounce re-ordering of the live slots (LBVH scenes, ptb_sort.cuh)
    int cap;               // paths per batch
    unsigned work_slot;    // host side: next of the 32 `work` counters to hand to a persistent kernel
    bool count_nodes;      // host side: instrumented run (PTB_FLAG_COUNT_NODES)
};

template <typename T> inline size_t queue_bytes(size_t cap, int bounces) {
    size_t v = sizeof(typename VecT<T>::v4);
    size_t per = 4 * v + sizeof(T) + sizeof(int) + v + 3 * v + sizeof(int2) + v + sizeof(int) + 3 * sizeof(int);
    per += 4 * sizeof(unsigned);                                    // sort keys / values, ping-pong
    per += sizeof(T) + sizeof(int);                                 // second hit record set
    per += 6 * sizeof(float2);                                      // physical mode: kgb x 2, sh_w x 4
    const size_t tiles = (cap + PS_TILE - 1) / PS_TILE;
    return per * cap + 256 * 40 + (size_t)(bounces + 8) * sizeof(unsigned) + tiles * 256 * sizeof(unsigned) + 2048 + (size_t)(bounces + 8) * sizeof(unsigned) + 1024;
}

// carve a workspace into queues; returns bytes used (0 if it does not fit)
template <typename T> inline size_t queue_carve(void *ws, size_t bytes, size_t cap, int bounces, Queues<T> &q) {
    typedef typename VecT<T>::v4 v4;
    char *p = (char *)ws;
    char *end = p + bytes;
    auto take = [&](size_t n) { char *r = p; p += (n + 255) & ~(size_t)255; return (void *)r; };
    q.qo[0] = (v4 *)take(cap * sizeof(v4)); q.qd[0] = (v4 *)take(cap * sizeof(v4));
    q.qo[1] = (v4 *)take(cap * sizeof(v4)); q.qd[1] = (v4 *)take(cap * sizeof(v4));
    q.sh_p = (v4 *)take(cap * sizeof(v4));
    q.sh_d = (v4 *)take(3 * cap * sizeof(v4));
    q.L = (v4 *)take(cap * sizeof(v4));
    q.hit_t = (T *)take(cap * sizeof(T));
    q.hit_tri = (int *)take(cap * sizeof(int));
    q.kgb[0] = (float2 *)take(cap * sizeof(float2)); q.kgb[1] = (float2 *)take(cap * sizeof(float2));
    q.sh_w = (float2 *)take(4 * cap * sizeof(float2));
    q.count_sh = (unsigned *)take((size_t)(bounces + 4) * sizeof(unsigned));
    q.hit_t2 = (T *)take(cap * sizeof(T));
    q.hit_tri2 = (int *)take(cap * sizeof(int));
    q.sh_aux = (int2 *)take(cap * sizeof(int2));
    q.unc_list = (int *)take(cap * sizeof(int));
    q.sh_res = (int *)take(3 * cap * sizeof(int));
    for (int i = 0; i < 2; ++i) { q.sort.keys[i] = (unsigned *)take(cap * sizeof(unsigned)); q.sort.vals[i] = (unsigned *)take(cap * sizeof(unsigned)); }
    q.sort.tiles_max = (int)((cap + PS_TILE - 1) / PS_TILE);
    q.sort.ghist = (unsigned *)take((size_t)q.sort.tiles_max * 256 * sizeof(unsigned));
    q.sort.totals = (unsigned *)take(256 * sizeof(unsigned));
    q.work = (unsigned *)take(256);
    q.count = (unsigned *)take((size_t)(bounces + 4) * sizeof(unsigned));
    q.unc_count = (unsigned *)take(4 * sizeof(unsigned));
    q.cap = (int)cap;
    q.work_slot = 0; q.count_nodes = false;
    return p <= end ? (size_t)(p - (char *)ws) : 0;
}

}  // namespace ptb
