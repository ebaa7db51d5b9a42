// ptb_common.cuh - shared device/host helpers: status handling, 3-vectors in float and double,
// Philox4x32-10.  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <atomic>
#include <mutex>

#include "../../include/ptb.h"

namespace ptb {

// ---------------------------------------------------------------- host-side status
void set_error(const char *fmt, ...);
extern std::atomic<unsigned long long> g_launches;   // kernels launched by this library (host counter)

// Per-device state of the library.  include/ptb.h lets any host thread call in on any device and stream, so nothing that
// CUDA keeps per device (kernel attributes, the SM count, the scene memory pool) may hide behind a process-wide flag.
#define PTB_MAX_DEVICES 64
#define PTB_POOL_KEEP_BYTES (16ull << 30)   // unused scene memory the private pool keeps across synchronisations
struct DeviceState {
    std::atomic<bool> ready;        // sms / pool filled in
    int sms;                        // multiprocessor count (persistent grids are multiples of it)
    std::atomic<bool> sort_optin;   // k_ps_scatter: 64 KB of dynamic shared memory asked for on this device
    std::atomic<bool> planes_optin; // plane-list kernels: more than 48 KB asked for on this device
    cudaMemPool_t pool;             // private stream-ordered pool of the scene memory (null: creation failed, plain cudaMallocAsync)
    size_t l2_persist_bytes;        // L2 set aside for persisting lines (PTB_L2_PERSIST_MB; 0: the feature is off)
    size_t l2_window_max;           // largest access-policy window the device takes
};
DeviceState *device_state();        // of the CURRENT device (ptb_api.cu)
extern std::mutex g_state_mu;       // guards first-use initialisation of a DeviceState and the profiler's lists

#define PTB_CUDA(call)                                                                        \
    do {                                                                                      \
        cudaError_t e_ = (call);                                                              \
        if (e_ != cudaSuccess) {                                                              \
            ptb::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
            return PTB_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)

#define PTB_LAUNCHED()                                                                        \
    do {                                                                                      \
        ++ptb::g_launches;                                                                    \
        cudaError_t e_ = cudaGetLastError();                                                  \
        if (e_ != cudaSuccess) {                                                              \
            ptb::set_error("%s:%d launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); \
            return PTB_ERR_CUDA;                                                              \
        }                                                                                     \
    } while (0)

// ---------------------------------------------------------------- 3-vectors
template <typename T> struct V3 { T x, y, z; };
typedef V3<float> f3;
typedef V3<double> d3;

template <typename T> __host__ __device__ __forceinline__ V3<T> mk3(T x, T y, T z) { V3<T> r; r.x = x; r.y = y; r.z = z; return r; }
template <typename T> __host__ __device__ __forceinline__ V3<T> operator+(V3<T> a, V3<T> b) { return mk3<T>(a.x + b.x, a.y + b.y, a.z + b.z); }
template <typename T> __host__ __device__ __forceinline__ V3<T> operator-(V3<T> a, V3<T> b) { return mk3<T>(a.x - b.x, a.y - b.y, a.z - b.z); }
template <typename T> __host__ __device__ __forceinline__ V3<T> operator*(V3<T> a, T s) { return mk3<T>(a.x * s, a.y * s, a.z * s); }

// FP32: free contraction is fine (results are judged by tolerance).
__device__ __forceinline__ float dot(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ f3 cross(f3 a, f3 b) {
    return mk3<float>(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}

// FP64 "exact" flavour: one IEEE rounding per reference operation, never contracted by the compiler.
// dot follows the BLAS ddot evaluation numpy uses for 3-vectors: fma(a2,b2, fma(a1,b1, a0*b0)).
__device__ __forceinline__ double xmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double xadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double xsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double xdiv(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double xsqrt(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ double xdot(d3 a, d3 b) { return __fma_rn(a.z, b.z, __fma_rn(a.y, b.y, __dmul_rn(a.x, b.x))); }
__device__ __forceinline__ double xnorm(d3 a) { return xsqrt(xdot(a, a)); }
__device__ __forceinline__ d3 xsub3(d3 a, d3 b) { return mk3<double>(xsub(a.x, b.x), xsub(a.y, b.y), xsub(a.z, b.z)); }
__device__ __forceinline__ d3 xdiv3(d3 a, double s) { return mk3<double>(xdiv(a.x, s), xdiv(a.y, s), xdiv(a.z, s)); }
__device__ __forceinline__ d3 xcross(d3 a, d3 b) {   // np.cross: two products, one subtraction each
    return mk3<double>(xsub(xmul(a.y, b.z), xmul(a.z, b.y)), xsub(xmul(a.z, b.x), xmul(a.x, b.z)),
                       xsub(xmul(a.x, b.y), xmul(a.y, b.x)));
}
// utils.py:48-49 squared_dist: sum from 0, left to right
__device__ __forceinline__ double xsqdist(d3 a, d3 b) {
    double dx = xsub(a.x, b.x), dy = xsub(a.y, b.y), dz = xsub(a.z, b.z);
    return xadd(xadd(xmul(dx, dx), xmul(dy, dy)), xmul(dz, dz));
}

// ---------------------------------------------------------------- Philox4x32-10
// key = (seed, 0), counter = (pixel, sample, bounce, dim/4); uniform u = (word >> 8) * 2^-24.
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
        uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
        c = make_uint4(hi1 ^ c.y ^ k0, lo1, hi0 ^ c.w ^ k1, lo0);
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    return c;
}
__device__ __forceinline__ uint4 rng4(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce, uint32_t quad) {
    return philox4x32_10(make_uint4(pixel, sample, bounce, quad), seed, 0u);
}
template <typename T> __device__ __forceinline__ T u01(uint32_t w) { return (T)(w >> 8) * (T)5.9604644775390625e-08; }

__device__ __forceinline__ float rcp_approx(float x) {
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
// id / d and id % d for a warp-uniform divisor without the integer-division sequence: float estimate
// + one correction step (exact for id < 2^31 whenever id / d < 2^23, which the batch limits guarantee)
__device__ __forceinline__ void fast_divmod(int id, int d, float inv_d, int &q, int &r) {
    q = (int)((float)id * inv_d);
    r = id - q * d;
    if (r < 0) { r += d; --q; }
    else if (r >= d) { r -= d; ++q; }
}

// ---------------------------------------------------------------- pixel-set mapping (ptb_pixels)
struct PixelMap {
    int W, H, shard, n_shards, tile, n;
    const int32_t *list;
    __host__ __device__ __forceinline__ int ray_index(int slot) const {
        if (list) {
#ifdef __CUDA_ARCH__
            return __ldg(list + slot);
#else
            return -1;
#endif
        }
        if (n_shards == 1) return slot;
        int c = slot / tile, w = slot - c * tile;
        return (c * n_shards + shard) * tile + w;
    }
};

inline int64_t pixel_count(const ptb_pixels *px) {
    if (px->d_pixels) return px->n_pixels;
    int64_t total = (int64_t)px->width * px->height;
    if (px->n_shards <= 1) return total;
    int64_t tile = px->tile, stride = tile * px->n_shards;
    int64_t full = total / stride, rem = total - full * stride;
    int64_t mine = full * tile;
    int64_t lo = (int64_t)px->shard * tile;
    if (rem > lo) mine += (rem - lo < tile) ? (rem - lo) : tile;
    return mine;
}

inline PixelMap make_pixel_map(const ptb_pixels *px) {
    PixelMap m;
    m.W = px->width; m.H = px->height;
    m.shard = px->shard; m.n_shards = px->n_shards < 1 ? 1 : px->n_shards;
    m.tile = px->tile < 1 ? 1 : px->tile;
    m.list = px->d_pixels;
    m.n = (int)pixel_count(px);
    return m;
}

}  // namespace ptb
