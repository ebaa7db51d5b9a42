// ptb_sort.cuh - per-bounce ray re-ordering for the LBVH kernels (north_star (2), (4)).
//
// After a bounce the surviving paths sit in the queue in an order that has nothing to do with where
// they are in the scene, so the 32 rays of a warp fetch 32 unrelated node chains.  The shade kernel
// therefore emits a 24-bit Morton key of the new path origin P (8 bits per axis inside the scene box)
// for every compacted slot, and the slots are sorted by that key with a stable LSD radix sort
// (3 passes x 8 bits).  The traversal kernels then walk `order[]`: neighbouring lanes get rays that
// start in the same ~1/256-of-the-scene cell, and shadow rays of one cell all run towards the same small
// light, so a warp shares its node fetches (same 128 B lines) and its control flow.  Only the visiting
// order changes: every result is still written to the ray's own slot, so the image is bit-identical.
//
// The element count lives on the device (the compaction counter), so every kernel reads it from memory
// and the grid is sized for the batch capacity: no host synchronisation between bounces.
#pragma once
#include "ptb_common.cuh"

namespace ptb {

#define PS_TILE 8192       // elements per block tile
#define PS_WARPS 8
#define PS_SUB (PS_TILE / PS_WARPS)     // consecutive elements per warp: keeps the scatter stable

// Morton key of a point, PS_KEY_BITS per axis; lo / inv = scene box origin and 256 / extent per axis
#ifndef PS_KEY_BITS
#define PS_KEY_BITS 8
#endif
#define PS_PASSES ((3 * PS_KEY_BITS + 7) / 8)
__device__ __forceinline__ unsigned ps_expand10(unsigned v) {  // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ unsigned ps_key(float x, float y, float z, const float *lo, const float *inv) {
    const float s = (float)(1 << PS_KEY_BITS) / 256.0f, top = (float)((1 << PS_KEY_BITS) - 1);
    const unsigned ix = (unsigned)fminf(fmaxf((x - lo[0]) * (inv[0] * s), 0.f), top);
    const unsigned iy = (unsigned)fminf(fmaxf((y - lo[1]) * (inv[1] * s), 0.f), top);
    const unsigned iz = (unsigned)fminf(fmaxf((z - lo[2]) * (inv[2] * s), 0.f), top);
    return (ps_expand10(ix) << 2) | (ps_expand10(iy) << 1) | ps_expand10(iz);
}

// per-tile digit histogram, bin-major: ghist[bin * tiles_max + tile]; tiles past the live count write zeros
__global__ void __launch_bounds__(PS_WARPS * 32)
k_ps_hist(const unsigned *__restrict__ keys, const unsigned *__restrict__ d_n, int shift, int tiles_max, unsigned *__restrict__ ghist) {
    __shared__ unsigned s_h[PS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x;
    const unsigned n = *d_n;
    for (int b = lane; b < 256; b += 32) s_h[warp][b] = 0;
    __syncwarp();
    const unsigned beg = (unsigned)tile * PS_TILE;
    if (beg < n) {
        const unsigned end = min(n, beg + PS_TILE);
        for (unsigned i = beg + threadIdx.x; i < end; i += PS_WARPS * 32) atomicAdd(&s_h[warp][(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    unsigned c = 0;
#pragma unroll
    for (int w = 0; w < PS_WARPS; ++w) c += s_h[w][threadIdx.x];
    ghist[(size_t)threadIdx.x * tiles_max + tile] = c;
}

// one block per digit value: exclusive scan of its row of tile counts in place, row total to totals[bin]
__global__ void __launch_bounds__(1024) k_ps_rowscan(unsigned *ghist, int tiles_max, unsigned *totals) {
    __shared__ unsigned s_part[1024];
    unsigned *row = ghist + (size_t)blockIdx.x * tiles_max;
    const int per = (tiles_max + 1023) / 1024;
    const int beg = min(tiles_max, per * (int)threadIdx.x), end = min(tiles_max, beg + per);
    unsigned sum = 0;
    for (int i = beg; i < end; ++i) sum += row[i];
    s_part[threadIdx.x] = sum;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
        const unsigned v = threadIdx.x >= (unsigned)o ? s_part[threadIdx.x - o] : 0u;
        __syncthreads();
        s_part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = threadIdx.x ? s_part[threadIdx.x - 1] : 0u;
    for (int i = beg; i < end; ++i) { const unsigned v = row[i]; row[i] = run; run += v; }
    if (threadIdx.x == 1023) totals[blockIdx.x] = s_part[1023];
}

// Stable scatter of one block tile; vals == nullptr means the identity permutation (first pass), keys_out == nullptr
// skips the key write (last pass).  The tile is first ordered by digit in shared memory (warp sub-tiles in order,
// iterations in order, lanes in order), so each digit leaves the block as ONE contiguous run (PS_TILE / 256 = 32
// elements on average): full 128 B lines instead of a 4 B write per element into 256 places.
#define PS_SCATTER_SMEM (2 * PS_TILE * sizeof(unsigned))
__global__ void __launch_bounds__(PS_WARPS * 32)
k_ps_scatter(const unsigned *__restrict__ keys, const unsigned *__restrict__ vals, const unsigned *__restrict__ d_n, int shift,
             int tiles_max, const unsigned *__restrict__ ghist, const unsigned *__restrict__ totals,
             unsigned *__restrict__ keys_out, unsigned *__restrict__ vals_out) {
    extern __shared__ __align__(16) unsigned s_stage[];           // [PS_TILE] keys, [PS_TILE] values, in digit order
    __shared__ unsigned s_cnt[PS_WARPS][256];                     // per warp and digit: count, then running local position
    __shared__ unsigned s_gbase[256];                             // digit -> global position of local position 0
    __shared__ unsigned s_w[PS_WARPS];
    unsigned *s_k = s_stage, *s_v = s_stage + PS_TILE;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x;
    const unsigned n = *d_n;
    const unsigned beg = (unsigned)tile * PS_TILE;
    if (beg >= n) return;                                         // whole block past the live count
    const unsigned end = min(n, beg + PS_TILE);
    const unsigned sub_beg = beg + (unsigned)warp * PS_SUB, sub_end = min(end, sub_beg + PS_SUB);
    for (int b = lane; b < 256; b += 32) s_cnt[warp][b] = 0;
    __syncwarp();
    for (unsigned i = sub_beg + lane; i < sub_end; i += 32) atomicAdd(&s_cnt[warp][(keys[i] >> shift) & 255u], 1u);
    // global start of every digit: exclusive scan of the 256 totals (one per thread, PS_WARPS * 32 == 256)
    const unsigned tot = totals[threadIdx.x];
    unsigned x = tot;
    for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) s_w[warp] = x;
    __syncthreads();
    unsigned gstart = x - tot;
    for (int w = 0; w < warp; ++w) gstart += s_w[w];
    // this thread's digit: positions of the warps' sub-runs inside the tile's run, and the tile's run inside the digit
    unsigned mine = 0;
#pragma unroll
    for (int w = 0; w < PS_WARPS; ++w) { const unsigned c = s_cnt[w][threadIdx.x]; s_cnt[w][threadIdx.x] = mine; mine += c; }
    __syncthreads();                                              // s_w is reused
    unsigned y = mine;
    for (int o = 1; o < 32; o <<= 1) { const unsigned z = __shfl_up_sync(0xffffffffu, y, o); if (lane >= o) y += z; }
    if (lane == 31) s_w[warp] = y;
    __syncthreads();
    unsigned lstart = y - mine;                                   // local position of the digit's first element
    for (int w = 0; w < warp; ++w) lstart += s_w[w];
#pragma unroll
    for (int w = 0; w < PS_WARPS; ++w) s_cnt[w][threadIdx.x] += lstart;
    s_gbase[threadIdx.x] = gstart + ghist[(size_t)threadIdx.x * tiles_max + tile] - lstart;
    __syncthreads();
    for (unsigned base = sub_beg; base < sub_end; base += 32) {
        const unsigned i = base + lane;
        const bool live = i < sub_end;
        const unsigned k = live ? keys[i] : 0u;
        const unsigned v = live ? (vals ? vals[i] : i) : 0u;
        const unsigned d = live ? ((k >> shift) & 255u) : 256u + lane;   // dead lanes match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rank = __popc(peers & ((1u << lane) - 1u));
        unsigned pos = 0;
        if (live) pos = s_cnt[warp][d] + rank;
        __syncwarp();
        if (live && rank == 0) s_cnt[warp][d] += __popc(peers);
        __syncwarp();
        if (live) { s_k[pos] = k; s_v[pos] = v; }
    }
    __syncthreads();
    const unsigned m = end - beg;
    for (unsigned p = threadIdx.x; p < m; p += PS_WARPS * 32) {
        const unsigned k = s_k[p];
        const unsigned g = s_gbase[(k >> shift) & 255u] + p;
        if (keys_out) keys_out[g] = k;
        vals_out[g] = s_v[p];
    }
}

// one pass of the LSD sort on `st`: n is read from *d_n on the device, tiles = ceil(capacity / PS_TILE)
static inline void ps_sort_pass(const unsigned *keys, const unsigned *vals, const unsigned *d_n, int shift, int tiles,
                                unsigned *ghist, unsigned *totals, unsigned *keys_out, unsigned *vals_out, cudaStream_t st);

struct SortBufs {
    unsigned *keys[2], *vals[2];   // keys[0] is written by the shade kernel; the sorted order ends in vals[1]
    unsigned *ghist, *totals;
    int tiles_max;
};

static inline void ps_sort_pass(const unsigned *keys, const unsigned *vals, const unsigned *d_n, int shift, int tiles,
                                unsigned *ghist, unsigned *totals, unsigned *keys_out, unsigned *vals_out, cudaStream_t st) {
    DeviceState *ds = device_state();      // the attribute is per device: a second GPU in the same process needs it too
    if (!ds->sort_optin.load(std::memory_order_acquire)) {
        std::lock_guard<std::mutex> lk(g_state_mu);
        cudaFuncSetAttribute(k_ps_scatter, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)PS_SCATTER_SMEM);
        ds->sort_optin.store(true, std::memory_order_release);
    }
    k_ps_hist<<<tiles, PS_WARPS * 32, 0, st>>>(keys, d_n, shift, tiles, ghist);
    k_ps_rowscan<<<256, 1024, 0, st>>>(ghist, tiles, totals);
    k_ps_scatter<<<tiles, PS_WARPS * 32, PS_SCATTER_SMEM, st>>>(keys, vals, d_n, shift, tiles, ghist, totals, keys_out, vals_out);
}

}  // namespace ptb
