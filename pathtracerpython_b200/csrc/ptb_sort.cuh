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

#define PS_TILE 2048       // elements per warp tile
#define PS_WARPS 8

__device__ __forceinline__ unsigned ps_expand8(unsigned v) {   // 8 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
// 24-bit Morton key of a point; lo / inv = scene box origin and 256 / extent per axis
__device__ __forceinline__ unsigned ps_key(float x, float y, float z, const float *lo, const float *inv) {
    const unsigned ix = (unsigned)fminf(fmaxf((x - lo[0]) * inv[0], 0.f), 255.f);
    const unsigned iy = (unsigned)fminf(fmaxf((y - lo[1]) * inv[1], 0.f), 255.f);
    const unsigned iz = (unsigned)fminf(fmaxf((z - lo[2]) * inv[2], 0.f), 255.f);
    return (ps_expand8(ix) << 2) | (ps_expand8(iy) << 1) | ps_expand8(iz);
}

// per-tile digit histogram, bin-major: ghist[bin * tiles_max + tile]; tiles past the live count write zeros
__global__ void __launch_bounds__(PS_WARPS * 32)
k_ps_hist(const unsigned *__restrict__ keys, const unsigned *__restrict__ d_n, int shift, int tiles_max, unsigned *__restrict__ ghist) {
    __shared__ unsigned s_h[PS_WARPS][256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * PS_WARPS + warp;
    if (tile >= tiles_max) return;
    const unsigned n = *d_n;
    for (int b = lane; b < 256; b += 32) s_h[warp][b] = 0;
    __syncwarp();
    const unsigned beg = (unsigned)tile * PS_TILE, end = min(n, beg + PS_TILE);
    for (unsigned i = beg + lane; i < end; i += 32) atomicAdd(&s_h[warp][(keys[i] >> shift) & 255u], 1u);
    __syncwarp();
    for (int b = lane; b < 256; b += 32) ghist[(size_t)b * tiles_max + tile] = s_h[warp][b];
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

// stable scatter; vals == nullptr means the identity permutation (first pass)
__global__ void __launch_bounds__(PS_WARPS * 32)
k_ps_scatter(const unsigned *__restrict__ keys, const unsigned *__restrict__ vals, const unsigned *__restrict__ d_n, int shift,
             int tiles_max, const unsigned *__restrict__ ghist, const unsigned *__restrict__ totals,
             unsigned *__restrict__ keys_out, unsigned *__restrict__ vals_out) {
    __shared__ unsigned s_off[PS_WARPS][256];
    __shared__ unsigned s_base[256];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * PS_WARPS + warp;
    const unsigned n = *d_n;
    if ((unsigned)blockIdx.x * PS_WARPS * PS_TILE >= n) return;    // whole block past the live count
    {   // exclusive scan of the 256 digit totals (one value per thread, PS_WARPS * 32 == 256)
        const unsigned v = totals[threadIdx.x];
        unsigned x = v;
        for (int o = 1; o < 32; o <<= 1) { const unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        __shared__ unsigned s_w[PS_WARPS];
        if (lane == 31) s_w[warp] = x;
        __syncthreads();
        unsigned before = 0;
        for (int w = 0; w < warp; ++w) before += s_w[w];
        s_base[threadIdx.x] = before + x - v;
        __syncthreads();
    }
    if (tile >= tiles_max) return;
    for (int b = lane; b < 256; b += 32) s_off[warp][b] = s_base[b] + ghist[(size_t)b * tiles_max + tile];
    __syncwarp();
    const unsigned beg = (unsigned)tile * PS_TILE, end = min(n, beg + PS_TILE);
    for (unsigned base = beg; base < end; base += 32) {
        const unsigned i = base + lane;
        const bool live = i < end;
        const unsigned k = live ? keys[i] : 0u;
        const unsigned v = live ? (vals ? vals[i] : i) : 0u;
        const unsigned d = live ? ((k >> shift) & 255u) : 256u + lane;   // dead lanes match nobody
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rank = __popc(peers & ((1u << lane) - 1u));
        unsigned pos = 0;
        if (live) pos = s_off[warp][d] + rank;
        __syncwarp();
        if (live && rank == 0) s_off[warp][d] += __popc(peers);
        __syncwarp();
        if (live) { keys_out[pos] = k; vals_out[pos] = v; }
    }
}

struct SortBufs {
    unsigned *keys[2], *vals[2];   // keys[0] is written by the shade kernel; the sorted order ends in vals[1]
    unsigned *ghist, *totals;
    int tiles_max;
};

}  // namespace ptb
