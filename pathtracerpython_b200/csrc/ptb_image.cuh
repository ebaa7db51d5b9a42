// ptb_image.cuh - make_image (utils.py:150-161) on the device: mean over samples (main.py:274-280),
// one global min and one global max over all pixels AND channels, (x - min) / max * 255, truncating
// cast to u8, pixel (ix, iy) -> row H-1-iy, column ix.
#pragma once
#include "ptb_common.cuh"

namespace ptb {

__device__ __forceinline__ unsigned long long dbl_order(double v) {
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b & 0x8000000000000000ull) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dbl_unorder(unsigned long long o) {
    const unsigned long long b = (o & 0x8000000000000000ull) ? (o & 0x7fffffffffffffffull) : ~o;
    return __longlong_as_double((long long)b);
}

__global__ void k_minmax_init(unsigned long long *mm) { mm[0] = ~0ull; mm[1] = 0ull; }

// mm[0] = ordered min, mm[1] = ordered max of the per-pixel means over r, g, b
__global__ void __launch_bounds__(256) k_minmax(const float4 *__restrict__ accum, int n, unsigned long long *mm) {
    double lo = 1.0e300, hi = -1.0e300;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = accum[i];
        const double c = a.w > 0.f ? (double)a.w : 1.0;
        const double r = (double)a.x / c, g = (double)a.y / c, b = (double)a.z / c;
        lo = fmin(lo, fmin(r, fmin(g, b)));
        hi = fmax(hi, fmax(r, fmax(g, b)));
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = fmin(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmax(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(mm, dbl_order(lo));
        atomicMax(mm + 1, dbl_order(hi));
    }
}

__global__ void __launch_bounds__(256) k_quantise(const float4 *__restrict__ accum, int W, int H,
                                                  const unsigned long long *__restrict__ mm, uint8_t *__restrict__ rgb) {
    const double mn = dbl_unorder(mm[0]);
    const double mx = dbl_unorder(mm[1]) - mn;     // max of (mat - min)
    const int n = W * H;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const float4 a = accum[r];
        const double c = a.w > 0.f ? (double)a.w : 1.0;
        const int ix = r / H, iy = r - ix * H;     // utils.py:66-68 (== utils.py:153-157 when W == H)
        uint8_t *o = rgb + ((size_t)(H - 1 - iy) * W + ix) * 3;
        const double v[3] = {(double)a.x / c, (double)a.y / c, (double)a.z / c};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const double q = ((v[k] - mn) / mx) * 255.0;
            o[k] = (uint8_t)(int)q;                // astype('uint8') truncates
        }
    }
}

// Opt-in display transform from two scene keywords the reference parses and never reads (scene_reader.py:158-159
// `background`, :167-168 `tonemapping`): v / (v + tm) on the per-pixel mean, clamped to [0, 1], * 255 truncated;
// pixels whose primary ray left the scene show the background colour.  Same pixel mapping as k_quantise.
__global__ void __launch_bounds__(256) k_quantise_scene(const float4 *__restrict__ accum, const int *__restrict__ primary_tri,
                                                        int W, int H, float tm, float bgr, float bgg, float bgb,
                                                        uint8_t *__restrict__ rgb) {
    const int n = W * H;
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
        const float4 a = accum[r];
        const float c = a.w > 0.f ? a.w : 1.f;
        const int ix = r / H, iy = r - ix * H;
        uint8_t *o = rgb + ((size_t)(H - 1 - iy) * W + ix) * 3;
        const bool miss = primary_tri && primary_tri[r] < 0;
        const float v[3] = {a.x / c, a.y / c, a.z / c}, bg[3] = {bgr, bgg, bgb};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            float q = miss ? bg[k] : (tm > 0.f ? v[k] / (v[k] + tm) : v[k]);
            q = q > 0.f ? (q < 1.f ? q : 1.f) : 0.f;            // also sends NaN to 0
            o[k] = (uint8_t)(int)(q * 255.f);
        }
    }
}

__global__ void __launch_bounds__(256) k_scatter_accum(PixelMap pm, const float4 *__restrict__ src, float4 *__restrict__ dst) {
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < pm.n; p += gridDim.x * blockDim.x)
        dst[pm.ray_index(p)] = src[p];
}

}  // namespace ptb
