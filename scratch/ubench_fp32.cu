// scratch/ubench_fp32.cu - issue-rate microbenchmarks behind the fused Cornell kernel's optimisation:
// FFMA with three distinct register sources vs constant operand, packed FFMA2 (sm_100 fma.rn.f32x2), FMNMX (ALU pipe),
// MUFU.RCP, and FFMA + FMNMX interleaved.  Prints warp-instructions per clock per SM and, for FMA forms, TFLOP/s.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o scratch/ubench_fp32 scratch/ubench_fp32.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define ITERS 4096
#define UNROLL 16

__global__ void __launch_bounds__(256) k_ffma_rrr(float *out, const float *in, int iters) {
    float a[8], b[8], c[8];
    for (int i = 0; i < 8; ++i) { a[i] = in[threadIdx.x + 32 * i]; b[i] = in[threadIdx.x + 32 * i + 7]; c[i] = in[threadIdx.x + 32 * i + 13]; }
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], b[i], c[i]);     // three distinct registers per FFMA
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_ffma_rrc(float *out, const float *in, int iters, float kb, float kc) {
    float a[8];
    for (int i = 0; i < 8; ++i) a[i] = in[threadIdx.x + 32 * i];
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], kb, kc);        // constant-bank operands
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_ffma_rr_shared(float *out, const float *in, int iters) {
    // a = fma(a, b, c) with b, c shared by all 8 chains (2 distinct sources + the accumulator): the plane-test shape
    float a[8];
    const float b = in[threadIdx.x + 3], c = in[threadIdx.x + 5];
    for (int i = 0; i < 8; ++i) a[i] = in[threadIdx.x + 32 * i];
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = fmaf(a[i], b, c);
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_ffma2(float *out, const float *in, int iters) {
    float2 a[8], b[8], c[8];
    for (int i = 0; i < 8; ++i) {
        a[i] = make_float2(in[threadIdx.x + 32 * i], in[threadIdx.x + 32 * i + 1]);
        b[i] = make_float2(in[threadIdx.x + 32 * i + 7], in[threadIdx.x + 32 * i + 8]);
        c[i] = make_float2(in[threadIdx.x + 32 * i + 13], in[threadIdx.x + 32 * i + 14]);
    }
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = __ffma2_rn(a[i], b[i], c[i]);
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_ffma2_shared(float *out, const float *in, int iters) {
    float2 a[8];
    const float2 b = make_float2(in[threadIdx.x + 3], in[threadIdx.x + 4]), c = make_float2(in[threadIdx.x + 5], in[threadIdx.x + 6]);
    for (int i = 0; i < 8; ++i) a[i] = make_float2(in[threadIdx.x + 32 * i], in[threadIdx.x + 32 * i + 1]);
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = __ffma2_rn(a[i], b, c);
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_fmnmx(float *out, const float *in, int iters) {
    float a[8], b[8];
    for (int i = 0; i < 8; ++i) { a[i] = in[threadIdx.x + 32 * i]; b[i] = in[threadIdx.x + 32 * i + 7]; }
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = fmaxf(a[i], b[i]); b[i] = fminf(b[i], a[i]) ; }
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + b[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_mufu(float *out, const float *in, int iters) {
    float a[8];
    for (int i = 0; i < 8; ++i) a[i] = in[threadIdx.x + 32 * i];
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_mix(float *out, const float *in, int iters) {
    // one FFMA (3 registers) + one FMNMX per pair: do the two pipes issue in the same cycle budget?
    float a[8], b[8], c[8], m[8];
    for (int i = 0; i < 8; ++i) { a[i] = in[threadIdx.x + 32 * i]; b[i] = in[threadIdx.x + 32 * i + 7]; c[i] = in[threadIdx.x + 32 * i + 13]; m[i] = in[threadIdx.x + 32 * i + 21]; }
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = fmaf(a[i], b[i], c[i]); m[i] = fmaxf(m[i], b[i]); }
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i] + m[i];
    if (s == 12345.678f) out[0] = s;
}
__global__ void __launch_bounds__(256) k_mix2(float *out, const float *in, int iters) {
    // one FFMA2 + two FMNMX: the packed form of the same work
    float2 a[8], b[8], c[8]; float m[8], n[8];
    for (int i = 0; i < 8; ++i) {
        a[i] = make_float2(in[threadIdx.x + 32 * i], in[threadIdx.x + 32 * i + 1]);
        b[i] = make_float2(in[threadIdx.x + 32 * i + 7], in[threadIdx.x + 32 * i + 8]);
        c[i] = make_float2(in[threadIdx.x + 32 * i + 13], in[threadIdx.x + 32 * i + 14]);
        m[i] = in[threadIdx.x + 32 * i + 21]; n[i] = in[threadIdx.x + 32 * i + 22];
    }
    for (int k = 0; k < iters; ++k) {
#pragma unroll
        for (int u = 0; u < UNROLL / 8; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) { a[i] = __ffma2_rn(a[i], b[i], c[i]); m[i] = fmaxf(m[i], b[i].x); n[i] = fmaxf(n[i], b[i].y); }
    }
    float s = 0; for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y + m[i] + n[i];
    if (s == 12345.678f) out[0] = s;
}

template <typename F> static double time_ms(F launch) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    launch(); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        cudaEventRecord(a); launch(); cudaEventRecord(b); cudaEventSynchronize(b);
        float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
    }
    cudaEventDestroy(a); cudaEventDestroy(b);
    return best;
}

int main() {
    int dev = 0, sms = 0, khz = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    float *in, *out;
    cudaMalloc(&in, 4096 * 4); cudaMalloc(&out, 64);
    float h[4096]; for (int i = 0; i < 4096; ++i) h[i] = 0.5f + 0.0001f * (i % 97);
    cudaMemcpy(in, h, sizeof(h), cudaMemcpyHostToDevice);
    const int grid = sms * 8, threads = 256;
    const double warps = (double)grid * threads / 32, per_iter = UNROLL;
    auto report = [&](const char *name, double ms, double instr_per_iter, double fma_per_instr) {
        const double winstr = warps * ITERS * instr_per_iter;
        const double clk = ms * 1e-3 * khz * 1e3;
        printf("%-28s %8.3f ms  %6.3f warp-instr/clk/SM  %7.2f TFLOP/s (at max clock %d MHz)\n", name, ms, winstr / clk / sms,
               fma_per_instr * 2.0 * 32 * winstr / (ms * 1e-3) / 1e12, khz / 1000);
    };
    report("FFMA  r,r,r", time_ms([&] { k_ffma_rrr<<<grid, threads>>>(out, in, ITERS); }), per_iter, 1);
    report("FFMA  r,c,c", time_ms([&] { k_ffma_rrc<<<grid, threads>>>(out, in, ITERS, 0.999f, 0.001f); }), per_iter, 1);
    report("FFMA  r,r,r shared b c", time_ms([&] { k_ffma_rr_shared<<<grid, threads>>>(out, in, ITERS); }), per_iter, 1);
    report("FFMA2 r,r,r", time_ms([&] { k_ffma2<<<grid, threads>>>(out, in, ITERS); }), per_iter, 2);
    report("FFMA2 r,r,r shared b c", time_ms([&] { k_ffma2_shared<<<grid, threads>>>(out, in, ITERS); }), per_iter, 2);
    report("FMNMX", time_ms([&] { k_fmnmx<<<grid, threads>>>(out, in, ITERS); }), 2 * per_iter, 0);
    report("MUFU.RCP", time_ms([&] { k_mufu<<<grid, threads>>>(out, in, ITERS); }), per_iter, 0);
    report("FFMA + FMNMX", time_ms([&] { k_mix<<<grid, threads>>>(out, in, ITERS); }), 2 * per_iter, 0.5);
    report("FFMA2 + 2 FMNMX", time_ms([&] { k_mix2<<<grid, threads>>>(out, in, ITERS); }), 3 * per_iter, 2.0 / 3.0);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e == cudaSuccess ? 0 : 1;
}
