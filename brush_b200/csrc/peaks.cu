// Roofline denominators measured on the device the bench runs on (SURVEY.md sections 6 and 8d: "FP32 and MUFU
// peaks measured on the box by micro-kernel before being used as denominators").  Two kernels with nothing but
// the instruction under test in their inner loop:
//   ffma_peak_kernel   16 independent FFMA chains per thread   -> FP32 TFLOP/s (FMA = 2 flop)
//   mufu_peak_kernel   8 independent MUFU.EX2 chains per thread -> 1e12 ex2 / s
// The rasterizers are bounded by these two pipes (rasterize.wgsl:76-99, rasterize_backwards.wgsl:194-272: 26 / 64
// flop and one exp per evaluated pair), not by HBM, so their roofline fractions are taken against these numbers.
#include "internal.h"

namespace brush {

constexpr int PEAK_THREADS = 256;
constexpr int PEAK_CHAINS = 16;

__global__ void __launch_bounds__(PEAK_THREADS)
ffma_peak_kernel(float a, float b, int iters, float* __restrict__ sink) {
    float x[PEAK_CHAINS];
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS; i++) x[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < PEAK_CHAINS; i++) x[i] = fmaf(x[i], a, b);  // |a| < 1: bounded
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS; i++) s += x[i];
    if (s == 12345.678f) sink[0] = s;  // keeps the chains alive; never true
}

// Three REGISTER operands per FFMA (the form real kernels issue: ffma_peak_kernel's multiplier and addend sit in the
// constant bank), 16 chains per thread.
__global__ void __launch_bounds__(PEAK_THREADS)
ffma3_peak_kernel(float a, float b, int iters, float* __restrict__ sink) {
    float x[PEAK_CHAINS], y[4], z[4];
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS; i++) x[i] = (float)(threadIdx.x + i) * 1e-3f;
#pragma unroll
    for (int i = 0; i < 4; i++) { y[i] = a - 1e-4f * (float)(threadIdx.x & (i + 1)); z[i] = b + 1e-5f * (float)(threadIdx.x & (i + 3)); }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < PEAK_CHAINS; i++) x[i] = fmaf(x[i], y[(i + u) & 3], z[i & 3]);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS; i++) s += x[i];
    if (s == 12345.678f) sink[0] = s;
}

// Packed FP32 (fma.rn.f32x2 -> FFMA2, new on sm_100): two FMAs per issued instruction, 8 packed chains per thread.
__global__ void __launch_bounds__(PEAK_THREADS)
ffma2_peak_kernel(float a, float b, int iters, float* __restrict__ sink) {
    unsigned long long x[PEAK_CHAINS / 2], y[2], z[2];
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS / 2; i++) {
        const float lo = (float)(threadIdx.x + i) * 1e-3f, hi = (float)(threadIdx.x + i) * 2e-3f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(x[i]) : "f"(lo), "f"(hi));
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const float ya = a - 1e-4f * (float)(threadIdx.x & (i + 1)), zb = b + 1e-5f * (float)(threadIdx.x & (i + 3));
        asm("mov.b64 %0, {%1, %2};" : "=l"(y[i]) : "f"(ya), "f"(ya * 0.999f));
        asm("mov.b64 %0, {%1, %2};" : "=l"(z[i]) : "f"(zb), "f"(zb * 1.001f));
    }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < PEAK_CHAINS / 2; i++)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(x[i]) : "l"(y[(i + u) & 1]), "l"(z[i & 1]));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < PEAK_CHAINS / 2; i++) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(x[i]));
        s += lo + hi;
    }
    if (s == 12345.678f) sink[0] = s;
}

__global__ void __launch_bounds__(PEAK_THREADS)
mufu_peak_kernel(int iters, float* __restrict__ sink) {
    float x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++) {
#pragma unroll
            for (int i = 0; i < 8; i++)  // x -> 2^-x maps [0, 1] into [0.5, 1]: bounded, the negation is an operand modifier
                asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(x[i]) : "f"(-x[i]));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i];
    if (s == 12345.678f) sink[0] = s;
}

// Best of `reps` timed launches each; synchronises `st`.  out[0] = scalar FFMA with constant-bank operands,
// out[1] = scalar FFMA with three register operands, out[2] = packed FFMA2, all in FP32 TFLOP/s (FMA = 2 flop);
// out[3] = 1e12 MUFU.EX2 per second.  Returns 0 or a cudaError_t.
int measure_peaks(float* out4, float* sink, cudaStream_t st) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int blocks = sms * 8;  // 8 CTAs x 8 warps = the SM's 64 warp slots
    cudaEvent_t e0, e1;
    cudaError_t err = cudaEventCreate(&e0);
    if (err != cudaSuccess) return (int)err;
    err = cudaEventCreate(&e1);
    if (err != cudaSuccess) return (int)err;
    const int ffma_iters = 4096, mufu_iters = 1024, reps = 4;
    float best[4] = {1e30f, 1e30f, 1e30f, 1e30f};
    for (int r = 0; r < reps + 1; r++) {  // first launch of each is the warm-up
        for (int k = 0; k < 4; k++) {
            g_launch_count++;
            cudaEventRecord(e0, st);
            if (k == 0) ffma_peak_kernel<<<blocks, PEAK_THREADS, 0, st>>>(0.999f, 1e-3f, ffma_iters, sink);
            else if (k == 1) ffma3_peak_kernel<<<blocks, PEAK_THREADS, 0, st>>>(0.999f, 1e-3f, ffma_iters, sink);
            else if (k == 2) ffma2_peak_kernel<<<blocks, PEAK_THREADS, 0, st>>>(0.999f, 1e-3f, ffma_iters, sink);
            else mufu_peak_kernel<<<blocks, PEAK_THREADS, 0, st>>>(mufu_iters, sink);
            cudaEventRecord(e1, st);
            cudaEventSynchronize(e1);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < best[k]) best[k] = ms;
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    err = cudaGetLastError();
    if (err != cudaSuccess) return (int)err;
    const double threads = (double)blocks * PEAK_THREADS;
    const double flop = threads * (double)ffma_iters * 4.0 * PEAK_CHAINS * 2.0;  // the same for all three FFMA kernels
    const double ex2 = threads * (double)mufu_iters * 4.0 * 8.0;
    for (int k = 0; k < 3; k++) out4[k] = (float)(flop / (best[k] * 1e-3) / 1e12);
    out4[3] = (float)(ex2 / (best[3] * 1e-3) / 1e12);
    return 0;
}

}  // namespace brush
