// rte_peak.cu — register-resident FFMA micro-benchmark: the measured FP32-FMA roofline denominator.
//
// MEASURED_PEAKS.json carries no FP32 figure, so bench.py measures it on the same box, in
// the same process and under the same clocks as the render kernel: every thread runs 16 independent FFMA
// chains (no memory traffic, no bank conflicts: each FFMA reads three distinct registers), the grid fills
// every SM with 2048 threads.  flop = 2 per FFMA.
#include "rte_kernels.h"

namespace rte {

constexpr int kPeakThreads = 256;
constexpr int kPeakChains = 16;
constexpr int kPeakInner = 64;  // FFMAs per chain per outer iteration

__global__ void __launch_bounds__(kPeakThreads) ffma_peak_kernel(float* out, int outer, float a, float b) {
    float acc[kPeakChains];
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) acc[k] = (float)(threadIdx.x + k) * 1e-3f;
#pragma unroll 1
    for (int it = 0; it < outer; ++it) {
#pragma unroll
        for (int j = 0; j < kPeakInner; ++j) {
#pragma unroll
            for (int k = 0; k < kPeakChains; ++k) acc[k] = __fmaf_rn(acc[k], a, b);
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) s += acc[k];
    if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // keeps the chains alive
}

// The same with the packed fma.rn.f32x2 (SASS FFMA2, sm_100+): 2 FMAs per lane per instruction.
__global__ void __launch_bounds__(kPeakThreads) ffma2_peak_kernel(float* out, int outer, float a, float b) {
    unsigned long long acc[kPeakChains], a2, b2;
    asm("mov.b64 %0, {%1, %1};" : "=l"(a2) : "f"(a));
    asm("mov.b64 %0, {%1, %1};" : "=l"(b2) : "f"(b));
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) {
        const float v = (float)(threadIdx.x + k) * 1e-3f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(acc[k]) : "f"(v), "f"(-v));
    }
#pragma unroll 1
    for (int it = 0; it < outer; ++it) {
#pragma unroll
        for (int j = 0; j < kPeakInner / 2; ++j) {
#pragma unroll
            for (int k = 0; k < kPeakChains; ++k) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(acc[k]) : "l"(a2), "l"(b2));
        }
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kPeakChains; ++k) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(acc[k]));
        s += lo + hi;
    }
    if (s == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

cudaError_t launch_ffma_peak(float* scratch, int outer, int sm_count, cudaStream_t stream, double* flop_out, int packed) {
    const int blocks = sm_count * 8;
    if (packed) ffma2_peak_kernel<<<blocks, kPeakThreads, 0, stream>>>(scratch, outer, 0.999f, 1e-3f);
    else ffma_peak_kernel<<<blocks, kPeakThreads, 0, stream>>>(scratch, outer, 0.999f, 1e-3f);
    *flop_out = 2.0 * (double)blocks * kPeakThreads * (double)outer * kPeakInner * kPeakChains;
    return cudaGetLastError();
}

}  // namespace rte
