// peak.cu -- FP32 FMA issue-peak microkernel: the measured roofline denominator for the gather
// (MEASURED_PEAKS.json has no FP32 entry).  16 independent FFMA chains per thread keep
// both FMA pipes of every SMSP saturated; __fmaf_rn forces FFMA regardless of -fmad=false.
#include "vrlb_kernels.h"

namespace vrlb {

__global__ void __launch_bounds__(256) fp32_peak_kernel(float *out, int iters, float a, float b) {
    float x[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) x[k] = (float) (threadIdx.x + k) * 1e-3f;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 4; ++r)
#pragma unroll
            for (int k = 0; k < 16; ++k) x[k] = __fmaf_rn(x[k], a, b);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += x[k];
    if (s == 123.456f) out[0] = s;     // never true; keeps the chains alive
}

cudaError_t launch_fp32_peak(float *scratch, int blocks, int iters, cudaStream_t stream) {
    fp32_peak_kernel<<<blocks, 256, 0, stream>>>(scratch, iters, 0.999f, 1e-3f);
    return cudaGetLastError();
}

}  // namespace vrlb
