// diag.cu -- measurement helpers of the C ABI (include/hmb200.h, "Diagnostics"): the FP64 issue rates of this GPU, which are
// the roofline denominators of the multi-RHS path (MEASURED_PEAKS.json carries no FP64 figure).
#include "hmb200.h"
#include "internal.h"

#include <cuda_runtime.h>
#include <stdexcept>
#include <string>

namespace {

__device__ __forceinline__ void diag_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// 8 independent accumulator pairs per warp: the DMMA pipe is the only limiter
__global__ void __launch_bounds__(256) diag_dmma_kernel(double *out, int iters, double a, double b) {
    double c[8][2];
#pragma unroll
    for (int i = 0; i < 8; i++) c[i][0] = c[i][1] = threadIdx.x;
    for (int it = 0; it < iters; it++)
#pragma unroll
        for (int i = 0; i < 8; i++) diag_dmma(c[i][0], c[i][1], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) diag_dfma_kernel(double *out, int iters, double a, double b) {
    double c[8];
#pragma unroll
    for (int i = 0; i < 8; i++) c[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++)
#pragma unroll
        for (int i = 0; i < 8; i++) c[i] = fma(c[i], a, b);
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

#define DCHECK(expr)                                                                                                   \
    do {                                                                                                               \
        cudaError_t _e = (expr);                                                                                       \
        if (_e != cudaSuccess) throw std::runtime_error(std::string(#expr) + " failed: " + cudaGetErrorString(_e));   \
    } while (0)

} // namespace

extern "C" int hmb200_fp64_rate(int device, double *dmma_tflops, double *dfma_tflops) {
    try {
        DCHECK(cudaSetDevice(device));
        cudaDeviceProp p;
        DCHECK(cudaGetDeviceProperties(&p, device));
        const int warps = 8, iters = 20000, ctas = p.multiProcessorCount;
        double *out = nullptr;
        DCHECK(cudaMalloc(&out, sizeof(double) * (size_t)ctas * warps * 32));
        cudaEvent_t e0, e1;
        DCHECK(cudaEventCreate(&e0));
        DCHECK(cudaEventCreate(&e1));
        float best[2] = {1e30f, 1e30f};
        for (int rep = 0; rep < 3; rep++)
            for (int k = 0; k < 2; k++) {
                DCHECK(cudaEventRecord(e0));
                if (k == 0) diag_dmma_kernel<<<ctas, warps * 32>>>(out, iters, 1.0000001, 0.9999999);
                else diag_dfma_kernel<<<ctas, warps * 32>>>(out, iters, 1.0000001, 0.9999999);
                DCHECK(cudaEventRecord(e1));
                DCHECK(cudaEventSynchronize(e1));
                float ms = 0;
                DCHECK(cudaEventElapsedTime(&ms, e0, e1));
                if (rep > 0 && ms < best[k]) best[k] = ms;
            }
        DCHECK(cudaGetLastError());
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(out);
        if (dmma_tflops) *dmma_tflops = 2.0 * 256 * 8 * (double)iters * warps * ctas / best[0] / 1e9;
        if (dfma_tflops) *dfma_tflops = 2.0 * 32 * 8 * (double)iters * warps * ctas / best[1] / 1e9;
        return 0;
    } catch (const std::exception &e) {
        hmb200i_set_error(e.what());
        return 1;
    }
}
