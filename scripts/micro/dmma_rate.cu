// micro-benchmark: FP64 DMMA (mma.sync.m8n8k4.f64) and DFMA issue rates on this GPU (roofline denominator of the matmat path)
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
template <int NACC> __global__ void k_dmma(double *out, int iters, double a, double b) {
    double c[NACC][2];
    for (int i = 0; i < NACC; i++) c[i][0] = c[i][1] = threadIdx.x;
    for (int it = 0; it < iters; it++)
#pragma unroll
        for (int i = 0; i < NACC; i++) dmma(c[i][0], c[i][1], a, b);
    double s = 0;
    for (int i = 0; i < NACC; i++) s += c[i][0] + c[i][1];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int NACC> __global__ void k_dfma(double *out, int iters, double a, double b) {
    double c[NACC];
    for (int i = 0; i < NACC; i++) c[i] = threadIdx.x + i;
    for (int it = 0; it < iters; it++)
#pragma unroll
        for (int i = 0; i < NACC; i++) c[i] = fma(c[i], a, b);
    double s = 0;
    for (int i = 0; i < NACC; i++) s += c[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main() {
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    double *out; cudaMalloc(&out, sizeof(double) * 148 * 8 * 1024);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 20000;
    for (int warps : {4, 8, 16, 32}) {
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0); k_dmma<8><<<p.multiProcessorCount, warps * 32>>>(out, iters, 1.0000001, 0.9999999); cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double flops = 2.0 * 256 * 8 * (double)iters * warps * p.multiProcessorCount;
            if (rep) printf("DMMA m8n8k4 warps/SM=%2d: %.2f TFLOP/s  (%.1f FMA/clk/SM at %d MHz)\n", warps, flops / ms / 1e9, flops / 2 / (ms * 1e-3) / p.multiProcessorCount / (p.clockRate * 1e3), p.clockRate / 1000);
            cudaEventRecord(e0); k_dfma<8><<<p.multiProcessorCount, warps * 32>>>(out, iters, 1.0000001, 0.9999999); cudaEventRecord(e1); cudaEventSynchronize(e1);
            cudaEventElapsedTime(&ms, e0, e1);
            flops = 2.0 * 32 * 8 * (double)iters * warps * p.multiProcessorCount;
            if (rep) printf("DFMA          warps/SM=%2d: %.2f TFLOP/s\n", warps, flops / ms / 1e9);
        }
    }
    printf("err=%s\n", cudaGetErrorString(cudaGetLastError()));
}
