// FP64 pipe ceilings measured in-process (kc_measure_fp64_peak): the roofline
// denominators of this path.  MEASURED_PEAKS.json only carries HBM and bf16.
#include "kc_internal.h"

namespace kc {

__global__ void __launch_bounds__(256) peak_dfma_kernel(double* out, int iters, double a, double b)
{
    double x[8];
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += x[i];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(256) peak_dmma_kernel(double* out, int iters)
{
    double c0[8], c1[8];
    double a = threadIdx.x * 1e-3, b = 1.0 - threadIdx.x * 1e-4;
#pragma unroll
    for (int i = 0; i < 8; i++) { c0[i] = i; c1[i] = -i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++)
            asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                         : "+d"(c0[i]), "+d"(c1[i]) : "d"(a), "d"(b));
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += c0[i] + c1[i];
    if (s == 123.456) out[0] = s;
}

cudaError_t measure_fp64_peak(int reps, double* dfma_tflops, double* dmma_tflops, cudaStream_t s)
{
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    double* d_out = nullptr;
    cudaError_t e = cudaMalloc(&d_out, 64);
    if (e != cudaSuccess) return e;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters = 4096, blocks = sms * 8, threads = 256;
    if (reps < 1) reps = 1;
    double best[2] = {1e30, 1e30};
    for (int which = 0; which < 2; which++) {
        for (int r = 0; r < reps + 1; r++) {
            cudaEventRecord(e0, s);
            if (which == 0) peak_dfma_kernel<<<blocks, threads, 0, s>>>(d_out, iters, 1.0000001, 1e-9);
            else peak_dmma_kernel<<<blocks, threads, 0, s>>>(d_out, iters);
            cudaEventRecord(e1, s);
            cudaEventSynchronize(e1);
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            if (r > 0 && ms < best[which]) best[which] = ms;
        }
    }
    double f0 = 2.0 * 8 * iters * (double)blocks * threads;
    double f1 = 2.0 * 8 * 8 * 4 * 8 * iters * (double)blocks * (threads / 32);
    if (dfma_tflops) *dfma_tflops = f0 / best[0] * 1e-9;
    if (dmma_tflops) *dmma_tflops = f1 / best[1] * 1e-9;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(d_out);
    return cudaGetLastError();
}

}  // namespace kc
