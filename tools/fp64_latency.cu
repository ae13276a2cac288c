// Microbenchmark: dependent-issue latency of FP64 operations on one warp, and throughput
// with k independent chains per thread (how much ILP a lone warp needs to fill the pipe).
#include <cstdio>
#include <cuda_runtime.h>

template <int CH>
__global__ void chain(double* out, long long* cyc, int iters, double a, double b)
{
    double x[CH];
#pragma unroll
    for (int k = 0; k < CH; k++) x[k] = a + k + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < CH; k++) x[k] = fma(x[k], b, a);
    }
    long long t1 = clock64();
    double s = 0;
#pragma unroll
    for (int k = 0; k < CH; k++) s += x[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

__global__ void mufu_chain(double* out, long long* cyc, int iters, double a)
{
    double x = a + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        double r;
        asm volatile("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
        x = r + 1.0;
    }
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

__global__ void shfl_chain(double* out, long long* cyc, int iters, double a)
{
    double x = a + threadIdx.x;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) x = __shfl_sync(0xffffffffu, x, (i + 1) & 31);
    long long t1 = clock64();
    out[threadIdx.x] = x;
    if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int CH>
void run(int warps, double* d_out, long long* d_cyc)
{
    const int iters = 4096;
    chain<CH><<<1, 32 * warps>>>(d_out, d_cyc, iters, 1.0000001, 0.999999);
    chain<CH><<<1, 32 * warps>>>(d_out, d_cyc, iters, 1.0000001, 0.999999);
    long long c;
    cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
    printf("DFMA chains/thread=%d warps/SM=%d: %.2f cycles per round, %.2f cycles per warp-DFMA\n", CH, warps,
           (double)c / iters, (double)c / iters / CH / ((warps + 3) / 4));
}

int main()
{
    double* d_out; long long* d_cyc;
    cudaMalloc(&d_out, sizeof(double) * 4096);
    cudaMalloc(&d_cyc, sizeof(long long));
    run<1>(1, d_out, d_cyc); run<2>(1, d_out, d_cyc); run<4>(1, d_out, d_cyc); run<8>(1, d_out, d_cyc); run<16>(1, d_out, d_cyc);
    run<1>(4, d_out, d_cyc); run<2>(4, d_out, d_cyc); run<4>(4, d_out, d_cyc);
    run<1>(12, d_out, d_cyc); run<2>(12, d_out, d_cyc); run<4>(12, d_out, d_cyc); run<8>(12, d_out, d_cyc);
    run<2>(16, d_out, d_cyc); run<4>(16, d_out, d_cyc);
    long long c;
    mufu_chain<<<1, 32>>>(d_out, d_cyc, 4096, 1.5);
    cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
    printf("rsqrt.approx.f64 + DADD chain: %.2f cycles per step\n", (double)c / 4096);
    shfl_chain<<<1, 32>>>(d_out, d_cyc, 4096, 1.5);
    cudaMemcpy(&c, d_cyc, sizeof(c), cudaMemcpyDeviceToHost);
    printf("64-bit shuffle chain: %.2f cycles per step\n", (double)c / 4096);
    return 0;
}
