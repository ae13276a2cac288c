// FP64 denominators for the roofline: MEASURED_PEAKS.json has HBM and bf16 only,
// and this path is FP64-bound.  Measures, on the current device:
//   * vector DFMA throughput (independent chains, full occupancy)
//   * tensor DMMA throughput, mma.sync m8n8k4 and m16n8k16 (f64)
//   * per-call cost of the fp64 math functions the Kepler chain needs
// Prints one JSON object.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <string>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { \
    fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1);} } while (0)

template <int ILP>
__global__ void __launch_bounds__(256) dfma_kernel(double* out, int iters, double a, double b)
{
    double x[ILP];
#pragma unroll
    for (int i = 0; i < ILP; i++) x[i] = threadIdx.x * 1e-9 + i;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) x[i] = fma(x[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += x[i];
    if (s == 123.456) out[0] = s;
}

__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ void dmma16816(double (&c)[4], const double (&a)[8], const double (&b)[4])
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, "
                 "{%4,%5,%6,%7,%8,%9,%10,%11}, {%12,%13,%14,%15}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(a[4]), "d"(a[5]), "d"(a[6]), "d"(a[7]),
                   "d"(b[0]), "d"(b[1]), "d"(b[2]), "d"(b[3]));
}
__device__ __forceinline__ void dmma1688(double (&c)[4], const double (&a)[4], const double (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f64.f64.f64.f64 {%0,%1,%2,%3}, "
                 "{%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+d"(c[0]), "+d"(c[1]), "+d"(c[2]), "+d"(c[3])
                 : "d"(a[0]), "d"(a[1]), "d"(a[2]), "d"(a[3]), "d"(b[0]), "d"(b[1]));
}

template <int ILP>
__global__ void __launch_bounds__(256) dmma884_kernel(double* out, int iters)
{
    double c0[ILP], c1[ILP];
    double a = threadIdx.x * 1e-3, b = 1.0 - threadIdx.x * 1e-4;
#pragma unroll
    for (int i = 0; i < ILP; i++) { c0[i] = i; c1[i] = -i; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) dmma884(c0[i], c1[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c0[i] + c1[i];
    if (s == 123.456) out[0] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) dmma16816_kernel(double* out, int iters)
{
    double c[ILP][4];
    double a[8], b[4];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
    for (int i = 0; i < 4; i++) b[i] = 1.0 - threadIdx.x * 1e-4 * i;
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = i; c[i][1] = -i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) dmma16816(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 123.456) out[0] = s;
}

template <int ILP>
__global__ void __launch_bounds__(256) dmma1688_kernel(double* out, int iters)
{
    double c[ILP][4];
    double a[4], b[2];
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] = threadIdx.x * 1e-3 + i;
#pragma unroll
    for (int i = 0; i < 2; i++) b[i] = 1.0 - threadIdx.x * 1e-4 * i;
#pragma unroll
    for (int i = 0; i < ILP; i++) { c[i][0] = i; c[i][1] = -i; c[i][2] = 1; c[i][3] = 2; }
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int i = 0; i < ILP; i++) dmma1688(c[i], a, b);
    }
    double s = 0;
#pragma unroll
    for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
    if (s == 123.456) out[0] = s;
}

// ---- math-function cost: ns per call per thread at full occupancy ----------
enum Fn { F_FMA, F_DIV, F_SQRT, F_RCP, F_SIN, F_COS, F_SINCOS, F_ACOS, F_POW23, F_CBRT, F_EXP, F_LOG,
          F_FMOD, F_LGAMMA, F_ATAN2, F_RSQRT, F_COUNT };
static const char* fn_names[] = {"fma", "div", "sqrt", "rcp", "sin", "cos", "sincos", "acos", "pow_2_3",
                                 "cbrt", "exp", "log", "fmod_2pi", "lgamma", "atan2", "rsqrt"};

template <int FN>
__global__ void __launch_bounds__(256) fn_kernel(double* out, int iters, double seed)
{
    double x = seed + (threadIdx.x + blockIdx.x * blockDim.x) * 1e-7;
    double y = 0.37 + threadIdx.x * 1e-6;
    double acc = 0;
    for (int it = 0; it < iters; it++) {
        double r;
        if (FN == F_FMA) r = fma(x, y, 0.5);
        else if (FN == F_DIV) r = y / x;
        else if (FN == F_SQRT) r = sqrt(x);
        else if (FN == F_RCP) r = 1.0 / x;
        else if (FN == F_SIN) r = sin(x);
        else if (FN == F_COS) r = cos(x);
        else if (FN == F_SINCOS) { double s, c; sincos(x, &s, &c); r = s + c; }
        else if (FN == F_ACOS) r = acos(y);
        else if (FN == F_POW23) r = pow(x, 2.0 / 3);
        else if (FN == F_CBRT) r = cbrt(x);
        else if (FN == F_EXP) r = exp(-x);
        else if (FN == F_LOG) r = log(x);
        else if (FN == F_FMOD) r = fmod(x * 1000.0, 6.283185307179586);
        else if (FN == F_LGAMMA) r = lgamma(x);
        else if (FN == F_ATAN2) r = atan2(y, x);
        else r = rsqrt(x);
        acc += r;
        x += 1e-3;  // keep arguments moving so nothing is hoisted
        y = fma(y, 0.999, 1e-4);
    }
    if (acc == 123.456) out[0] = acc;
}

template <typename F>
static double time_ms(F launch, int reps)
{
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    launch(); launch(); launch();
    CK(cudaDeviceSynchronize());
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        CK(cudaEventRecord(e0));
        launch();
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

template <int FN>
static double fn_cost(double* d_out, int sms, int reps, double base_ns)
{
    const int iters = 2048, blocks = sms * 8, threads = 256;
    double ms = time_ms([&] { fn_kernel<FN><<<blocks, threads>>>(d_out, iters, 1.3); }, reps);
    double calls = (double)iters * blocks * threads;
    return ms * 1e6 / calls - base_ns;  // ns per call, chip-wide throughput view
}

int main(int argc, char** argv)
{
    int reps = argc > 1 ? atoi(argv[1]) : 10;
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    int sms = prop.multiProcessorCount;
    double* d_out; CK(cudaMalloc(&d_out, 64));
    printf("{\"gpu\": \"%s\", \"sms\": %d, \"clock_khz\": %d", prop.name, sms, prop.clockRate);

    {   // DFMA: 8 CTAs/SM x 256 thr, ILP 8
        const int iters = 4096, blocks = sms * 8, threads = 256;
        double ms = time_ms([&] { dfma_kernel<8><<<blocks, threads>>>(d_out, iters, 1.0000001, 1e-9); }, reps);
        double flops = 2.0 * 8 * iters * (double)blocks * threads;
        printf(", \"dfma_tflops\": %.3f", flops / ms * 1e-9);
        double ms1 = time_ms([&] { dfma_kernel<1><<<blocks, threads>>>(d_out, iters * 8, 1.0000001, 1e-9); }, reps);
        printf(", \"dfma_ilp1_tflops\": %.3f", flops / ms1 * 1e-9);
        // latency: one warp per SM, ILP 1
        double msl = time_ms([&] { dfma_kernel<1><<<sms, 32>>>(d_out, iters * 8, 1.0000001, 1e-9); }, reps);
        printf(", \"dfma_dep_ns\": %.3f", msl * 1e6 / (iters * 8.0));
    }
    {
        const int iters = 4096, blocks = sms * 8, threads = 256;
        double ms = time_ms([&] { dmma884_kernel<8><<<blocks, threads>>>(d_out, iters); }, reps);
        double flops = 2.0 * 8 * 8 * 4 * 8 * iters * (double)blocks * (threads / 32);
        printf(", \"dmma_m8n8k4_tflops\": %.3f", flops / ms * 1e-9);
        double ms2 = time_ms([&] { dmma16816_kernel<4><<<blocks, threads>>>(d_out, iters); }, reps);
        double flops2 = 2.0 * 16 * 8 * 16 * 4 * iters * (double)blocks * (threads / 32);
        printf(", \"dmma_m16n8k16_tflops\": %.3f", flops2 / ms2 * 1e-9);
        double ms3 = time_ms([&] { dmma1688_kernel<4><<<blocks, threads>>>(d_out, iters); }, reps);
        double flops3 = 2.0 * 16 * 8 * 8 * 4 * iters * (double)blocks * (threads / 32);
        printf(", \"dmma_m16n8k8_tflops\": %.3f", flops3 / ms3 * 1e-9);
        double msl = time_ms([&] { dmma884_kernel<1><<<sms, 32>>>(d_out, iters * 8); }, reps);
        printf(", \"dmma_m8n8k4_dep_ns\": %.3f", msl * 1e6 / (iters * 8.0));
    }
    {
        double base = 0;
        printf(", \"fn_ns_per_call_chipwide\": {");
        double c[F_COUNT];
        c[F_FMA] = fn_cost<F_FMA>(d_out, sms, reps, 0);
        base = 0;
        c[F_DIV] = fn_cost<F_DIV>(d_out, sms, reps, base);
        c[F_SQRT] = fn_cost<F_SQRT>(d_out, sms, reps, base);
        c[F_RCP] = fn_cost<F_RCP>(d_out, sms, reps, base);
        c[F_SIN] = fn_cost<F_SIN>(d_out, sms, reps, base);
        c[F_COS] = fn_cost<F_COS>(d_out, sms, reps, base);
        c[F_SINCOS] = fn_cost<F_SINCOS>(d_out, sms, reps, base);
        c[F_ACOS] = fn_cost<F_ACOS>(d_out, sms, reps, base);
        c[F_POW23] = fn_cost<F_POW23>(d_out, sms, reps, base);
        c[F_CBRT] = fn_cost<F_CBRT>(d_out, sms, reps, base);
        c[F_EXP] = fn_cost<F_EXP>(d_out, sms, reps, base);
        c[F_LOG] = fn_cost<F_LOG>(d_out, sms, reps, base);
        c[F_FMOD] = fn_cost<F_FMOD>(d_out, sms, reps, base);
        c[F_LGAMMA] = fn_cost<F_LGAMMA>(d_out, sms, reps, base);
        c[F_ATAN2] = fn_cost<F_ATAN2>(d_out, sms, reps, base);
        c[F_RSQRT] = fn_cost<F_RSQRT>(d_out, sms, reps, base);
        for (int i = 0; i < F_COUNT; i++)
            printf("%s\"%s\": %.6f", i ? ", " : "", fn_names[i], c[i]);
        printf("}");
    }
    printf("}\n");
    return 0;
}
