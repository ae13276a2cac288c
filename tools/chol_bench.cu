// Microbenchmark: cycles of chol32_panels (kc_gp_common.cuh) for one warp alone on an SM, and of its parts.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I kima_b200/csrc tools/chol_bench.cu -o tools/_build/chol_bench
#include <cstdio>
#include <vector>
#include "kc_gp_common.cuh"
using namespace kc;

constexpr int LD = 36;  // 4 mod 16

__global__ void bench(const double* A, long long* cyc, double* out, int reps, int nwarps_busy)
{
    __shared__ double S[32 * LD];
    __shared__ double dinv[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (warp > 0) {
        // background load: FP64 chains as the generation warps would issue
        double x = lane * 1e-3, y = 1.0;
        for (int i = 0; i < reps * 4000; i++) { x = kfma(x, 1.0000001, 1e-9); y = kfma(y, 0.9999999, x); }
        if (x + y == 123.456) out[1] = x;
        return;
    }
    long long tot = 0;
    double acc = 0.0;
    for (int r = 0; r < reps; r++) {
        for (int i = lane; i < 32 * 32; i += 32) S[(i >> 5) * LD + (i & 31)] = A[(i >> 5) * 32 + (i & 31)];
        __syncwarp();
        double logdet = 0.0;
        long long t0 = clock64();
        bool ok = chol32_panels<false>(S, LD, lane, 32, logdet, dinv, nullptr, 0);
        long long t1 = clock64();
        tot += t1 - t0;
        acc += logdet + (ok ? 0.0 : 1e300);
    }
    if (lane == 0) { cyc[0] = tot / reps; out[0] = acc; }
}

int main()
{
    std::vector<double> A(32 * 32);
    for (int c = 0; c < 32; c++)
        for (int r = 0; r < 32; r++) A[c * 32 + r] = (r == c ? 40.0 : 0.0) + 1.0 / (1.0 + abs(r - c));
    double *dA, *dout; long long* dc;
    cudaMalloc(&dA, sizeof(double) * 1024); cudaMalloc(&dout, 16); cudaMalloc(&dc, 8);
    cudaMemcpy(dA, A.data(), sizeof(double) * 1024, cudaMemcpyHostToDevice);
    for (int warps : {1, 2, 5, 9, 16}) {
        bench<<<1, 32 * warps>>>(dA, dc, dout, 50, warps);
        cudaDeviceSynchronize();
        long long c; double o[2];
        cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(o, dout, 16, cudaMemcpyDeviceToHost);
        printf("warps in CTA %2d: chol32_panels %lld cycles per 32x32 block (%.1f per column), logdet sum %.6f, err %s\n", warps, c, c / 32.0, o[0],
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
