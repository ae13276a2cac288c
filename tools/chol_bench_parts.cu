// Microbenchmark: cycles of chol32_panels (kc_gp_common.cuh) for one warp alone on an SM, and of its parts.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I kima_b200/csrc tools/chol_bench.cu -o tools/_build/chol_bench
#include <cstdio>
#include <vector>
#include "kc_gp_common.cuh"
using namespace kc;
namespace kc {
template <bool MASK, int SKIP>
__device__ __forceinline__ bool chol32_var(double* S, int ldS, int lane, int nvalid, double& logdet,
                                              double* dinv_out, double* __restrict__ Lout, size_t ld,
                                              int* progress = nullptr, int progress_base = 0)
{
    constexpr unsigned FULL = 0xffffffffu;
    const int fr = lane >> 2, fk = lane & 3;
    bool ok = true;
    double lmine = 1.0;  // lane c keeps L(c,c) for the log-determinant
    double d = 1.0, inv = 1.0, l = 1.0;
#pragma unroll 1
    for (int p = 0; p < 4; p++) {
        const int c0 = CHOL_PANEL * p;
        // ---- the panel: 8 columns of row `lane` ----
        double a[CHOL_PANEL];
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            double v = (lane == c) ? 1.0 : 0.0;  // identity padding of a partial block; zeros above the diagonal
            if (lane < nvalid && c <= lane) v = S[c * ldS + lane];
            a[k] = v;
        }
        d = __shfl_sync(FULL, a[0], c0);
        chol_pivot(d, inv, l);
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (!(d > 0.0)) ok = false;  // NaN or non-positive pivot; the NaNs that follow are discarded
            if (lane == c) lmine = l;
            if (lane == k) dinv_out[c] = inv;
            const double lrc = a[k] * inv;  // L(lane, c) below the diagonal
            const double lcc = l;
            if (k + 1 < CHOL_PANEL) {
                // The next pivot is known to lane c+1 before the column is passed round:
                // S(c+1,c+1) - L(c+1,c)^2, the same FMA the update below performs.  Its rsqrt
                // chain (latency-bound) overlaps the shuffles + FMAs of the rest of the panel.
                d = __shfl_sync(FULL, kfma(-lrc, lrc, a[k + 1]), c + 1);
                if (SKIP & 2) { inv = d * 0.024; l = d * 1.3; } else chol_pivot(d, inv, l);
                if (!(SKIP & 4)) {
#pragma unroll
                for (int k2 = k + 1; k2 < CHOL_PANEL; k2++)
                    a[k2] = kfma(-lrc, __shfl_sync(FULL, lrc, c0 + k2), a[k2]);
                }
            }
            a[k] = (lane == c) ? lcc : lrc;
        }
#pragma unroll
        for (int k = 0; k < CHOL_PANEL; k++) {
            const int c = c0 + k;
            if (lane >= c && lane < nvalid) {
                S[c * ldS + lane] = a[k];
                if (Lout) Lout[(size_t)c * ld + lane] = a[k];
            }
        }
        __syncwarp();
        if (progress && lane == 0) st_release_cta(progress, progress_base + c0 + CHOL_PANEL);
        // ---- trailing update of the block on the tensor cores: S(i,j) -= sum_k L(i,k) L(j,k) ----
        if (p < 3 && !(SKIP & 1)) {
            double af[4][2];  // fragments of L[8I .. 8I+7, panel]: A operand of row block I, B operand of column block I
#pragma unroll
            for (int I = 1; I < 4; I++) {
                af[I][0] = af[I][1] = 0.0;
                const int row = 8 * I + fr;
                if (I > p && (!MASK || row < nvalid)) {
                    af[I][0] = S[(c0 + fk) * ldS + row];
                    af[I][1] = S[(c0 + 4 + fk) * ldS + row];
                }
            }
#pragma unroll
            for (int J = 1; J < 4; J++) {
#pragma unroll
                for (int I = J; I < 4; I++) {
                    if (J > p) {  // warp-uniform
                        double u0 = 0.0, u1 = 0.0;
                        dmma(u0, u1, af[I][0], af[J][0]);
                        dmma(u0, u1, af[I][1], af[J][1]);
                        const int row = 8 * I + fr;
                        double* q = S + (8 * J + 2 * fk) * ldS + row;
                        if (!MASK || row < nvalid) {
                            q[0] -= u0;
                            q[ldS] -= u1;
                        }
                    }
                }
            }
            __syncwarp();
        }
    }
    double lg = lane < nvalid ? log(lmine) : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) lg += __shfl_xor_sync(FULL, lg, o);
    logdet += lg;
    return ok;
}

}

constexpr int LD = 36;  // 4 mod 16

template <int SKIP>
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
        bool ok = chol32_var<false, SKIP>(S, LD, lane, 32, logdet, dinv, nullptr, 0);
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
    for (int v = 0; v < 8; v++) {
        const int warps = 1;
        switch (v) {
            case 0: bench<0><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 1: bench<1><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 2: bench<2><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 3: bench<3><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 4: bench<4><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 5: bench<5><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            case 6: bench<6><<<1, 32>>>(dA, dc, dout, 50, 1); break;
            default: bench<7><<<1, 32>>>(dA, dc, dout, 50, 1); break;
        }
        cudaDeviceSynchronize();
        long long c; double o[2];
        cudaMemcpy(&c, dc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(o, dout, 16, cudaMemcpyDeviceToHost);
        printf("skip mask %d (1 = no trailing DMMA update, 2 = no rsqrt chain, 4 = no in-panel updates), warps %2d: chol32_panels %lld cycles per 32x32 block (%.1f per column), logdet sum %.6f, err %s\n", v, warps, c, c / 32.0, o[0],
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
