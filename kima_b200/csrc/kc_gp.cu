// Kernel family 2, test hook: the full quasi-periodic covariance matrix of one particle
// (RVmodel::calculate_C, /root/reference/src/RVmodel.cpp:170-216).  The product path never
// materialises C (kc_gp_blocked.cu generates each entry inside the tile that consumes it);
// this kernel exists so tests can compare the entries themselves with the reference's.
#include "kc_model.cuh"
#include "kc_gp_common.cuh"

namespace kc {

// ---- test hook: the full covariance matrix of one particle (calculate_C) ------------
__global__ void __launch_bounds__(CTA_THREADS) gp_cov_kernel(const KParams p, int b, double* C)
{
    __shared__ ParticleScalars sc;
    __shared__ PlanetInv dummy[1];
    __shared__ GpConsts g;
    if (threadIdx.x < 32) {
        // only the scalars are needed; np is forced to 0 planets worth of work
        const kc_layout& L = p.L;
        if (threadIdx.x == 0) {
            sc.eta1 = ldp(p, L.row_eta, b); sc.eta2 = ldp(p, L.row_eta + 1, b);
            sc.eta3 = ldp(p, L.row_eta + 2, b); sc.eta4 = ldp(p, L.row_eta + 3, b);
            g.eta1sq = kmul(sc.eta1, sc.eta1);
            g.eta2 = sc.eta2; g.r_eta2 = kdiv(1.0, sc.eta2);
            g.eta3 = sc.eta3; g.r_eta3 = kdiv(1.0, sc.eta3);
            g.eta4 = sc.eta4; g.r_eta4 = kdiv(1.0, sc.eta4);
        }
        if (threadIdx.x < KC_MAX_INST) {
            double j2 = 0.0;
            if (threadIdx.x < L.n_jit) { double jt = ldp(p, L.row_jit + threadIdx.x, b); j2 = kmul(jt, jt); }
            sc.jit2[threadIdx.x] = j2;
        }
    }
    (void)dummy;
    __syncthreads();
    const int N = p.N;
    const bool multi = p.F.multi_instrument != 0;
    for (size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x; e < (size_t)N * N;
         e += (size_t)gridDim.x * blockDim.x) {
        int i = (int)(e % N), j = (int)(e / N);  // column-major like Eigen::MatrixXd
        int lo = i < j ? i : j, hi = i < j ? j : i;
        double v;
        if (i == j) {
            int in = multi ? p.obsi[i] - 1 : 0;
            v = kadd(g.eta1sq, kadd(p.sig2[i], sc.jit2[in]));
        } else {
            v = qp_kernel(g, kadd(p.t[lo], -p.t[hi]));  // the reference fills (i<j) and mirrors
        }
        C[e] = v;
    }
}

cudaError_t launch_gp_cov(const KParams& p, int lane, double* C, cudaStream_t s)
{
    int grid = (int)(((size_t)p.N * p.N + CTA_THREADS - 1) / CTA_THREADS);
    if (grid > 4096) grid = 4096;
    gp_cov_kernel<<<grid, CTA_THREADS, 0, s>>>(p, lane, C);
    return cudaGetLastError();
}

}  // namespace kc
