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
            g = gp_consts(sc);
        }
        if (threadIdx.x < KC_MAX_INST) {
            double j2 = 0.0;
            if (threadIdx.x < L.n_jit) { double jt = ldp(p, L.row_jit + threadIdx.x, b); j2 = kmul(jt, jt); }
            sc.jit2[threadIdx.x] = j2;
        }
    }
    (void)dummy;
    __shared__ double etab[64];
    exp_tab_load(etab, threadIdx.x);
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
            // the reference fills (i<j) and mirrors; the generator is the product path's (qp_entry)
            double si, ci, sj, cj;
            gp_phase(g, p.t[lo], si, ci);
            gp_phase(g, p.t[hi], sj, cj);
            v = qp_entry(g, etab, kadd(p.t[lo], -p.t[hi]), si, ci, sj, cj);
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
