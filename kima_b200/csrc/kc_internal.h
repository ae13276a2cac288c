// Internal declarations shared by the libkima_cuda translation units.
#pragma once
#include <cuda_runtime.h>
#include <cstddef>
#include <cstdint>
#include <string>
#include "../../include/kima_cuda.h"

namespace kc {

constexpr int MAX_PLANETS = KC_MAX_NPMAX + KC_MAX_KO;  // per particle, in shared memory
constexpr int CTA_THREADS = 256;
constexpr int CTA_WARPS = CTA_THREADS / 32;

// Everything a kernel needs, passed by value (lives in the constant bank).
struct KParams {
    // Data singleton, uploaded once (src/Data.h:14-16)
    const double* t;
    const double* y;
    const double* sig2;     // sig[i]*sig[i], the product the reference forms each call
    const int32_t* obsi;    // 1-based
    const double* actind;   // [n_ind][N]
    int N, n_inst, n_ind;
    double M0_epoch, t_middle;
    // model
    kc_model_flags F;
    kc_layout L;
    // batch
    const double* params;   // [n_rows][ld]
    const int32_t* np;      // [B]
    int ld, B;
    double* logL;           // [B]
    int* counter;           // dynamic work counter (zeroed before launch)
    const int* order;       // small GP path: lane run by CTA i (lanes in order of planet count), null = i
};

// Launchers (kc_keplerian.cu, kc_gp.cu).  All asynchronous on `stream`.
cudaError_t launch_keplerian(const KParams& p, int kepler_mode, int sm_count, cudaStream_t stream,
                             int* n_launches);
cudaError_t launch_kepler_E(int n, const double* M, const double* e, double* E, cudaStream_t s);
cudaError_t launch_true_anomaly(int n, const double* t, const double* P, const double* e,
                                const double* Tp, double* f, cudaStream_t s);
cudaError_t launch_model_mu(const KParams& p, int kepler_mode, int lane, double* mu, cudaStream_t s);

struct GpLane;  // kc_gp_common.cuh
struct GpWorkspace {
    double* ws;          // per-slot factor storage (lower triangle of L + the residual row)
    size_t slot_doubles; // doubles per slot
    int slots;           // lanes in flight per wave
    GpLane* lanes;       // per-slot state
    // lane groups factorised concurrently: group 0 on the context stream, the others on aux
    // streams, so that one group's latency-bound diagonal blocks overlap another's updates
    int nstreams;               // groups (1..4)
    cudaStream_t aux[3];
    cudaEvent_t fork, join[3];
    // TMA descriptors (CUtensorMap, 128 bytes each) over ws as a (row, column, slot) tensor: boxes of
    // the update kernel's A and B panels (kc_gp_blocked.cu: gpb_make_tensor_maps)
    alignas(64) unsigned char tmap_a[128];
    alignas(64) unsigned char tmap_b[128];
    bool tma_ok;
};
// posterior-predictive GP mean, computed wave by wave while the factors are still in HBM
// (kc_predict.cu)
struct GpPredict {
    const double* t_pred;  // device [nt], null = the data epochs
    int nt;
    double* alpha;         // device [B][N]: C^-1 r per lane
    double* gp_out;        // device [B][nt], null = no mean
    double* std_out;       // device [B][nt], null = no predictive standard deviation
};
size_t gpb_ld(int N);
size_t gpb_slot_doubles(int N);
size_t gpb_lane_bytes();
cudaError_t gpb_make_tensor_maps(GpWorkspace& w, int N);
cudaError_t launch_gp_blocked(const KParams& p, const GpWorkspace& w, int kepler_mode, int sm_count,
                              cudaStream_t stream, int* n_launches, const GpPredict* pred = nullptr);
cudaError_t launch_gp_predict_wave(const KParams& p, const double* ws, size_t slot_doubles,
                                   const GpLane* lanes, int lane0, int n, const GpPredict& pr,
                                   cudaStream_t s, int* n_launches);
cudaError_t launch_predict_model(const KParams& p, int kepler_mode, const double* d_tpred, int nt,
                                 double* d_out, cudaStream_t s, int* n_launches);
// small-N path (kc_gp_small.cu): shared-memory-resident, one CTA per particle
size_t gps_smem_bytes(int N);  // 0 = N does not fit
// Factor cache of the small path (kc_gp_small.cu, "factor cache"): two entries per slot -- a slot is whatever
// the caller says carries one particle from evaluation to evaluation (the samplers: the particle index) -- each
// the factor of the last covariance matrices seen in that slot, keyed by the parameters C depends on.
struct GpsCacheMeta {
    double key[4 + KC_MAX_INST];  // eta1..eta4 and the jitters, as given
    double logdet;
    int valid, stamp;             // stamp: launch counter of the last use (the older entry is replaced)
};
struct GpsCache {
    double* img;                  // [n_slots][2][img_doubles]: the factor as it sits in shared memory + 1/diag; null = no cache
    GpsCacheMeta* meta;           // [n_slots][2]
    const int32_t* slots;         // [B]: slot of every lane of this launch (< 0: not cached)
    size_t img_doubles;
    int stamp;
    // per-launch work lists (gps_classify_kernel), each [max_batch]
    int* cls;                     // lane -> entry that holds its factor (0 / 1), -1: not cached, -2 - e: a miss that fills entry e
    int* hit_list;
    int* miss_list;
    int* cnt;                     // [0]: hits, [1]: misses
    int debug;                    // developer switch (KIMA_CUDA_GP_CACHE_DEBUG): 1 = every cached lane is treated as a miss
};
size_t gps_cache_img_doubles(int N);
void launch_gp_cache_compare(const KParams& p, const double* plain, const GpsCache& gc, cudaStream_t s);  // developer check
cudaError_t launch_gp_small(const KParams& p, int kepler_mode, cudaStream_t stream, int* n_launches,
                            const GpsCache* cache = nullptr, int* order_scratch = nullptr);
// kernel = celerite (kc_celerite.cu): resid is [N][ldr] doubles of scratch
cudaError_t launch_gp_celerite(const KParams& p, double* resid, int ldr, int kepler_mode,
                               cudaStream_t stream, int* n_launches);
cudaError_t launch_celerite_predict(const KParams& p, double* resid, double* fac, int ldr, int kepler_mode,
                                    const double* d_tpred, int nt, double* d_out, cudaStream_t stream,
                                    int* n_launches);
cudaError_t launch_gp_cov(const KParams& p, int lane, double* C, cudaStream_t s);

// The communicator state of a context, as kc_comm.cu sees it.
struct CommView {
    int device;
    cudaStream_t stream;
    void** nccl_lib;
    void** nccl_comm;
    int* rank;
    int* nranks;
    void** d_buf;
    size_t* buf_bytes;
};

cudaError_t measure_fp64_peak(int reps, double* dfma_tflops, double* dmma_tflops, cudaStream_t s);

}  // namespace kc

struct kc_ctx;
namespace kc {
CommView comm_view(kc_ctx* c);                       // kc_api.cu
int comm_fail(int code, const std::string& msg);     // kc_api.cu
// kc_comm.cu: device-to-device collectives on the context's stream
int comm_ranks(kc_ctx* c);  // 1 without a communicator
int comm_allgather_dev(kc_ctx* c, const void* d_send, void* d_recv, size_t nbytes);
int comm_allreduce_u64_dev(kc_ctx* c, uint64_t* d, size_t n);
}  // namespace kc
