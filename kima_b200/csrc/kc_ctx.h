// The context behind the C ABI, shared by the translation units that implement
// include/kima_cuda.h (kc_api.cu) and include/kima_sampler.h (kc_sampler_api.cu).
#pragma once
#include "kc_internal.h"

namespace kc { struct SamplerState; }

struct kc_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    // data
    int N = 0, n_inst = 1, n_ind = 0;
    double M0_epoch = 0, t_middle = 0;
    double *d_t = nullptr, *d_y = nullptr, *d_sig2 = nullptr, *d_actind = nullptr;
    int32_t* d_obsi = nullptr;
    bool have_data = false;
    // model
    kc_model_flags F{};
    kc_layout L{};
    bool configured = false;
    int max_batch = 0, ld = 0;
    double *h_params = nullptr, *d_params = nullptr;
    int32_t *h_np = nullptr, *d_np = nullptr;
    double *d_logL = nullptr, *h_logL = nullptr;
    int* d_counter = nullptr;
    // GP
    kc::GpWorkspace gp{};
    int gp_streams_opt = 2;
    double* d_resid = nullptr;   // kernel = celerite: residuals, [N][ld]
    // factor cache of the shared-memory-resident GP path (kc_gp_cache_enable)
    kc::GpsCache gpc{};
    int gpc_nslots = 0;
    int32_t *d_slots = nullptr, *h_slots = nullptr;  // slots of the lanes of the next evaluation
    int slots_B = 0;                                  // lanes kc_set_slots named (0: none pending)
    cudaEvent_t slots_ev = nullptr;                   // the last upload of h_slots
    double* d_chk = nullptr;                          // developer check (KIMA_CUDA_GP_CACHE_DEBUG=2): uncached results
    // options / instrumentation
    int kepler_mode = 0;
    int gp_slots_opt = 0;
    int eval_chunk_opt = 32768;  // kc_eval: batches of >= 2 chunks overlap H2D with the kernels (0 = off)
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_ev = nullptr;
    int gp_order_opt = 1;   // small GP path: batches of >= 1024 lanes run in order of planet count (keeps the SMs in step)
    int* d_order = nullptr;
    bool np_uniform = false; // set by kc_eval_async around its launches: every lane of the batch has the same planet count
    int gp_cache_opt = 1;   // kc_sampler_create: one factor-cache slot per particle where the cache applies
    int gp_small_opt = 1;   // 1 = shared-memory-resident kernel whenever N fits, 0 = always blocked
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool timed = false;
    float last_ms = 0.f;
    int64_t launches = 0;
    double* pending_out = nullptr;
    int pending_B = 0;
    // comm
    void* nccl_lib = nullptr;
    void* nccl_comm = nullptr;
    int rank = 0, nranks = 1;
    void* d_comm_buf = nullptr;
    size_t comm_buf_bytes = 0;
    void* d_flush = nullptr;
    // device-side sampler (kc_sampler_api.cu), null until kc_sampler_create
    kc::SamplerState* sampler = nullptr;
    // posterior-predictive scratch, grown on demand
    double* d_pred = nullptr;
    size_t pred_doubles = 0;
};

namespace kc {
// kc_api.cu: thread-local error message + the launch of the likelihood kernels of the context
// over `B` columns of a device parameter block of the context's shape (params [n_rows][ld],
// np [B]); asynchronous on the context stream.
int fail(int code, const std::string& msg);
int fail_cuda(cudaError_t e, const char* what);
int run_kernels(kc_ctx* c, int B, double* d_logL, const double* d_params, const int32_t* d_np,
                bool timed = true, const int32_t* d_slots = nullptr);
void gp_cache_free(kc_ctx* c);
int check_ready(kc_ctx* c, int B);
void sampler_free(kc_ctx* c);  // kc_sampler_api.cu
}  // namespace kc
