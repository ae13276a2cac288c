// libkima_cuda: the C ABI (include/kima_cuda.h) over the two kernel families.
// Context, buffers, copies and launch order only — no arithmetic lives here.
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "kc_ctx.h"

namespace {
thread_local std::string g_err;
}

namespace kc {
int fail(int code, const std::string& msg)
{
    g_err = msg;
    return code;
}
int fail_cuda(cudaError_t e, const char* what)
{
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return KC_ERR_CUDA;
}
}  // namespace kc
using kc::fail;
using kc::fail_cuda;
namespace {
#define CU(call)                                                  \
    do {                                                          \
        cudaError_t e_ = (call);                                  \
        if (e_ != cudaSuccess) return fail_cuda(e_, #call);       \
    } while (0)

}  // namespace


namespace {

void free_model(kc_ctx* c)
{
    kc::sampler_free(c);
    kc::gp_cache_free(c);
    if (c->h_params) cudaFreeHost(c->h_params);
    if (c->h_np) cudaFreeHost(c->h_np);
    if (c->h_logL) cudaFreeHost(c->h_logL);
    cudaFree(c->d_params); cudaFree(c->d_np); cudaFree(c->d_logL);
    cudaFree(c->gp.ws); cudaFree(c->gp.lanes); cudaFree(c->d_resid);
    cudaFree(c->d_order);
    c->d_order = nullptr;
    c->d_resid = nullptr;
    for (int i = 0; i < 3; i++) {
        if (c->gp.aux[i]) cudaStreamDestroy(c->gp.aux[i]);
        if (c->gp.join[i]) cudaEventDestroy(c->gp.join[i]);
    }
    if (c->gp.fork) cudaEventDestroy(c->gp.fork);
    c->h_params = nullptr; c->h_np = nullptr; c->h_logL = nullptr;
    c->d_params = nullptr; c->d_np = nullptr; c->d_logL = nullptr;
    c->gp = kc::GpWorkspace{};
    c->configured = false;
}
void free_data(kc_ctx* c)
{
    kc::gp_cache_free(c);  // the cached factors are those of the old epochs
    cudaFree(c->d_t); cudaFree(c->d_y); cudaFree(c->d_sig2); cudaFree(c->d_actind);
    cudaFree(c->d_obsi);
    c->d_t = c->d_y = c->d_sig2 = c->d_actind = nullptr;
    c->d_obsi = nullptr;
    c->have_data = false;
}

kc::KParams make_params(kc_ctx* c, int B, double* d_logL, const double* d_params = nullptr,
                        const int32_t* d_np = nullptr)
{
    kc::KParams p{};
    p.t = c->d_t; p.y = c->d_y; p.sig2 = c->d_sig2; p.obsi = c->d_obsi; p.actind = c->d_actind;
    p.N = c->N; p.n_inst = c->n_inst; p.n_ind = c->n_ind;
    p.M0_epoch = c->M0_epoch; p.t_middle = c->t_middle;
    p.F = c->F; p.L = c->L;
    p.params = d_params ? d_params : c->d_params; p.np = d_np ? d_np : c->d_np; p.ld = c->ld; p.B = B;
    p.logL = d_logL; p.counter = c->d_counter;
    return p;
}

}  // namespace

namespace kc {
void gp_cache_free(kc_ctx* c)
{
    cudaFree(c->gpc.img);
    cudaFree(c->gpc.meta);
    cudaFree(c->gpc.cls);  // one allocation: cls, hit_list, miss_list, cnt
    cudaFree(c->d_chk);
    c->d_chk = nullptr;
    cudaFree(c->d_slots);
    if (c->h_slots) cudaFreeHost(c->h_slots);
    if (c->slots_ev) cudaEventDestroy(c->slots_ev);
    c->slots_ev = nullptr;
    c->gpc = GpsCache{};
    c->gpc_nslots = 0;
    c->d_slots = c->h_slots = nullptr;
    c->slots_B = 0;
}

int run_kernels(kc_ctx* c, int B, double* d_logL, const double* d_params, const int32_t* d_np, bool timed,
                const int32_t* d_slots)
{
    kc::KParams p = make_params(c, B, d_logL, d_params, d_np);
    int n = 0;
    if (timed) CU(cudaEventRecord(c->ev0, c->stream));
    cudaError_t e;
    if (c->F.GP && c->F.kernel == 1)
        e = kc::launch_gp_celerite(p, c->d_resid, c->ld, c->kepler_mode, c->stream, &n);
    else if (c->F.GP && c->gp_small_opt && kc::gps_smem_bytes(c->N))
    {
        kc::GpsCache gc = c->gpc;
        gc.slots = d_slots;
        gc.stamp = ++c->gpc.stamp;
        if (!c->d_order && c->gp_order_opt) CU(cudaMalloc(&c->d_order, sizeof(int) * (size_t)c->max_batch));
        e = kc::launch_gp_small(p, c->kepler_mode, c->stream, &n, gc.img && d_slots ? &gc : nullptr,
                                c->gp_order_opt && !c->np_uniform ? c->d_order : nullptr);
        if (e == cudaSuccess && gc.img && d_slots && gc.debug >= 2) {
            // developer check (KIMA_CUDA_GP_CACHE_DEBUG=2): the same lanes through the uncached kernel, compared on the device
            if (!c->d_chk) cudaMalloc(&c->d_chk, sizeof(double) * c->max_batch);
            double* d_chk = c->d_chk;
            kc::KParams q = p;
            q.logL = d_chk;
            e = kc::launch_gp_small(q, c->kepler_mode, c->stream, &n, nullptr);
            if (e == cudaSuccess) kc::launch_gp_cache_compare(p, d_chk, gc, c->stream);
        }
    }
    else if (c->F.GP) e = kc::launch_gp_blocked(p, c->gp, c->kepler_mode, c->sm_count, c->stream, &n);
    else e = kc::launch_keplerian(p, c->kepler_mode, c->sm_count, c->stream, &n);
    if (e != cudaSuccess) return fail_cuda(e, "kernel launch");
    if (timed) {
        CU(cudaEventRecord(c->ev1, c->stream));
        c->timed = true;
    }
    c->launches += n;
    return KC_OK;
}

int check_ready(kc_ctx* c, int B)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    if (!c->have_data) return fail(KC_ERR_ARG, "kc_upload_data has not been called");
    if (!c->configured) return fail(KC_ERR_ARG, "kc_configure has not been called");
    if (B < 0 || B > c->max_batch) return fail(KC_ERR_ARG, "B outside [0, max_batch]");
    return KC_OK;
}
}  // namespace kc
using kc::check_ready;
namespace {
int run_kernels(kc_ctx* c, int B, double* d_logL) { return kc::run_kernels(c, B, d_logL, nullptr, nullptr); }
}  // namespace

namespace kc {
CommView comm_view(kc_ctx* c)
{
    return {c->device, c->stream, &c->nccl_lib, &c->nccl_comm, &c->rank, &c->nranks,
            &c->d_comm_buf, &c->comm_buf_bytes};
}
int comm_fail(int code, const std::string& msg) { return fail(code, msg); }
}  // namespace kc

extern "C" {

const char* kc_last_error(void) { return g_err.c_str(); }
const char* kc_version(void) { return "kima_cuda 0.1 sm_100a"; }

int kc_create(int device, kc_ctx** out)
{
    if (!out) return fail(KC_ERR_ARG, "out is null");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(KC_ERR_CUDA, std::string("no CUDA device: libkima_cuda has no CPU fallback (") +
                                     cudaGetErrorString(e) + ")");
    if (device < 0 || device >= ndev) return fail(KC_ERR_ARG, "device index out of range");
    CU(cudaSetDevice(device));
    kc_ctx* c = new kc_ctx();
    c->device = device;
    // a failure half-way leaves nothing behind: kc_destroy tolerates the null members
    auto bail = [&](cudaError_t err, const char* what) {
        kc_destroy(c);
        return fail_cuda(err, what);
    };
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return bail(e, "cudaGetDeviceProperties");
    c->sm_count = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaMalloc(&c->d_counter, 64)) != cudaSuccess) return bail(e, "cudaMalloc");
    *out = c;
    return KC_OK;
}

int kc_destroy(kc_ctx* c)
{
    if (!c) return KC_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    kc_comm_destroy(c);
    free_model(c);
    free_data(c);
    cudaFree(c->d_counter);
    cudaFree(c->d_comm_buf);
    cudaFree(c->d_flush);
    cudaFree(c->d_pred);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->copy_ev) cudaEventDestroy(c->copy_ev);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return KC_OK;
}

int kc_upload_data(kc_ctx* c, int N, const double* t, const double* y, const double* sig,
                   const int32_t* obsi, int n_inst, int n_ind, const double* actind,
                   double M0_epoch, double t_middle)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    if (N <= 0 || !t || !y || !sig) return fail(KC_ERR_ARG, "N <= 0 or null data array");
    if (n_inst < 1 || n_inst > KC_MAX_INST) return fail(KC_ERR_ARG, "n_inst outside [1, KC_MAX_INST]");
    if (n_ind < 0 || n_ind > KC_MAX_IND) return fail(KC_ERR_ARG, "n_ind outside [0, KC_MAX_IND]");
    if (n_ind > 0 && !actind) return fail(KC_ERR_ARG, "n_ind > 0 but actind is null");
    if (n_inst > 1 && !obsi) return fail(KC_ERR_ARG, "n_inst > 1 but obsi is null");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    free_model(c);  // capacity depends on N
    free_data(c);
    std::vector<double> sig2(N);
    std::vector<int32_t> ob(N);
    for (int i = 0; i < N; i++) {
        sig2[i] = sig[i] * sig[i];  // the product every reference call forms (:201,:900,:921)
        ob[i] = obsi ? obsi[i] : 1;
        if (ob[i] < 1 || ob[i] > n_inst) return fail(KC_ERR_ARG, "obsi must be 1-based ids <= n_inst");
    }
    size_t nb = sizeof(double) * N;
    CU(cudaMalloc(&c->d_t, nb));
    CU(cudaMalloc(&c->d_y, nb));
    CU(cudaMalloc(&c->d_sig2, nb));
    CU(cudaMalloc(&c->d_obsi, sizeof(int32_t) * N));
    CU(cudaMemcpy(c->d_t, t, nb, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_y, y, nb, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_sig2, sig2.data(), nb, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(c->d_obsi, ob.data(), sizeof(int32_t) * N, cudaMemcpyHostToDevice));
    if (n_ind > 0) {
        CU(cudaMalloc(&c->d_actind, nb * n_ind));
        CU(cudaMemcpy(c->d_actind, actind, nb * n_ind, cudaMemcpyHostToDevice));
    }
    c->N = N; c->n_inst = n_inst; c->n_ind = n_ind;
    c->M0_epoch = M0_epoch; c->t_middle = t_middle;
    c->have_data = true;
    return KC_OK;
}

int kc_configure(kc_ctx* c, const kc_model_flags* f, int max_batch)
{
    if (!c || !f) return fail(KC_ERR_ARG, "null argument");
    if (!c->have_data) return fail(KC_ERR_ARG, "kc_upload_data must come before kc_configure");
    if (max_batch < 1) return fail(KC_ERR_ARG, "max_batch < 1");
    if (f->kernel != 0 && f->kernel != 1) return fail(KC_ERR_ARG, "kernel must be 0 (standard) or 1 (celerite)");
    // the reference's celerite branch reads extra_sigma, which multi_instrument leaves unset
    // (src/RVmodel.cpp:258 against :124-127)
    if (f->GP && f->kernel == 1 && f->multi_instrument)
        return fail(KC_ERR_UNSUPPORTED, "kernel=celerite with multi_instrument: the reference reads an unset jitter");
    if (f->npmax < 0 || f->npmax > KC_MAX_NPMAX) return fail(KC_ERR_ARG, "npmax outside [0, KC_MAX_NPMAX]");
    if (f->known_object && (f->n_known_object < 0 || f->n_known_object > KC_MAX_KO))
        return fail(KC_ERR_ARG, "n_known_object outside [0, KC_MAX_KO]");
    if (f->trend && (f->degree < 1 || f->degree > 3))
        return fail(KC_ERR_ARG, "trend=true needs degree in 1..3 (src/RVmodel.cpp:43-47)");
    if (f->multi_instrument && c->n_inst < 1) return fail(KC_ERR_ARG, "multi_instrument without instruments");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    free_model(c);
    c->F = *f;
    kc_layout_from_flags(f, c->n_inst, c->n_ind, &c->L);
    c->max_batch = max_batch;
    c->ld = (max_batch + 3) & ~3;
    size_t pb = sizeof(double) * (size_t)c->L.n_rows * c->ld;
    CU(cudaMallocHost(&c->h_params, pb));
    CU(cudaMallocHost(&c->h_np, sizeof(int32_t) * c->ld));
    CU(cudaMallocHost(&c->h_logL, sizeof(double) * c->ld));
    std::memset(c->h_params, 0, pb);
    std::memset(c->h_np, 0, sizeof(int32_t) * c->ld);
    CU(cudaMalloc(&c->d_params, pb));
    CU(cudaMalloc(&c->d_np, sizeof(int32_t) * c->ld));
    CU(cudaMalloc(&c->d_logL, sizeof(double) * c->ld));
    CU(cudaMemset(c->d_params, 0, pb));
    CU(cudaMemset(c->d_np, 0, sizeof(int32_t) * c->ld));
    if (f->GP && f->kernel == 1) {
        CU(cudaMalloc(&c->d_resid, sizeof(double) * (size_t)c->N * c->ld));
    } else if (f->GP) {
        // factor storage for as many lanes in flight as fit the budget; more lanes than
        // slots run in waves
        c->gp.slot_doubles = kc::gpb_slot_doubles(c->N);
        // budget: 64 GB, or 80% of what the device has free right now if that is less
        size_t free_b = 0, total_b = 0;
        CU(cudaMemGetInfo(&free_b, &total_b));
        size_t cap = (size_t)64 << 30;
        if (cap > free_b / 5 * 4) cap = free_b / 5 * 4;
        size_t fit = cap / (sizeof(double) * c->gp.slot_doubles);
        if (fit < 1) fit = 1;
        int slots = c->gp_slots_opt > 0 ? c->gp_slots_opt : (int)(fit > (size_t)max_batch ? (size_t)max_batch : fit);
        if (slots > max_batch) slots = max_batch;
        c->gp.slots = slots;
        CU(cudaMalloc(&c->gp.ws, sizeof(double) * c->gp.slot_doubles * slots));
        CU(cudaMalloc((void**)&c->gp.lanes, kc::gpb_lane_bytes() * slots));
        {
            cudaError_t te = kc::gpb_make_tensor_maps(c->gp, c->N);
            if (te != cudaSuccess) return fail_cuda(te, "cuTensorMapEncodeTiled over the GP factor storage (TMA descriptors)");
            c->gp.tma_ok = true;
        }
        c->gp.nstreams = c->gp_streams_opt;
        CU(cudaEventCreateWithFlags(&c->gp.fork, cudaEventDisableTiming));
        for (int i = 0; i + 1 < c->gp.nstreams; i++) {
            CU(cudaStreamCreateWithFlags(&c->gp.aux[i], cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&c->gp.join[i], cudaEventDisableTiming));
        }
    }
    c->configured = true;
    return KC_OK;
}

int kc_get_layout(kc_ctx* c, kc_layout* out)
{
    if (!c || !out) return fail(KC_ERR_ARG, "null argument");
    if (!c->configured) return fail(KC_ERR_ARG, "kc_configure has not been called");
    *out = c->L;
    return KC_OK;
}

int kc_host_params(kc_ctx* c, double** params, int32_t** np, int* ld)
{
    if (!c || !c->configured) return fail(KC_ERR_ARG, "context not configured");
    if (params) *params = c->h_params;
    if (np) *np = c->h_np;
    if (ld) *ld = c->ld;
    return KC_OK;
}

int kc_device_params(kc_ctx* c, double** d_params, int32_t** d_np, int* ld)
{
    if (!c || !c->configured) return fail(KC_ERR_ARG, "context not configured");
    if (d_params) *d_params = c->d_params;
    if (d_np) *d_np = c->d_np;
    if (ld) *ld = c->ld;
    return KC_OK;
}

int kc_eval_async(kc_ctx* c, int B, double* logL_host)
{
    int rc = check_ready(c, B);
    if (rc) return rc;
    if (!logL_host) return fail(KC_ERR_ARG, "logL_host is null");
    if (c->pending_out) return fail(KC_ERR_ARG, "an asynchronous evaluation is already in flight");
    CU(cudaSetDevice(c->device));
    if (B == 0) return KC_OK;
    // slots named by kc_set_slots for exactly this batch: consumed here
    const int32_t* slots = nullptr;
    if (c->slots_B == B && c->gpc.img) {
        CU(cudaMemcpyAsync(c->d_slots, c->h_slots, sizeof(int32_t) * B, cudaMemcpyHostToDevice, c->stream));
        CU(cudaEventRecord(c->slots_ev, c->stream));
        slots = c->d_slots;
    }
    c->slots_B = 0;
    // the planet counts are at hand here: a batch in which they are all equal needs no ordering (gp_order)
    c->np_uniform = true;
    for (int i = 1; i < B && c->np_uniform; i++) c->np_uniform = c->h_np[i] == c->h_np[0];
    struct Reset { bool& f; ~Reset() { f = false; } } reset_hint{c->np_uniform};
    const int chunk = c->eval_chunk_opt;
    if (chunk > 0 && B >= 2 * chunk) {
        // large batches: the columns travel in chunks on a copy stream while the kernels of the
        // previous chunk run, so that PCIe time hides behind compute (one event, re-recorded:
        // cudaStreamWaitEvent binds to the record that precedes it)
        if (!c->copy_stream) {
            CU(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
            CU(cudaEventCreateWithFlags(&c->copy_ev, cudaEventDisableTiming));
        }
        CU(cudaEventRecord(c->copy_ev, c->stream));       // earlier readers of the device block
        CU(cudaStreamWaitEvent(c->copy_stream, c->copy_ev, 0));
        CU(cudaEventRecord(c->ev0, c->stream));
        for (int c0 = 0; c0 < B; c0 += chunk) {
            const int nb = B - c0 < chunk ? B - c0 : chunk;
            CU(cudaMemcpy2DAsync(c->d_params + c0, sizeof(double) * c->ld, c->h_params + c0,
                                 sizeof(double) * c->ld, sizeof(double) * nb, c->L.n_rows,
                                 cudaMemcpyHostToDevice, c->copy_stream));
            CU(cudaMemcpyAsync(c->d_np + c0, c->h_np + c0, sizeof(int32_t) * nb, cudaMemcpyHostToDevice,
                               c->copy_stream));
            CU(cudaEventRecord(c->copy_ev, c->copy_stream));
            CU(cudaStreamWaitEvent(c->stream, c->copy_ev, 0));
            rc = kc::run_kernels(c, nb, c->d_logL + c0, c->d_params + c0, c->d_np + c0, false, slots ? slots + c0 : nullptr);
            if (rc) return rc;
        }
        CU(cudaEventRecord(c->ev1, c->stream));
        c->timed = true;
    } else {
        // one strided copy moves columns [0,B) of every row
        CU(cudaMemcpy2DAsync(c->d_params, sizeof(double) * c->ld, c->h_params, sizeof(double) * c->ld,
                             sizeof(double) * B, c->L.n_rows, cudaMemcpyHostToDevice, c->stream));
        CU(cudaMemcpyAsync(c->d_np, c->h_np, sizeof(int32_t) * B, cudaMemcpyHostToDevice, c->stream));
        rc = kc::run_kernels(c, B, c->d_logL, nullptr, nullptr, true, slots);
        if (rc) return rc;
    }
    CU(cudaMemcpyAsync(c->h_logL, c->d_logL, sizeof(double) * B, cudaMemcpyDeviceToHost, c->stream));
    c->pending_out = logL_host;
    c->pending_B = B;
    return KC_OK;
}

int kc_wait(kc_ctx* c)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    if (c->pending_out) {
        std::memcpy(c->pending_out, c->h_logL, sizeof(double) * c->pending_B);
        c->pending_out = nullptr;
        c->pending_B = 0;
    }
    return KC_OK;
}

int kc_eval(kc_ctx* c, int B, double* logL_host)
{
    int rc = kc_eval_async(c, B, logL_host);
    if (rc) return rc;
    return kc_wait(c);
}

int kc_eval_device(kc_ctx* c, int B, double* d_logL)
{
    int rc = check_ready(c, B);
    if (rc) return rc;
    CU(cudaSetDevice(c->device));
    if (B == 0) return KC_OK;
    const int32_t* slots = nullptr;
    if (c->slots_B == B && c->gpc.img) {
        CU(cudaMemcpyAsync(c->d_slots, c->h_slots, sizeof(int32_t) * B, cudaMemcpyHostToDevice, c->stream));
        CU(cudaEventRecord(c->slots_ev, c->stream));
        slots = c->d_slots;
    }
    c->slots_B = 0;
    return kc::run_kernels(c, B, d_logL ? d_logL : c->d_logL, nullptr, nullptr, true, slots);
}

int kc_gp_cache_enable(kc_ctx* c, int n_slots)
{
    if (!c || !c->configured) return fail(KC_ERR_ARG, "context not configured");
    if (n_slots < 0) return fail(KC_ERR_ARG, "n_slots is negative");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    kc::gp_cache_free(c);
    // only the shared-memory-resident dense path keeps factors (and not with the MA(1) term, whose residual
    // recurrence borrows the factor storage): elsewhere the call succeeds and the cache stays inactive
    if (n_slots == 0 || !c->F.GP || c->F.kernel != 0 || c->F.MA || !c->gp_small_opt || !kc::gps_smem_bytes(c->N))
        return KC_OK;
    const size_t img = kc::gps_cache_img_doubles(c->N);
    const size_t bytes = (size_t)n_slots * 2 * img * sizeof(double);
    size_t free_b = 0, total_b = 0;
    CU(cudaMemGetInfo(&free_b, &total_b));
    if (bytes > free_b / 10 * 9) return fail(KC_ERR_ARG, "factor cache: n_slots x 2 factors do not fit the free device memory");
    cudaError_t e = cudaMalloc(&c->gpc.img, bytes);
    if (e != cudaSuccess) { kc::gp_cache_free(c); return fail_cuda(e, "cudaMalloc (factor cache)"); }
    e = cudaMalloc(&c->gpc.meta, sizeof(kc::GpsCacheMeta) * 2 * (size_t)n_slots);
    if (e == cudaSuccess) e = cudaMemsetAsync(c->gpc.meta, 0, sizeof(kc::GpsCacheMeta) * 2 * (size_t)n_slots, c->stream);
    if (e == cudaSuccess) e = cudaMalloc(&c->d_slots, sizeof(int32_t) * (size_t)c->max_batch);
    if (e == cudaSuccess) e = cudaMalloc(&c->gpc.cls, sizeof(int) * (3 * (size_t)c->max_batch + 2));
    if (e == cudaSuccess) {
        c->gpc.hit_list = c->gpc.cls + c->max_batch;
        c->gpc.miss_list = c->gpc.hit_list + c->max_batch;
        c->gpc.cnt = c->gpc.miss_list + c->max_batch;
    }
    if (e == cudaSuccess) e = cudaMallocHost(&c->h_slots, sizeof(int32_t) * (size_t)c->max_batch);
    if (e == cudaSuccess) e = cudaEventCreateWithFlags(&c->slots_ev, cudaEventDisableTiming);
    if (e != cudaSuccess) { kc::gp_cache_free(c); return fail_cuda(e, "factor cache allocation"); }
    c->gpc.img_doubles = img;
    c->gpc_nslots = n_slots;
    if (const char* dbg = std::getenv("KIMA_CUDA_GP_CACHE_DEBUG")) c->gpc.debug = std::atoi(dbg);
    return KC_OK;
}

int kc_gp_cache_slots(kc_ctx* c, int* n_slots)
{
    if (!c || !n_slots) return fail(KC_ERR_ARG, "null argument");
    *n_slots = c->gpc_nslots;
    return KC_OK;
}

int kc_set_slots(kc_ctx* c, int B, const int32_t* slots)
{
    int rc = check_ready(c, B);
    if (rc) return rc;
    if (!slots) return fail(KC_ERR_ARG, "slots is null");
    if (!c->gpc.img) return KC_OK;  // inactive cache: nothing to name
    if (c->pending_out) return fail(KC_ERR_ARG, "an asynchronous evaluation is in flight");
    for (int i = 0; i < B; i++)
        if (slots[i] >= c->gpc_nslots) return fail(KC_ERR_ARG, "slot index outside the cache");
    CU(cudaEventSynchronize(c->slots_ev));  // the previous upload has left the staging buffer
    std::memcpy(c->h_slots, slots, sizeof(int32_t) * (size_t)B);
    c->slots_B = B;
    return KC_OK;
}

int kc_sync(kc_ctx* c)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

int kc_stream(kc_ctx* c, void** stream)
{
    if (!c || !stream) return fail(KC_ERR_ARG, "null argument");
    *stream = (void*)c->stream;
    return KC_OK;
}

int kc_set_option(kc_ctx* c, const char* key, int value)
{
    if (!c || !key) return fail(KC_ERR_ARG, "null argument");
    std::string k(key);
    if (k == "kepler_mode") {
        if (value != 0 && value != 1) return fail(KC_ERR_ARG, "kepler_mode must be 0 or 1");
        c->kepler_mode = value;

    } else if (k == "gp_small") {
        c->gp_small_opt = value ? 1 : 0;
    } else if (k == "gp_streams") {
        if (value < 1 || value > 4) return fail(KC_ERR_ARG, "gp_streams must be 1..4");
        c->gp_streams_opt = value;  // takes effect at the next kc_configure
    } else if (k == "eval_chunk") {
        if (value < 0) return fail(KC_ERR_ARG, "eval_chunk must be >= 0 (0 = one copy, no overlap)");
        c->eval_chunk_opt = value;
    } else if (k == "gp_order") {
        c->gp_order_opt = value ? 1 : 0;  // small GP path: lanes handed to the SMs in order of planet count
    } else if (k == "gp_cache") {
        c->gp_cache_opt = value ? 1 : 0;  // kc_sampler_create attaches a factor cache (takes effect at the next call)
    } else if (k == "gp_slots") {
        if (value < 0) return fail(KC_ERR_ARG, "gp_slots must be >= 0");
        c->gp_slots_opt = value;  // takes effect at the next kc_configure
    } else {
        return fail(KC_ERR_ARG, "unknown option: " + k);
    }
    return KC_OK;
}

int kc_last_kernel_ms(kc_ctx* c, float* keplerian_ms, float* gp_ms)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    if (!c->timed) return fail(KC_ERR_ARG, "no evaluation has been timed yet");
    CU(cudaSetDevice(c->device));
    CU(cudaEventSynchronize(c->ev1));
    float ms = 0.f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    if (keplerian_ms) *keplerian_ms = c->F.GP ? 0.f : ms;
    if (gp_ms) *gp_ms = c->F.GP ? ms : 0.f;
    return KC_OK;
}

int kc_flush_l2(kc_ctx* c)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    const size_t bytes = (size_t)256 << 20;  // 2x the 126 MB L2
    if (!c->d_flush) CU(cudaMalloc(&c->d_flush, bytes));
    CU(cudaMemsetAsync(c->d_flush, 0, bytes, c->stream));
    return KC_OK;
}

int64_t kc_launch_count(kc_ctx* c) { return c ? c->launches : 0; }

int kc_measure_fp64_peak(kc_ctx* c, int reps, double* dfma_tflops, double* dmma_tflops)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    CU(cudaSetDevice(c->device));
    cudaError_t e = kc::measure_fp64_peak(reps, dfma_tflops, dmma_tflops, c->stream);
    if (e != cudaSuccess) return fail_cuda(e, "measure_fp64_peak");
    c->launches += 2 * (reps + 1);
    return KC_OK;
}

// ---- test hooks ------------------------------------------------------------------------
static int hook_arrays(kc_ctx* c, int n, int n_in, const double* const* in, double* out, int which)
{
    if (!c) return fail(KC_ERR_ARG, "null context");
    if (n <= 0) return KC_OK;
    CU(cudaSetDevice(c->device));
    double* d = nullptr;
    size_t nb = sizeof(double) * n;
    CU(cudaMalloc(&d, nb * (n_in + 1)));
    for (int k = 0; k < n_in; k++) CU(cudaMemcpyAsync(d + (size_t)k * n, in[k], nb, cudaMemcpyHostToDevice, c->stream));
    cudaError_t e;
    if (which == 0) e = kc::launch_kepler_E(n, d, d + n, d + 2 * (size_t)n, c->stream);
    else e = kc::launch_true_anomaly(n, d, d + n, d + 2 * (size_t)n, d + 3 * (size_t)n, d + 4 * (size_t)n, c->stream);
    if (e != cudaSuccess) { cudaFree(d); return fail_cuda(e, "hook launch"); }
    c->launches += 1;
    CU(cudaMemcpyAsync(out, d + (size_t)n_in * n, nb, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    return KC_OK;
}

int kc_kepler_E(kc_ctx* c, int n, const double* M, const double* ecc, double* E)
{
    const double* in[2] = {M, ecc};
    return hook_arrays(c, n, 2, in, E, 0);
}

int kc_true_anomaly(kc_ctx* c, int n, const double* t, const double* P, const double* ecc,
                    const double* Tp, double* f)
{
    const double* in[4] = {t, P, ecc, Tp};
    return hook_arrays(c, n, 4, in, f, 1);
}

static int stage_lane(kc_ctx* c, int lane)
{
    int rc = check_ready(c, lane + 1);
    if (rc) return rc;
    if (lane < 0) return fail(KC_ERR_ARG, "lane < 0");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpy2DAsync(c->d_params + lane, sizeof(double) * c->ld, c->h_params + lane,
                         sizeof(double) * c->ld, sizeof(double), c->L.n_rows,
                         cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_np + lane, c->h_np + lane, sizeof(int32_t), cudaMemcpyHostToDevice, c->stream));
    return KC_OK;
}

int kc_model_mu(kc_ctx* c, int lane, double* mu)
{
    int rc = stage_lane(c, lane);
    if (rc) return rc;
    if (!mu) return fail(KC_ERR_ARG, "mu is null");
    double* d = nullptr;
    CU(cudaMalloc(&d, sizeof(double) * c->N));
    kc::KParams p = make_params(c, lane + 1, c->d_logL);
    cudaError_t e = kc::launch_model_mu(p, c->kepler_mode, lane, d, c->stream);
    if (e != cudaSuccess) { cudaFree(d); return fail_cuda(e, "model_mu launch"); }
    c->launches += 1;
    CU(cudaMemcpyAsync(mu, d, sizeof(double) * c->N, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    return KC_OK;
}

int kc_gp_cov(kc_ctx* c, int lane, double* C)
{
    int rc = stage_lane(c, lane);
    if (rc) return rc;
    if (!C) return fail(KC_ERR_ARG, "C is null");
    if (!c->F.GP || c->F.kernel != 0) return fail(KC_ERR_ARG, "kc_gp_cov needs GP=true and the standard kernel");
    double* d = nullptr;
    size_t nb = sizeof(double) * (size_t)c->N * c->N;
    CU(cudaMalloc(&d, nb));
    kc::KParams p = make_params(c, lane + 1, c->d_logL);
    cudaError_t e = kc::launch_gp_cov(p, lane, d, c->stream);
    if (e != cudaSuccess) { cudaFree(d); return fail_cuda(e, "gp_cov launch"); }
    c->launches += 1;
    CU(cudaMemcpyAsync(C, d, nb, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    CU(cudaFree(d));
    return KC_OK;
}

static int predict_impl(kc_ctx* c, int B, int nt, const double* t_pred, int what, double* model_out,
                        double* gp_out, double* std_out)
{
    int rc = check_ready(c, B);
    if (rc) return rc;
    const bool want_model = (what & KC_PRED_MODEL) != 0, want_gp = (what & KC_PRED_GP) != 0;
    const bool want_std = (what & KC_PRED_GP_STD) != 0;
    if (!want_model && !want_gp && !want_std)
        return fail(KC_ERR_ARG, "what must name KC_PRED_MODEL, KC_PRED_GP and/or KC_PRED_GP_STD");
    if (what & ~(KC_PRED_MODEL | KC_PRED_GP | KC_PRED_GP_STD)) return fail(KC_ERR_ARG, "unknown bits in what");
    if (nt <= 0) return fail(KC_ERR_ARG, "nt <= 0");
    if (!t_pred && nt != c->N) return fail(KC_ERR_ARG, "t_pred is null (the data epochs) but nt != N");
    if (want_model && !model_out) return fail(KC_ERR_ARG, "model_out is null");
    if (want_gp && !gp_out) return fail(KC_ERR_ARG, "gp_out is null");
    if (want_std && !std_out) return fail(KC_ERR_ARG, "gp_std_out is null");
    if ((want_gp || want_std) && !c->F.GP) return fail(KC_ERR_UNSUPPORTED, "KC_PRED_GP / KC_PRED_GP_STD need GP=true");
    const bool cel = c->F.GP && c->F.kernel == 1;
    if (want_std && cel)
        return fail(KC_ERR_UNSUPPORTED, "KC_PRED_GP_STD needs the standard kernel (kernel=celerite predicts the mean only)");
    if (want_gp && cel && t_pred)  // celerite's predict walks the grid and the epochs together (cholesky.h:613-687)
        for (int k = 1; k < nt; k++)
            if (!(t_pred[k] >= t_pred[k - 1])) return fail(KC_ERR_ARG, "kernel=celerite: t_pred must be sorted ascending");
    if (c->pending_out) return fail(KC_ERR_ARG, "an asynchronous evaluation is in flight");
    CU(cudaSetDevice(c->device));
    if (B == 0) return KC_OK;
    const size_t out_d = (size_t)B * nt;
    // GP scratch: the standard kernel keeps alpha [B][N]; celerite its factor [N][8][ld]
    const size_t gp_scratch = cel ? (size_t)8 * c->N * c->ld : (size_t)B * c->N;
    const size_t need = (size_t)nt + (want_model ? out_d : 0) + (want_gp ? out_d + gp_scratch : 0) +
                        (want_std ? out_d : 0);
    if (need > c->pred_doubles) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_pred);
        c->d_pred = nullptr;
        c->pred_doubles = 0;
        CU(cudaMalloc(&c->d_pred, sizeof(double) * need));
        c->pred_doubles = need;
    }
    double* d_t = c->d_pred;
    double* d_model = d_t + nt;
    double* d_gp = d_model + (want_model ? out_d : 0);
    double* d_alpha = d_gp + (want_gp ? out_d : 0);
    double* d_std = d_alpha + (want_gp ? gp_scratch : 0);
    if (t_pred) CU(cudaMemcpyAsync(d_t, t_pred, sizeof(double) * nt, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpy2DAsync(c->d_params, sizeof(double) * c->ld, c->h_params, sizeof(double) * c->ld,
                         sizeof(double) * B, c->L.n_rows, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->d_np, c->h_np, sizeof(int32_t) * B, cudaMemcpyHostToDevice, c->stream));
    kc::KParams p = make_params(c, B, c->d_logL);
    int n = 0;
    CU(cudaEventRecord(c->ev0, c->stream));
    cudaError_t e = cudaSuccess;
    if (want_model)
        e = kc::launch_predict_model(p, c->kepler_mode, t_pred ? d_t : nullptr, nt, d_model, c->stream, &n);
    if (e == cudaSuccess && want_gp && cel) {
        e = kc::launch_celerite_predict(p, c->d_resid, d_alpha, c->ld, c->kepler_mode, t_pred ? d_t : nullptr, nt, d_gp,
                                        c->stream, &n);
    } else if (e == cudaSuccess && (want_gp || want_std)) {
        kc::GpPredict pr{t_pred ? d_t : nullptr, nt, d_alpha, want_gp ? d_gp : nullptr, want_std ? d_std : nullptr};
        e = kc::launch_gp_blocked(p, c->gp, c->kepler_mode, c->sm_count, c->stream, &n, &pr);
    }
    if (e != cudaSuccess) return fail_cuda(e, "predict launch");
    CU(cudaEventRecord(c->ev1, c->stream));
    c->timed = true;
    c->launches += n;
    if (want_model) CU(cudaMemcpyAsync(model_out, d_model, sizeof(double) * out_d, cudaMemcpyDeviceToHost, c->stream));
    if (want_gp) CU(cudaMemcpyAsync(gp_out, d_gp, sizeof(double) * out_d, cudaMemcpyDeviceToHost, c->stream));
    if (want_std) CU(cudaMemcpyAsync(std_out, d_std, sizeof(double) * out_d, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return KC_OK;
}

int kc_predict(kc_ctx* c, int B, int nt, const double* t_pred, int what, double* model_out,
               double* gp_out)
{
    if (what & KC_PRED_GP_STD) return fail(KC_ERR_ARG, "KC_PRED_GP_STD goes through kc_predict_std");
    return predict_impl(c, B, nt, t_pred, what, model_out, gp_out, nullptr);
}

int kc_predict_std(kc_ctx* c, int B, int nt, const double* t_pred, double* gp_out, double* gp_std_out)
{
    return predict_impl(c, B, nt, t_pred, (gp_out ? KC_PRED_GP : 0) | KC_PRED_GP_STD, nullptr, gp_out, gp_std_out);
}

}  // extern "C"
