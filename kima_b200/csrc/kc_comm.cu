// Multi-GPU exchange at DNest4's level-creation barrier
// (/root/reference/DNest4/code/SamplerImpl.h:306-329): every rank's `above`
// list is gathered so each rank can sort and place the new level identically,
// and the per-level accept/try/visit/exceed deltas are summed.  One process per
// GPU; NCCL over NVLink.  Payloads are KB-sized, so this is latency-bound and
// there is nothing to fuse it with: the likelihood kernels have no data-path
// collective at all (particles are independent).
//
// NCCL is bound at run time (dlopen) so that libkima_cuda.so loads in
// single-GPU processes that never initialise a communicator.
#include <dlfcn.h>
#include <nccl.h>
#include <cstring>
#include <string>
#include <vector>
#include "kc_internal.h"


namespace {

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

bool load_nccl(std::string& err)
{
    if (g_nccl.lib) return true;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    void* lib = nullptr;
    for (const char* n : names) {
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    if (!lib) { err = std::string("cannot dlopen libnccl: ") + dlerror(); return false; }
    auto sym = [&](const char* s) { return dlsym(lib, s); };
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))sym("ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))sym("ncclCommInitRank");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))sym("ncclCommDestroy");
    g_nccl.AllGather = (decltype(g_nccl.AllGather))sym("ncclAllGather");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))sym("ncclAllReduce");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))sym("ncclGetErrorString");
    if (!g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllGather ||
        !g_nccl.AllReduce || !g_nccl.GetErrorString) {
        err = "libnccl lacks a required symbol";
        return false;
    }
    g_nccl.lib = lib;
    return true;
}

#define NC(call)                                                                         \
    do {                                                                                 \
        ncclResult_t r_ = (call);                                                        \
        if (r_ != ncclSuccess)                                                           \
            return kc::comm_fail(KC_ERR_COMM, std::string(#call) + ": " + g_nccl.GetErrorString(r_)); \
    } while (0)
#define CUC(call)                                                                        \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess)                                                           \
            return kc::comm_fail(KC_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

int ensure_buf(kc::CommView& v, size_t bytes)
{
    if (*v.buf_bytes >= bytes) return KC_OK;
    if (*v.d_buf) cudaFree(*v.d_buf);
    *v.d_buf = nullptr;
    *v.buf_bytes = 0;
    CUC(cudaMalloc(v.d_buf, bytes));
    *v.buf_bytes = bytes;
    return KC_OK;
}

}  // namespace

// Collectives on buffers that already sit in HBM, enqueued on the context's stream (kc_sampler.cu: the
// `above` lists and the level counters meet on the GPUs, SamplerImpl.h:306-329, without touching the hosts).
namespace kc {
int comm_ranks(kc_ctx* c)
{
    CommView v = comm_view(c);
    return *v.nccl_comm ? *v.nranks : 1;
}
int comm_allgather_dev(kc_ctx* c, const void* d_send, void* d_recv, size_t nbytes)
{
    CommView v = comm_view(c);
    if (!*v.nccl_comm) return comm_fail(KC_ERR_ARG, "kc_comm_init has not been called");
    NC(g_nccl.AllGather(d_send, d_recv, nbytes, ncclChar, (ncclComm_t)*v.nccl_comm, v.stream));
    return KC_OK;
}
int comm_allreduce_u64_dev(kc_ctx* c, uint64_t* d, size_t n)
{
    CommView v = comm_view(c);
    if (!*v.nccl_comm) return comm_fail(KC_ERR_ARG, "kc_comm_init has not been called");
    NC(g_nccl.AllReduce(d, d, n, ncclUint64, ncclSum, (ncclComm_t)*v.nccl_comm, v.stream));
    return KC_OK;
}
}  // namespace kc

extern "C" {

int kc_comm_unique_id(void* id128)
{
    static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
    if (!id128) return kc::comm_fail(KC_ERR_ARG, "id128 is null");
    std::string err;
    if (!load_nccl(err)) return kc::comm_fail(KC_ERR_COMM, err);
    ncclUniqueId id;
    NC(g_nccl.GetUniqueId(&id));
    std::memcpy(id128, &id, 128);
    return KC_OK;
}

int kc_comm_init(kc_ctx* c, int rank, int nranks, const void* id128)
{
    if (!c || !id128) return kc::comm_fail(KC_ERR_ARG, "null argument");
    if (nranks < 1 || rank < 0 || rank >= nranks) return kc::comm_fail(KC_ERR_ARG, "bad rank / nranks");
    std::string err;
    if (!load_nccl(err)) return kc::comm_fail(KC_ERR_COMM, err);
    kc::CommView v = kc::comm_view(c);
    if (*v.nccl_comm) return kc::comm_fail(KC_ERR_ARG, "communicator already initialised");
    CUC(cudaSetDevice(v.device));
    ncclUniqueId id;
    std::memcpy(&id, id128, 128);
    ncclComm_t comm;
    NC(g_nccl.CommInitRank(&comm, nranks, id, rank));
    *v.nccl_comm = comm;
    *v.rank = rank;
    *v.nranks = nranks;
    return KC_OK;
}

int kc_comm_destroy(kc_ctx* c)
{
    if (!c) return KC_OK;
    kc::CommView v = kc::comm_view(c);
    if (*v.nccl_comm && g_nccl.CommDestroy) {
        g_nccl.CommDestroy((ncclComm_t)*v.nccl_comm);
        *v.nccl_comm = nullptr;
    }
    return KC_OK;
}

int kc_allgather_above(kc_ctx* c, const double* pairs, int n, double* out, int max_total,
                       int* counts)
{
    if (!c || !out || !counts || n < 0 || (n > 0 && !pairs))
        return kc::comm_fail(KC_ERR_ARG, "bad argument");
    kc::CommView v = kc::comm_view(c);
    const int R = *v.nranks;
    if (!*v.nccl_comm) {  // single rank without a communicator: identity
        if (R != 1) return kc::comm_fail(KC_ERR_ARG, "kc_comm_init has not been called");
        if (n > max_total) return kc::comm_fail(KC_ERR_ARG, "max_total too small");
        std::memcpy(out, pairs, sizeof(double) * 2 * n);
        counts[0] = n;
        return KC_OK;
    }
    ncclComm_t comm = (ncclComm_t)*v.nccl_comm;
    CUC(cudaSetDevice(v.device));
    // 1) counts
    int rc = ensure_buf(v, sizeof(int) * (R + 1));
    if (rc) return rc;
    int* d_cnt = (int*)*v.d_buf;
    CUC(cudaMemcpyAsync(d_cnt + R, &n, sizeof(int), cudaMemcpyHostToDevice, v.stream));
    NC(g_nccl.AllGather(d_cnt + R, d_cnt, 1, ncclInt32, comm, v.stream));
    CUC(cudaMemcpyAsync(counts, d_cnt, sizeof(int) * R, cudaMemcpyDeviceToHost, v.stream));
    CUC(cudaStreamSynchronize(v.stream));
    int mx = 0;
    long total = 0;
    for (int r = 0; r < R; r++) { mx = counts[r] > mx ? counts[r] : mx; total += counts[r]; }
    if (total > max_total) return kc::comm_fail(KC_ERR_ARG, "max_total too small");
    if (mx == 0) return KC_OK;
    // 2) payload, padded to the largest list
    size_t per = sizeof(double) * 2 * (size_t)mx;
    rc = ensure_buf(v, per * (R + 1));
    if (rc) return rc;
    char* d = (char*)*v.d_buf;
    char* d_send = d + per * R;
    CUC(cudaMemsetAsync(d_send, 0, per, v.stream));
    if (n) CUC(cudaMemcpyAsync(d_send, pairs, sizeof(double) * 2 * n, cudaMemcpyHostToDevice, v.stream));
    NC(g_nccl.AllGather(d_send, d, 2 * (size_t)mx, ncclFloat64, comm, v.stream));
    std::vector<double> host((size_t)2 * mx * R);
    CUC(cudaMemcpyAsync(host.data(), d, per * R, cudaMemcpyDeviceToHost, v.stream));
    CUC(cudaStreamSynchronize(v.stream));
    size_t o = 0;
    for (int r = 0; r < R; r++) {
        std::memcpy(out + o, host.data() + (size_t)2 * mx * r, sizeof(double) * 2 * counts[r]);
        o += (size_t)2 * counts[r];
    }
    return KC_OK;
}

int kc_allgather_bytes(kc_ctx* c, const void* send, size_t nbytes, void* recv)
{
    if (!c || (nbytes > 0 && (!send || !recv))) return kc::comm_fail(KC_ERR_ARG, "bad argument");
    kc::CommView v = kc::comm_view(c);
    const int R = *v.nranks;
    if (!*v.nccl_comm) {  // single rank without a communicator: identity
        if (R != 1) return kc::comm_fail(KC_ERR_ARG, "kc_comm_init has not been called");
        if (nbytes && recv != send) std::memcpy(recv, send, nbytes);
        return KC_OK;
    }
    if (nbytes == 0) return KC_OK;
    CUC(cudaSetDevice(v.device));
    int rc = ensure_buf(v, nbytes * (size_t)(R + 1));
    if (rc) return rc;
    char* d = (char*)*v.d_buf;
    char* d_send = d + nbytes * (size_t)R;
    CUC(cudaMemcpyAsync(d_send, send, nbytes, cudaMemcpyHostToDevice, v.stream));
    NC(g_nccl.AllGather(d_send, d, nbytes, ncclChar, (ncclComm_t)*v.nccl_comm, v.stream));
    CUC(cudaMemcpyAsync(recv, d, nbytes * (size_t)R, cudaMemcpyDeviceToHost, v.stream));
    CUC(cudaStreamSynchronize(v.stream));
    return KC_OK;
}

int kc_allreduce_levels(kc_ctx* c, uint64_t* counters, int n)
{
    if (!c || n < 0 || (n > 0 && !counters)) return kc::comm_fail(KC_ERR_ARG, "bad argument");
    kc::CommView v = kc::comm_view(c);
    if (!*v.nccl_comm) {
        if (*v.nranks != 1) return kc::comm_fail(KC_ERR_ARG, "kc_comm_init has not been called");
        return KC_OK;
    }
    if (n == 0) return KC_OK;
    CUC(cudaSetDevice(v.device));
    int rc = ensure_buf(v, sizeof(uint64_t) * n);
    if (rc) return rc;
    CUC(cudaMemcpyAsync(*v.d_buf, counters, sizeof(uint64_t) * n, cudaMemcpyHostToDevice, v.stream));
    NC(g_nccl.AllReduce(*v.d_buf, *v.d_buf, n, ncclUint64, ncclSum, (ncclComm_t)*v.nccl_comm, v.stream));
    CUC(cudaMemcpyAsync(counters, *v.d_buf, sizeof(uint64_t) * n, cudaMemcpyDeviceToHost, v.stream));
    CUC(cudaStreamSynchronize(v.stream));
    return KC_OK;
}

}  // extern "C"
