// TEST INFRASTRUCTURE, not a product path: an Evaluator for host/BatchedSampler.h that scores the staged
// particles with the reference's own RVmodel::log_likelihood() (src/RVmodel.cpp:827-943) on the host cores.
// With it the lock-step sampler -- phase A / phase B around the likelihood call, the two lane groups, the lane
// pool, the level bookkeeping -- runs on a machine without a GPU, and must write the files of the reference's
// sampler byte for byte (tests/test_host_sampler_cpu.py).  Built only by that test, from the reference's
// UNWEAKENED objects: perturb() keeps mu and C up to date as in the reference, nothing is stripped.
//
// With RANK / WORLD_SIZE / MASTER_PORT in the environment several processes share a run as the GPU ranks do
// (lanes sharded, BatchedSampler::exchange at every barrier); the collectives go over TCP on 127.0.0.1.
//
// The Evaluator concept is described at the top of host/BatchedSampler.h.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <type_traits>
#include <vector>
#include <arpa/inet.h>
#include <netinet/in.h>
#include <netinet/tcp.h>
#include <sys/socket.h>
#include <unistd.h>
#include <chrono>
#include <cstring>
#include <stdexcept>
#include <thread>
#include "kima_cuda.hpp"   // declarations only: BatchedSampler::run_device names these types (never instantiated here)
#include "LanePool.h"
#include "LanePriors.h"
#include "ParticlePack.h"

extern const bool hyperpriors;

namespace kima_b200 {

class RefEvaluator {
public:
    static constexpr int GROUPS = 2;
    typedef std::false_type has_device_sampler;   // host proposals only: run_device() is never instantiated

    explicit RefEvaluator(int max_batch) : cap_(max_batch)
    {
        for (auto& s : staged_) s.assign((size_t)max_batch, nullptr);
        // one process per "rank" as a torchrun-style launcher starts them (RANK / WORLD_SIZE / MASTER_PORT):
        // the barrier exchange of host/BatchedSampler.h::exchange over TCP on 127.0.0.1, rank 0 as the hub
        if (const char* e = std::getenv("WORLD_SIZE")) world_ = std::max(1, std::atoi(e));
        if (const char* e = std::getenv("RANK")) rank_ = std::atoi(e);
        if (world_ > 1) connect_ranks();
    }
    ~RefEvaluator()
    {
        for (int fd : peers_)
            if (fd >= 0) ::close(fd);
        if (hub_ >= 0) ::close(hub_);
    }
    RefEvaluator(const RefEvaluator&) = delete;
    RefEvaluator& operator=(const RefEvaluator&) = delete;
    int max_batch() const { return cap_; }
    int max_batch(int) const { return cap_; }
    int rank() const { return rank_; }
    int world() const { return world_; }
    long long launches() const { return 0; }
    void set_host_pool(std::shared_ptr<LanePool> p) { pool_ = std::move(p); }

    // the CPU likelihood reads mu, C and the data: nothing is dropped
    void strip(RVmodel&) const {}
    void slim_host_data() {}

    void stage(int lane, const RVmodel& m) { stage(0, lane, m); }
    void stage(int g, int lane, const RVmodel& m) { staged_[g][(size_t)lane] = &m; }
    void stage_slot(int, int, int) {}
    void enable_factor_cache(int, int) {}
    bool check_staged(int, int, const RVmodel&, int) { return true; }

    void evaluate(int B, double* logL) { score(0, B, logL); }
    void submit(int g, int B, double* logL) { score(g, B, logL); }   // blocking: the result is there on return
    void wait(int) {}

    void reserve_lane_priors(size_t lanes) { lane_priors_.reserve(hyperpriors, lanes); }
    void own_lane_priors(RVmodel& m, size_t lane) const
    {
        // KIMA_SHARED_PRIORS=1: leave the reference's shared objects in place (to show the race they carry)
        static const bool shared = std::getenv("KIMA_SHARED_PRIORS") != nullptr;
        if (!shared) lane_priors_.own(m, lane);
    }

    // ---- a particle for the exchange: the parameters (host/ParticlePack.h) and, because the CPU likelihood reads
    // them and perturb() updates them incrementally, the model vector mu and the covariance matrix C as well
    size_t state_doubles() const
    {
        RVmodel probe;
        return particle_state_doubles() + 2 * probe.mu.size() + (size_t)probe.C.size();
    }
    void pack(const RVmodel& m, double* out) const
    {
        particle_pack(m, out);
        double* o = out + particle_state_doubles();
        std::memset(o, 0, sizeof(double) * 2 * m.mu.size());   // a long double is 10 bytes in a 16-byte slot
        for (size_t i = 0; i < m.mu.size(); i++) std::memcpy(o + 2 * i, &m.mu[i], 10);
        o += 2 * m.mu.size();
        for (Eigen::Index i = 0; i < m.C.size(); i++) o[i] = m.C.data()[i];
    }
    void unpack(RVmodel& m, const double* in) const
    {
        particle_unpack(m, in);
        const double* o = in + particle_state_doubles();
        for (size_t i = 0; i < m.mu.size(); i++) {
            long double v = 0;
            std::memcpy(&v, o + 2 * i, 10);
            m.mu[i] = v;
        }
        o += 2 * m.mu.size();
        for (Eigen::Index i = 0; i < m.C.size(); i++) m.C.data()[i] = o[i];
    }

    // ---- the collectives of the exchange (kc_allgather_bytes / kc_allgather_above / kc_allreduce_levels) ----
    void allgather_bytes(const void* send, size_t nbytes, void* recv)
    {
        char* out = static_cast<char*>(recv);
        if (world_ == 1) { if (recv != send) std::memcpy(out, send, nbytes); return; }
        if (rank_ == 0) {
            std::memcpy(out, send, nbytes);
            for (int r = 1; r < world_; r++) read_all(peers_[(size_t)r], out + nbytes * (size_t)r, nbytes);
            for (int r = 1; r < world_; r++) write_all(peers_[(size_t)r], out, nbytes * (size_t)world_);
        } else {
            write_all(hub_, send, nbytes);
            read_all(hub_, out, nbytes * (size_t)world_);
        }
    }
    // pairs (value, tiebreaker) of all ranks, concatenated in rank order; counts[r] = pairs of rank r
    void allgather_above(const double* pairs, int n, double* out, int max_total, int* counts)
    {
        std::vector<int> c((size_t)world_);
        allgather_bytes(&n, sizeof(int), c.data());
        int most = 0;
        for (int r = 0; r < world_; r++) most = std::max(most, c[(size_t)r]);
        std::vector<double> mine(2 * (size_t)std::max(most, 1), 0.0), all(mine.size() * (size_t)world_);
        std::memcpy(mine.data(), pairs, sizeof(double) * 2 * (size_t)n);
        allgather_bytes(mine.data(), sizeof(double) * mine.size(), all.data());
        int total = 0;
        for (int r = 0; r < world_; r++) {
            counts[r] = c[(size_t)r];
            for (int q = 0; q < c[(size_t)r] && total < max_total; q++, total++) {
                out[2 * (size_t)total] = all[mine.size() * (size_t)r + 2 * (size_t)q];
                out[2 * (size_t)total + 1] = all[mine.size() * (size_t)r + 2 * (size_t)q + 1];
            }
        }
    }
    void allreduce_levels(uint64_t* counters, int n)
    {
        std::vector<uint64_t> all((size_t)n * (size_t)world_);
        allgather_bytes(counters, sizeof(uint64_t) * (size_t)n, all.data());
        for (int j = 0; j < n; j++) {
            uint64_t s = 0;
            for (int r = 0; r < world_; r++) s += all[(size_t)r * (size_t)n + (size_t)j];
            counters[j] = s;
        }
    }

private:
    static void fail(const char* what)
    {
        std::fprintf(stderr, "RefEvaluator: %s\n", what);
        std::abort();
    }
    static void read_all(int fd, void* buf, size_t n)
    {
        char* p = static_cast<char*>(buf);
        while (n > 0) {
            const ssize_t k = ::read(fd, p, n);
            if (k <= 0) fail("peer closed the connection");
            p += k;
            n -= (size_t)k;
        }
    }
    static void write_all(int fd, const void* buf, size_t n)
    {
        const char* p = static_cast<const char*>(buf);
        while (n > 0) {
            const ssize_t k = ::write(fd, p, n);
            if (k <= 0) fail("write to a peer failed");
            p += k;
            n -= (size_t)k;
        }
    }
    void connect_ranks()
    {
        const int port = std::getenv("MASTER_PORT") ? std::atoi(std::getenv("MASTER_PORT")) : 29533;
        sockaddr_in addr{};
        addr.sin_family = AF_INET;
        addr.sin_port = htons((uint16_t)port);
        addr.sin_addr.s_addr = htonl(INADDR_LOOPBACK);
        const int one = 1;
        if (rank_ == 0) {
            const int ls = ::socket(AF_INET, SOCK_STREAM, 0);
            ::setsockopt(ls, SOL_SOCKET, SO_REUSEADDR, &one, sizeof(one));
            if (::bind(ls, (sockaddr*)&addr, sizeof(addr)) != 0 || ::listen(ls, world_) != 0) fail("cannot listen on MASTER_PORT");
            peers_.assign((size_t)world_, -1);
            for (int k = 1; k < world_; k++) {
                const int fd = ::accept(ls, nullptr, nullptr);
                if (fd < 0) fail("accept failed");
                ::setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
                int32_t r = -1;
                read_all(fd, &r, sizeof(r));
                if (r < 1 || r >= world_ || peers_[(size_t)r] >= 0) fail("unexpected rank at the hub");
                peers_[(size_t)r] = fd;
            }
            ::close(ls);
        } else {
            for (int tries = 0;; tries++) {
                hub_ = ::socket(AF_INET, SOCK_STREAM, 0);
                if (::connect(hub_, (sockaddr*)&addr, sizeof(addr)) == 0) break;
                ::close(hub_);
                hub_ = -1;
                if (tries > 3000) fail("rank 0 does not answer on MASTER_PORT");
                std::this_thread::sleep_for(std::chrono::milliseconds(10));
            }
            ::setsockopt(hub_, IPPROTO_TCP, TCP_NODELAY, &one, sizeof(one));
            const int32_t r = rank_;
            write_all(hub_, &r, sizeof(r));
        }
    }
    void score(int g, int B, double* logL)
    {
        const std::vector<const RVmodel*>& s = staged_[g];
        auto body = [&](int a, int b) {
            for (int i = a; i < b; ++i) logL[i] = s[(size_t)i]->log_likelihood();
        };
        if (pool_) pool_->run(0, B, body);
        else body(0, B);
    }
    int cap_;
    int rank_ = 0, world_ = 1;
    int hub_ = -1;               // ranks > 0: the connection to rank 0
    std::vector<int> peers_;     // rank 0: the connections to the others
    std::vector<const RVmodel*> staged_[GROUPS];
    std::shared_ptr<LanePool> pool_;
    LanePriors lane_priors_;
};

}  // namespace kima_b200
