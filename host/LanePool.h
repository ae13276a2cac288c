// Host-side worker threads for the lock-step sampler: the per-lane halves of
// Sampler::update_particle (DNest4/code/SamplerImpl.h:168-221) are independent across lanes
// (own RNG, own particles, own copy of the levels), exactly as they are across the reference's
// threads (SamplerImpl.h:134-165), so they can run on all host cores while the GPU scores the
// other half of the lanes.  Static contiguous chunks: lane -> thread is fixed, results do not
// depend on the number of threads.
#pragma once
#include <atomic>
#include <functional>
#include <thread>
#include <vector>

namespace kima_b200 {

class LanePool {
public:
    explicit LanePool(int nthreads) : n_(nthreads < 1 ? 1 : nthreads)
    {
        for (int t = 1; t < n_; ++t) workers_.emplace_back([this, t] { loop(t); });
    }
    ~LanePool()
    {
        stop_.store(true, std::memory_order_release);
        gen_.fetch_add(1, std::memory_order_acq_rel);
        for (auto& w : workers_) w.join();
    }
    LanePool(const LanePool&) = delete;
    LanePool& operator=(const LanePool&) = delete;

    int threads() const { return n_; }

    // f(begin, end) over [lo, hi) in n_ contiguous chunks; returns when all are done.
    void run(int lo, int hi, const std::function<void(int, int)>& f)
    {
        const int n = hi - lo;
        if (n_ == 1 || n < 2 * n_) {
            if (n > 0) f(lo, hi);
            return;
        }
        job_ = &f;
        lo_ = lo;
        hi_ = hi;
        pending_.store(n_ - 1, std::memory_order_release);
        gen_.fetch_add(1, std::memory_order_acq_rel);
        chunk(0);
        int spins = 0;
        while (pending_.load(std::memory_order_acquire) != 0)
            if (++spins > 4096) std::this_thread::yield();
    }

private:
    void chunk(int t)
    {
        const long n = hi_ - lo_;
        const int a = lo_ + (int)(n * t / n_), b = lo_ + (int)(n * (t + 1) / n_);
        if (b > a) (*job_)(a, b);
    }
    void loop(int t)
    {
        unsigned long seen = 0;
        while (true) {
            int spins = 0;
            unsigned long g;
            while ((g = gen_.load(std::memory_order_acquire)) == seen)
                if (++spins > 4096) std::this_thread::yield();
            seen = g;
            if (stop_.load(std::memory_order_acquire)) return;
            chunk(t);
            pending_.fetch_sub(1, std::memory_order_acq_rel);
        }
    }

    int n_;
    std::vector<std::thread> workers_;
    const std::function<void(int, int)>* job_ = nullptr;
    int lo_ = 0, hi_ = 0;
    std::atomic<unsigned long> gen_{0};
    std::atomic<int> pending_{0};
    std::atomic<bool> stop_{false};
};

}  // namespace kima_b200
