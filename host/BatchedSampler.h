// DNest4::BatchedSampler — the reference's Sampler with its particle loop turned into
// lock-step lanes, so that every MCMC step scores all lanes' proposals with ONE batched
// likelihood call instead of one ModelType::log_likelihood() per thread.
//
// What is reused unchanged from the reference (it is the base class): the level table,
// do_bookkeeping / kill_lagging_particles / save_particle / save_levels, the output files,
// update_level_assignment, log_push, print/read, the Options/CommandLineOptions handling
// (DNest4/code/Sampler.h, SamplerImpl.h:224-265, 377-637).
// What is new: initialise() and run(), i.e. the restatement of
//     Sampler::initialise      SamplerImpl.h:52-79
//     Sampler::run/run_thread  SamplerImpl.h:82-131, 268-335
//     Sampler::mcmc_thread     SamplerImpl.h:134-165
//     Sampler::update_particle SamplerImpl.h:168-221   (split around the likelihood call)
//
// A lane is exactly one DNest4 "thread" (-t): its own RNG (seed + lane), its own
// num_particles particles, its own copy of the level table between barriers.  Since
// log_likelihood() draws no random numbers, splitting update_particle around it leaves every
// lane's RNG stream identical to the reference thread with the same seed: `-t 4 -s 43` on
// the CI OPTIONS reproduces the reference's golden sample.txt (tests/test_host_cpp.py).
//
// The Evaluator concept (host/RVmodelDevice.h is the RVmodel instance):
//     explicit Evaluator(int max_batch);  int max_batch() const;
//     void stage(int slot, const ModelType&);  void evaluate(int B, double* logL);   (group 0, blocking)
//     void strip(ModelType&);   (drop per-particle host state only the CPU likelihood needs)
//     void slim_host_data();    (after initialise: drop host copies of the data the sampler no longer reads)
//     void stage(int g, int slot, const ModelType&);  void submit(int g, int B, double* logL);  void wait(int g);
//     typedef std::true_type / std::false_type has_device_sampler;   (the sampler_* members run_device() uses)
// tests/cpp/RefEvaluator.h is a second instance: the reference's own log_likelihood() on the host cores, so
// that the lock-step sampler itself is tested on a machine without a GPU.
//
// This header must be included after Sampler.h has been included with its private section
// opened (host/overlay/kima.h does that).
#pragma once
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <cstring>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>
#include <thread>
#include <type_traits>
#include <vector>
#include "LanePool.h"

namespace DNest4 {

template <class ModelType, class Evaluator>
class BatchedSampler : public Sampler<ModelType> {
    using Base = Sampler<ModelType>;

    // evaluator shared so that the sampler stays copyable (setup() returns it by value)
    std::shared_ptr<Evaluator> evaluator;

    struct LaneStep {
        unsigned int which = 0;
        bool particle_first = true;  // the coin of mcmc_thread (SamplerImpl.h:151)
        bool scored = false;         // passed the prior-ratio test, proposal staged
        int slot = -1;
        ModelType proposal;
    };
    std::vector<LaneStep> lane_steps;
    std::vector<double> logl_buf;            // initialise(): one batch
    std::vector<double> logl_grp[2];         // run(): one buffer per lane group
    std::shared_ptr<kima_b200::LanePool> pool;
    std::vector<char> xbuf_send, xbuf_recv;  // exchange() buffers, kept between rounds
    // wall-clock split of run(), printed at the end when KIMA_PROFILE is set
    bool cache_tried = false;  // the evaluator's factor cache is asked for once, at the first step
    double prof[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // perturb, stage, submit, wait, phase B, bookkeeping, exchange, level copies
    static double now()
    {
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }

    void ensure_evaluator()
    {
        if (evaluator) return;
        const size_t total = this->particles.size();
        const int cap = (int)std::max<size_t>(this->num_threads, std::min<size_t>(total, 8192));
        evaluator = std::make_shared<Evaluator>(cap);
        logl_buf.resize(cap);
        for (auto& v : logl_grp) v.resize(cap);
        int nt = (int)std::thread::hardware_concurrency();
        if (const char* e = std::getenv("KIMA_HOST_THREADS")) nt = std::atoi(e);
        pool = std::make_shared<kima_b200::LanePool>(std::max(1, std::min(nt, 64)));
        evaluator->set_host_pool(pool);
    }

    // Sampler::print (SamplerImpl.h:590-612) with the particles formatted by all host threads: RVmodel::print
    // sets its own format flags before the first number (src/RVmodel.cpp:947-948), so a run of particles
    // prints the same into a stream of its own; what follows them inherits fixed / precision 8 as it does there.
    void print_state(std::ostream& out)
    {
        const size_t n = this->particles.size();
        if (n < 4096 || std::getenv("KIMA_SERIAL_STATE")) { this->print(out); return; }
        out << this->save_to_disk << ' ';
        out << this->num_threads << ' ';
        out << this->compression << ' ';
        out << this->options << ' ';
        const size_t chunk = 2048, nchunks = (n + chunk - 1) / chunk;
        std::vector<std::string> parts(nchunks);
        pool->run(0, (int)nchunks, [&](int a, int b) {
            for (int c = a; c < b; ++c) {
                std::ostringstream os;
                const size_t i1 = std::min(n, ((size_t)c + 1) * chunk);
                for (size_t i = (size_t)c * chunk; i < i1; ++i) this->particles[i].print(os);
                parts[c] = os.str();
            }
        });
        for (const std::string& s : parts) out << s;
        out.setf(std::ios::fixed, std::ios::floatfield);
        out.precision(8);
        // the likelihood pairs and the level assignments inherit those flags: chunk streams start from them
        pool->run(0, (int)nchunks, [&](int a, int b) {
            for (int c = a; c < b; ++c) {
                std::ostringstream os;
                os.copyfmt(out);
                const size_t i1 = std::min(n, ((size_t)c + 1) * chunk);
                for (size_t i = (size_t)c * chunk; i < i1; ++i) this->log_likelihoods[i].print(os);
                parts[c] = os.str();
            }
        });
        for (const std::string& s : parts) out << s;
        pool->run(0, (int)nchunks, [&](int a, int b) {
            for (int c = a; c < b; ++c) {
                std::ostringstream os;
                const size_t i1 = std::min(n, ((size_t)c + 1) * chunk);
                for (size_t i = (size_t)c * chunk; i < i1; ++i) os << this->level_assignments[i] << ' ';
                parts[c] = os.str();
            }
        });
        for (const std::string& s : parts) out << s;
        for (const auto& l : this->levels) l.print(out);
        out << this->count_saves << ' ';
        out << this->count_mcmc_steps << ' ';
    }

    // update_particle up to its likelihood call, for the lanes [k0, k1) of group g:
    // choose the particle, the coin, perturb, prior-ratio test; then stage the survivors
    // in lane order (so the batch does not depend on the number of host threads).
    int phase_a(int g, unsigned int k0, unsigned int k1)
    {
        const unsigned int P = this->options.num_particles;
        double t0 = now();
        pool->run((int)k0, (int)k1, [&](int a, int b) {
            for (int k = a; k < b; ++k) {
                RNG& rng = this->rngs[k];
                LaneStep& st = lane_steps[k];
                st.which = k * P + rng.rand_int(P);
                st.particle_first = (rng.rand() <= 0.5);
                if (!st.particle_first) this->update_level_assignment(k, st.which);
                st.proposal = this->particles[st.which];
                evaluator->own_lane_priors(st.proposal, (size_t)k);  // hyperpriors: no prior object shared between threads
                double log_H = st.proposal.perturb(rng);
                if (log_H > 0.0) log_H = 0.0;
                st.scored = (rng.rand() <= exp(log_H));
            }
        });
        double t1 = now();
        prof[0] += t1 - t0;
        int nb = 0;
        for (unsigned int k = k0; k < k1; ++k) lane_steps[k].slot = lane_steps[k].scored ? nb++ : -1;
        pool->run((int)k0, (int)k1, [&](int a, int b) {
            for (int k = a; k < b; ++k)
                if (lane_steps[k].scored) {
                    evaluator->stage(g, lane_steps[k].slot, lane_steps[k].proposal);
                    // the particle behind the lane, numbered within the group (factor cache)
                    evaluator->stage_slot(g, lane_steps[k].slot, (int)(lane_steps[k].which - k0 * P));
                }
        });
        prof[1] += now() - t1;
        static const bool verify_stage = std::getenv("KIMA_VERIFY_STAGE") != nullptr;
        if (verify_stage)  // developer check: the block the pool staged against a serial restaging
            for (unsigned int k = k0; k < k1; ++k)
                if (lane_steps[k].scored) evaluator->check_staged(g, lane_steps[k].slot, lane_steps[k].proposal, (int)k);
        return nb;
    }

    // the rest of update_particle, then the level assignment and the `above` list
    void phase_b(int g, unsigned int k0, unsigned int k1)
    {
        const double* logl = logl_grp[g].data();
        const double t0 = now();
        pool->run((int)k0, (int)k1, [&](int a, int b) {
            for (int k = a; k < b; ++k) {
                RNG& rng = this->rngs[k];
                LaneStep& st = lane_steps[k];
                std::vector<Level>& lv = this->copies_of_levels[k];
                Level& level = lv[this->level_assignments[st.which]];
                if (st.scored) {
                    LikelihoodType cand(logl[st.slot], this->log_likelihoods[st.which].get_tiebreaker());
                    cand.perturb(rng);  // tiebreaker move: draws from the lane's stream
                    if (level.get_log_likelihood() < cand) {
                        this->particles[st.which] = st.proposal;
                        this->log_likelihoods[st.which] = cand;
                        level.increment_accepts(1);
                    }
                }
                level.increment_tries(1);
                for (unsigned int j = this->level_assignments[st.which]; j + 1 < lv.size(); ++j) {
                    lv[j].increment_visits(1);
                    if (lv[j + 1].get_log_likelihood() < this->log_likelihoods[st.which])
                        lv[j].increment_exceeds(1);
                    else
                        break;
                }
                if (st.particle_first) this->update_level_assignment(k, st.which);
                if (!this->enough_levels(lv) &&
                    lv.back().get_log_likelihood() < this->log_likelihoods[st.which])
                    this->above[k].push_back(this->log_likelihoods[st.which]);
            }
        });
        prof[4] += now() - t0;
    }

    void timed_wait(int g)
    {
        const double t0 = now();
        evaluator->wait(g);
        prof[3] += now() - t0;
    }

    // ---- one process per GPU -----------------------------------------------------------------
    // Rank r owns the lanes [r T/R, (r+1) T/R) and is the only one to step them, but every rank
    // keeps the whole sampler state (all particles, likelihoods, level assignments, levels), so
    // that after the exchange below each rank runs the same do_bookkeeping() on the same data
    // and no decision has to be broadcast.  With R ranks the run is, lane for lane and draw for
    // draw, the single-process run with the same -t and -s.
    int rank() const { return evaluator ? evaluator->rank() : 0; }
    int world() const { return evaluator ? evaluator->world() : 1; }
    unsigned int lane_lo() const { return (unsigned int)((unsigned long long)this->num_threads * rank() / world()); }
    unsigned int lane_hi() const { return (unsigned int)((unsigned long long)this->num_threads * (rank() + 1) / world()); }

    // The barrier of run_thread (SamplerImpl.h:306-329) between processes: the owners' lanes
    // (particles, likelihoods, level assignments) go to everybody, rank 0's RNG 0 (the one
    // do_bookkeeping draws from) too; `above` lists are concatenated in lane order and the
    // level counters summed.
    void exchange(const std::vector<Level>& before, const bool with_particles)
    {
        const int R = world();
        const unsigned int T = this->num_threads, P = this->options.num_particles;
        // with_particles = false: likelihoods and level assignments only; particles then travel when the
        // bookkeeping asks for one (bookkeeping_ranks below)
        const size_t S = with_particles ? evaluator->state_doubles() : 0, per = S + 3;
        const unsigned int lanes_per_rank = T / R;
        constexpr size_t RNG_BYTES = 16384;
        const size_t blob = sizeof(double) * per * P * lanes_per_rank + RNG_BYTES;
        std::vector<char>& send = xbuf_send;
        std::vector<char>& recv = xbuf_recv;
        if (send.size() != blob) send.assign(blob, 0);
        if (recv.size() != blob * R) recv.assign(blob * R, 0);
        double* sd = reinterpret_cast<double*>(send.data());
        const unsigned int my_lo = lane_lo();
        pool->run((int)lane_lo(), (int)lane_hi(), [&](int a, int b) {
            for (int k = a; k < b; ++k)
                for (unsigned int q = 0; q < P; ++q) {
                    const size_t i = (size_t)k * P + q;
                    double* o = sd + ((size_t)(k - my_lo) * P + q) * per;
                    if (with_particles) evaluator->pack(this->particles[i], o);
                    o[S] = this->log_likelihoods[i].get_value();
                    o[S + 1] = this->log_likelihoods[i].get_tiebreaker();
                    o[S + 2] = (double)this->level_assignments[i];
                }
        });
        if (rank() == 0) {
            std::ostringstream os;
            os << this->rngs[0].twister << ' ' << this->rngs[0].uniform << ' ' << this->rngs[0].normal;
            const std::string st = os.str();
            if (st.size() + 1 > RNG_BYTES) throw std::runtime_error("RNG state larger than its slot");
            std::memcpy(send.data() + blob - RNG_BYTES, st.c_str(), st.size() + 1);
        }
        evaluator->allgather_bytes(send.data(), blob, recv.data());
        for (int r = 0; r < R; ++r) {
            if (r == rank()) continue;
            const double* rd = reinterpret_cast<const double*>(recv.data() + blob * r);
            const unsigned int lo = (unsigned int)((unsigned long long)T * r / R);
            pool->run(0, (int)lanes_per_rank, [&](int a, int b) {
                for (int k = a; k < b; ++k)
                    for (unsigned int q = 0; q < P; ++q) {
                        const size_t i = (size_t)(lo + k) * P + q;
                        const double* o = rd + ((size_t)k * P + q) * per;
                        if (with_particles) evaluator->unpack(this->particles[i], o);
                        this->log_likelihoods[i] = LikelihoodType(o[S], o[S + 1]);
                        this->level_assignments[i] = (unsigned int)o[S + 2];
                    }
            });
        }
        if (rank() != 0) {
            std::istringstream is(std::string(recv.data() + blob - RNG_BYTES));
            is >> this->rngs[0].twister >> this->rngs[0].uniform >> this->rngs[0].normal;
        }
        // level counters: this rank's lanes' increments, summed over ranks
        const size_t L = this->levels.size();
        std::vector<uint64_t> cnt(4 * L, 0);
        for (unsigned int k = lane_lo(); k < lane_hi(); ++k) {
            const std::vector<Level>& lv = this->copies_of_levels[k];
            for (size_t j = 0; j < L; ++j) {
                cnt[4 * j + 0] += lv[j].get_accepts() - before[j].get_accepts();
                cnt[4 * j + 1] += lv[j].get_tries() - before[j].get_tries();
                cnt[4 * j + 2] += lv[j].get_visits() - before[j].get_visits();
                cnt[4 * j + 3] += lv[j].get_exceeds() - before[j].get_exceeds();
            }
        }
        evaluator->allreduce_levels(cnt.data(), (int)cnt.size());
        for (size_t j = 0; j < L; ++j) {
            this->levels[j].increment_accepts(cnt[4 * j + 0]);
            this->levels[j].increment_tries(cnt[4 * j + 1]);
            this->levels[j].increment_visits(cnt[4 * j + 2]);
            this->levels[j].increment_exceeds(cnt[4 * j + 3]);
        }
        // `above`: this rank's lists in lane order, then all ranks' in rank (= lane) order
        std::vector<double> mine;
        for (unsigned int k = lane_lo(); k < lane_hi(); ++k) {
            for (const LikelihoodType& a : this->above[k]) {
                mine.push_back(a.get_value());
                mine.push_back(a.get_tiebreaker());
            }
            this->above[k].clear();
        }
        const int max_total = (int)(T * this->options.thread_steps) + 16;
        std::vector<double> all(2 * (size_t)max_total);
        std::vector<int> counts(R);
        evaluator->allgather_above(mine.data(), (int)(mine.size() / 2), all.data(), max_total, counts.data());
        size_t tot = 0;
        for (int c : counts) tot += (size_t)c;
        for (size_t i = 0; i < tot; ++i) this->all_above.push_back(LikelihoodType(all[2 * i], all[2 * i + 1]));
    }

    // ---- host proposals over several ranks: particles travel only when the bookkeeping asks for one ----
    // Every rank keeps likelihoods, level assignments, levels and rngs[0] of the whole run (exchange above), but
    // a particle itself only where its lane is stepped: the copies of another rank's particles go stale after
    // initialise().  do_bookkeeping (SamplerImpl.h:377-440) reads particles in three places -- the sources of
    // kill_lagging_particles' copies (:575), the saved particle (:512) and the record holder (:437) -- and those
    // are fetched from their owners; the end-of-run state file gathers everything once.  All-gathering every
    // particle at every barrier (KIMA_FULL_EXCHANGE=1 keeps that) was the largest part of the barrier.
    unsigned int deletions_ranks = 0;
    int owner_of_particle(size_t i) const
    {
        return (int)((i / this->options.num_particles) / (this->num_threads / (unsigned int)world()));
    }
    // particles `which[0..K)` from their owners into every rank's copy
    void fetch_host_particles(const std::vector<size_t>& which)
    {
        if (which.empty()) return;
        const size_t S = evaluator->state_doubles(), K = which.size();
        std::vector<double> mine(S * K, 0.0), all(S * K * (size_t)world());
        for (size_t q = 0; q < K; ++q)
            if (owner_of_particle(which[q]) == rank()) evaluator->pack(this->particles[which[q]], mine.data() + q * S);
        evaluator->allgather_bytes(mine.data(), sizeof(double) * S * K, all.data());
        for (size_t q = 0; q < K; ++q) {
            const int own = owner_of_particle(which[q]);
            if (own != rank()) evaluator->unpack(this->particles[which[q]], all.data() + ((size_t)own * K + q) * S);
        }
    }
    // kill_lagging_particles (SamplerImpl.h:534-587): same draws from rngs[0] on every rank, same messages
    void kill_lagging_ranks()
    {
        const size_t n = this->particles.size();
        std::vector<bool> good;
        std::vector<uint32_t> lagging;
        double max_log_push = 0;
        std::vector<double> push, kill;
        level_push_and_kill(push, kill);
        const unsigned int num_bad = flag_lagging(push, kill, good, lagging, max_log_push);
        if (num_bad >= n) return;
        std::vector<double> keep(push.size());
        for (size_t j = 0; j < push.size(); ++j) keep[j] = exp(push[j] - max_log_push);
        std::vector<size_t> dst, src;
        for (const uint32_t i : lagging) {
            int i_copy;
            do {
                i_copy = this->rngs[0].rand_int((int)n);
            } while (!good[i_copy] || this->rngs[0].rand() >= keep[this->level_assignments[i_copy]]);
            dst.push_back(i);
            src.push_back((size_t)i_copy);
            this->log_likelihoods[i] = this->log_likelihoods[i_copy];
            this->level_assignments[i] = this->level_assignments[i_copy];
            ++deletions_ranks;
            if (rank() == 0)
                std::cout << "# Replacing lagging particle. This has happened " << deletions_ranks << " times." << std::endl;
        }
        // sources are good particles and never overwritten: the copies are independent of their order.  A source
        // on another rank than its destination travels first.
        std::vector<size_t> remote;
        for (size_t k = 0; k < dst.size(); ++k)
            if (owner_of_particle(dst[k]) != owner_of_particle(src[k])) remote.push_back(src[k]);
        fetch_host_particles(remote);
        for (size_t k = 0; k < dst.size(); ++k)
            if (owner_of_particle(dst[k]) == rank()) this->particles[dst[k]] = this->particles[src[k]];
    }
    // save_particle (SamplerImpl.h:503-531): every rank draws `which` (rngs[0] is replicated), rank 0 prints
    void save_particle_ranks()
    {
        ++this->count_saves;
        this->count_mcmc_steps_since_save = 0;
        if (rank() == 0 && this->save_to_disk && this->count_saves % this->thin_print == 0)
            std::cout << "# Saving particle to disk. N = " << this->count_saves << "." << std::endl;
        const int which = this->rngs[0].rand_int((int)this->particles.size());
        fetch_host_particles(std::vector<size_t>(1, (size_t)which));
        if (!this->save_to_disk) return;
        std::fstream fout;
        fout.open(this->options.sample_file, std::ios::out | std::ios::app);
        this->particles[which].print(fout);
        fout << std::endl;
        fout.close();
        fout.open(this->options.sample_info_file, std::ios::out | std::ios::app);
        fout << std::setprecision(12);
        fout << this->level_assignments[which] << ' ';
        fout << this->log_likelihoods[which].get_value() << ' ';
        fout << this->log_likelihoods[which].get_tiebreaker() << ' ';
        fout << which << std::endl;
        fout.close();
    }
    // do_bookkeeping (SamplerImpl.h:377-440)
    void bookkeeping_ranks()
    {
        bool created_level = false;
        if (!this->enough_levels(this->levels) && (this->all_above.size() >= this->options.new_level_interval)) {
            std::sort(this->all_above.begin(), this->all_above.end());
            const int index = static_cast<int>((1. - 1. / this->compression) * this->all_above.size());
            if (rank() == 0) {
                std::cout << "# Creating level " << this->levels.size() << " with log likelihood = ";
                std::cout << std::setprecision(12);
                std::cout << this->all_above[index].get_value() << "." << std::endl;
            }
            this->levels.push_back(Level(this->all_above[index]));
            this->all_above.erase(this->all_above.begin(), this->all_above.begin() + index + 1);
            for (auto& a : this->above) a.clear();
            if (this->enough_levels(this->levels)) {
                const double reg = this->options.new_level_interval * sqrt(this->options.lambda);
                Level::renormalise_visits(this->levels, static_cast<int>(reg));
                this->all_above.clear();
                this->options.max_num_levels = this->levels.size();
                if (rank() == 0) std::cout << "# Done creating levels." << std::endl;
            } else {
                kill_lagging_ranks();
            }
            created_level = true;
        }
        Level::recalculate_log_X(this->levels, this->compression,
                                 this->options.new_level_interval * sqrt(this->options.lambda));
        if (created_level) this->save_levels();
        if (this->count_mcmc_steps_since_save >= this->options.save_interval) {
            save_particle_ranks();
            if (!created_level) this->save_levels();
        }
        // record holder (:433-439): the largest likelihood; its particle comes from its owner
        size_t best = 0;
        for (size_t i = 1; i < this->log_likelihoods.size(); ++i)
            if (this->log_likelihoods[best] < this->log_likelihoods[i]) best = i;
        if (this->best_ever_log_likelihood < this->log_likelihoods[best]) {
            fetch_host_particles(std::vector<size_t>(1, best));
            this->best_ever_particle = this->particles[best];
            this->best_ever_log_likelihood = this->log_likelihoods[best];
        }
    }
    // the end-of-run state file holds every particle: the owners' lanes go to everybody, once
    void gather_particles()
    {
        const int R = world();
        const unsigned int T = this->num_threads, P = this->options.num_particles;
        const size_t S = evaluator->state_doubles();
        const unsigned int lanes_per_rank = T / R;
        const size_t blob = S * P * lanes_per_rank;
        std::vector<double> send(blob), recv(blob * (size_t)R);
        const unsigned int my_lo = lane_lo();
        pool->run((int)lane_lo(), (int)lane_hi(), [&](int a, int b) {
            for (int k = a; k < b; ++k)
                for (unsigned int q = 0; q < P; ++q)
                    evaluator->pack(this->particles[(size_t)k * P + q], send.data() + ((size_t)(k - my_lo) * P + q) * S);
        });
        evaluator->allgather_bytes(send.data(), sizeof(double) * blob, recv.data());
        for (int r = 0; r < R; ++r) {
            if (r == rank()) continue;
            const double* rd = recv.data() + blob * (size_t)r;
            const unsigned int lo = (unsigned int)((unsigned long long)T * r / R);
            pool->run(0, (int)lanes_per_rank, [&](int a, int b) {
                for (int k = a; k < b; ++k)
                    for (unsigned int q = 0; q < P; ++q)
                        evaluator->unpack(this->particles[(size_t)(lo + k) * P + q], rd + ((size_t)k * P + q) * S);
            });
        }
    }

    // thread_steps lock-step MCMC steps of all lanes (mcmc_thread for every thread at once).
    // The lanes form two groups; while the GPU scores one group's proposals the host finishes
    // the previous step of the other group and prepares its next one.  Within a lane the
    // order of operations, and therefore its RNG stream, is that of the reference thread.
    void mcmc_lockstep()
    {
        const unsigned int T = this->num_threads;
        if (lane_steps.size() != T) {
            lane_steps.resize(T);
            for (auto& st : lane_steps) evaluator->strip(st.proposal);
            evaluator->reserve_lane_priors(T);
        }
        // this rank's lanes [k0, k1), in two groups
        const unsigned int k0 = lane_lo(), k1 = lane_hi(), n = k1 - k0;
        const unsigned int split[3] = {k0, n < 2 ? k1 : k0 + (n + 1) / 2, k1};
        const int ngroups = n < 2 ? 1 : 2;
        if (!cache_tried) {
            // one factor-cache slot per particle of each group (inactive where it does not apply)
            cache_tried = true;
            for (int g = 0; g < ngroups; ++g)
                evaluator->enable_factor_cache(g, (int)((split[g + 1] - split[g]) * this->options.num_particles));
        }
        int nb[2] = {0, 0};
        for (unsigned int i = 0; i < this->options.thread_steps; ++i) {
            for (int g = 0; g < ngroups; ++g) {
                if (i > 0) {
                    if (nb[g] > 0) timed_wait(g);
                    phase_b(g, split[g], split[g + 1]);
                }
                nb[g] = phase_a(g, split[g], split[g + 1]);
                const double t0 = now();
                if (nb[g] > 0) evaluator->submit(g, nb[g], logl_grp[g].data());
                prof[2] += now() - t0;
            }
        }
        for (int g = 0; g < ngroups; ++g) {
            if (nb[g] > 0) timed_wait(g);
            phase_b(g, split[g], split[g + 1]);
        }
    }


    // ---- device-side proposals (KIMA_DEVICE_PROPOSALS=1) ----------
    // mcmc_thread of every lane runs on the GPU (include/kima_sampler.h): the particles live in
    // HBM, a block of thread_steps steps is one asynchronous enqueue, and the host only does the
    // thread-0 duties of run_thread (SamplerImpl.h:296-331) and do_bookkeeping (:377-440) between
    // blocks.  The algorithm is the reference's; the lanes' random streams are counter-based
    // (include/kc_philox.h), so the chains are not the reference's mt19937 chains: this path is
    // opt-in, the default stays byte-identical.  rngs[0] (bookkeeping draws) stays the host's.
    unsigned int seed0 = 0;
    unsigned int deletions_dev = 0;
    // one process: all_above stays in HBM (kc_sampler_collect_pool / kc_sampler_pool_select), the host
    // only knows its size; several processes: the lists are gathered on the host as before
    double book_prof[3] = {0, 0, 0};  // selecting the new level, lagging particles, saving
    bool pool_mode = false;
    int64_t pool_n = 0;

    static bool device_proposals_requested()
    {
        const char* e = std::getenv("KIMA_DEVICE_PROPOSALS");
        return e && std::atoi(e) != 0;
    }

    // One process per GPU: rank r owns the lanes [lane_lo, lane_hi) and keeps their particles in
    // its GPU's HBM; likelihoods, level assignments, levels and rngs[0] are replicated (every rank
    // runs the same bookkeeping on the same data), particle columns travel only when asked for.
    size_t part_lo() const { return (size_t)lane_lo() * this->options.num_particles; }
    size_t part_hi() const { return (size_t)lane_hi() * this->options.num_particles; }
    int owner_of(size_t i) const { return (int)(i / (this->particles.size() / (size_t)world())); }

    // column of global particle `which` on every rank (its owner fetches it from HBM)
    void fetch_particle(size_t which, std::vector<double>& col, int32_t& np)
    {
        const int S = evaluator->sampler_state_rows();
        std::vector<double> mine((size_t)S + 1, 0.0);
        const int own = owner_of(which);
        if (own == rank()) {
            int32_t k = 0;
            evaluator->sampler_ctx().sampler_get_particle((int)(which - part_lo()), mine.data(), &k);
            mine[S] = (double)k;
        }
        if (world() > 1) {
            std::vector<double> all(((size_t)S + 1) * world());
            evaluator->allgather_bytes(mine.data(), sizeof(double) * (S + 1), all.data());
            mine.assign(all.begin() + (size_t)own * (S + 1), all.begin() + (size_t)(own + 1) * (S + 1));
        }
        np = (int32_t)mine[S];
        col.assign(mine.begin(), mine.begin() + S);
    }

    // Engine outputs one rngs[0].rand() consumes (1 with mt19937_64, DNest4/code/RNG.h:17-20), found by trying:
    // 0 = not a fixed number this code recognises, the caller keeps the serial loop.
    int engine_outputs_per_rand()
    {
        static int k = -1;
        if (k >= 0) return k;
        DNest4::RNG probe = this->rngs[0];
        auto stepped = probe.twister;
        probe.rand();
        k = 0;
        for (int t = 1; t <= 2 && k == 0; ++t) {
            stepped.discard(1);
            if (stepped == probe.twister) k = t;
        }
        return k;
    }

    // The first loop of kill_lagging_particles (SamplerImpl.h:543-558) for both proposal paths: which particles
    // lag (good[], and their indices in particle order), and the largest log_push.  push / kill: log_push and the
    // kill probability per LEVEL (they depend on the level alone: one evaluation per level instead of four per
    // particle; same values, same draws from rngs[0] in the same order).  Returns the number of lagging particles.
    unsigned int flag_lagging(const std::vector<double>& push, const std::vector<double>& kill, std::vector<bool>& good,
                              std::vector<uint32_t>& lagging, double& max_log_push)
    {
        const size_t n = this->particles.size();
        good.assign(n, true);
        lagging.clear();
        max_log_push = -std::numeric_limits<double>::max();
        unsigned int num_bad = 0;
        const int per_draw = world() > 1 ? engine_outputs_per_rand() : 0;
        if (per_draw > 0) {
            // Several ranks: the loop is one draw per particle in particle order, i.e. a fixed number of engine
            // outputs per particle, so each rank draws for the particles it owns and steps the (replicated)
            // engine over the others' -- mt19937_64::discard skips the tempering and the conversion, a tenth of
            // a draw.  Same good[] and the same engine state afterwards as the serial loop; with millions of
            // particles that loop was most of the replicated bookkeeping.
            const size_t a = part_lo(), b = part_hi();
            auto& tw = this->rngs[0].twister;
            const auto check_engine = tw;  // KIMA_CHECK_SHARDED_KILL: the serial loop on a copy, below
            tw.discard((unsigned long long)per_draw * a);
            std::vector<uint32_t> bad;
            double mx = max_log_push;
            for (size_t i = a; i < b; ++i) {
                const unsigned int lv = this->level_assignments[i];
                if (push[lv] > mx) mx = push[lv];
                if (this->rngs[0].rand() <= kill[lv]) bad.push_back((uint32_t)i);
            }
            tw.discard((unsigned long long)per_draw * (n - b));
            struct Head { double mx; uint64_t cnt; } mine{mx, (uint64_t)bad.size()};
            std::vector<Head> heads((size_t)world());
            evaluator->allgather_bytes(&mine, sizeof(Head), heads.data());
            size_t most = 0;
            for (const Head& h : heads) {
                if (h.mx > max_log_push) max_log_push = h.mx;
                num_bad += (unsigned int)h.cnt;
                most = std::max(most, (size_t)h.cnt);
            }
            if (most > 0) {
                std::vector<uint32_t> send(most, 0u), recv(most * (size_t)world());
                std::copy(bad.begin(), bad.end(), send.begin());
                evaluator->allgather_bytes(send.data(), sizeof(uint32_t) * most, recv.data());
                for (int r = 0; r < world(); ++r)  // rank order is particle order
                    for (uint64_t q = 0; q < heads[(size_t)r].cnt; ++q) {
                        good[recv[(size_t)r * most + q]] = false;
                        lagging.push_back(recv[(size_t)r * most + q]);
                    }
            }
            static const bool check = std::getenv("KIMA_CHECK_SHARDED_KILL") != nullptr;
            if (check) {
                // developer check: the reference's loop on a copy of the engine
                DNest4::RNG serial = this->rngs[0];
                serial.twister = check_engine;
                std::vector<bool> good_serial(n, true);
                double mx_serial = -std::numeric_limits<double>::max();
                for (size_t i = 0; i < n; ++i) {
                    const unsigned int lv = this->level_assignments[i];
                    if (push[lv] > mx_serial) mx_serial = push[lv];
                    if (serial.rand() <= kill[lv]) good_serial[i] = false;
                }
                if (good_serial != good || mx_serial != max_log_push || !(serial.twister == tw)) {
                    std::cerr << "# SHARDED KILL MISMATCH on rank " << rank() << std::endl;
                    std::abort();
                }
            }
        } else {
            for (size_t i = 0; i < n; ++i) {
                const unsigned int lv = this->level_assignments[i];
                if (push[lv] > max_log_push) max_log_push = push[lv];
                if (this->rngs[0].rand() <= kill[lv]) {
                    good[i] = false;
                    lagging.push_back((uint32_t)i);
                    ++num_bad;
                }
            }
        }
        return num_bad;
    }
    // log_push and the kill probability of every level (SamplerImpl.h:547-550)
    void level_push_and_kill(std::vector<double>& push, std::vector<double>& kill) const
    {
        const size_t L = this->levels.size();
        push.resize(L);
        kill.resize(L);
        for (size_t j = 0; j < L; ++j) {
            push[j] = this->log_push((unsigned int)j);
            kill[j] = pow(1.0 - 1.0 / (1.0 + exp(-push[j] - 4.0)), 3);
        }
    }

    // kill_lagging_particles (SamplerImpl.h:534-587) with the copies done in HBM: same draws from
    // rngs[0], same messages; a replacement taken from another rank's lane travels as one column
    void kill_lagging_device()
    {
        const size_t n = this->particles.size();
        std::vector<bool> good;
        std::vector<uint32_t> lagging;  // the lagging particles in particle order
        double max_log_push = 0;
        std::vector<double> push, kill;
        level_push_and_kill(push, kill);
        const size_t L = push.size();
        const unsigned int num_bad = flag_lagging(push, kill, good, lagging, max_log_push);
        if (num_bad >= n) return;
        std::vector<double> keep(L);
        for (size_t j = 0; j < L; ++j) keep[j] = exp(push[j] - max_log_push);
        std::vector<size_t> dst, src;
        std::ostringstream msg;
        for (const uint32_t i : lagging) {
            int i_copy;
            do {
                i_copy = this->rngs[0].rand_int((int)n);
            } while (!good[i_copy] || this->rngs[0].rand() >= keep[this->level_assignments[i_copy]]);
            dst.push_back(i);
            src.push_back((size_t)i_copy);
            this->log_likelihoods[i] = this->log_likelihoods[i_copy];
            this->level_assignments[i] = this->level_assignments[i_copy];
            ++deletions_dev;
            if (rank() == 0) msg << "# Replacing lagging particle. This has happened " << deletions_dev << " times.\n";
        }
        if (rank() == 0 && !dst.empty()) std::cout << msg.str() << std::flush;  // the reference's lines, one write
        // sources are good particles and never overwritten: the copies are independent
        std::vector<int32_t> ldst, lsrc;
        std::vector<size_t> cross;
        for (size_t k = 0; k < dst.size(); ++k) {
            const int od = owner_of(dst[k]), os = owner_of(src[k]);
            if (od == os) {
                if (od == rank()) {
                    ldst.push_back((int32_t)(dst[k] - part_lo()));
                    lsrc.push_back((int32_t)(src[k] - part_lo()));
                }
            } else {
                cross.push_back(k);
            }
        }
        kimacuda::Context& ctx = evaluator->sampler_ctx();
        ctx.sampler_copy_particles((int)ldst.size(), ldst.data(), lsrc.data());
        if (cross.empty()) return;
        const int S = evaluator->sampler_state_rows();
        const size_t slot = (size_t)S + 1, K = cross.size();
        std::vector<double> mine(slot * K, 0.0), all(slot * K * world());
        for (size_t q = 0; q < K; ++q)
            if (owner_of(src[cross[q]]) == rank()) {
                int32_t np = 0;
                ctx.sampler_get_particle((int)(src[cross[q]] - part_lo()), mine.data() + q * slot, &np);
                mine[q * slot + S] = (double)np;
            }
        evaluator->allgather_bytes(mine.data(), sizeof(double) * slot * K, all.data());
        for (size_t q = 0; q < K; ++q) {
            const size_t d = dst[cross[q]];
            if (owner_of(d) != rank()) continue;
            const double* col = all.data() + ((size_t)owner_of(src[cross[q]]) * K + q) * slot;
            ctx.sampler_set_particle((int)(d - part_lo()), col, (int32_t)col[S], this->log_likelihoods[d].get_value(),
                                     this->log_likelihoods[d].get_tiebreaker(), (int32_t)this->level_assignments[d]);
        }
    }

    // save_particle (SamplerImpl.h:503-531): the one particle that is printed comes back from HBM
    void save_particle_device()
    {
        ++this->count_saves;
        this->count_mcmc_steps_since_save = 0;
        if (rank() == 0 && this->save_to_disk && this->count_saves % this->thin_print == 0)
            std::cout << "# Saving particle to disk. N = " << this->count_saves << "." << std::endl;
        if (world() == 1 && !this->save_to_disk) return;  // the reference returns before drawing `which`
        // every rank draws `which` (rngs[0] is replicated) and takes part in the fetch
        const int which = this->rngs[0].rand_int((int)this->particles.size());
        std::vector<double> col;
        int32_t np = 0;
        fetch_particle((size_t)which, col, np);
        if (!this->save_to_disk) return;
        evaluator->column_into(this->particles[which], col.data(), np);
        std::fstream fout;
        fout.open(this->options.sample_file, std::ios::out | std::ios::app);
        this->particles[which].print(fout);
        fout << std::endl;
        fout.close();
        fout.open(this->options.sample_info_file, std::ios::out | std::ios::app);
        fout << std::setprecision(12);
        fout << this->level_assignments[which] << ' ';
        fout << this->log_likelihoods[which].get_value() << ' ';
        fout << this->log_likelihoods[which].get_tiebreaker() << ' ';
        fout << which << std::endl;
        fout.close();
    }

    // do_bookkeeping (SamplerImpl.h:377-440), particles in HBM
    void bookkeeping_device()
    {
        bool created_level = false;
        const size_t n_all = pool_mode ? (size_t)pool_n : this->all_above.size();
        if (!this->enough_levels(this->levels) && (n_all >= this->options.new_level_interval)) {
            // the reference sorts all_above and keeps what lies above position `index`; only that
            // element and the SET above it matter afterwards (the next creation orders again)
            const int index = static_cast<int>((1. - 1. / this->compression) * n_all);
            LikelihoodType new_level;
            const double ts = now();
            if (pool_mode) {
                // ordered and cut on the GPU, where the lists already are: one pair comes back
                double v = 0, t = 0;
                evaluator->sampler_ctx().sampler_pool_select(index, &v, &t, &pool_n);
                new_level = LikelihoodType(v, t);
            } else {
                // a selection does the same in O(n): with millions of lanes-times-steps entries per
                // barrier the full sort was most of the bookkeeping
                std::nth_element(this->all_above.begin(), this->all_above.begin() + index, this->all_above.end());
                new_level = this->all_above[index];
                this->all_above.erase(this->all_above.begin(), this->all_above.begin() + index + 1);
            }
            book_prof[0] += now() - ts;
            if (rank() == 0) {
                std::cout << "# Creating level " << this->levels.size() << " with log likelihood = ";
                std::cout << std::setprecision(12);
                std::cout << new_level.get_value() << "." << std::endl;
            }
            this->levels.push_back(Level(new_level));
            if (this->enough_levels(this->levels)) {
                const double reg = this->options.new_level_interval * sqrt(this->options.lambda);
                Level::renormalise_visits(this->levels, static_cast<int>(reg));
                this->all_above.clear();
                if (pool_mode) {
                    evaluator->sampler_ctx().sampler_pool_clear();
                    pool_n = 0;
                }
                this->options.max_num_levels = this->levels.size();
                if (rank() == 0) std::cout << "# Done creating levels." << std::endl;
            } else {
                const double tk = now();
                kill_lagging_device();
                book_prof[1] += now() - tk;
            }
            created_level = true;
        }
        Level::recalculate_log_X(this->levels, this->compression,
                                 this->options.new_level_interval * sqrt(this->options.lambda));
        if (created_level) this->save_levels();
        const double tv = now();
        if (this->count_mcmc_steps_since_save >= this->options.save_interval) {
            save_particle_device();
            if (!created_level) this->save_levels();
        }
        book_prof[2] += now() - tv;
        // record holder (:433-439)
        size_t best = 0;
        if (world() > 1) {
            // each rank searches the particles it owns; the first maximum in particle order, as the serial loop
            size_t mine = part_lo();
            for (size_t i = part_lo() + 1; i < part_hi(); ++i)
                if (this->log_likelihoods[mine] < this->log_likelihoods[i]) mine = i;
            const uint64_t send = (uint64_t)mine;
            std::vector<uint64_t> all((size_t)world());
            evaluator->allgather_bytes(&send, sizeof(uint64_t), all.data());
            best = (size_t)all[0];
            for (int r = 1; r < world(); ++r)
                if (this->log_likelihoods[best] < this->log_likelihoods[(size_t)all[(size_t)r]]) best = (size_t)all[(size_t)r];
        } else {
            for (size_t i = 1; i < this->log_likelihoods.size(); ++i)
                if (this->log_likelihoods[best] < this->log_likelihoods[i]) best = i;
        }
        if (this->best_ever_log_likelihood < this->log_likelihoods[best]) {
            std::vector<double> col;
            int32_t np = 0;
            fetch_particle(best, col, np);
            evaluator->column_into(this->particles[best], col.data(), np);
            this->best_ever_particle = this->particles[best];
            this->best_ever_log_likelihood = this->log_likelihoods[best];
        }
    }

    // an evaluator without the device-side sampler (has_device_sampler = std::false_type: the CPU evaluator of
    // tests/cpp/) never instantiates run_device() and what it calls
    bool run_device_if(std::true_type) { return run_device(); }
    bool run_device_if(std::false_type) { return false; }

    // false = the model has a prior without a device counterpart: the caller runs host proposals
    bool run_device()
    {
        const double t_begin = now();
        const unsigned int T = this->num_threads, P = this->options.num_particles;
        const int W = world();
        if (T % (unsigned int)W != 0)
            throw std::invalid_argument("the number of lanes (-t) must be a multiple of the number of ranks");
        const size_t n = this->particles.size(), p0 = part_lo(), p1 = part_hi(), nl = p1 - p0;
        const unsigned int Tl = lane_hi() - lane_lo();
        kc_sampler_config cfg{};
        cfg.lanes = (int32_t)Tl;
        cfg.num_particles = (int32_t)P;
        cfg.max_steps = (int32_t)this->options.thread_steps;
        cfg.max_levels = this->options.max_num_levels ? (int32_t)this->options.max_num_levels : 1024;
        cfg.lambda = this->options.lambda;
        cfg.beta = this->options.beta;
        cfg.seed = seed0;
        cfg.lane0 = (int32_t)lane_lo();  // lane k of the run draws from stream seed + k whatever rank steps it
        // all_above lives in HBM (with several ranks: replicated, the ranks' lists exchanged GPU to GPU)
        pool_mode = !(std::getenv("KIMA_HOST_ABOVE") && std::atoi(std::getenv("KIMA_HOST_ABOVE")));
        pool_n = 0;
        cfg.pool_reserve = pool_mode ? (int32_t)this->options.new_level_interval : 0;
        const std::string bad = evaluator->sampler_create(this->particles[0], cfg);
        if (!bad.empty()) {
            std::cerr << "# KIMA_DEVICE_PROPOSALS: " << bad << " has no device counterpart; proposals stay on the host"
                      << std::endl;
            return false;
        }
        std::vector<double> logl(n), tb(n);
        std::vector<int32_t> level(n);
        for (size_t i = 0; i < n; ++i) {
            logl[i] = this->log_likelihoods[i].get_value();
            tb[i] = this->log_likelihoods[i].get_tiebreaker();
            level[i] = (int32_t)this->level_assignments[i];
        }
        const double t_created = now();
        evaluator->sampler_upload(this->particles, p0, p1, logl, tb, level);
        kimacuda::Context& ctx = evaluator->sampler_ctx();
        const int max_local = (int)(Tl * this->options.thread_steps), max_pairs = max_local * W;
        std::vector<double> above(pool_mode ? 0 : 2 * (size_t)max_pairs),
            above_local(W > 1 && !pool_mode ? 2 * (size_t)max_local : 0);
        std::vector<double> lv_logl, lv_tb, lv_logX, xs(W > 1 ? 3 * nl : 0), xr(W > 1 ? 3 * n : 0);
        std::vector<uint64_t> lv_tries, counters;
        std::vector<int> counts(W);
        double t_dev = 0, t_book = 0, t_xfer = 0, t_fold = 0;
        const double t_loop0 = now();
        while (true) {
            const size_t L = this->levels.size();
            lv_logl.resize(L); lv_tb.resize(L); lv_logX.resize(L); lv_tries.resize(L); counters.resize(4 * L);
            for (size_t j = 0; j < L; ++j) {
                lv_logl[j] = this->levels[j].get_log_likelihood().get_value();
                lv_tb[j] = this->levels[j].get_log_likelihood().get_tiebreaker();
                lv_logX[j] = this->levels[j].get_log_X();
                lv_tries[j] = this->levels[j].tries;
            }
            if (this->options.max_num_saves != 0 && this->count_saves != 0 &&
                (this->count_saves % this->options.max_num_saves == 0))
                break;
            const double t0 = now();
            ctx.sampler_set_levels((int)L, lv_logl.data(), lv_tb.data(), lv_logX.data(), lv_tries.data(),
                                   this->enough_levels(this->levels), L == this->options.max_num_levels);
            ctx.sampler_run((int)this->options.thread_steps);
            int n_above = 0;
            if (pool_mode) {
                int64_t n_new = 0;
                ctx.sampler_collect_pool(counters.data(), &n_new, &pool_n);
            } else {
                n_above = ctx.sampler_collect(counters.data(), W > 1 ? above_local.data() : above.data(), max_local);
            }
            const double tc = now();
            ctx.sampler_download(nullptr, nullptr, logl.data() + p0, tb.data() + p0, level.data() + p0);
            if (W > 1) {
                // the barrier between processes (SamplerImpl.h:306-329): counters summed, `above` lists
                // concatenated in rank (= lane) order, likelihoods and level assignments to everybody
                if (!pool_mode) {  // (with the pool, kc_sampler_collect_pool has done both between the GPUs)
                    evaluator->allreduce_levels(counters.data(), (int)counters.size());
                    evaluator->allgather_above(above_local.data(), std::min(n_above, max_local), above.data(), max_pairs,
                                               counts.data());
                    n_above = 0;
                    for (int c : counts) n_above += c;
                }
                for (size_t i = 0; i < nl; ++i) {
                    xs[3 * i] = logl[p0 + i];
                    xs[3 * i + 1] = tb[p0 + i];
                    xs[3 * i + 2] = (double)level[p0 + i];
                }
                evaluator->allgather_bytes(xs.data(), sizeof(double) * 3 * nl, xr.data());
            }
            const double t1 = now();
            t_dev += tc - t0;
            t_xfer += t1 - tc;
            const unsigned long long done = (unsigned long long)T * this->options.thread_steps;
            this->count_mcmc_steps += done;
            this->count_mcmc_steps_since_save += done;
            for (size_t j = 0; j < L; ++j) {
                // (the members directly: the increment_* methods take an int, Level.h:56-59)
                this->levels[j].accepts += counters[4 * j + 0];
                this->levels[j].tries += counters[4 * j + 1];
                this->levels[j].visits += counters[4 * j + 2];
                this->levels[j].exceeds += counters[4 * j + 3];
            }
            {
                const size_t base = this->all_above.size(), take = (size_t)std::min(n_above, max_pairs);
                this->all_above.resize(base + take);
                LikelihoodType* dst = this->all_above.data() + base;
                pool->run(0, (int)take, [&](int a, int b) {
                    for (int i = a; i < b; ++i) dst[i] = LikelihoodType(above[2 * (size_t)i], above[2 * (size_t)i + 1]);
                });
            }
            if (W > 1)  // everybody's triples, straight from the exchange buffer
                pool->run(0, (int)n, [&](int a, int b) {
                    for (int i = a; i < b; ++i) {
                        this->log_likelihoods[i] = LikelihoodType(xr[3 * (size_t)i], xr[3 * (size_t)i + 1]);
                        this->level_assignments[i] = (unsigned int)xr[3 * (size_t)i + 2];
                    }
                });
            else
                pool->run(0, (int)n, [&](int a, int b) {
                    for (int i = a; i < b; ++i) {
                        this->log_likelihoods[i] = LikelihoodType(logl[i], tb[i]);
                        this->level_assignments[i] = (unsigned int)level[i];
                    }
                });
            const double t2 = now();
            t_fold += t2 - t1;
            bookkeeping_device();
            t_book += now() - t2;
        }
        const double t_loop1 = now();
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# sampling loop " << t_loop1 - t_loop0 << " s (before it: sampler state in HBM allocated "
                      << t_created - t_begin << " s, particles converted and uploaded " << t_loop0 - t_created << " s)"
                      << std::endl;
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# device proposals: steps + collect " << t_dev << " s, particle arrays back / exchange " << t_xfer
                      << " s, folding counters / above / likelihoods in " << t_fold << " s, bookkeeping " << t_book
                      << " s (new level " << book_prof[0] << ", lagging particles " << book_prof[1] << ", saves "
                      << book_prof[2] << ")" << std::endl;
        // the end-of-run state file holds every particle: the ranks' columns go to everybody
        {
            std::vector<double> st;
            std::vector<int32_t> np;
            evaluator->sampler_local_state(st, np);
            if (W == 1) {
                evaluator->columns_into(this->particles, 0, st, np);
            } else {
                const size_t S = (size_t)evaluator->sampler_state_rows(), blob = (S + 1) * nl;
                std::vector<double> send(blob), recv(blob * W);
                std::copy(st.begin(), st.end(), send.begin());
                for (size_t i = 0; i < nl; ++i) send[S * nl + i] = (double)np[i];
                evaluator->allgather_bytes(send.data(), sizeof(double) * blob, recv.data());
                for (int r = 0; r < W; ++r) {
                    const double* b = recv.data() + blob * r;
                    st.assign(b, b + S * nl);
                    for (size_t i = 0; i < nl; ++i) np[i] = (int32_t)b[S * nl + i];
                    evaluator->columns_into(this->particles, (size_t)r * nl, st, np);
                }
            }
        }
        const double t_state = now();
        if (rank() == 0) {
            std::fstream fout("sampler_state.txt", std::ios::out);
            print_state(fout);
            fout.close();
        }
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# after the loop: particles back to the host " << t_state - t_loop1 << " s, sampler_state.txt "
                      << now() - t_state << " s" << std::endl;
        return true;
    }

public:
    BatchedSampler() {}
    BatchedSampler(unsigned int num_threads, double compression, const Options& options,
                   bool save_to_disk = true)
        : Base(num_threads, compression, options, save_to_disk)
    {
    }

    // Same console output and RNG consumption as Sampler::initialise; the likelihoods of
    // the prior draws are scored in batches afterwards.
    void initialise(unsigned int first_seed)
    {
        RNG& rng = this->rngs[0];
        seed0 = first_seed;
        this->all_above.reserve(2 * this->options.new_level_interval);
        for (auto& a : this->above) a.reserve(2 * this->options.new_level_interval);
        std::cout << "# Seeding random number generators. First seed = ";
        std::cout << first_seed << "." << std::endl;
        for (RNG& r : this->rngs) r.set_seed(first_seed++);
        const size_t n = this->particles.size();
        std::cout << "# Generating " << n;
        std::cout << " particle" << ((n > 1) ? ("s") : (""));
        std::cout << " from the prior..." << std::flush;
        std::vector<double> tiebreakers(n);
        const double ti0 = now();
        ensure_evaluator();
        const double ti1 = now();
        for (size_t i = 0; i < n; ++i) {
            this->particles[i].from_prior(rng);
            tiebreakers[i] = rng.rand();
            evaluator->strip(this->particles[i]);  // host-side state the GPU path never reads
        }
        const size_t cap = (size_t)evaluator->max_batch();
        for (size_t i0 = 0; i0 < n; i0 += cap) {
            const int nb = (int)std::min(cap, n - i0);
            for (int b = 0; b < nb; ++b) evaluator->stage(b, this->particles[i0 + b]);
            evaluator->evaluate(nb, logl_buf.data());
            for (int b = 0; b < nb; ++b)
                this->log_likelihoods[i0 + b] = LikelihoodType(logl_buf[b], tiebreakers[i0 + b]);
        }
        std::cout << "done." << std::endl;
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# initialise: evaluator " << ti1 - ti0 << " s, prior draws + first likelihoods "
                      << now() - ti1 << " s" << std::endl;
        this->initialise_output_files();
        evaluator->slim_host_data();  // nothing on the host reads the data arrays from here on
    }

    // Alternate lock-step MCMC and bookkeeping until max_num_saves is reached
    // (run_thread with every "thread == 0" duty done by the one host thread).
    void run(unsigned int thin = 1)
    {
        this->thin_print = thin;
        const double t_run0 = now();
        ensure_evaluator();
        if (device_proposals_requested() && run_device_if(typename Evaluator::has_device_sampler())) {
            if (std::getenv("KIMA_PROFILE")) std::cerr << "# run() " << now() - t_run0 << " s (device proposals)" << std::endl;
            return;
        }
        const unsigned int T = this->num_threads;
        if (T % (unsigned int)world() != 0)
            throw std::invalid_argument("the number of lanes (-t) must be a multiple of the number of ranks");
        // several ranks: every particle to everybody at every barrier (the first version), or only those the
        // bookkeeping reads (bookkeeping_ranks)
        const bool full_exchange = std::getenv("KIMA_FULL_EXCHANGE") != nullptr;
        while (true) {
            {
                // each lane's own copy of the levels (SamplerImpl.h:283-284); only the lanes this
                // rank steps need one
                const double tc = now();
                pool->run((int)lane_lo(), (int)lane_hi(), [&](int a, int b) {
                    for (int k = a; k < b; ++k) this->copies_of_levels[k] = this->levels;
                });
                prof[7] += now() - tc;
            }
            if (this->options.max_num_saves != 0 && this->count_saves != 0 &&
                (this->count_saves % this->options.max_num_saves == 0))
                break;
            mcmc_lockstep();
            const unsigned long long done = (unsigned long long)T * this->options.thread_steps;
            this->count_mcmc_steps += done;
            this->count_mcmc_steps_since_save += done;
            // fold every lane's counter increments back into the master level table
            const std::vector<Level> before = this->levels;
            if (world() > 1) {
                const double te = now();
                exchange(before, full_exchange);
                prof[6] += now() - te;
            } else {
                for (const auto& lv : this->copies_of_levels)
                    for (size_t j = 0; j < this->levels.size(); ++j) {
                        this->levels[j].increment_accepts(lv[j].get_accepts() - before[j].get_accepts());
                        this->levels[j].increment_tries(lv[j].get_tries() - before[j].get_tries());
                        this->levels[j].increment_visits(lv[j].get_visits() - before[j].get_visits());
                        this->levels[j].increment_exceeds(lv[j].get_exceeds() - before[j].get_exceeds());
                    }
                for (auto& a : this->above) {
                    this->all_above.insert(this->all_above.end(), a.begin(), a.end());
                    a.clear();
                }
            }
            const double tb = now();
            if (world() > 1 && !full_exchange) bookkeeping_ranks();
            else this->do_bookkeeping();
            prof[5] += now() - tb;
        }
        if (world() > 1 && !full_exchange) gather_particles();
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# run() " << now() - t_run0 << " s" << std::endl;
        if (std::getenv("KIMA_PROFILE"))
            std::cerr << "# host threads " << pool->threads() << "  perturb " << prof[0] << " s, stage " << prof[1]
                      << " s, submit " << prof[2] << " s, wait " << prof[3] << " s, phase B " << prof[4]
                      << " s, bookkeeping " << prof[5] << " s, exchange " << prof[6] << " s, level copies " << prof[7]
                      << " s" << std::endl;
        if (rank() == 0) {
            std::fstream fout("sampler_state.txt", std::ios::out);
            this->print(fout);
            fout.close();
        }
    }

    long long device_launches() const { return evaluator ? evaluator->launches() : 0; }
};

}  // namespace DNest4
