/*
 * Ensemble of independent daecpp::solve() instances (parameter sweeps) -- an ADDITION to the dae-cpp
 * API, not a change of it: the reference has no ensemble entry point.
 *
 * Members are independent DAE systems; nothing is exchanged between them.  Each member is integrated
 * by the unmodified daecpp::solve() of whichever solver.hpp is on the include path (the B200 overlay
 * or the reference's own), so step counts, exit codes and tolerances are exactly those of a single
 * solve.  With the overlay every in-flight member owns one device context and one CUDA stream
 * (include/daecpp_b200.h: "one context = one DAE system ... an ensemble may own many contexts"); a pool
 * of host threads keeps `threads` members in flight per GPU, which overlaps the small kernels of
 * different members on the device and the host functors of some members with the device work of
 * others.  One process drives one GPU (DAECPP_B200_DEVICE); across GPUs the members are partitioned
 * by the launcher (bench.py: one rank per GPU, contiguous blocks of members, one gather at the end).
 *
 * Usage:
 *   auto result = daecpp::solve_ensemble(4096, [&](std::size_t m) {            // called on a pool thread
 *       MyMass mass{...}; MyRhs rhs{param[m]}; MyJac jac{param[m]};
 *       daecpp::state_vector x0 = ...; daecpp::ensemble_capture out;            // keeps the last state
 *       int status = daecpp::solve(mass, rhs, jac, x0, t_end, out.observer(), opt);
 *       return daecpp::ensemble_member{status, out.steps(), std::move(out.x)};
 *   }, threads);
 */
#pragma once

#include <dae-cpp/solver.hpp>

#include <atomic>
#include <cstddef>
#include <cstdlib>
#include <thread>
#include <utility>
#include <vector>

namespace daecpp
{

struct ensemble_member
{
    int status{-1};      // exit code of daecpp::solve for this member
    long steps{0};       // accepted time steps
    state_vector x;      // final state
};

// Solution-manager helper: remembers the last state handed to it (solution-manager.hpp:175-189 semantics).
struct ensemble_capture
{
    state_vector x;
    double t{0.0};
    long calls{0};
    struct Observer
    {
        ensemble_capture *self;
        int operator()(const state_vector &xk, const double tk)
        {
            self->x = xk;
            self->t = tk;
            self->calls++;
            return 0;
        }
    };
    Observer observer() { return Observer{this}; }
    long steps() const { return calls > 0 ? calls - 1 : 0; } // the initial condition is reported first
};

struct ensemble_result
{
    std::vector<ensemble_member> members;
    int failed{0};
};

// Runs member(m) for m in [first, first + count) on `threads` host threads (0: hardware concurrency).
template <class MemberFn>
ensemble_result solve_ensemble(std::size_t first, std::size_t count, MemberFn member, unsigned threads = 0)
{
    ensemble_result R;
    R.members.resize(count);
    if (threads == 0) threads = std::thread::hardware_concurrency();
    if (threads == 0) threads = 1;
    if (threads > count) threads = (unsigned)count;
#ifdef DAECPP_B200_H
    {   // waiting threads sleep: the cores belong to the functors of the other members in flight
        int device = 0;
        if (const char *env = std::getenv("DAECPP_B200_DEVICE")) device = std::atoi(env);
        if (threads > 2 * std::thread::hardware_concurrency()) daecpp_b200_set_blocking_sync(device, 1);
        daecpp_b200_set_pooling(1); // contexts and pinned buffers are recycled from member to member
    }
#endif
    std::atomic<std::size_t> next{0};
    std::atomic<int> failed{0};
    auto work = [&]() {
        for (;;)
        {
            const std::size_t k = next.fetch_add(1);
            if (k >= count) return;
            R.members[k] = member(first + k);
            if (R.members[k].status != 0) failed.fetch_add(1);
        }
    };
    std::vector<std::thread> pool;
    for (unsigned t = 1; t < threads; ++t) pool.emplace_back(work);
    work();
    for (auto &th : pool) th.join();
#ifdef DAECPP_B200_H
    daecpp_b200_set_pooling(0);
#endif
    R.failed = failed.load();
    return R;
}

template <class MemberFn>
ensemble_result solve_ensemble(std::size_t count, MemberFn member, unsigned threads = 0)
{
    return solve_ensemble(0, count, member, threads);
}

} // namespace daecpp
