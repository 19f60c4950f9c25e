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
#ifdef DAECPP_B200_H
#include <dae-cpp/device-functor.hpp>
#endif

#include <algorithm>
#include <atomic>
#include <chrono>
#include <future>
#include <cmath>
#include <cstddef>
#include <cstdio>
#include <cstdlib>
#include <memory>
#include <thread>
#include <type_traits>
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

#ifdef DAECPP_B200_H
/*
 * ---- lock-step batched ensembles (B200 only) ---------------------------------------------------------
 *
 * All members of a sweep share one sparsity pattern, so instead of one device context per member the
 * members become LANES of one batch context (include/daecpp_b200.h: daecpp_b200_batch_*): every Newton
 * round is one launch sequence for all lanes in flight.  Each lane keeps its own integrator state and makes
 * its own step-size / order / refactorisation decisions -- `lane_control` below restates the decisions of
 * the reference's time loop (dae-cpp/solver.hpp:402-797) as a state machine that is advanced once per
 * round -- so exit code, accepted steps and final state of a member are those of a single daecpp::solve.
 *
 * Restrictions of this path (the thread-pool solve_ensemble above has none of them): one common pattern
 * recognised as a block chain (banded systems), a mass matrix that does not change in time, and a Solution
 * Manager that only observes: solve_ensemble_device_observed hands it the states of the output times (a lane's
 * state is downloaded only then -- with device functors nothing else on the host needs it) and honours
 * solver_command::stop_intergration; the two "decrease the time step" commands are not available.
 */
namespace core
{
namespace detail
{

class lane_control
{
    const SolverOptions &m_opt;
    std::vector<double> m_t_out;
    std::size_t m_out{0};
    bool m_in_time_loop{false};
    double m_t1{0.0};

    double m_dt[DAECPP_MAX_ORDER];
    double m_t{0.0}, m_t_prev{0.0};
    unsigned int m_order{1};
    int m_iter{0};
    bool m_fact{false};
    int m_dt_increase_threshold{0}, m_dt_decrease_threshold{0}, m_max_Newton_iter{0};
    uint64_t m_n_steps{0};
    uint32_t m_n_iter_failed{0};
    bool m_delay_timestep_inc{false};
    int m_pending_pre{DAECPP_B200_LANE_KEEP};
    bool m_done{false};
    bool m_last_was_accept{false};
    int m_prev_iters{0};        // Newton iterations the last finished Newton loop of this lane took
    bool m_event_accept{false}; // a step was accepted in the round just consumed, at time m_event_t
    double m_event_t{0.0};

    double &dt() { return m_dt[0]; }

    // dae-cpp/solver.hpp:383-404 and the head of the time loop :402-408
    void next_step()
    {
        for (;;)
        {
            if (!m_in_time_loop)
            {
                if (m_out >= m_t_out.size())
                {
                    status = exit_code::success;
                    m_done = true;
                    return;
                }
                m_t1 = m_t_out[m_out];
                if (m_t1 <= 0.0) { ++m_out; continue; }
                if (dt() > m_t1 - m_t)
                {
                    dt() = m_t1 - m_t;
                    ASSERT(dt() > 0.0, "Negative or zero time step: dt = " << dt() << ".");
                }
                m_delay_timestep_inc = false;
                m_in_time_loop = true;
            }
            if (m_t < m_t1)
            {
                m_t_prev = m_t;
                m_t += dt();
                m_n_steps++;
                m_iter = 0;
                return;
            }
            m_in_time_loop = false;
            ++m_out;
        }
    }

    // dae-cpp/solver.hpp:636-651
    void shrink_and_rewind()
    {
        m_delay_timestep_inc = true;
        m_t = m_t_prev;
        m_n_steps--;
        dt() /= m_opt.dt_decrease_factor;
        m_pending_pre = DAECPP_B200_LANE_ROLLBACK;
        m_last_was_accept = false;
        if (dt() < m_opt.dt_min)
        {
            status = exit_code::diverged;
            m_done = true;
            return;
        }
        next_step();
    }

    // the Newton loop of this step has ended at iteration index `iter` (dae-cpp/solver.hpp:622-797)
    void end_newton(bool diverged, double variability)
    {
        m_prev_iters = std::min(m_iter + 1, m_max_Newton_iter);
        if (m_iter == m_max_Newton_iter) diverged = true; // "couldn't converge"
        if (diverged)
        {
            m_n_iter_failed++;
            shrink_and_rewind();
            return;
        }
        // no Solution Manager on this path; the step is accepted
        m_n_iter_failed = 0;
        accepted_steps++;
        m_pending_pre = DAECPP_B200_LANE_ACCEPT;
        m_last_was_accept = true;
        m_event_accept = true;
        m_event_t = m_t;
        for (int k = DAECPP_MAX_ORDER - 1; k > 0; --k) m_dt[k] = m_dt[k - 1];
        if ((m_iter >= m_dt_decrease_threshold) || (variability > m_opt.variability_threshold_high))
        {
            dt() /= m_opt.dt_decrease_factor;
            if (dt() < m_opt.dt_min)
            {
                status = exit_code::diverged;
                m_done = true;
                return;
            }
        }
        else if ((m_iter <= m_dt_increase_threshold - 1) && !m_delay_timestep_inc && (variability <= m_opt.variability_threshold_low))
        {
            dt() *= m_opt.dt_increase_factor;
            if (dt() > m_opt.dt_max) dt() = m_opt.dt_max;
        }
        m_delay_timestep_inc = false;
        if (dt() > m_t1 - m_t) dt() = m_t1 - m_t;
        if (m_order < m_opt.BDF_order) m_order++;
        if (dt() < m_opt.dt_min)
        {
            if (m_dt[2] > m_opt.dt_min) dt() = m_dt[2];
            else if (m_dt[1] > m_opt.dt_min) dt() = m_dt[1];
            else dt() = m_opt.dt_init;
            m_in_time_loop = false; // leave the time loop of this output time
            ++m_out;
        }
        next_step();
    }

public:
    int status{exit_code::unknown};
    long accepted_steps{0};
    Counters count;

    lane_control(const SolverOptions &opt, const std::vector<double> &t_out) : m_opt(opt), m_t_out(t_out)
    {
        std::fill(m_dt, m_dt + DAECPP_MAX_ORDER, 0.0);
        m_dt_increase_threshold = 2 * (m_opt.Newton_scheme + 1) + m_opt.dt_increase_threshold_delta;
        m_dt_decrease_threshold = 4 * (m_opt.Newton_scheme + 1) + m_opt.dt_decrease_threshold_delta;
        ASSERT(m_dt_increase_threshold > 0, "Too small delta `dt_increase_threshold_delta`: " << m_opt.dt_increase_threshold_delta);
        ASSERT(m_dt_decrease_threshold > 0, "Too small delta `dt_decrease_threshold_delta`: " << m_opt.dt_decrease_threshold_delta);
        ASSERT(m_dt_decrease_threshold > m_dt_increase_threshold, "Adaptive time stepping thresholds are not consistent with each other.");
        m_max_Newton_iter = m_opt.max_Jacobian_updates * (m_opt.Newton_scheme + 1);
        dt() = m_opt.dt_init;
        next_step();
    }

    bool done() const { return m_done; }
    // How many iterations the Newton loop in progress will probably still run: what the last one took (3 before there is a
    // last one), less those already done.  A multi-iteration round is planned this long -- iterations queued beyond the
    // end of the loop are launches that find nothing to do, a loop that needs more simply goes on in the next round.
    int expected_iterations() const { return std::max(1, (m_prev_iters > 0 ? m_prev_iters : 3) - m_iter); }
    // did the round just consumed end with an accepted step (the point where the reference calls the Solution Manager,
    // dae-cpp/solver.hpp:659)?  Its time goes to `t`; the lane's Newton iterate is the accepted state until the next round.
    bool take_accept(double &t)
    {
        const bool e = m_event_accept;
        m_event_accept = false;
        t = m_event_t;
        return e;
    }
    // solver_command::stop_intergration from an observer (dae-cpp/solver.hpp:683-689): the member ends here, successfully
    void stop()
    {
        status = exit_code::success;
        m_done = true;
    }
    double time() const { return m_done && !m_last_was_accept ? m_t_prev : m_t; }
    // the final state of a finished lane is its Newton iterate when the last event was an accepted step,
    // the last accepted state x[0] otherwise
    bool final_state_is_iterate() const { return m_last_was_accept; }

    // what this lane does in the coming round (dae-cpp/solver.hpp:431-537)
    void plan(daecpp_b200_lane_cmd &cmd, int max_iterations = 1)
    {
        cmd = daecpp_b200_lane_cmd{};
        cmd.pre = m_pending_pre;
        m_pending_pre = DAECPP_B200_LANE_KEEP;
        if (m_done)
        {
            cmd.mode = DAECPP_B200_LANE_IDLE;
            return;
        }
        m_fact = (m_n_iter_failed < m_opt.max_Newton_failed_attempts) ? !(m_iter % (m_opt.Newton_scheme + 1)) : true;
        cmd.mode = m_fact ? DAECPP_B200_LANE_REFACTOR : DAECPP_B200_LANE_SOLVE;
        cmd.order = static_cast<int32_t>(m_order);
        cmd.t = m_t;
        for (int k = 0; k < DAECPP_MAX_ORDER; ++k) cmd.dt[k] = m_dt[k];
        // should the driver let the lane run several iterations in one round: which of them refactorise (:533) and how
        // many the Newton loop has left
        cmd.period = (m_n_iter_failed < m_opt.max_Newton_failed_attempts) ? static_cast<int32_t>(m_opt.Newton_scheme + 1) : 1;
        cmd.phase = m_iter;
        cmd.max_iterations = std::min(max_iterations, m_max_Newton_iter - m_iter);
    }

    // the outcome of that round (dae-cpp/solver.hpp:541-627)
    void consume(const daecpp_b200_lane_result &res)
    {
        if (m_done) return;
        // iterations the lane ran ahead of the last one of the round ended "not converged, not diverged"
        for (int k = 1; k < res.iterations; ++k)
        {
            if (m_fact) count.n_fact_calls++;
            count.n_lin_calls++;
            ++m_iter;
            m_fact = (m_n_iter_failed < m_opt.max_Newton_failed_attempts) ? !(m_iter % (m_opt.Newton_scheme + 1)) : true;
        }
        if (m_fact) count.n_fact_calls++;
        if (res.status == DAECPP_B200_EDECOMP || res.status == DAECPP_B200_ESOLVE)
        {
            if (m_opt.recover_from_linsolver_failure)
            {
                if (res.status == DAECPP_B200_ESOLVE) count.n_lin_calls++;
                end_newton(true, 0.0);
                return;
            }
            status = res.status == DAECPP_B200_EDECOMP ? exit_code::linsolver_failed_decomposition : exit_code::linsolver_failed_solving;
            m_done = true;
            return;
        }
        count.n_lin_calls++;
        const bool diverged = (res.flags & DAECPP_B200_DIVERGED) != 0;
        const bool converged = (res.flags & DAECPP_B200_NOT_CONVERGED) == 0;
        const double variability = m_opt.solution_variability_control ? res.variability : 0.0;
        if (converged || diverged)
        {
            end_newton(diverged, variability);
            return;
        }
        if (++m_iter == m_max_Newton_iter) end_newton(false, variability);
    }
};

} // namespace detail
} // namespace core

struct lockstep_stats
{
    long rounds{0};
    double ms_device{0.0}; // CUDA-event time of all rounds
    double ms_wall{0.0};
    long lane_iterations{0}, lane_refactorisations{0};
    double kernels_launched{0.0};
    double refine_rounds{0.0}, refined_lanes{0.0}; // guarded refinement passes (include/daecpp_b200.h) and the lanes they touched
    double ms_setup{0.0}; // batch creation, pattern analysis, uploads (before the first round)
    long observed_states{0}; // lane states downloaded for the observer (solve_ensemble_device_observed)
    double ms_rounds_wall{0.0}, ms_download{0.0}, ms_destroy{0.0}; // host wall clock of the round loop, of the final-state download, of the teardown
};

/*
 * Integrates eqs.size() members on [0, t_end] (or through every time of t_output) in lock step on one GPU.
 *   Eq     -- device functor type of the equations (dae-cpp/device-functor.hpp), one object per member
 *   shape  -- the non-zero elements of the Jacobian, shared by all members
 *   mass   -- the usual host mass-matrix functor `mass(sparse_matrix &M, double t)`; evaluated once, at t = 0
 *   x0     -- one initial state shared by all members, or one per member
 * Needs nvcc for the translation unit (the equations are instantiated as kernels).
 *
 * solve_ensemble_device_observed: the same with an observer `int observer(std::size_t member, const state_vector &x, double t)`
 * called for the initial state (if 0 is an output time) and for every accepted step that lands on a time of t_output --
 * what `daecpp::Solution(holder, t_output)` records (dae-cpp/solution-manager.hpp:175-189).  Only then is the lane's
 * state brought to the host.  A return value of solver_command::stop_intergration ends that member.
 */
struct no_observer
{
    int operator()(std::size_t, const state_vector &, double) const { return 0; }
};

template <class Eq, class Mass, class Observer>
ensemble_result solve_ensemble_device_observed(const std::vector<Eq> &eqs, const device::Shape &shape, Mass mass, const std::vector<state_vector> &x0,
                                               const std::vector<double> &t_output, Observer &&observer, const SolverOptions &opt = SolverOptions(),
                                               lockstep_stats *stats = nullptr)
{
    constexpr bool observe = !std::is_same<typename std::decay<Observer>::type, no_observer>::value;
    const std::size_t members = eqs.size();
    ensemble_result R;
    R.members.resize(members);
    if (members == 0) return R;
    ASSERT(x0.size() == 1 || x0.size() == members, "solve_ensemble_device: one initial state, or one per member.");
    const std::size_t n = x0[0].size();
    ASSERT(n > 0 && n <= 0x7fffffffu, "solve_ensemble_device: bad system size.");
    opt.check();
    std::vector<double> t_out(t_output);
    std::sort(t_out.begin(), t_out.end());
    t_out.erase(std::unique(t_out.begin(), t_out.end()), t_out.end());
    ASSERT(!t_out.empty() && t_out.back() >= 0.0, "solve_ensemble_device: no (or a negative) target time.");

    const auto wall0 = std::chrono::steady_clock::now();
    const bool timing = std::getenv("DAECPP_B200_TIMING") != nullptr; // debug: host-side phases to stderr
    auto stamp = [&](const char *what) {
        if (timing) std::fprintf(stderr, "  [solve_ensemble_device] %-18s at %8.3f ms\n", what, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count());
    };
    double st_setup_ms = 0.0;
    std::vector<char> initial_stop;
    int dev = 0;
    if (const char *env = std::getenv("DAECPP_B200_DEVICE")) dev = std::atoi(env);
    // The batch's device memory (gigabytes) and its analysed pattern are parked at the end for the next sweep of this
    // process instead of going through cudaFree; daecpp_b200_set_pooling(0) hands everything back.
    daecpp_b200_set_pooling(1);
    daecpp_b200_batch *b = nullptr;
    int rc = daecpp_b200_batch_create(static_cast<int32_t>(n), static_cast<int32_t>(members), dev, &b);
    if (rc) { ERROR("daecpp_b200: cannot create the batch context (code " << rc << "): " << daecpp_b200_batch_last_error(nullptr)); }
    auto fatal = [&](const char *what, int code) { ERROR("daecpp_b200: " << what << " failed (code " << code << "): " << daecpp_b200_batch_last_error(b)); };

    stamp("batch_create");
    sparse_matrix M;
    mass(M, 0.0);
    M.check();
    if ((rc = daecpp_b200_batch_set_pattern(b, shape.rows().data(), shape.cols().data(), static_cast<int64_t>(shape.size()), M.i.data(), M.j.data(),
                                            static_cast<int64_t>(M.A.size()))))
        fatal("batch_set_pattern", rc);
    stamp("set_pattern");
    if ((rc = daecpp_b200_batch_set_mass_values(b, M.A.data(), 0, 0))) fatal("batch_set_mass_values", rc);
    {
        if (x0.size() == 1 && members > 1)
        {   // every member starts from the same state: it travels once
            if ((rc = daecpp_b200_batch_set_state(b, x0[0].data(), 0, 0))) fatal("batch_set_state", rc);
        }
        else
        {
            std::unique_ptr<double[]> slab(new double[members * n]);
            for (std::size_t m = 0; m < members; ++m)
            {
                const state_vector &x = x0.size() == 1 ? x0[0] : x0[m];
                ASSERT(x.size() == n, "solve_ensemble_device: initial states differ in size.");
                std::copy(x.begin(), x.end(), slab.get() + m * n);
            }
            if ((rc = daecpp_b200_batch_set_state(b, slab.get(), 0, static_cast<int32_t>(members)))) fatal("batch_set_state", rc);
        }
    }
    stamp("set_state");
    if (observe && std::binary_search(t_out.begin(), t_out.end(), 0.0))
    {   // the initial condition is reported first (dae-cpp/solver.hpp:358)
        for (std::size_t m = 0; m < members; ++m) initial_stop.push_back(observer(m, x0.size() == 1 ? x0[0] : x0[m], 0.0) == 1);
    }
    device::DeviceFunctors<Eq> functors(eqs, shape);
    daecpp_b200_device_functors fn = functors.binding();
    if ((rc = daecpp_b200_batch_set_device_functors(b, &fn))) fatal("batch_set_device_functors", rc);

    st_setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    std::vector<core::detail::lane_control> lanes;
    lanes.reserve(members);
    for (std::size_t m = 0; m < members; ++m) lanes.emplace_back(opt, t_out);
    for (std::size_t m = 0; m < initial_stop.size(); ++m)
        if (initial_stop[m]) lanes[m].stop();
    state_vector observed; // one lane's state on its way to the observer
    std::vector<daecpp_b200_lane_cmd> cmd(members);
    std::vector<daecpp_b200_lane_result> res(members);
    lockstep_stats st;
    const bool trace_rounds = std::getenv("DAECPP_B200_TRACE_ROUNDS") != nullptr; // debug: lanes in flight and wall clock per round
    long trace_lane = -1; // debug: DAECPP_B200_TRACE_LANE=<lane> prints that lane's rounds to stderr
    if (const char *env = std::getenv("DAECPP_B200_TRACE_LANE")) trace_lane = std::atol(env);
    // With many lanes in flight a round is one Newton iteration (the device is busy, the host round trip is small
    // change); when only a few stragglers are left the round trip dominates, and they run several iterations per round.
    std::size_t in_flight = members;
    std::size_t straggler_threshold = 64; // (DAECPP_B200_STRAGGLERS=<lanes> overrides: tests force either regime)
    if (const char *env = std::getenv("DAECPP_B200_STRAGGLERS")) straggler_threshold = (std::size_t)std::atol(env);
    for (;;)
    {
        int per_round = 1;
        if (in_flight < straggler_threshold)
        {
            for (std::size_t m = 0; m < members; ++m)
                if (!lanes[m].done()) per_round = std::max(per_round, lanes[m].expected_iterations());
            per_round = std::min(per_round, DAECPP_B200_MAX_ROUND_ITERATIONS);
        }
        std::size_t active = 0;
        for (std::size_t m = 0; m < members; ++m)
        {
            lanes[m].plan(cmd[m], per_round);
            if (cmd[m].mode != DAECPP_B200_LANE_IDLE) ++active;
        }
        in_flight = active;
        if (!active) break;
        if ((rc = daecpp_b200_batch_round(b, cmd.data(), opt.atol, opt.rtol, opt.max_err_abs, opt.solution_variability_control ? 1 : 0,
                                          opt.variability_tolerance, res.data())))
            fatal("batch_round", rc);
        st.rounds++;
        if (trace_rounds) std::fprintf(stderr, "round %ld: %zu lanes, %d iteration(s), done at %.3f ms\n", st.rounds, active, per_round, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count());
        if (trace_lane >= 0 && (std::size_t)trace_lane < members && cmd[trace_lane].mode != DAECPP_B200_LANE_IDLE)
            std::fprintf(stderr, "lane %ld round %ld: t=%.10g dt=%.6g order=%d mode=%d pre=%d -> flags=%d status=%d var=%.3g\n", trace_lane, st.rounds,
                         cmd[trace_lane].t, cmd[trace_lane].dt[0], cmd[trace_lane].order, cmd[trace_lane].mode, cmd[trace_lane].pre, res[trace_lane].flags,
                         res[trace_lane].status, res[trace_lane].variability);
        if (trace_lane >= 0 && (std::size_t)trace_lane < members && cmd[trace_lane].mode != DAECPP_B200_LANE_IDLE && std::getenv("DAECPP_B200_TRACE_DX"))
        {
            std::vector<double> dx(n), xa(n);
            daecpp_b200_batch_get_vector(b, 1, dx.data(), (int32_t)trace_lane, 1);
            daecpp_b200_batch_get_vector(b, 4, xa.data(), (int32_t)trace_lane, 1);
            std::size_t ia = 0, ir = 0;
            double ma = 0, mr = 0;
            for (std::size_t k = 0; k < n; ++k)
            {
                const double e = std::fabs(dx[k]), a = std::fabs(xa[k]);
                if (e > ma) { ma = e; ia = k; }
                if (a > 1e-14 && e / a > mr) { mr = e / a; ir = k; }
            }
            std::fprintf(stderr, "    max|dx| = %.3e at %zu (x0 = %.3e)   max rel = %.3e at %zu (x0 = %.3e, dx = %.3e)\n", ma, ia, xa[ia], mr, ir, xa[ir], dx[ir]);
        }
        for (std::size_t m = 0; m < members; ++m)
            if (cmd[m].mode != DAECPP_B200_LANE_IDLE)
            {
                for (int k = 0; k < res[m].iterations; ++k)
                {
                    const bool refac = k == 0 ? cmd[m].mode == DAECPP_B200_LANE_REFACTOR : (cmd[m].period <= 1 || (cmd[m].phase + k) % cmd[m].period == 0);
                    st.lane_iterations++;
                    if (refac) st.lane_refactorisations++;
                }
                lanes[m].consume(res[m]);
                if (observe)
                {
                    double ta = 0.0;
                    if (lanes[m].take_accept(ta) && std::binary_search(t_out.begin(), t_out.end(), ta))
                    {   // the lane's iterate is the accepted state until its next round: the only download of this step
                        observed.resize(n);
                        if ((rc = daecpp_b200_batch_get_vector(b, 3, observed.data(), static_cast<int32_t>(m), 1))) fatal("batch_get_vector", rc);
                        st.observed_states++;
                        if (observer(m, observed, ta) == 1) lanes[m].stop();
                    }
                }
            }
    }
    stamp("rounds");
    auto since0 = [&] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count(); };
    st.ms_rounds_wall = since0() - st_setup_ms;
    {   // final states: the iterate of lanes that ended on an accepted step, x[0] of the others
        // downloaded in chunks of lanes through a few small buffers; the copy into the members' vectors (first touch of
        // fresh pages: the slow part) runs on helper threads while the next chunk is on its way
        constexpr int RING = 4;
        const std::size_t chunk = std::max<std::size_t>(1, (std::size_t(4) << 20) / (n * sizeof(double)));
        std::unique_ptr<double[]> buf[RING];
        std::future<void> busy[RING];
        for (auto &p : buf) p.reset(new double[2 * chunk * n]);
        int slot = 0;
        for (std::size_t m0 = 0; m0 < members; m0 += chunk, slot = (slot + 1) % RING)
        {
            const std::size_t cnt = std::min(chunk, members - m0);
            if (busy[slot].valid()) busy[slot].get();
            double *xk = buf[slot].get(), *xa = xk + chunk * n;
            if ((rc = daecpp_b200_batch_get_vector(b, 3, xk, static_cast<int32_t>(m0), static_cast<int32_t>(cnt)))) fatal("batch_get_vector", rc);
            bool need_x0 = false;
            for (std::size_t m = m0; m < m0 + cnt; ++m) need_x0 = need_x0 || !lanes[m].final_state_is_iterate();
            if (need_x0 && (rc = daecpp_b200_batch_get_vector(b, 4, xa, static_cast<int32_t>(m0), static_cast<int32_t>(cnt)))) fatal("batch_get_vector", rc);
            busy[slot] = std::async(std::launch::async, [&R, &lanes, xk, xa, m0, cnt, n] {
                for (std::size_t m = m0; m < m0 + cnt; ++m)
                {
                    const double *src = (lanes[m].final_state_is_iterate() ? xk : xa) + (m - m0) * n;
                    R.members[m].x.assign(src, src + n);
                }
            });
        }
        for (auto &f : busy)
            if (f.valid()) f.get();
        for (std::size_t m = 0; m < members; ++m)
        {
            R.members[m].status = lanes[m].status;
            R.members[m].steps = lanes[m].accepted_steps;
            if (R.members[m].status != 0) R.failed++;
        }
    }
    double v = 0.0;
    if (daecpp_b200_batch_get_stat(b, "ms_device", &v) == DAECPP_B200_OK) st.ms_device = v;
    if (daecpp_b200_batch_get_stat(b, "kernels_launched", &v) == DAECPP_B200_OK) st.kernels_launched = v;
    if (daecpp_b200_batch_get_stat(b, "refine_rounds", &v) == DAECPP_B200_OK) st.refine_rounds = v;
    if (daecpp_b200_batch_get_stat(b, "refined_lanes", &v) == DAECPP_B200_OK) st.refined_lanes = v;
    st.ms_setup = st_setup_ms;
    st.ms_download = since0() - st_setup_ms - st.ms_rounds_wall;
    stamp("final states");
    daecpp_b200_batch_destroy(b);
    stamp("destroy");
    st.ms_destroy = since0() - st_setup_ms - st.ms_rounds_wall - st.ms_download;
    st.ms_wall = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - wall0).count();
    if (stats) *stats = st;
    return R;
}

template <class Eq, class Mass>
ensemble_result solve_ensemble_device(const std::vector<Eq> &eqs, const device::Shape &shape, Mass mass, const std::vector<state_vector> &x0,
                                      const std::vector<double> &t_output, const SolverOptions &opt = SolverOptions(),
                                      lockstep_stats *stats = nullptr)
{
    return solve_ensemble_device_observed(eqs, shape, mass, x0, t_output, no_observer{}, opt, stats);
}

template <class Eq, class Mass>
ensemble_result solve_ensemble_device(const std::vector<Eq> &eqs, const device::Shape &shape, Mass mass, const std::vector<state_vector> &x0,
                                      const double t_end, const SolverOptions &opt = SolverOptions(), lockstep_stats *stats = nullptr)
{
    return solve_ensemble_device(eqs, shape, mass, x0, std::vector<double>{t_end}, opt, stats);
}
#endif // DAECPP_B200_H

} // namespace daecpp
