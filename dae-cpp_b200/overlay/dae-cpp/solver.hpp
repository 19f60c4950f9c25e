/*
 * dae-cpp solver front end, B200 edition.
 *
 * Drop-in replacement for dae-cpp/solver.hpp of dae-cpp 2.2.0: the same `daecpp::solve`
 * overloads, `daecpp::System`, `core::detail::SolverState`, `core::detail::time_derivative_approx`
 * and exit codes, but the Newton linear-solve path -- triplet assembly, `J - alpha*M`, sparse LU
 * refactorisation, triangular solves, BDF residual, error control, history shift -- runs in
 * hand-written sm_100a kernels behind the C ABI of include/daecpp_b200.h.  User functors (mass,
 * RHS, Jacobian, Solution Manager) still run on the host, exactly as before.
 *
 * Usage: put this directory BEFORE the dae-cpp installation on the include path and link
 * libdaecpp_b200.so:
 *     g++ -std=c++17 -I dae-cpp_b200/overlay -I include -I <dae-cpp root> app.cpp -ldaecpp_b200
 * Every other dae-cpp header (sparse-matrix.hpp, jacobian-matrix.hpp, solver-options.hpp, ...) is
 * the user's own, unmodified.
 *
 * Reference regions replaced (paths relative to the dae-cpp tree):
 *     core::detail::solve            dae-cpp/solver.hpp:237-815
 *       Eigen members                dae-cpp/solver.hpp:325-337   -> one device context
 *       M.convert / J.convert        dae-cpp/solver.hpp:481,502   -> set_mass / set_jacobian
 *       b, Jb, compute, solve        dae-cpp/solver.hpp:512-582   -> residual / refactor / solve_update
 *       error loop, history shift    dae-cpp/solver.hpp:590-619,698-721 -> solve_update / accept_step
 * There is no CPU fallback: without a CUDA device the first call reports a fatal ERROR.
 */
#ifndef DAECPP_SOLVER_H
#define DAECPP_SOLVER_H

#include <algorithm>
#include <array>
#include <cstdint>
#include <cstdio>
#include <iomanip>
#include <sstream>
#include <string>
#include <type_traits>
#include <vector>

#include <dae-cpp/jacobian-matrix.hpp>
#include <dae-cpp/mass-matrix.hpp>
#include <dae-cpp/solution-manager.hpp>
#include <dae-cpp/solver-options.hpp>
#include <dae-cpp/timer.hpp>
#include <dae-cpp/vector-function.hpp>
#include <dae-cpp/version.hpp>

#include <daecpp_b200.h>

namespace daecpp_namespace_name
{
namespace core
{
namespace detail
{

/*
 * Integrator state: step sizes, times, solution history and BDF order.  Same public layout as
 * the reference's (dae-cpp/solver.hpp:38-69); tests construct and fill it directly.  During a
 * solve() the history vectors live on the device -- only `dt`, `t`, `t_prev` and `order` are
 * used on the host.
 */
struct SolverState
{
    double dt[DAECPP_MAX_ORDER];
    double t{0.0};
    double t_prev{0.0};
    std::array<rvec, DAECPP_MAX_ORDER> x;
    unsigned int order{1};

    explicit SolverState(const std::size_t size)
    {
        std::fill(dt, dt + DAECPP_MAX_ORDER, 0.0);
        try
        {
            for (auto &v : x)
                v.resize(size);
        }
        catch (const std::exception &e)
        {
            ERROR("Failed to allocate memory for the solver state.\n"
                  << e.what());
        }
    }
};

struct Counters
{
    uint64_t n_fact_calls{0}; // Jacobian updates + refactorisations
    uint64_t n_lin_calls{0};  // linear solves
};

/*
 * Variable-step BDF weights of orders 1-4: w[0] multiplies xk (and is d(dxdt)/d(xk)),
 * w[1..4] multiply x[0..3] with alternating signs.  Every product and quotient keeps the
 * association of dae-cpp/solver.hpp:160-207 so the values are bit-identical to the reference's.
 * The device uses the same weights (daecpp_b200_residual).
 */
inline bool bdf_weights(unsigned int order, const double *dt, double *w)
{
    const double a = dt[0], b = dt[1], c = dt[2], d = dt[3];
    const double ab = a + b, bc = b + c, cd = c + d;
    const double abc = a + b + c, bcd = b + c + d, abcd = a + b + c + d;
    std::fill(w, w + 5, 0.0);
    switch (order)
    {
    case 1:
        w[0] = 1.0 / a;
        w[1] = w[0];
        return true;
    case 2:
        w[0] = (2.0 * a + b) / (a * ab);
        w[1] = ab / (a * b);
        w[2] = a / (b * ab);
        return true;
    case 3:
        w[0] = (3.0 * a * a + b * bc + 2.0 * a * (2.0 * b + c)) / (a * ab * abc);
        w[1] = ab * abc / (a * b * bc);
        w[2] = a * abc / (b * c * ab);
        w[3] = a * ab / (c * bc * abc);
        return true;
    case 4:
        w[0] = (4.0 * a * a * a + b * bc * bcd + 3.0 * a * a * (3.0 * b + 2.0 * c + d) +
                2.0 * a * (3.0 * b * b + c * cd + 2.0 * b * (2.0 * c + d))) /
               (a * ab * abc * abcd);
        w[1] = ab * abc * abcd / (a * b * bc * bcd);
        w[2] = a * abc * abcd / (b * ab * c * cd);
        w[3] = a * ab * abcd / (c * bc * abc * d);
        w[4] = a * ab * abc / (d * cd * bcd * abcd);
        return true;
    default:
        return false;
    }
}

/*
 * Host evaluation of the BDF time derivative and of alpha = d(dxdt)/d(xk).  Kept with the
 * reference's signature (dae-cpp/solver.hpp:141) because user code and tests call it; the
 * integrator below evaluates the same expression on the device, fused with b = M*dxdt - f.
 */
inline double time_derivative_approx(eivec &dxdt, const rvec &xk, const SolverState &state, const std::size_t size)
{
    double w[5];
    if (!bdf_weights(state.order, state.dt, w))
    {
        ERROR("Unsupported time integration order.");
    }
    if (state.order == 1)
    {
        for (std::size_t i = 0; i < size; ++i)
            dxdt[i] = (xk[i] - state.x[0][i]) * w[0];
        return w[0];
    }
    for (std::size_t i = 0; i < size; ++i)
    {
        double v = w[0] * xk[i] - w[1] * state.x[0][i] + w[2] * state.x[1][i];
        if (state.order >= 3)
            v = v - w[3] * state.x[2][i];
        if (state.order >= 4)
            v = v + w[4] * state.x[3][i];
        dxdt[i] = v;
    }
    return w[0];
}

/*
 * Owner of the device context: replaces the Eigen locals of the reference's solve()
 * (linsolver, M_, Jb, f_, b, dx, dxdt: dae-cpp/solver.hpp:325-337) and the history vectors.
 * Fatal (CUDA / usage) errors go through ERROR like every other fatal condition of dae-cpp;
 * numerical failures are returned to the integrator, which handles them the way the reference
 * handles `linsolver.info() != Eigen::Success`.
 */
// Jacobian functors whose triplet (i, j) stream cannot change during a solve
template <class T> struct is_jacobian_matrix_shape : std::false_type {};
template <class R> struct is_jacobian_matrix_shape<JacobianMatrixShape<R>> : std::true_type {};

class DeviceNewton
{
    daecpp_b200_ctx *m_ctx{nullptr};
    std::size_t m_n{0};
    double *m_f{nullptr};  // page-locked RHS values
    double *m_xk{nullptr}; // page-locked copy of the Newton iterate

    // Page-locked registration of the Jacobian VALUE array of a JacobianMatrixShape functor.  Only that array
    // is ever registered: the functor reserves and fills exactly its fixed element list on every call
    // (dae-cpp/jacobian-matrix.hpp:66-92), so after the first call the storage behind J.A can never be reallocated
    // -- and therefore never freed while registered (CUDA requires the unregister to come first).  Every other
    // triplet array (mass matrix, index arrays, user-written Jacobian functors whose element count may change)
    // is handed over as ordinary pageable memory.
    struct Pinned
    {
        const void *p{nullptr};
        std::size_t bytes{0};
    };
    Pinned m_pin;
    std::size_t m_nT_pattern{0}; // length of the (i, j) stream the device has cached for J

    void fatal(int rc, const char *what) const
    {
        ERROR("daecpp_b200: " << what << " failed (code " << rc << "): " << daecpp_b200_last_error(m_ctx));
    }

    void unpin()
    {
        if (m_pin.p && daecpp_b200_host_unregister(const_cast<void *>(m_pin.p)) != DAECPP_B200_OK)
        {
            WARNING("daecpp_b200: could not release the page lock on the Jacobian values.");
        }
        m_pin = Pinned();
    }
    // Page-lock the storage once it is large and has stopped moving.
    template <class Vec>
    void pin(const Vec &v)
    {
        const void *p = v.data();
        const std::size_t bytes = v.capacity() * sizeof(typename Vec::value_type);
        if (p == m_pin.p && bytes == m_pin.bytes)
            return;
        unpin();
        if (bytes >= (std::size_t(1) << 22) && daecpp_b200_host_register(const_cast<void *>(p), bytes) == DAECPP_B200_OK)
        {
            m_pin.p = p;
            m_pin.bytes = bytes;
        }
    }

public:
    explicit DeviceNewton(std::size_t n) : m_n(n)
    {
        ASSERT(n <= 0x7fffffffu, "System is too large for 32-bit indices.");
        int device = 0;
        if (const char *env = std::getenv("DAECPP_B200_DEVICE"))
            device = std::atoi(env);
        const int rc = daecpp_b200_create(static_cast<int32_t>(n), device, &m_ctx);
        if (rc != DAECPP_B200_OK)
        {
            ERROR("daecpp_b200: cannot create the device context (code " << rc << "): " << daecpp_b200_last_error(nullptr));
        }
        m_f = static_cast<double *>(daecpp_b200_host_alloc(n * sizeof(double)));
        m_xk = static_cast<double *>(daecpp_b200_host_alloc(n * sizeof(double)));
        if (!m_f || !m_xk)
        {
            ERROR("daecpp_b200: cannot allocate page-locked host buffers.");
        }
        daecpp_b200_set_option(m_ctx, "pinned_io", 1.0);
    }
    DeviceNewton(const DeviceNewton &) = delete;
    DeviceNewton &operator=(const DeviceNewton &) = delete;
    ~DeviceNewton()
    {
        unpin(); // (the integrator declares its sparse matrices before this object: their storage is still alive here)
        daecpp_b200_host_free(m_f);
        daecpp_b200_host_free(m_xk);
        daecpp_b200_destroy(m_ctx);
    }

    // Device-side phase stopwatch (CUDA events between the phases of every iteration; costs a stream synchronisation per
    // iteration, so it is only switched on when the caller wants the timing report: verbosity >= 1, systems above the
    // one-launch tiny path).  phase_totals: milliseconds spent in {assembly, residual, refactor, solve, update} so far.
    void enable_phase_timing() { daecpp_b200_set_option(m_ctx, "phase_timing", 1.0); }
    void phase_totals(double ms[5]) const
    {
        static const char *keys[5] = {"ms_phase_assembly", "ms_phase_residual", "ms_phase_refactor", "ms_phase_solve", "ms_phase_update"};
        double samples = 0.0;
        daecpp_b200_get_stat(m_ctx, "n_phase_samples", &samples);
        for (int k = 0; k < 5; ++k)
        {
            double avg = 0.0;
            daecpp_b200_get_stat(m_ctx, keys[k], &avg);
            ms[k] = avg * samples;
        }
    }
    double *f_host() { return m_f; }
    bool jacobian_values_page_locked() const { return m_pin.p != nullptr; }
    const double *xk_host() const { return m_xk; }

    void set_state(const state_vector &x0)
    {
        std::copy(x0.begin(), x0.end(), m_xk);
        const int rc = daecpp_b200_set_state(m_ctx, m_xk);
        if (rc)
            fatal(rc, "set_state");
    }
    // M_ = M.convert(size)   (dae-cpp/solver.hpp:481)
    void set_mass(const sparse_matrix &M)
    {
        const int rc = daecpp_b200_set_mass(m_ctx, M.A.data(), M.i.data(), M.j.data(), static_cast<int64_t>(M.A.size()));
        ASSERT(rc != DAECPP_B200_EINDEX, "Mass matrix: sparse matrix index is out of boundaries, matrix size is " << m_n << "x" << m_n);
        if (rc)
            fatal(rc, "set_mass");
    }
    // Jb = J.convert(size)   (dae-cpp/solver.hpp:502).  fixed_pattern: the functor is a
    // JacobianMatrixShape, whose (i, j) stream is fixed at construction (jacobian-matrix.hpp:66-92):
    // after the first hand-off only the values travel.
    void set_jacobian(const sparse_matrix &J, bool fixed_pattern)
    {
        if (fixed_pattern)
            pin(J.A);
        if (fixed_pattern && m_nT_pattern == J.A.size() && m_nT_pattern > 0)
        {
            const int rc = daecpp_b200_set_jacobian_values(m_ctx, J.A.data(), static_cast<int64_t>(J.A.size()));
            if (rc)
                fatal(rc, "set_jacobian_values");
            return;
        }
        m_nT_pattern = J.A.size();
        const int rc = daecpp_b200_set_jacobian(m_ctx, J.A.data(), J.i.data(), J.j.data(), static_cast<int64_t>(J.A.size()));
        ASSERT(rc != DAECPP_B200_EINDEX, "Jacobian matrix: sparse matrix index is out of boundaries, matrix size is " << m_n << "x" << m_n);
        if (rc)
            fatal(rc, "set_jacobian");
    }
    // dxdt, alpha and b = M*dxdt - f   (dae-cpp/solver.hpp:439,516-518); f is read from f_host()
    double residual(const SolverState &state)
    {
        double alpha = 0.0;
        const int rc = daecpp_b200_residual(m_ctx, m_f, static_cast<int32_t>(state.order), state.dt, &alpha);
        if (rc)
            fatal(rc, "residual");
        return alpha;
    }
    // Jb -= M_*alpha; linsolver.compute(Jb)   (dae-cpp/solver.hpp:523,537).  false: decomposition failed.
    bool refactor(double alpha)
    {
        const int rc = daecpp_b200_refactor(m_ctx, alpha);
        if (rc == DAECPP_B200_EDECOMP)
            return false;
        if (rc)
            fatal(rc, "refactor");
        return true;
    }
    // dx = linsolver.solve(b), error checks, xk += dx   (dae-cpp/solver.hpp:558,590-619)
    bool solve_update(const SolverOptions &opt, bool &diverged, bool &converged)
    {
        int32_t flags = 0;
        const int rc = daecpp_b200_solve_update(m_ctx, opt.atol, opt.rtol, opt.max_err_abs, &flags);
        if (rc == DAECPP_B200_ESOLVE)
            return false;
        if (rc)
            fatal(rc, "solve_update");
        diverged = (flags & DAECPP_B200_DIVERGED) != 0;
        converged = (flags & DAECPP_B200_NOT_CONVERGED) == 0;
        return true;
    }
    // One whole Newton iteration in one submission and one synchronisation: f upload, b, (refactor,)
    // solves, checks, xk += dx, download of xk   (dae-cpp/solver.hpp:439-619).
    // Returns 0, or 1 when the decomposition failed (the caller rolls back like the reference, :541-551).
    int iterate(const SolverState &state, const SolverOptions &opt, bool do_refactor, bool &diverged, bool &converged,
                state_vector &xk)
    {
        int32_t flags = 0;
        double alpha = 0.0;
        const int rc = daecpp_b200_iteration_host(m_ctx, m_f, do_refactor ? 1 : 0, static_cast<int32_t>(state.order), state.dt,
                                                  opt.atol, opt.rtol, opt.max_err_abs, m_xk, &alpha, &flags);
        if (rc == DAECPP_B200_EDECOMP)
            return 1;
        if (rc)
            fatal(rc, "iteration_host");
        diverged = (flags & DAECPP_B200_DIVERGED) != 0;
        converged = (flags & DAECPP_B200_NOT_CONVERGED) == 0;
        if (!diverged)
            std::copy(m_xk, m_xk + m_n, xk.begin());
        return 0;
    }
    void fetch_xk(state_vector &xk)
    {
        const int rc = daecpp_b200_get_xk(m_ctx, m_xk);
        if (rc)
            fatal(rc, "get_xk");
        std::copy(m_xk, m_xk + m_n, xk.begin());
    }
    // xk = x[0]   (dae-cpp/solver.hpp:643,671)
    void rollback()
    {
        const int rc = daecpp_b200_rollback(m_ctx);
        if (rc)
            fatal(rc, "rollback");
    }
    // history shift + maximum relative variability   (dae-cpp/solver.hpp:698-721)
    double accept_step(const SolverOptions &opt)
    {
        double variability = 0.0;
        const int rc = daecpp_b200_accept_step(m_ctx, opt.solution_variability_control ? 1 : 0, opt.variability_tolerance, &variability);
        if (rc)
            fatal(rc, "accept_step");
        return variability;
    }
};

inline void print_char(bool condition, char ch) noexcept
{
    if (condition)
        std::putchar(ch);
}

/*
 * One line of the timing report: value, unit (ms, or s for long runs) and share of the total.
 */
inline std::string print_time(double val, double t_total)
{
    const bool seconds = t_total > 1e4;
    std::ostringstream oss;
    oss << std::setw(11) << (seconds ? val * 1e-3 : val) << ' ' << (seconds ? "s" : "ms");
    oss << std::setprecision(3) << "  (" << static_cast<int>(val / t_total * 1e5) / 1000.0 << "%)";
    return oss.str();
}

/*
 * Timing report, same buckets and layout as the reference (dae-cpp/solver.hpp:116-136).  Device
 * phases are synchronous inside their C-ABI calls, so the host timers around them stay truthful.
 */
inline void finalize(const timer::Time &time, const int v, const Counters c)
{
    if (v < 1)
        return;
    const auto &t = time.timers;
    const std::string rule(64, '-');
    const std::pair<const char *, double> lines[] = {
        {"  Initialization:       ", t[timer::init]},
        {"  Time derivative:      ", t[timer::time_derivative]},
        {"  RHS:                  ", t[timer::rhs]},
        {"  Mass matrix:          ", t[timer::mass]},
        {"  Jacobian matrix:      ", t[timer::jacobian]},
        {"  Linear algebra:       ", t[timer::linear_algebra]},
    };
    PRINT(true, "\nComputation time:" << std::right);
    PRINT(true, rule);
    for (const auto &ln : lines)
        PRINT(true, ln.first << print_time(ln.second, time.total));
    PRINT(true, "  Matrix factorization: " << print_time(t[timer::factorization], time.total) << "  <--  " << c.n_fact_calls << " calls");
    PRINT(true, "  Linear solver:        " << print_time(t[timer::linear_solver], time.total) << "  <--  " << c.n_lin_calls << " calls");
    PRINT(true, "  Error control:        " << print_time(t[timer::error_check], time.total));
    PRINT(true, "  Solution Manager:     " << print_time(t[timer::manager], time.total));
    PRINT(true, "  Other calculations:   " << print_time(time.other(), time.total));
    PRINT(true, rule);
    PRINT(true, "Total time:             " << print_time(time.total, time.total) << std::left);
    PRINT(true, rule << '\n');
}

/*
 * The integrator: variable-step BDF-1..4 with a quasi-Newton iteration.  Step acceptance,
 * rejection and step-size control follow the reference's rules (dae-cpp/solver.hpp:402-797)
 * decision for decision; everything that touches vectors of length N or the matrices is a call
 * into the device context.
 */
template <class Mass, class RHS, class Jacobian, class Manager>
class Integrator
{
    Mass &m_mass;
    RHS &m_rhs;
    Jacobian &m_jac;
    Manager &m_mgr;
    const SolverOptions &m_opt;
    const std::size_t m_size;

    Counters m_count;
    timer::Time m_time;
    SolverState m_state;
    sparse_matrix m_M, m_J; // declared BEFORE the device object: it may hold a page lock on m_J.A and must be destroyed first
    DeviceNewton m_dev;

    state_vector m_xk;     // host copy of the Newton iterate (what the functors see)
    state_vector m_x_prev; // host copy of the last accepted state, x[0]
    state_type m_f_ad, m_xk_ad; // autodiff-typed views handed to the RHS functor

    int m_dt_increase_threshold{0}, m_dt_decrease_threshold{0}, m_max_Newton_iter{0};
    uint64_t m_n_steps{0};
    uint32_t m_n_iter_failed{0};
    bool m_delay_timestep_inc{false};
    bool m_phase_timing{false}, m_phase_folded{false};

    enum class StepResult { accepted, redo, leave_time_loop, finished };
    exit_code::status m_status{exit_code::unknown};

    double &dt() { return m_state.dt[0]; }
    bool verbose(int level) const { return m_opt.verbosity >= level; }

    // Undo the time step, shrink dt, restore xk = x[0].  False: dt fell below dt_min.
    bool shrink_and_rewind()
    {
        m_delay_timestep_inc = true;
        m_state.t = m_state.t_prev;
        m_n_steps--;
        dt() /= m_opt.dt_decrease_factor;
        m_dev.rollback();
        m_xk = m_x_prev;
        if (dt() < m_opt.dt_min)
        {
            PRINT(verbose(1), "The time step was reduced to `t_min` but the scheme failed to converge.");
            m_status = exit_code::diverged;
            return false;
        }
        return true;
    }

    void eval_rhs()
    {
        try
        {
            Timer timer(&m_time.timers[timer::rhs]);
            for (std::size_t k = 0; k < m_size; ++k)
                m_xk_ad[k] = m_xk[k];
            m_rhs(m_f_ad, m_xk_ad, m_state.t);
            ASSERT(static_cast<std::size_t>(m_f_ad.size()) == m_size, "The RHS vector size (" << m_f_ad.size() << ") does not match the initial condition vector size (" << m_size << ").");
            double *f = m_dev.f_host();
            for (std::size_t k = 0; k < m_size; ++k)
                f[k] = m_f_ad[k].val();
        }
        catch (const std::exception &e)
        {
            ERROR("Failed to compute and convert the vector function (RHS).\n"
                  << e.what());
        }
    }

    void eval_mass()
    {
        try
        {
            Timer timer(&m_time.timers[timer::mass]);
            m_M.clear();
            m_mass(m_M, m_state.t);
            m_M.check();
            m_dev.set_mass(m_M);
        }
        catch (const std::exception &e)
        {
            ERROR("Failed to compute and convert the Mass matrix.\n"
                  << e.what());
        }
    }

    void eval_jacobian()
    {
        try
        {
            Timer timer(&m_time.timers[timer::jacobian]);
            m_J.clear();
            m_jac(m_J, m_xk, m_state.t);
            m_J.check();
            m_dev.set_jacobian(m_J, is_jacobian_matrix_shape<Jacobian>::value);
        }
        catch (const std::exception &e)
        {
            ERROR("Failed to compute and convert the Jacobian matrix.\n"
                  << e.what());
        }
    }

    // Returns false when the whole integration has to stop (m_status is set).
    // On return `iter` is the Newton iteration index the loop stopped at, `diverged` the outcome.
    bool newton(int &iter, bool &diverged)
    {
        auto &t = m_time.timers;
        diverged = false;
        for (iter = 0; iter < m_max_Newton_iter; ++iter)
        {
            eval_rhs();

            if (!m_opt.is_mass_matrix_static || (!iter && !m_count.n_lin_calls))
                eval_mass();

            // Refactorise on every (Newton_scheme+1)-th iteration, or always after repeated failures
            const bool is_fact_enabled = (m_n_iter_failed < m_opt.max_Newton_failed_attempts) ? !(iter % (m_opt.Newton_scheme + 1)) : true;
            if (is_fact_enabled)
                eval_jacobian();

            bool converged = true;
            {
                // residual, (refactorisation,) solves, error checks and the download of the new iterate
                // are one device submission; its time goes to the bucket of its dominant part
                Timer timer(&t[is_fact_enabled ? timer::factorization : timer::linear_solver]);
                const int rc = m_dev.iterate(m_state, m_opt, is_fact_enabled, diverged, converged, m_xk);
                if (is_fact_enabled)
                    m_count.n_fact_calls++;
                if (rc == 1)
                {
                    PRINT(verbose(2), " <- decomposition failed");
                    if (m_opt.recover_from_linsolver_failure)
                    {
                        diverged = true;
                        return true;
                    }
                    m_status = exit_code::linsolver_failed_decomposition;
                    return false;
                }
                m_count.n_lin_calls++;
                print_char(verbose(2), is_fact_enabled ? '#' : '*');
                if (diverged)
                {
                    PRINT(verbose(2), " <- diverged");
                }
            }
            if (converged || diverged)
                return true;
        }
        return true;
    }

    StepResult step(const double t1)
    {
        m_state.t_prev = m_state.t;
        m_state.t += dt();
        m_n_steps++;

        if (verbose(2))
        {
            std::cout << std::left << "Step " << std::setw(8) << m_n_steps << " :: t = " << std::setw(11) << m_state.t << " :: "
                      << "BDF-" << m_state.order << ": dt=" << std::setw(11) << dt() << " :: ";
        }

        int iter = 0;
        bool is_diverged = false;
        if (!newton(iter, is_diverged))
            return StepResult::finished;
        if (iter == m_max_Newton_iter)
        {
            PRINT(verbose(2), " <- couldn't converge");
            is_diverged = true;
        }
        if (is_diverged)
        {
            m_n_iter_failed++;
            return shrink_and_rewind() ? StepResult::redo : StepResult::finished;
        }

        bool decrease_time_step = false;
        try
        {
            Timer timer(&m_time.timers[timer::manager]);
            const auto command = m_mgr(m_xk, m_state.t);
            if (command == solver_command::decrease_time_step)
                decrease_time_step = true;
            else if (command == solver_command::decrease_time_step_and_redo)
            {
                PRINT(verbose(2), " <- decrease_time_step_and_redo");
                return shrink_and_rewind() ? StepResult::redo : StepResult::finished;
            }
            else if (command) // stop_intergration
            {
                print_char(verbose(2), '\n');
                PRINT(verbose(1), "Stop event in Solution Manager triggered.");
                m_status = exit_code::success;
                return StepResult::finished;
            }
        }
        catch (const std::exception &e)
        {
            ERROR("Solution Manager functor call failed.\n"
                  << e.what());
        }

        // ---- the step is accepted
        m_n_iter_failed = 0;
        const double variability = m_dev.accept_step(m_opt); // x[3]<-x[2]<-x[1]<-x[0]<-xk on the device
        m_x_prev = m_xk;
        for (int k = DAECPP_MAX_ORDER - 1; k > 0; --k)
            m_state.dt[k] = m_state.dt[k - 1];

        // ---- next step size
        if ((iter >= m_dt_decrease_threshold) || (variability > m_opt.variability_threshold_high) || decrease_time_step)
        {
            dt() /= m_opt.dt_decrease_factor;
            print_char(verbose(2), '<');
            if (dt() < m_opt.dt_min)
            {
                PRINT(verbose(2), " <- reached dt_min");
                PRINT(verbose(1), "The time step was reduced to `t_min` but the scheme failed to converge.");
                m_status = exit_code::diverged;
                return StepResult::finished;
            }
        }
        else if ((iter <= m_dt_increase_threshold - 1) && !m_delay_timestep_inc && (variability <= m_opt.variability_threshold_low))
        {
            dt() *= m_opt.dt_increase_factor;
            const bool capped = dt() > m_opt.dt_max;
            if (capped)
                dt() = m_opt.dt_max;
            print_char(verbose(2), capped ? '|' : '>');
        }
        m_delay_timestep_inc = false;

        if (dt() > t1 - m_state.t)
            dt() = t1 - m_state.t; // land exactly on the output time
        if (m_state.order < m_opt.BDF_order)
            m_state.order++;
        print_char(verbose(2), '\n');

        if (dt() < m_opt.dt_min)
        { // the output time has been reached: restart from a sane step size
            if (m_state.dt[2] > m_opt.dt_min)
                dt() = m_state.dt[2];
            else if (m_state.dt[1] > m_opt.dt_min)
                dt() = m_state.dt[1];
            else
                dt() = m_opt.dt_init;
            return StepResult::leave_time_loop;
        }
        return StepResult::accepted;
    }

public:
    Integrator(Mass &mass, RHS &rhs, Jacobian &jac, Manager &mgr, const state_vector &x0, const SolverOptions &opt)
        : m_mass(mass), m_rhs(rhs), m_jac(jac), m_mgr(mgr), m_opt(opt), m_size(x0.size()), m_state(x0.size()), m_dev(x0.size()),
          m_xk(x0), m_x_prev(x0), m_f_ad(x0.size()), m_xk_ad(x0.size())
    {
    }

    const Counters &counters() const { return m_count; }
    // The reference's timer buckets (dae-cpp/timer.hpp:59-71, filled at dae-cpp/solver.hpp:438,514,535-539,556-560,588).
    // Here one device submission does the work of five of them, so the host stopwatch around it can only feed
    // "factorization" / "linear_solver"; with the device-side phase stopwatch on, its split is handed on:
    //   linear_algebra <- assembly of Jb + residual b (which forms dxdt on the fly: time_derivative stays with it)
    //   linear_solver  <- triangular solves          error_check <- error check + update
    //   factorization  <- everything else of the submissions: numeric refactorisation, transfers, host turn-around
    // (kernels that fuse two phases report the fused time under the phase that hosts them).
    timer::Time &time()
    {
        if (m_phase_timing && !m_phase_folded)
        {
            m_phase_folded = true;
            double ms[5];
            m_dev.phase_totals(ms);
            auto &t = m_time.timers;
            const double submissions = t[timer::factorization] + t[timer::linear_solver];
            const double moved = ms[0] + ms[1] + ms[3] + ms[4];
            if (moved > 0.0 && moved <= submissions)
            {
                t[timer::linear_algebra] += ms[0] + ms[1];
                t[timer::error_check] += ms[4];
                t[timer::linear_solver] = ms[3];
                t[timer::factorization] = submissions - moved;
            }
        }
        return m_time;
    }

    exit_code::status run(const state_vector &x0, const std::vector<double> &t_out, bool is_jac_auto)
    {
        {
            Timer timer_init(&m_time.timers[timer::init]);

            PRINT(verbose(1), "Starting dae-cpp version " << version_major << '.' << version_minor << '.' << version_patch << " (B200 device path: " << daecpp_b200_version() << ")...");
            PRINT(verbose(1) && is_jac_auto, "NOTE: Using automatic Jacobian...");

            m_dt_increase_threshold = 2 * (m_opt.Newton_scheme + 1) + m_opt.dt_increase_threshold_delta;
            ASSERT(m_dt_increase_threshold > 0, "Too small delta `dt_increase_threshold_delta`: " << m_opt.dt_increase_threshold_delta);
            m_dt_decrease_threshold = 4 * (m_opt.Newton_scheme + 1) + m_opt.dt_decrease_threshold_delta;
            ASSERT(m_dt_decrease_threshold > 0, "Too small delta `dt_decrease_threshold_delta`: " << m_opt.dt_decrease_threshold_delta);
            ASSERT(m_dt_decrease_threshold > m_dt_increase_threshold, "Adaptive time stepping thresholds are not consistent with each other.");
            m_max_Newton_iter = m_opt.max_Jacobian_updates * (m_opt.Newton_scheme + 1);

            dt() = m_opt.dt_init;
            m_dev.set_state(x0);
            m_phase_timing = verbose(1) && m_size > 32;
            if (m_phase_timing)
                m_dev.enable_phase_timing();

            PRINT(verbose(2), "Float size:      " << 8 * sizeof(float_type) << " bit");
            PRINT(verbose(2), "Integer size:    " << 8 * sizeof(int_type) << " bit");
            PRINT(verbose(2), "BDF max order:   " << m_opt.BDF_order);
            PRINT(verbose(2), "Newton scheme:   " << m_opt.Newton_scheme);
            PRINT(verbose(2), "Max time step:   " << m_opt.dt_max);
            PRINT(verbose(1), "DAE system size: " << m_size << " equations");
            PRINT(verbose(1), "Calculating...");

            try
            {
                Timer timer(&m_time.timers[timer::manager]);
                if (m_mgr(x0, 0.0) == solver_command::stop_intergration)
                {
                    PRINT(verbose(1), "Stop event in Solution Manager triggered.");
                    return exit_code::success;
                }
            }
            catch (const std::exception &e)
            {
                ERROR("Solution Manager functor call failed.\n"
                      << e.what());
            }
        }

        for (const double t1 : t_out)
        {
            PRINT(verbose(2), "\n-- Integration time t = " << t1 << ":");
            if (t1 <= 0.0)
            {
                WARNING("Negative or zero integration time t = " << t1 << ". Skipped.");
                continue;
            }
            if (dt() > t1 - m_state.t)
            {
                dt() = t1 - m_state.t;
                ASSERT(dt() > 0.0, "Negative or zero time step: dt = " << dt() << ".\n"
                                       << "This assertion triggered after adjusting the time step to match the output time t_output = " << t1 << ".");
            }
            m_delay_timestep_inc = false;
            while (m_state.t < t1)
            {
                const StepResult r = step(t1);
                if (r == StepResult::finished)
                    return m_status;
                if (r == StepResult::leave_time_loop)
                    break;
            }
        }
        return exit_code::success;
    }
};

/*
 * The main solver: integrates M(t) dx/dt = f(x,t) from 0 to the last output time.  Same
 * signature and semantics as the reference's core::detail::solve (dae-cpp/solver.hpp:237-238).
 */
template <class Mass, class RHS, class Jacobian, class Manager>
inline exit_code::status solve(Mass mass, RHS rhs, Jacobian jac, Manager mgr, const state_vector &x0, const double t_end, const std::vector<double> &t_output, const SolverOptions &opt, bool is_jac_auto)
{
    exit_code::status status = exit_code::unknown;
    Counters counters;
    timer::Time time;
    {
        Timer timer_global(&time.total);

        std::vector<double> t_out(t_output);
        if (t_out.empty())
            t_out.push_back(t_end);
        else
        {
            std::sort(t_out.begin(), t_out.end());
            t_out.erase(std::unique(t_out.begin(), t_out.end()), t_out.end());
        }
        ASSERT(t_out.back() >= 0.0, "Target time t_end cannot be negative. The solver integrates from 0 to t_end.");
        opt.check();
        ASSERT(x0.size() > 0, "Initial condition vector x is empty.");

        Integrator<Mass, RHS, Jacobian, Manager> integrator(mass, rhs, jac, mgr, x0, opt);
        status = integrator.run(x0, t_out, is_jac_auto);
        counters = integrator.counters();
        for (std::size_t k = 0; k < time.timers.size(); ++k)
            time.timers[k] = integrator.time().timers[k];

        PRINT(opt.verbosity >= 1, "...done");
    }
    finalize(time, opt.verbosity, counters);
    return status;
}

} // namespace detail
} // namespace core

/*
 * Public entry points.  Signatures are those of dae-cpp 2.2.0 (dae-cpp/solver.hpp:835-902):
 * functors are taken by value, the Solution Manager and the options are optional trailing
 * arguments, and leaving the Jacobian out selects automatic differentiation of the RHS.
 *
 *   solve(mass, rhs, jac, x0, t_end,    mgr, opt)      user Jacobian, integrate on [0, t_end]
 *   solve(mass, rhs,      x0, t_end,    mgr, opt)      automatic Jacobian
 *   solve(mass, rhs, jac, x0, t_output, mgr, opt)      stop exactly at every time in t_output
 *   solve(mass, rhs,      x0, t_output, mgr, opt)
 *
 * All return daecpp::exit_code::success or an error code.
 */
template <class Mass, class RHS, class Jacobian, class Manager = SolutionManager>
inline exit_code::status solve(Mass mass, RHS rhs, Jacobian jac, const state_vector &x0, const double t_end, Manager mgr = SolutionManager(), const SolverOptions &opt = SolverOptions())
{
    const std::vector<double> no_output_times;
    return core::detail::solve(mass, rhs, jac, mgr, x0, t_end, no_output_times, opt, false);
}

template <class Mass, class RHS, class Manager = SolutionManager>
inline exit_code::status solve(Mass mass, RHS rhs, const state_vector &x0, const double t_end, Manager mgr = SolutionManager(), const SolverOptions &opt = SolverOptions())
{
    const std::vector<double> no_output_times;
    return core::detail::solve(mass, rhs, JacobianAutomatic(rhs), mgr, x0, t_end, no_output_times, opt, true);
}

template <class Mass, class RHS, class Jacobian, class Manager = SolutionManager>
inline exit_code::status solve(Mass mass, RHS rhs, Jacobian jac, const state_vector &x0, const std::vector<double> &t_output, Manager mgr = SolutionManager(), const SolverOptions &opt = SolverOptions())
{
    return core::detail::solve(mass, rhs, jac, mgr, x0, 0.0, t_output, opt, false);
}

template <class Mass, class RHS, class Manager = SolutionManager>
inline exit_code::status solve(Mass mass, RHS rhs, const state_vector &x0, const std::vector<double> &t_output, Manager mgr = SolutionManager(), const SolverOptions &opt = SolverOptions())
{
    return core::detail::solve(mass, rhs, JacobianAutomatic(rhs), mgr, x0, 0.0, t_output, opt, true);
}

/*
 * Convenience holder of a DAE system (mass matrix + RHS) with its options, solution and status
 * (dae-cpp/solver.hpp:909-1001).  `sol` receives the solution through a `Solution` manager.
 */
template <class Mass, class RHS>
class System
{
    Mass m_mass;
    RHS m_rhs;

    template <class Jacobian>
    exit_code::status integrate(Jacobian jac, bool automatic, const state_vector &x0, double t_end, const std::vector<double> &t_output)
    {
        if (t_output.empty())
            status = core::detail::solve(m_mass, m_rhs, jac, Solution(sol), x0, t_end, t_output, opt, automatic);
        else
            status = core::detail::solve(m_mass, m_rhs, jac, Solution(sol, t_output), x0, 0.0, t_output, opt, automatic);
        return status;
    }

public:
    SolverOptions opt = SolverOptions();          // Solver options
    SolutionHolder sol;                           // Solution holder
    exit_code::status status{exit_code::unknown}; // Exit code of the last solve()

    System(Mass mass, RHS rhs) : m_mass(mass), m_rhs(rhs) {}

    // Integrates on [0, t_end] with a user-defined Jacobian
    template <class Jacobian>
    exit_code::status solve(const state_vector &x0, const double t_end, Jacobian jac)
    {
        return integrate(jac, false, x0, t_end, {});
    }

    // Integrates on [0, t_end] with the automatic Jacobian
    exit_code::status solve(const state_vector &x0, const double t_end)
    {
        return integrate(JacobianAutomatic(m_rhs), true, x0, t_end, {});
    }

    // Integrates through every output time with a user-defined Jacobian
    template <class Jacobian>
    exit_code::status solve(const state_vector &x0, const std::vector<double> &t_output, Jacobian jac)
    {
        if (t_output.empty())
            status = core::detail::solve(m_mass, m_rhs, jac, Solution(sol, t_output), x0, 0.0, t_output, opt, false);
        else
            integrate(jac, false, x0, 0.0, t_output);
        return status;
    }

    // Integrates through every output time with the automatic Jacobian
    exit_code::status solve(const state_vector &x0, const std::vector<double> &t_output)
    {
        if (t_output.empty())
            status = core::detail::solve(m_mass, m_rhs, JacobianAutomatic(m_rhs), Solution(sol, t_output), x0, 0.0, t_output, opt, true);
        else
            integrate(JacobianAutomatic(m_rhs), true, x0, 0.0, t_output);
        return status;
    }
};

} // namespace daecpp_namespace_name

#endif // DAECPP_SOLVER_H
