// Trajectory driver for the BASELINE configurations, written against the public dae-cpp API only
// (daecpp::solve, sparse_matrix, state_type, JacobianMatrixShape, SolverOptions).  The same file
// is compiled twice (hostcpp/Makefile):
//   _bin/traj_dump      overlay solver.hpp + libdaecpp_b200.so  (the B200 path)
//   _bin/traj_dump_ref  the unmodified reference headers        (Eigen::SparseLU on the host)
// so that step counts, final states and wall times can be compared run against run.
//
// Usage: traj_dump <workload> [key=value ...]
//   workloads: flame | perovskite | flame1d | grid2d | ensemble
//   ensemble keys: members (size of the whole sweep), first, count (this process's block), threads, n
//   keys: n, lambda, t_end, out=<file for the final state, raw f64>, shape=0|1, repeat
// Prints one JSON line.
#include <dae-cpp/solver.hpp>
#include <dae-cpp/ensemble.hpp>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>

using namespace daecpp;

namespace
{

struct Tally
{
    long rhs = 0, jac = 0, mass = 0, mgr = 0;
    double ms_rhs = 0, ms_jac = 0;
    state_vector last;
    double t_last = 0;
};

struct Stopwatch
{
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    double ms() const { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
};

struct Observer
{
    Tally *tally;
    int operator()(const state_vector &x, const double t)
    {
        tally->mgr++;
        tally->last = x;
        tally->t_last = t;
        return 0;
    }
};

// ---- C1: examples/flame_propagation, scalar y' = y^2 - y^3 ------------------------------------
struct FlameScalarRHS
{
    Tally *tally;
    void operator()(state_type &f, const state_type &x, const double) const
    {
        tally->rhs++;
        const state_value y = x[0];
        f[0] = y * y - y * y * y;
    }
};

// ---- C2 / C5 member: examples/perovskite_model (Poisson-Nernst-Planck, singular mass matrix) ---
struct Perovskite
{
    int_type N;
    double L, lambda;
    double inv_h2() const { const double h = L / double(N - 1); return 1.0 / (h * h); }
    double inv_l2() const { return 1.0 / (lambda * lambda); }
};
struct PerovskiteMass
{
    Perovskite p; Tally *tally;
    void operator()(sparse_matrix &M, const double) const
    {
        tally->mass++;
        M.reserve(p.N);
        for (int_type k = 0; k < p.N; ++k) M(k, k, 1.0);
    }
};
struct PerovskiteRHS
{
    Perovskite p; Tally *tally;
    void operator()(state_type &f, const state_type &x, const double t) const
    {
        tally->rhs++;
        Stopwatch sw;
        const int_type N = p.N;
        const double ih2 = p.inv_h2(), il2 = p.inv_l2();
        // ion concentration: divergence of the drift-diffusion flux
        f[0] = (x[1] - x[0] + 0.5 * (x[1] + x[0]) * (x[N + 1] - x[N])) * ih2;
        for (int_type k = 1; k < N - 1; ++k)
            f[k] = (x[k + 1] - 2.0 * x[k] + x[k - 1] +
                    0.5 * ((x[k + 1] + x[k]) * (x[N + k + 1] - x[N + k]) - (x[k] + x[k - 1]) * (x[N + k] - x[N + k - 1]))) * ih2;
        f[N - 1] = -(x[N - 1] - x[N - 2] + 0.5 * (x[N - 1] + x[N - 2]) * (x[2 * N - 1] - x[2 * N - 2])) * ih2;
        // potential: Poisson equation with Dirichlet ends -t / +t
        f[N] = x[N] + t;
        for (int_type k = 1; k < N - 1; ++k)
            f[N + k] = (x[N + k + 1] - 2.0 * x[N + k] + x[N + k - 1]) * ih2 - (1.0 - x[k]) * il2;
        f[2 * N - 1] = x[2 * N - 1] - t;
        tally->ms_rhs += sw.ms();
    }
};
struct PerovskiteElements : VectorFunctionElements // the same equations, one at a time, for JacobianMatrixShape
{
    Perovskite p; Tally *tally;
    PerovskiteElements(Perovskite p_, Tally *t_) : p(p_), tally(t_) {}
    state_value equations(const state_type &x, const double t, const int_type k) const override
    {
        const int_type N = p.N;
        const double ih2 = p.inv_h2(), il2 = p.inv_l2();
        if (k == 0) return (x[1] - x[0] + 0.5 * (x[1] + x[0]) * (x[N + 1] - x[N])) * ih2;
        if (k < N - 1)
            return (x[k + 1] - 2.0 * x[k] + x[k - 1] +
                    0.5 * ((x[k + 1] + x[k]) * (x[N + k + 1] - x[N + k]) - (x[k] + x[k - 1]) * (x[N + k] - x[N + k - 1]))) * ih2;
        if (k == N - 1) return -(x[N - 1] - x[N - 2] + 0.5 * (x[N - 1] + x[N - 2]) * (x[2 * N - 1] - x[2 * N - 2])) * ih2;
        if (k == N) return x[N] + t;
        if (k < 2 * N - 1) return (x[k + 1] - 2.0 * x[k] + x[k - 1]) * ih2 - (1.0 - x[k - N]) * il2;
        return x[2 * N - 1] - t;
    }
};
struct PerovskiteJac
{
    Perovskite p; Tally *tally;
    void operator()(sparse_matrix &J, const state_vector &x, const double) const
    {
        tally->jac++;
        Stopwatch sw;
        const int_type N = p.N;
        const double ih2 = p.inv_h2(), il2 = p.inv_l2();
        J.reserve(12 * N);
        for (int_type r = 0; r < N; ++r)
        {
            if (r == 0)
            {
                const double dphi = x[N + 1] - x[N], ps = x[0] + x[1];
                J(r, 0, (-1.0 + 0.5 * dphi) * ih2);
                J(r, 1, (1.0 + 0.5 * dphi) * ih2);
                J(r, N, -0.5 * ps * ih2);
                J(r, N + 1, 0.5 * ps * ih2);
            }
            else if (r < N - 1)
            {
                J(r, r - 1, (1.0 - 0.5 * (x[N + r] - x[N + r - 1])) * ih2);
                J(r, r, (-2.0 + 0.5 * (x[N + r + 1] - 2.0 * x[N + r] + x[N + r - 1])) * ih2);
                J(r, r + 1, (1.0 + 0.5 * (x[N + r + 1] - x[N + r])) * ih2);
                J(r, N + r - 1, 0.5 * (x[r] + x[r - 1]) * ih2);
                J(r, N + r, -0.5 * (x[r + 1] + 2.0 * x[r] + x[r - 1]) * ih2);
                J(r, N + r + 1, 0.5 * (x[r + 1] + x[r]) * ih2);
            }
            else
            {
                const double dphi = x[2 * N - 1] - x[2 * N - 2], ps = x[N - 1] + x[N - 2];
                J(r, r - 1, (1.0 - 0.5 * dphi) * ih2);
                J(r, r, (-1.0 - 0.5 * dphi) * ih2);
                J(r, N + r - 1, 0.5 * ps * ih2);
                J(r, N + r, -0.5 * ps * ih2);
            }
        }
        J(N, N, 1.0);
        for (int_type r = N + 1; r < 2 * N - 1; ++r)
        {
            J(r, r - N, il2);
            J(r, r - 1, ih2);
            J(r, r, -2.0 * ih2);
            J(r, r + 1, ih2);
        }
        J(2 * N - 1, 2 * N - 1, 1.0);
        tally->ms_jac += sw.ms();
    }
};

// ---- C3: 1-D flame front u_t = D u_xx + u^2 - u^3, zero-flux ends -------------
struct Flame1D
{
    int_type n;
    double D = 1e-3;
    double inv_h2() const { const double h = 1.0 / double(n - 1); return 1.0 / (h * h); }
};
struct Flame1DRHS
{
    Flame1D p; Tally *tally;
    void operator()(state_type &f, const state_type &u, const double) const
    {
        tally->rhs++;
        Stopwatch sw;
        const int_type n = p.n;
        const double c = p.D * p.inv_h2();
        f[0] = c * (u[1] - u[0]) + u[0] * u[0] - u[0] * u[0] * u[0];
        for (int_type k = 1; k < n - 1; ++k)
            f[k] = c * (u[k + 1] - 2.0 * u[k] + u[k - 1]) + u[k] * u[k] - u[k] * u[k] * u[k];
        f[n - 1] = c * (u[n - 2] - u[n - 1]) + u[n - 1] * u[n - 1] - u[n - 1] * u[n - 1] * u[n - 1];
        tally->ms_rhs += sw.ms();
    }
};
struct Flame1DElements : VectorFunctionElements // RHS as equations(x,t,i) + JacobianMatrixShape
{
    Flame1D p;
    explicit Flame1DElements(Flame1D p_) : p(p_) {}
    state_value equations(const state_type &u, const double, const int_type k) const override
    {
        const int_type n = p.n;
        const double c = p.D * p.inv_h2();
        const state_value react = u[k] * u[k] - u[k] * u[k] * u[k];
        if (k == 0) return c * (u[1] - u[0]) + react;
        if (k == n - 1) return c * (u[n - 2] - u[n - 1]) + react;
        return c * (u[k + 1] - 2.0 * u[k] + u[k - 1]) + react;
    }
};
struct Flame1DJac // analytic tridiagonal Jacobian, row by row
{
    Flame1D p; Tally *tally;
    void operator()(sparse_matrix &J, const state_vector &u, const double) const
    {
        tally->jac++;
        Stopwatch sw;
        const int_type n = p.n;
        const double c = p.D * p.inv_h2();
        J.reserve(3 * n);
        for (int_type k = 0; k < n; ++k)
        {
            const double react = 2.0 * u[k] - 3.0 * u[k] * u[k];
            if (k > 0) J(k, k - 1, c);
            J(k, k, ((k == 0 || k == n - 1) ? -c : -2.0 * c) + react);
            if (k < n - 1) J(k, k + 1, c);
        }
        tally->ms_jac += sw.ms();
    }
};

// ---- C4: 2-D u_t = Lap(u) + u^2 - u^3 on an m x m grid, 5-point stencil, Neumann edges ---------
struct Grid2D
{
    int_type m;
    double inv_h2() const { const double h = 1.0 / double(m - 1); return 1.0 / (h * h); }
    double D = 1e-4;
};
struct Grid2DRHS
{
    Grid2D p; Tally *tally;
    void operator()(state_type &f, const state_type &u, const double) const
    {
        tally->rhs++;
        Stopwatch sw;
        const int_type m = p.m;
        const double c = p.D * p.inv_h2();
        for (int_type r = 0; r < m; ++r)
            for (int_type q = 0; q < m; ++q)
            {
                const int_type k = r * m + q;
                state_value lap = 0.0;
                if (r > 0) lap += u[k - m] - u[k];
                if (q > 0) lap += u[k - 1] - u[k];
                if (q < m - 1) lap += u[k + 1] - u[k];
                if (r < m - 1) lap += u[k + m] - u[k];
                f[k] = c * lap + u[k] * u[k] - u[k] * u[k] * u[k];
            }
        tally->ms_rhs += sw.ms();
    }
};
struct Grid2DJac
{
    Grid2D p; Tally *tally;
    void operator()(sparse_matrix &J, const state_vector &u, const double) const
    {
        tally->jac++;
        Stopwatch sw;
        const int_type m = p.m;
        const double c = p.D * p.inv_h2();
        J.reserve(5 * m * m);
        for (int_type r = 0; r < m; ++r)
            for (int_type q = 0; q < m; ++q)
            {
                const int_type k = r * m + q;
                const int nb = (r > 0) + (q > 0) + (q < m - 1) + (r < m - 1);
                if (r > 0) J(k, k - m, c);
                if (q > 0) J(k, k - 1, c);
                J(k, k, -c * nb + 2.0 * u[k] - 3.0 * u[k] * u[k]);
                if (q < m - 1) J(k, k + 1, c);
                if (r < m - 1) J(k, k + m, c);
            }
        tally->ms_jac += sw.ms();
    }
};

double arg(const std::map<std::string, std::string> &kv, const char *key, double dflt)
{
    auto it = kv.find(key);
    return it == kv.end() ? dflt : std::atof(it->second.c_str());
}

} // namespace

int main(int argc, char **argv)
{
    if (argc < 2)
    {
        std::fprintf(stderr, "usage: traj_dump flame|perovskite|flame1d|grid2d [key=value ...]\n");
        return 64;
    }
    const std::string wl = argv[1];
    std::map<std::string, std::string> kv;
    for (int a = 2; a < argc; ++a)
    {
        const char *eq = std::strchr(argv[a], '=');
        if (eq) kv[std::string(argv[a], eq - argv[a])] = eq + 1;
    }
    Tally tally;
    SolverOptions opt;
    int status = -1;
    std::size_t n = 0;
    double t_end = 0;
    Stopwatch wall;
    // repeat=R: the same solve R times in this process, the LAST one reported (wall_ms) next to the
    // first (first_wall_ms, which carries the one-off costs of the process: CUDA context, module load)
    const int repeat = std::max(1, (int)arg(kv, "repeat", 1));
    double first_ms = 0.0;
    for (int rep = 0; rep < repeat; ++rep)
    {
    tally = Tally{};
    opt = SolverOptions{};
    wall = Stopwatch{};

    if (wl == "flame")
    {
        const double delta = arg(kv, "delta", 1e-4);
        t_end = arg(kv, "t_end", 2.0 / delta);
        opt.dt_init = t_end / 100.0;
        opt.dt_increase_threshold_delta = -2;
        opt.variability_threshold_low = 0.10;
        opt.variability_threshold_high = 0.10;
        n = 1;
        status = solve(MassMatrixIdentity(1), FlameScalarRHS{&tally}, state_vector{delta}, t_end, Observer{&tally}, opt);
    }
    else if (wl == "perovskite")
    {
        Perovskite p{(int_type)arg(kv, "n", 4000), 1.0, arg(kv, "lambda", 1.0)};
        n = 2 * (std::size_t)p.N;
        t_end = arg(kv, "t_end", 10.0);
        state_vector x0(n, 0.0);
        for (int_type k = 0; k < p.N; ++k) x0[k] = 1.0;
        opt.solution_variability_control = arg(kv, "variability", 0) != 0;
        if (arg(kv, "shape", 0) != 0)
        {
            PerovskiteElements rhs(p, &tally);
            JacobianMatrixShape<PerovskiteElements> jac(rhs);
            const int_type N = p.N;
            for (int_type r = 0; r < N; ++r)
            {
                std::vector<int_type> cols;
                for (int_type c : {r - 1, r, r + 1}) if (c >= 0 && c < N) { cols.push_back(c); }
                for (int_type c : {r - 1, r, r + 1}) if (c >= 0 && c < N) { cols.push_back(N + c); }
                jac.add_element(r, cols);
            }
            jac.add_element(N, N);
            for (int_type r = N + 1; r < 2 * N - 1; ++r) jac.add_element(r, std::vector<int_type>{r - N, r - 1, r, r + 1});
            jac.add_element(2 * N - 1, 2 * N - 1);
            status = solve(PerovskiteMass{p, &tally}, rhs, jac, x0, t_end, Observer{&tally}, opt);
        }
        else if (arg(kv, "outputs", 0) > 0)
        {   // daecpp::Solution with output times: what the lock-step driver's observer path is checked against
            const int K = (int)arg(kv, "outputs", 0);
            std::vector<double> t_output;
            for (int k = 1; k <= K; ++k) t_output.push_back(t_end * double(k) / double(K));
            SolutionHolder sol;
            status = solve(PerovskiteMass{p, &tally}, PerovskiteRHS{p, &tally}, PerovskiteJac{p, &tally}, x0, t_output, Solution(sol, t_output), opt);
            tally.mgr = (long)sol.t.size() + 1;
            if (!sol.x.empty()) { tally.last = sol.x.back(); tally.t_last = sol.t.back(); }
            auto ob = kv.find("observed");
            if (ob != kv.end())
            {
                FILE *fp = std::fopen(ob->second.c_str(), "wb");
                if (!fp) { std::perror("observed"); return 65; }
                for (std::size_t k = 0; k < sol.t.size(); ++k)
                {
                    std::fwrite(&sol.t[k], sizeof(double), 1, fp);
                    std::fwrite(sol.x[k].data(), sizeof(double), sol.x[k].size(), fp);
                }
                std::fclose(fp);
            }
        }
        else
            status = solve(PerovskiteMass{p, &tally}, PerovskiteRHS{p, &tally}, PerovskiteJac{p, &tally}, x0, t_end, Observer{&tally}, opt);
    }
    else if (wl == "flame1d")
    {
        Flame1D p{(int_type)arg(kv, "n", 1000000)};
        n = (std::size_t)p.n;
        t_end = arg(kv, "t_end", 50.0);
        const double delta = 1e-4;
        state_vector u0(n);
        for (std::size_t k = 0; k < n; ++k)
        {
            const double x = double(k) / double(n - 1);
            u0[k] = delta + (1.0 - delta) * 0.5 * (1.0 - std::tanh((x - 0.1) / 0.02));
        }
        opt.dt_init = arg(kv, "dt_init", t_end / 100.0);
        opt.dt_increase_threshold_delta = -2;
        opt.variability_threshold_low = 0.10;
        opt.variability_threshold_high = 0.10;
        opt.is_mass_matrix_static = true;
        opt.solution_variability_control = arg(kv, "variability", 1) != 0;
        Flame1DRHS rhs{p, &tally};
        if (arg(kv, "shape", 0) != 0)
        {
            Flame1DElements erhs(p);
            JacobianMatrixShape<Flame1DElements> jac(erhs);
            for (int_type k = 0; k < p.n; ++k)
            {
                std::vector<int_type> cols;
                for (int_type c : {k - 1, k, k + 1}) if (c >= 0 && c < p.n) { cols.push_back(c); }
                jac.add_element(k, cols);
            }
            status = solve(MassMatrixIdentity(n), erhs, jac, u0, t_end, Observer{&tally}, opt);
        }
        else
            status = solve(MassMatrixIdentity(n), rhs, Flame1DJac{p, &tally}, u0, t_end, Observer{&tally}, opt);
    }
    else if (wl == "grid2d")
    {
        Grid2D p{(int_type)arg(kv, "n", 1024)};
        n = (std::size_t)p.m * p.m;
        t_end = arg(kv, "t_end", 5.0);
        state_vector u0(n);
        for (int_type r = 0; r < p.m; ++r)
            for (int_type q = 0; q < p.m; ++q)
            {
                const double x = double(q) / double(p.m - 1) - 0.5, y = double(r) / double(p.m - 1) - 0.5;
                u0[(std::size_t)r * p.m + q] = (x * x + y * y < 0.01) ? 1.0 : 1e-4;
            }
        opt.dt_init = arg(kv, "dt_init", t_end / 100.0);
        opt.is_mass_matrix_static = true;
        opt.solution_variability_control = arg(kv, "variability", 1) != 0;
        status = solve(MassMatrixIdentity(n), Grid2DRHS{p, &tally}, Grid2DJac{p, &tally}, u0, t_end, Observer{&tally}, opt);
    }
    else if (wl == "ensemble")
    {
        // C5: perovskite members with lambda_m = 0.5 * 4^(m/(members-1)), same pattern,
        // different values and different adaptive step sequences; this process integrates the block
        // [first, first+count) with `threads` members in flight.
        const std::size_t members = (std::size_t)arg(kv, "members", 4096), first = (std::size_t)arg(kv, "first", 0);
        const std::size_t count = (std::size_t)arg(kv, "count", (double)(members - first));
        const unsigned threads = (unsigned)arg(kv, "threads", 0);
        const int_type N = (int_type)arg(kv, "n", 4000);
        n = 2 * (std::size_t)N;
        t_end = arg(kv, "t_end", 10.0);
        std::vector<Tally> tallies(count);
        Stopwatch sw;
        auto result = solve_ensemble(first, count, [&](std::size_t m) {
            Tally &tl = tallies[m - first];
            const double lam = 0.5 * std::pow(4.0, members > 1 ? double(m) / double(members - 1) : 0.0);
            Perovskite p{N, 1.0, lam};
            state_vector x0(n, 0.0);
            for (int_type k = 0; k < N; ++k) x0[k] = 1.0;
            SolverOptions o;
            o.solution_variability_control = false;
            ensemble_capture cap;
            const int st = solve(PerovskiteMass{p, &tl}, PerovskiteRHS{p, &tl}, PerovskiteJac{p, &tl}, x0, t_end, cap.observer(), o);
            return ensemble_member{st, cap.steps(), std::move(cap.x)};
        }, threads);
        const double ms_ens = sw.ms();
        long steps = 0, rhs_calls = 0, jac_calls = 0;
        double checksum = 0.0;
        for (std::size_t k = 0; k < count; ++k)
        {
            steps += result.members[k].steps;
            rhs_calls += tallies[k].rhs; jac_calls += tallies[k].jac;
            for (double v : result.members[k].x) checksum += v;
        }
        auto it = kv.find("out");
        if (it != kv.end())
        {
            FILE *fp = std::fopen(it->second.c_str(), "wb");
            if (!fp) { std::perror("out"); return 65; }
            for (std::size_t k = 0; k < count; ++k)
            {
                const double head[2] = {(double)result.members[k].status, (double)result.members[k].steps};
                std::fwrite(head, sizeof(double), 2, fp);
                std::fwrite(result.members[k].x.data(), sizeof(double), result.members[k].x.size(), fp);
            }
            std::fclose(fp);
        }
#ifdef TRAJ_REFERENCE
        const char *impl_e = "reference";
#else
        const char *impl_e = "b200";
#endif
        std::printf("{\"impl\": \"%s\", \"workload\": \"ensemble\", \"n\": %zu, \"members\": %zu, \"first\": %zu, \"count\": %zu, "
                    "\"threads\": %u, \"failed\": %d, \"steps\": %ld, \"rhs_calls\": %ld, \"jac_calls\": %ld, \"wall_ms\": %.3f, "
                    "\"systems_per_s\": %.6f, \"checksum\": %.17g}\n",
                    impl_e, n, members, first, count, threads ? threads : std::thread::hardware_concurrency(), result.failed, steps,
                    rhs_calls, jac_calls, ms_ens, 1e3 * double(count) / ms_ens, checksum);
        return result.failed ? 1 : 0;
    }
    else
    {
        std::fprintf(stderr, "unknown workload %s\n", wl.c_str());
        return 64;
    }
    if (rep == 0) first_ms = wall.ms();
    } // repeat
    const double ms = wall.ms();

    double s1 = 0, s2 = 0, amax = 0;
    for (double v : tally.last) { s1 += v; s2 += v * v; amax = std::max(amax, std::fabs(v)); }
    auto it = kv.find("out");
    if (it != kv.end())
    {
        FILE *fp = std::fopen(it->second.c_str(), "wb");
        if (!fp) { std::perror("out"); return 65; }
        std::fwrite(tally.last.data(), sizeof(double), tally.last.size(), fp);
        std::fclose(fp);
    }
#ifdef TRAJ_REFERENCE
    const char *impl = "reference";
#else
    const char *impl = "b200";
#endif
    std::printf("{\"impl\": \"%s\", \"workload\": \"%s\", \"n\": %zu, \"status\": %d, \"steps\": %ld, \"t_final\": %.17g, "
                "\"rhs_calls\": %ld, \"jac_calls\": %ld, \"mass_calls\": %ld, \"wall_ms\": %.3f, \"first_wall_ms\": %.3f, \"repeat\": %d, \"ms_rhs_functor\": %.3f, "
                "\"ms_jac_functor\": %.3f, \"sum\": %.17g, \"sum_sq\": %.17g, \"max_abs\": %.17g, \"first\": %.17g, \"last\": %.17g}\n",
                impl, wl.c_str(), n, status, tally.mgr - 1, tally.t_last, tally.rhs, tally.jac, tally.mass, ms, first_ms, repeat, tally.ms_rhs,
                tally.ms_jac, s1, s2, amax, tally.last.empty() ? 0.0 : tally.last.front(), tally.last.empty() ? 0.0 : tally.last.back());
    return status;
}
