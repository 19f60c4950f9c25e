// Driver of the device-functor / lock-step batched path (dae-cpp/ensemble.hpp: solve_ensemble_device).
// The same file is compiled twice (hostcpp/Makefile):
//   _bin/ens_device      nvcc, libdaecpp_b200.so          (the B200 path: equations instantiated as kernels)
//   _bin/ens_device_emu  g++,  tests/emu/libbatchemu.so   (TEST ONLY: host loops + the oracle's arithmetic,
//                                                          checks the host-side lane logic without a GPU)
// Usage: ens_device ensemble|flame1d [key=value ...]
//   ensemble: members (whole sweep), first, count (this process's block), n (perovskite N), t_end, out=<file>
//   flame1d : n, t_end, out=<file>           (one lane: a single system integrated on the device)
// Prints one JSON line.  Output file layout per member: {status, accepted steps, final state[n]} as f64.
#include <dae-cpp/ensemble.hpp>

#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>

#include "problems_device.hpp"

using namespace daecpp;

namespace
{
double arg(const std::map<std::string, std::string> &kv, const char *key, double dflt)
{
    auto it = kv.find(key);
    return it == kv.end() ? dflt : std::atof(it->second.c_str());
}
struct PerovskiteMass
{
    int N;
    void operator()(sparse_matrix &M, const double) const
    {
        M.reserve(N);
        for (int k = 0; k < N; ++k) M(k, k, 1.0);
    }
};
int dump(const std::map<std::string, std::string> &kv, const ensemble_result &R)
{
    auto it = kv.find("out");
    if (it == kv.end()) return 0;
    FILE *fp = std::fopen(it->second.c_str(), "wb");
    if (!fp) { std::perror("out"); return 65; }
    for (const ensemble_member &m : R.members)
    {
        const double head[2] = {(double)m.status, (double)m.steps};
        std::fwrite(head, sizeof(double), 2, fp);
        std::fwrite(m.x.data(), sizeof(double), m.x.size(), fp);
    }
    std::fclose(fp);
    return 0;
}
} // namespace

// In-process entry for bench.py (ctypes): integrates members first, first + stride, ... (count of them) of the C5 sweep
// (perovskite, lambda_m = 0.5 * 4^(m / (members - 1))) in lock step on the device named by DAECPP_B200_DEVICE.
// out: [count][2N + 2] = {status, accepted steps, final state} per member (may be NULL);
// stats[13]: {rounds, lane iterations, lane refactorisations, device ms, wall ms, failed members, accepted steps, checksum,
//             kernels launched, setup ms, wall ms of the round loop, of the final-state download, of the teardown}
extern "C" int ensdev_perovskite(int64_t members, int64_t first, int64_t count, int64_t stride, int32_t N, double t_end, int32_t variability,
                                 double *out, double *stats)
{
    const std::size_t n = 2 * (std::size_t)N;
    std::vector<problems::PerovskiteEq> eqs;
    for (int64_t k = 0; k < count; ++k)
    {   // members first, first + stride, ...: round-robin over the ranks spreads the slow (small lambda) members evenly
        const int64_t m = first + k * stride;
        eqs.push_back(problems::PerovskiteEq::make(N, 1.0, 0.5 * std::pow(4.0, members > 1 ? double(m) / double(members - 1) : 0.0)));
    }
    state_vector x0(n, 0.0);
    for (int k = 0; k < N; ++k) x0[k] = 1.0;
    SolverOptions o;
    o.solution_variability_control = (variability & 1) != 0;
    lockstep_stats st;
    ensemble_result R;
    if (variability & 2) // shape Jacobian by dual numbers (what JacobianMatrixShape does); default: the example's closed-form Jacobian
        R = solve_ensemble_device(eqs, problems::perovskite_shape(N), PerovskiteMass{N}, std::vector<state_vector>{x0}, t_end, o, &st);
    else
    {
        std::vector<problems::PerovskiteEqAnalytic> eqa(eqs.begin(), eqs.end());
        R = solve_ensemble_device(eqa, problems::perovskite_shape(N), PerovskiteMass{N}, std::vector<state_vector>{x0}, t_end, o, &st);
    }
    long steps = 0;
    double checksum = 0.0;
    {   // hand the members to the caller's array (and sum them up) on a few threads: this is harness work, not the solver's
        const std::size_t nm = R.members.size();
        const unsigned nt = (unsigned)std::max<std::size_t>(1, std::min<std::size_t>(4, nm / 64));
        std::vector<double> part(nt, 0.0);
        std::vector<std::thread> pool;
        for (unsigned w = 0; w < nt; ++w)
            pool.emplace_back([&, w] {
                double acc = 0.0;
                for (std::size_t k = w * nm / nt; k < (w + 1) * nm / nt; ++k)
                {
                    const ensemble_member &m = R.members[k];
                    for (double v : m.x) acc += v;
                    if (out)
                    {
                        double *row = out + k * (n + 2);
                        row[0] = (double)m.status; row[1] = (double)m.steps;
                        std::copy(m.x.begin(), m.x.end(), row + 2);
                    }
                }
                part[w] = acc;
            });
        for (std::thread &t : pool) t.join();
        for (double v : part) checksum += v;
        for (const ensemble_member &m : R.members) steps += m.steps;
    }
    if (stats)
    {
        stats[0] = (double)st.rounds; stats[1] = (double)st.lane_iterations; stats[2] = (double)st.lane_refactorisations; stats[3] = st.ms_device;
        stats[4] = st.ms_wall; stats[5] = (double)R.failed; stats[6] = (double)steps; stats[7] = checksum;
        stats[8] = st.kernels_launched; stats[9] = st.ms_setup;
        stats[10] = st.ms_rounds_wall; stats[11] = st.ms_download; stats[12] = st.ms_destroy;
    }
    return R.failed;
}

// In-process entry for bench.py: C3 (flame-1D, n unknowns) integrated to t_end through solve_ensemble_device with ONE lane --
// a single system whose equations and shape Jacobian are evaluated on the device, so that no state or triplet crosses
// PCIe per Newton iteration.  stats as in ensdev_perovskite.
extern "C" int ensdev_flame1d(int32_t n, double t_end, int32_t variability, double *stats)
{
    SolverOptions o;
    o.dt_init = t_end / 100.0;
    o.dt_increase_threshold_delta = -2;
    o.variability_threshold_low = 0.10;
    o.variability_threshold_high = 0.10;
    o.is_mass_matrix_static = true;
    o.solution_variability_control = variability != 0;
    lockstep_stats st;
    std::vector<problems::Flame1DEq> eqs{problems::Flame1DEq::make(n)};
    ensemble_result R = solve_ensemble_device(eqs, problems::flame1d_shape(n), MassMatrixIdentity((std::size_t)n),
                                              std::vector<state_vector>{problems::flame1d_initial(n)}, t_end, o, &st);
    double checksum = 0.0;
    for (double v : R.members[0].x) checksum += v;
    if (stats)
    {
        stats[0] = (double)st.rounds; stats[1] = (double)st.lane_iterations; stats[2] = (double)st.lane_refactorisations; stats[3] = st.ms_device;
        stats[4] = st.ms_wall; stats[5] = (double)R.failed; stats[6] = (double)R.members[0].steps; stats[7] = checksum;
        stats[8] = st.kernels_launched; stats[9] = st.ms_setup;
        stats[10] = st.ms_rounds_wall; stats[11] = st.ms_download; stats[12] = st.ms_destroy;
    }
    return R.failed;
}

#ifndef ENSDEV_NO_MAIN
int main(int argc, char **argv)
{
    if (argc < 2) { std::fprintf(stderr, "usage: ens_device ensemble|flame1d [key=value ...]\n"); return 64; }
    const std::string wl = argv[1];
    std::map<std::string, std::string> kv;
    for (int a = 2; a < argc; ++a)
    {
        const char *eq = std::strchr(argv[a], '=');
        if (eq) kv[std::string(argv[a], eq - argv[a])] = eq + 1;
    }
#if defined(__CUDACC__)
    const char *impl = "b200-lockstep";
#else
    const char *impl = "emulated-lockstep";
#endif
    const int repeat = std::max(1, (int)arg(kv, "repeat", 1));
    ensemble_result R;
    lockstep_stats st;
    std::size_t n = 0, members = 1, first = 0, count = 1;
    double first_ms = 0.0;
    long observed_total = -1;
    for (int rep = 0; rep < repeat; ++rep)
    {
        if (wl == "ensemble")
        {
            members = (std::size_t)arg(kv, "members", 4096); first = (std::size_t)arg(kv, "first", 0);
            count = (std::size_t)arg(kv, "count", (double)(members - first));
            const int N = (int)arg(kv, "n", 4000);
            n = 2 * (std::size_t)N;
            std::vector<problems::PerovskiteEq> eqs;
            for (std::size_t m = first; m < first + count; ++m)
                eqs.push_back(problems::PerovskiteEq::make(N, 1.0, 0.5 * std::pow(4.0, members > 1 ? double(m) / double(members - 1) : 0.0)));
            state_vector x0(n, 0.0);
            for (int k = 0; k < N; ++k) x0[k] = 1.0;
            SolverOptions o;
            o.solution_variability_control = arg(kv, "variability", 0) != 0;
            if (arg(kv, "outputs", 0) > 0)
            {   // observer path: every member's state at K output times; the first member's go to observed=<file> as {t, x} rows
                const int K = (int)arg(kv, "outputs", 0);
                const double t_end = arg(kv, "t_end", 10.0);
                std::vector<double> t_output;
                for (int k = 1; k <= K; ++k) t_output.push_back(t_end * double(k) / double(K));
                std::vector<SolutionHolder> sols(count);
                std::vector<problems::PerovskiteEqAnalytic> eqa(eqs.begin(), eqs.end());
                R = solve_ensemble_device_observed(eqa, problems::perovskite_shape(N), PerovskiteMass{N}, std::vector<state_vector>{x0}, t_output,
                                                   [&](std::size_t m, const state_vector &x, double t) { sols[m].x.push_back(x); sols[m].t.push_back(t); return 0; }, o, &st);
                observed_total = 0;
                for (const SolutionHolder &h : sols) observed_total += (long)h.t.size();
                auto ob = kv.find("observed");
                if (ob != kv.end())
                {
                    FILE *fp = std::fopen(ob->second.c_str(), "wb");
                    if (!fp) { std::perror("observed"); return 65; }
                    for (std::size_t k = 0; k < sols[0].t.size(); ++k)
                    {
                        std::fwrite(&sols[0].t[k], sizeof(double), 1, fp);
                        std::fwrite(sols[0].x[k].data(), sizeof(double), sols[0].x[k].size(), fp);
                    }
                    std::fclose(fp);
                }
            }
            else if (arg(kv, "dual", 0) != 0) // shape Jacobian by dual numbers; default: the example's closed-form Jacobian
                R = solve_ensemble_device(eqs, problems::perovskite_shape(N), PerovskiteMass{N}, std::vector<state_vector>{x0}, arg(kv, "t_end", 10.0), o, &st);
            else
            {
                std::vector<problems::PerovskiteEqAnalytic> eqa(eqs.begin(), eqs.end());
                R = solve_ensemble_device(eqa, problems::perovskite_shape(N), PerovskiteMass{N}, std::vector<state_vector>{x0}, arg(kv, "t_end", 10.0), o, &st);
            }
        }
        else if (wl == "flame1d")
        {
            const int nn = (int)arg(kv, "n", 1000000);
            n = (std::size_t)nn;
            const double t_end = arg(kv, "t_end", 50.0);
            SolverOptions o;
            o.dt_init = arg(kv, "dt_init", t_end / 100.0);
            o.dt_increase_threshold_delta = -2;
            o.variability_threshold_low = 0.10;
            o.variability_threshold_high = 0.10;
            o.is_mass_matrix_static = true;
            o.solution_variability_control = arg(kv, "variability", 1) != 0;
            std::vector<problems::Flame1DEq> eqs{problems::Flame1DEq::make(nn)};
            R = solve_ensemble_device(eqs, problems::flame1d_shape(nn), MassMatrixIdentity(n), std::vector<state_vector>{problems::flame1d_initial(nn)}, t_end, o, &st);
        }
        else { std::fprintf(stderr, "unknown workload %s\n", wl.c_str()); return 64; }
        if (rep == 0) first_ms = st.ms_wall;
    }
    long steps = 0;
    double checksum = 0.0;
    for (const ensemble_member &m : R.members)
    {
        steps += m.steps;
        for (double v : m.x) checksum += v;
    }
    if (int rc = dump(kv, R)) return rc;
    std::printf("{\"impl\": \"%s\", \"workload\": \"%s\", \"n\": %zu, \"members\": %zu, \"first\": %zu, \"count\": %zu, \"failed\": %d, "
                "\"steps\": %ld, \"rounds\": %ld, \"lane_iterations\": %ld, \"lane_refactorisations\": %ld, \"wall_ms\": %.3f, "
                "\"first_wall_ms\": %.3f, \"refine_rounds\": %.0f, \"refined_lanes\": %.0f, \"device_ms\": %.3f, \"systems_per_s\": %.6f, \"status0\": %d, \"observed\": %ld, \"observed_downloads\": %ld, \"checksum\": %.17g}\n",
                impl, wl.c_str(), n, members, first, count, R.failed, steps, st.rounds, st.lane_iterations, st.lane_refactorisations, st.ms_wall,
                first_ms, st.refine_rounds, st.refined_lanes, st.ms_device, 1e3 * double(count) / st.ms_wall, R.members.empty() ? -1 : R.members[0].status, observed_total, st.observed_states, checksum);
    return R.failed ? 1 : 0;
}
#endif // ENSDEV_NO_MAIN
