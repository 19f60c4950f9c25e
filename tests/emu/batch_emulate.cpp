// TEST INFRASTRUCTURE ONLY.  A CPU stand-in for the lock-step batch entry points of include/daecpp_b200.h
// (daecpp_b200_batch_*), built on the oracle's plain-C arithmetic (oracle/oracle.c), so that the HOST side of
// the batched ensemble -- dae-cpp/ensemble.hpp: lane_control, solve_ensemble_device; dae-cpp/device-functor.hpp
// in its host-loop mode -- can be exercised against the unmodified reference in the GPU-less container.
// Never linked into libdaecpp_b200.so; the product path has no CPU fallback.
#include <cmath>
#include <cstring>
#include <string>
#include <vector>

#include "../../include/daecpp_b200.h"

extern "C" {
int64_t oracle_assemble_jb(int32_t n, int64_t nJ, const double *JA, const int32_t *Ji, const int32_t *Jj, int64_t nM, const double *MA,
                           const int32_t *Mi, const int32_t *Mj, double alpha, int32_t *outer, int32_t *inner, double *vals);
int64_t oracle_convert(int32_t n, int64_t nT, const double *A, const int32_t *ti, const int32_t *tj, int32_t *outer, int32_t *inner, double *vals);
double oracle_time_derivative(int32_t n, int32_t order, const double *dt, const double *xk, const double *xh, double *dxdt);
void oracle_residual(int32_t n, const int32_t *mo, const int32_t *mi, const double *mv, const double *dxdt, const double *f, double *b);
int oracle_lu_solve(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals, const double *b, double *x);
}

struct daecpp_b200_batch
{
    int32_t n, lanes;
    std::vector<int32_t> Ji, Jj, Mi, Mj;
    std::vector<double> JA, MA; // [lanes][nT]
    std::vector<double> xk, f, b, dx, xh; // xh: [lanes][4][n]
    std::vector<std::vector<int32_t>> outer, inner; // factor "storage": the assembled Jb of every lane
    std::vector<std::vector<double>> vals;
    daecpp_b200_device_functors fn{};
    bool have_fn = false;
    double rounds = 0;
    std::string err;
};

extern "C" {

int daecpp_b200_batch_create(int32_t n, int32_t lanes, int, daecpp_b200_batch **out)
{
    if (!out || n <= 0 || lanes <= 0) return DAECPP_B200_EINVAL;
    daecpp_b200_batch *b = new daecpp_b200_batch();
    b->n = n; b->lanes = lanes;
    const size_t per = (size_t)n * lanes;
    b->xk.assign(per, 0); b->f.assign(per, 0); b->b.assign(per, 0); b->dx.assign(per, 0); b->xh.assign(4 * per, 0);
    b->outer.resize(lanes); b->inner.resize(lanes); b->vals.resize(lanes);
    *out = b;
    return 0;
}
void daecpp_b200_batch_destroy(daecpp_b200_batch *b) { delete b; }
const char *daecpp_b200_batch_last_error(const daecpp_b200_batch *b) { return b ? b->err.c_str() : ""; }

int daecpp_b200_batch_set_pattern(daecpp_b200_batch *b, const int32_t *Ji, const int32_t *Jj, int64_t nTJ, const int32_t *Mi, const int32_t *Mj, int64_t nTM)
{
    b->Ji.assign(Ji, Ji + nTJ); b->Jj.assign(Jj, Jj + nTJ);
    b->Mi.assign(Mi, Mi + nTM); b->Mj.assign(Mj, Mj + nTM);
    b->JA.assign((size_t)nTJ * b->lanes, 0.0); b->MA.assign((size_t)nTM * b->lanes, 0.0);
    return 0;
}
int daecpp_b200_batch_set_mass_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count)
{
    const size_t nT = b->Mi.size();
    if (count == 0) { for (int32_t l = 0; l < b->lanes; ++l) std::memcpy(b->MA.data() + l * nT, A, sizeof(double) * nT); return 0; }
    std::memcpy(b->MA.data() + (size_t)first * nT, A, sizeof(double) * nT * count);
    return 0;
}
int daecpp_b200_batch_set_jacobian_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count)
{
    const size_t nT = b->Ji.size();
    std::memcpy(b->JA.data() + (size_t)first * nT, A, sizeof(double) * nT * count);
    return 0;
}
int daecpp_b200_batch_set_rhs_values(daecpp_b200_batch *b, const double *f, int32_t first, int32_t count)
{
    std::memcpy(b->f.data() + (size_t)first * b->n, f, sizeof(double) * (size_t)b->n * count);
    return 0;
}
int daecpp_b200_batch_set_state(daecpp_b200_batch *b, const double *x0, int32_t first, int32_t count)
{
    const size_t n = b->n;
    const bool shared = count == 0;
    if (shared) { first = 0; count = b->lanes; }
    for (int32_t l = first; l < first + count; ++l)
    {
        const double *src = x0 + (shared ? 0 : (size_t)(l - first) * n);
        std::memcpy(b->xk.data() + l * n, src, sizeof(double) * n);
        std::memset(b->xh.data() + (size_t)l * 4 * n, 0, sizeof(double) * 4 * n);
        std::memcpy(b->xh.data() + (size_t)l * 4 * n, src, sizeof(double) * n);
    }
    return 0;
}
int daecpp_b200_batch_get_vector(daecpp_b200_batch *b, int32_t which, double *out, int32_t first, int32_t count)
{
    const size_t n = b->n;
    for (int32_t l = first; l < first + count; ++l)
    {
        const double *src = which == 0 ? b->b.data() + l * n : which == 1 ? b->dx.data() + l * n : which == 3 ? b->xk.data() + l * n
                                                                                              : b->xh.data() + ((size_t)l * 4 + (which - 4)) * n;
        std::memcpy(out + (size_t)(l - first) * n, src, sizeof(double) * n);
    }
    return 0;
}
int daecpp_b200_batch_set_device_functors(daecpp_b200_batch *b, const daecpp_b200_device_functors *fn)
{
    b->have_fn = fn && fn->rhs && fn->jacobian;
    if (b->have_fn) b->fn = *fn;
    return 0;
}

int daecpp_b200_batch_round(daecpp_b200_batch *b, const daecpp_b200_lane_cmd *cmd, double atol, double rtol, double max_err_abs,
                            int32_t variability_on, double var_tol, daecpp_b200_lane_result *res)
{
    const int32_t n = b->n, L = b->lanes;
    std::vector<int32_t> mode(L), running(L);
    std::vector<double> t(L);
    int K = 1;
    for (int32_t l = 0; l < L; ++l)
    {
        t[l] = cmd[l].t;
        running[l] = cmd[l].mode != DAECPP_B200_LANE_IDLE;
        res[l] = daecpp_b200_lane_result{0, 0, 0.0, 0, 0};
        if (running[l] && cmd[l].max_iterations > K) K = cmd[l].max_iterations < DAECPP_B200_MAX_ROUND_ITERATIONS ? cmd[l].max_iterations : DAECPP_B200_MAX_ROUND_ITERATIONS;
        double *xk = b->xk.data() + (size_t)l * n, *xh = b->xh.data() + (size_t)l * 4 * n;
        if (cmd[l].pre == DAECPP_B200_LANE_ACCEPT)
        {
            std::memmove(xh + n, xh, sizeof(double) * 3 * n);
            std::memcpy(xh, xk, sizeof(double) * n);
        }
        else if (cmd[l].pre == DAECPP_B200_LANE_ROLLBACK) std::memcpy(xk, xh, sizeof(double) * n);
    }
    const int64_t nTJ = (int64_t)b->Ji.size(), nTM = (int64_t)b->Mi.size();
    std::vector<double> dxdt(n);
    std::vector<int32_t> mo(n + 1), mi(nTM ? nTM : 1);
    std::vector<double> mv(nTM ? nTM : 1);
    for (int j = 0; j < K; ++j)
    {   // iteration j of the round: the lanes whose loop goes on (include/daecpp_b200.h: daecpp_b200_lane_cmd)
        for (int32_t l = 0; l < L; ++l)
        {
            mode[l] = DAECPP_B200_LANE_IDLE;
            if (!running[l] || j >= (cmd[l].max_iterations > 1 ? cmd[l].max_iterations : 1)) { running[l] = 0; continue; }
            mode[l] = j == 0 ? cmd[l].mode : ((cmd[l].period <= 1 || (cmd[l].phase + j) % cmd[l].period == 0) ? DAECPP_B200_LANE_REFACTOR : DAECPP_B200_LANE_SOLVE);
        }
        if (b->have_fn)
        {   // the functor callbacks, on host pointers (dae-cpp/device-functor.hpp compiled without nvcc)
            daecpp_b200_functor_args a{};
            a.n = n; a.lanes = L; a.x_stride = n; a.x = b->xk.data(); a.t = t.data(); a.mode = mode.data(); a.stream = nullptr;
            a.min_mode = DAECPP_B200_LANE_SOLVE; a.out = b->f.data(); a.out_stride = n;
            if (b->fn.rhs(b->fn.user, &a)) return DAECPP_B200_EINVAL;
            a.min_mode = DAECPP_B200_LANE_REFACTOR; a.out = b->JA.data(); a.out_stride = (int64_t)b->Ji.size();
            if (b->fn.jacobian(b->fn.user, &a)) return DAECPP_B200_EINVAL;
        }
        for (int32_t l = 0; l < L; ++l)
        {
            if (mode[l] == DAECPP_B200_LANE_IDLE) continue;
            res[l].iterations += 1;
            double *xk = b->xk.data() + (size_t)l * n, *xh = b->xh.data() + (size_t)l * 4 * n;
            const double alpha = oracle_time_derivative(n, cmd[l].order, cmd[l].dt, xk, xh, dxdt.data());
            oracle_convert(n, nTM, b->MA.data() + (size_t)l * nTM, b->Mi.data(), b->Mj.data(), mo.data(), mi.data(), mv.data());
            oracle_residual(n, mo.data(), mi.data(), mv.data(), dxdt.data(), b->f.data() + (size_t)l * n, b->b.data() + (size_t)l * n);
            if (mode[l] == DAECPP_B200_LANE_REFACTOR)
            {
                b->outer[l].assign(n + 1, 0); b->inner[l].assign(nTJ + nTM + 1, 0); b->vals[l].assign(nTJ + nTM + 1, 0.0);
                oracle_assemble_jb(n, nTJ, b->JA.data() + (size_t)l * nTJ, b->Ji.data(), b->Jj.data(), nTM, b->MA.data() + (size_t)l * nTM,
                                   b->Mi.data(), b->Mj.data(), alpha, b->outer[l].data(), b->inner[l].data(), b->vals[l].data());
            }
            double *dx = b->dx.data() + (size_t)l * n;
            if (oracle_lu_solve(n, b->outer[l].data(), b->inner[l].data(), b->vals[l].data(), b->b.data() + (size_t)l * n, dx))
            {
                res[l].status = DAECPP_B200_EDECOMP;
                running[l] = 0;
                continue;
            }
            int flags = 0;
            double v = 0.0;
            for (int32_t i = 0; i < n; ++i)
            {   // dae-cpp/solver.hpp:590-619 over ALL elements (the device does not stop at the first offender) + :703-717
                const double e = std::fabs(dx[i]), xa = std::fabs(xh[i]);
                if (e > max_err_abs || std::isnan(dx[i])) flags |= DAECPP_B200_DIVERGED;
                if (xa > 1e-14 && e / xa > rtol) flags |= DAECPP_B200_NOT_CONVERGED;
                if (e > atol) flags |= DAECPP_B200_NOT_CONVERGED;
                xk[i] += dx[i];
                if (variability_on && xa > var_tol && std::fabs(xk[i]) > var_tol)
                {
                    const double rel = std::fabs((xh[i] - xk[i]) / xh[i]);
                    if (rel > v) v = rel;
                }
            }
            res[l].flags = flags;
            res[l].variability = v;
            if (!(flags & DAECPP_B200_NOT_CONVERGED) || (flags & DAECPP_B200_DIVERGED)) running[l] = 0; // the Newton loop of this step is over
        }
    }
    b->rounds += 1;
    return 0;
}

int daecpp_b200_batch_get_stat(daecpp_b200_batch *b, const char *key, double *value)
{
    if (std::string(key) == "rounds") { *value = b->rounds; return 0; }
    *value = 0.0;
    return 0;
}

int daecpp_b200_set_pooling(int) { return 0; } // (nothing is pooled here)
} // extern "C"
