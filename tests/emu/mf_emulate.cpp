// TEST INFRASTRUCTURE ONLY.  A scalar CPU walk over the schedule produced by the product's
// symbolic analysis (dae-cpp_b200/csrc/symbolic.cpp), executing launch group by launch group
// what the CUDA kernels execute: subtree units through the SAME per-front code the kernels use
// (dae-cpp_b200/csrc/unit_ops.hpp, instantiated with one lane), big fronts front by front.  It
// exists so that ordering, unit partition, blobs, relative indices and slot maps can be
// validated in the GPU-less build container; it is never linked into libdaecpp_b200.so.
#include "../../dae-cpp_b200/csrc/symbolic.hpp"
#include "../../dae-cpp_b200/csrc/unit_ops.hpp"

#include <cmath>
#include <cstring>
#include <vector>

using namespace daecpp_b200;
using namespace daecpp_b200::sched;

namespace
{
struct OneLane
{
    int lane() const { return 0; }
    int width() const { return 1; }
    void sync() const {}
};
} // namespace

extern "C" int mfemu_solve(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals,
                           const double *b, double *x, int32_t leaf_size, int32_t unit_smem_bytes,
                           double *stats /* >= 12 */)
{
    Symbolic S;
    SymbolicOptions opt;
    if (leaf_size > 0) opt.leaf_size = leaf_size;
    if (unit_smem_bytes >= 0) opt.unit_smem_bytes = unit_smem_bytes;
    const char *err = nullptr;
    std::vector<double> mag(vals, vals + outer[n]); // what numeric.cu: upload_symbolic hands over when the diagonal has holes
    for (double &v : mag) v = v < 0 ? -v : v;
    opt.weights = mag.data();
    if (!analyse_pattern(n, outer, inner, opt, S, &err)) return 21;
    if (stats)
    {
        stats[0] = S.nsn; stats[1] = S.nlevels * 1000 + S.nstages; stats[2] = (double)S.nnz_factor; stats[3] = S.flops;
        stats[4] = S.max_front; stats[5] = (double)S.lu_size; stats[6] = (double)S.cb_size; stats[7] = S.max_pivot_block;
        stats[8] = S.nunit; stats[9] = S.nbig; stats[10] = (double)S.groups.size(); stats[11] = S.factor_smem_bytes;
    }
    // The CUDA path runs a launch of units beside the level of big fronts behind it when the schedule says they share a
    // stage of the tree (LaunchGroup::same_stage): neither may depend on the other.  The walk here takes such a pair in
    // the OPPOSITE order -- the level first, then the units -- so a schedule that flags a dependent pair gives wrong numbers.
    std::vector<LaunchGroup> order(S.groups.begin(), S.groups.end());
    int swapped = 0;
    for (size_t gi = 1; gi + 1 < order.size(); ++gi)
        if (order[gi].kind == 0 && order[gi].same_stage && order[gi + 1].kind != 0)
        {
            std::swap(order[gi], order[gi + 1]);
            ++gi; ++swapped;
        }
    if (stats) stats[10] = (double)S.groups.size() + 1e-3 * swapped;
    std::vector<double> LU((size_t)S.lu_size + 1, 0.0), CB((size_t)S.cb_size + 1, 0.0), upd((size_t)S.upd_size + 1, 0.0);
    std::vector<double> y((size_t)n);
    int fail = 0;
    OneLane P;

    // ---------------------------------------------------------------- factorisation
    for (size_t q = 0; q < S.big_u.size(); ++q) LU[S.big_dst[q]] += vals[S.big_u[q]];
    std::vector<double> W;
    for (const LaunchGroup &g : order)
    {
        if (g.kind == 0)
        {
            for (int32_t t = g.begin; t < g.end; ++t)
            {
                const int32_t *B = S.blob.data() + S.unit_blob[t];
                if ((int64_t)B[H_NINTS] * 4 + (int64_t)B[H_WDOUBLES] * 8 > S.factor_smem_bytes) return 30;
                W.assign((size_t)B[H_WDOUBLES] + 1, 0.0);
                for (int32_t e = 0; e < B[H_ENTCOUNT]; ++e) W[S.ent_dst[B[H_ENT0] + e]] = vals[S.ent_u[B[H_ENT0] + e]];
                const int32_t *lp = B + B[H_LEVELS];
                const int32_t *lg = lp + B[H_NLEV] + 1;
                for (int32_t l = 0; l < B[H_NLEV]; ++l)
                    for (int32_t fi = lp[l]; fi < lp[l + 1]; ++fi)
                    {
                        if (lg[l] == 1 && opt.serial_dim <= 4) // the kernels' register fast path
                            unit_ops::factor_front_small<4>(B, B + B[H_FRONTS] + (int64_t)fi * F_SIZE, W.data(),
                                                            LU.data() + (uint32_t)B[H_LU0], CB.data(), 0.0, &fail);
                        else
                            unit_ops::factor_front(P, B, B + B[H_FRONTS] + (int64_t)fi * F_SIZE, W.data(),
                                                   LU.data() + (uint32_t)B[H_LU0], CB.data(), 0.0, &fail);
                    }
            }
        }
        else
        {
            for (int32_t q = g.begin; q < g.end; ++q)
            {
                const int32_t f = S.big_sn[q];
                const int64_t s = S.s_of(f), r = S.r_of(f), ld = s + r;
                double *Lp = LU.data() + S.sn_lu[f], *Ut = Lp + ld * s;
                double *C = r ? CB.data() + S.sn_cb[f] : nullptr;
                for (int64_t e = 0; e < r * r; ++e) C[e] = 0.0;
                for (int32_t ci = S.child_ptr[f]; ci < S.child_ptr[f + 1]; ++ci)
                {
                    const int32_t c = S.child_idx[ci];
                    const int64_t rc_ = S.r_of(c);
                    if (!rc_) continue;
                    const double *Cc = CB.data() + S.sn_cb[c];
                    const int32_t *rel = S.sn_rel.data() + S.sn_rowptr[c];
                    for (int64_t bcol = 0; bcol < rc_; ++bcol)
                        for (int64_t a = 0; a < rc_; ++a)
                        {
                            const double v = Cc[bcol * rc_ + a];
                            const int64_t ra = rel[a], rb = rel[bcol];
                            if (rb < s) Lp[rb * ld + ra] += v;
                            else if (ra < s) Ut[ra * r + (rb - s)] += v;
                            else C[(rb - s) * r + (ra - s)] += v;
                        }
                }
                for (int64_t k = 0; k < s; ++k)
                {
                    const double d = Lp[k * ld + k];
                    if (d == 0.0 || !std::isfinite(d)) fail = 1;
                    const double inv = 1.0 / d;
                    for (int64_t i = k + 1; i < ld; ++i) Lp[k * ld + i] *= inv;
                    for (int64_t j = k + 1; j < s; ++j)
                    {
                        const double u = Lp[j * ld + k];
                        for (int64_t i = k + 1; i < ld; ++i) Lp[j * ld + i] -= Lp[k * ld + i] * u;
                        const double l = Lp[k * ld + j];
                        for (int64_t i = 0; i < r; ++i) Ut[j * r + i] -= Ut[k * r + i] * l;
                    }
                }
                for (int64_t bcol = 0; bcol < r; ++bcol)
                    for (int64_t k = 0; k < s; ++k)
                    {
                        const double u = Ut[k * r + bcol];
                        for (int64_t a = 0; a < r; ++a) C[bcol * r + a] -= Lp[k * ld + s + a] * u;
                    }
            }
        }
    }

    // ---------------------------------------------------------------- forward sweep
    for (int32_t k = 0; k < n; ++k) y[k] = b[S.row_src.empty() ? S.perm[k] : S.row_src[(size_t)S.perm[k]]]; // Q A x = Q b
    std::vector<double> Y, U, X;
    for (const LaunchGroup &g : order)
    {
        if (g.kind == 0)
        {
            for (int32_t t = g.begin; t < g.end; ++t)
            {
                const int32_t *B = S.blob.data() + S.unit_blob[t];
                const int32_t *fr = B + B[H_FRONTS];
                Y.assign((size_t)B[H_YDOUBLES] + 1, 0.0); U.assign((size_t)B[H_UDOUBLES] + 1, 0.0);
                for (int32_t fi = 0; fi < B[H_NF]; ++fi)
                    for (int32_t k = 0; k < fr[fi * F_SIZE + F_S]; ++k)
                        Y[fr[fi * F_SIZE + F_YOFF] + k] = y[fr[fi * F_SIZE + F_COL0] + k];
                const int32_t *lp = B + B[H_LEVELS];
                for (int32_t l = 0; l < B[H_NLEV]; ++l)
                    for (int32_t fi = lp[l]; fi < lp[l + 1]; ++fi)
                        unit_ops::forward_front(P, B, fr + (int64_t)fi * F_SIZE, Y.data(), U.data(),
                                                LU.data() + (uint32_t)B[H_LU0], upd.data());
                for (int32_t fi = 0; fi < B[H_NF]; ++fi)
                    for (int32_t k = 0; k < fr[fi * F_SIZE + F_S]; ++k)
                        y[fr[fi * F_SIZE + F_COL0] + k] = Y[fr[fi * F_SIZE + F_YOFF] + k];
            }
        }
        else
        {
            for (int32_t q = g.begin; q < g.end; ++q)
            {
                const int32_t f = S.big_sn[q];
                const int64_t s = S.s_of(f), r = S.r_of(f), ld = s + r;
                const int32_t first = S.sn_first[f];
                const double *Lp = LU.data() + S.sn_lu[f];
                double *u = upd.data() + S.sn_rowptr[f];
                for (int64_t a = 0; a < r; ++a) u[a] = 0.0;
                for (int32_t ci = S.child_ptr[f]; ci < S.child_ptr[f + 1]; ++ci)
                {
                    const int32_t c = S.child_idx[ci];
                    const int64_t rc_ = S.r_of(c);
                    const double *uc = upd.data() + S.sn_rowptr[c];
                    const int32_t *rel = S.sn_rel.data() + S.sn_rowptr[c];
                    for (int64_t a = 0; a < rc_; ++a)
                    {
                        if (rel[a] < s) y[first + rel[a]] += uc[a];
                        else u[rel[a] - s] += uc[a];
                    }
                }
                for (int64_t k = 0; k < s; ++k)
                {
                    const double yk = y[first + k];
                    for (int64_t i = k + 1; i < s; ++i) y[first + i] -= Lp[k * ld + i] * yk;
                    for (int64_t a = 0; a < r; ++a) u[a] -= Lp[k * ld + s + a] * yk;
                }
            }
        }
    }

    // ---------------------------------------------------------------- backward sweep
    for (size_t gi = order.size(); gi-- > 0;)
    {
        const LaunchGroup &g = order[gi];
        if (g.kind == 0)
        {
            for (int32_t t = g.end - 1; t >= g.begin; --t)
            {
                const int32_t *B = S.blob.data() + S.unit_blob[t];
                const int32_t *fr = B + B[H_FRONTS];
                const int32_t ny = B[H_YDOUBLES];
                X.assign((size_t)ny + B[H_NXR] + 1, 0.0);
                for (int32_t fi = 0; fi < B[H_NF]; ++fi)
                    for (int32_t k = 0; k < fr[fi * F_SIZE + F_S]; ++k)
                        X[fr[fi * F_SIZE + F_YOFF] + k] = y[fr[fi * F_SIZE + F_COL0] + k];
                for (int32_t a = 0; a < B[H_NXR]; ++a) X[ny + a] = y[B[B[H_XR] + a]];
                const int32_t *lp = B + B[H_LEVELS];
                for (int32_t l = B[H_NLEV] - 1; l >= 0; --l)
                    for (int32_t fi = lp[l]; fi < lp[l + 1]; ++fi)
                        unit_ops::backward_front(P, B, fr + (int64_t)fi * F_SIZE, X.data(), LU.data() + (uint32_t)B[H_LU0]);
                for (int32_t fi = 0; fi < B[H_NF]; ++fi)
                    for (int32_t k = 0; k < fr[fi * F_SIZE + F_S]; ++k)
                        y[fr[fi * F_SIZE + F_COL0] + k] = X[fr[fi * F_SIZE + F_YOFF] + k];
            }
        }
        else
        {
            for (int32_t q = g.end - 1; q >= g.begin; --q)
            {
                const int32_t f = S.big_sn[q];
                const int64_t s = S.s_of(f), r = S.r_of(f), ld = s + r;
                const int32_t first = S.sn_first[f];
                const double *Lp = LU.data() + S.sn_lu[f], *Ut = Lp + ld * s;
                const int32_t *rows = S.sn_rows.data() + S.sn_rowptr[f];
                for (int64_t k = s - 1; k >= 0; --k)
                {
                    double acc = y[first + k];
                    for (int64_t a = 0; a < r; ++a) acc -= Ut[k * r + a] * y[rows[a]];
                    for (int64_t j = k + 1; j < s; ++j) acc -= Lp[j * ld + k] * y[first + j];
                    y[first + k] = acc / Lp[k * ld + k];
                }
            }
        }
    }
    for (int32_t k = 0; k < n; ++k) x[S.perm[k]] = y[k];
    return fail ? 2 : 0;
}
