// TEST INFRASTRUCTURE ONLY.  A scalar CPU walk over the data structures produced by the
// product's symbolic analysis (dae-cpp_b200/csrc/symbolic.cpp), doing front by front what the
// CUDA kernels do level by level.  It exists so that the symbolic phase (ordering, supernodes,
// relative indices, slot map) can be validated in the GPU-less build container; it is never
// linked into libdaecpp_b200.so.
#include "../../dae-cpp_b200/csrc/symbolic.hpp"

#include <cmath>
#include <cstring>
#include <vector>

using namespace daecpp_b200;

extern "C" int mfemu_solve(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals,
                           const double *b, double *x, int32_t leaf_size, double *stats /* >= 8 */)
{
    Symbolic S;
    SymbolicOptions opt;
    if (leaf_size > 0) opt.leaf_size = leaf_size;
    const char *err = nullptr;
    if (!analyse_pattern(n, outer, inner, opt, S, &err)) return 21;
    if (stats)
    {
        stats[0] = S.nsn; stats[1] = S.nlevels; stats[2] = (double)S.nnz_factor; stats[3] = S.flops;
        stats[4] = S.max_front; stats[5] = (double)S.lu_size; stats[6] = (double)S.cb_size; stats[7] = S.max_pivot_block;
    }
    std::vector<double> LU((size_t)S.lu_size, 0.0), CB((size_t)S.cb_size, 0.0), upd((size_t)S.upd_size, 0.0);
    for (int64_t p = 0; p < outer[n]; ++p) LU[S.a_dest[p]] += vals[p];
    int rc = 0;
    for (int32_t f = 0; f < S.nsn; ++f)
    {
        const int64_t s = S.s_of(f), r = S.r_of(f), ld = s + r;
        double *Lp = LU.data() + S.sn_lu[f], *Ut = Lp + ld * s, *C = CB.data() + S.sn_cb[f];
        // extend-add children
        for (int32_t ci = S.child_ptr[f]; ci < S.child_ptr[f + 1]; ++ci)
        {
            const int32_t c = S.child_idx[ci];
            const int64_t rc_ = S.r_of(c);
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
        // factor: right-looking on the pivot block, panels follow
        for (int64_t k = 0; k < s; ++k)
        {
            const double d = Lp[k * ld + k];
            if (d == 0.0 || !std::isfinite(d)) rc = 2;
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
    // forward: y = L^-1 P b with per-front update vectors
    std::vector<double> y((size_t)n);
    for (int32_t k = 0; k < n; ++k) y[k] = b[S.perm[k]];
    for (int32_t f = 0; f < S.nsn; ++f)
    {
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
    // backward: U x = y
    for (int32_t f = S.nsn - 1; f >= 0; --f)
    {
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
    for (int32_t k = 0; k < n; ++k) x[S.perm[k]] = y[k];
    return rc;
}
