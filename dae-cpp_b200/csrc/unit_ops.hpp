// Per-front work of a subtree unit (schedule.hpp), written once for the CUDA kernels
// (subtree.cu: a front is handled by G cooperating lanes, G = 1, 8, 32 or the whole block) and for
// the scalar walk used by the GPU-less tests (tests/emu/: G = 1).  `P` is a policy type with
//   int lane() const, int width() const, void sync() const
// All pointers except LUg / CBg / UPDg point into the unit's (shared-memory) arenas.
//
// Role in the reference: this is what Eigen::SparseLU::factorize and _solve_impl do column by
// column (Eigen/src/SparseLU/SparseLU.h:611-814, 338-358), restated as a multifrontal sweep.
#pragma once
#include <cmath>
#include <cstdint>

#include "schedule.hpp"

#if defined(__CUDACC__)
#define DAE_HD __host__ __device__ __forceinline__
#else
#define DAE_HD inline
#endif
// Data written by OTHER thread blocks of the same launch (contribution blocks, update vectors,
// finished unknowns) must bypass the non-coherent L1.
#if defined(__CUDA_ARCH__)
#define DAE_LDCG(p) __ldcg(p)
#else
#define DAE_LDCG(p) (*(p))
#endif

#if defined(__CUDA_ARCH__)
#define DAE_UNROLL _Pragma("unroll")
#else
#define DAE_UNROLL
#endif

namespace daecpp_b200
{
namespace unit_ops
{
using namespace sched;

// Extend-add the children's contribution blocks, eliminate the s pivots (right-looking, static
// pivoting), store the L and transposed-U panels, and leave / export the contribution block.
template <class P>
DAE_HD void factor_front(const P &p, const int32_t *B, const int32_t *F, double *W, double *LUunit, double *CBg,
                         double pivot_tol, int *fail)
{
    const int lane = p.lane(), G = p.width();
    const int32_t s = F[F_S], r = F[F_R], d = s + r;
    double *A = W + F[F_WOFF];
    const int32_t *fronts = B + B[H_FRONTS];
    for (int32_t k = 0; k < F[F_NCHILD]; ++k)
    {
        const int32_t v = B[F[F_CHILD] + k];
        const double *Cc; const int32_t *rel; int32_t rc, ldc;
        if (v >= 0)
        {
            const int32_t *Fc = fronts + (int64_t)v * F_SIZE;
            const int32_t sc = Fc[F_S];
            rc = Fc[F_R]; ldc = sc + rc;
            Cc = W + Fc[F_WOFF] + (int64_t)sc * ldc + sc;
            rel = B + Fc[F_ROWS];
        }
        else
        {
            const int32_t *X = B - v;
            rc = X[X_R]; ldc = rc;
            Cc = CBg + (uint32_t)X[X_CB];
            rel = B + X[X_REL];
        }
        for (int32_t e = lane; e < rc * rc; e += G)
        {
            const int32_t a = e % rc, b = e / rc;
            const double cv = (v >= 0) ? Cc[(int64_t)b * ldc + a] : DAE_LDCG(Cc + (int64_t)b * ldc + a);
            A[(int64_t)rel[b] * d + rel[a]] += cv;
        }
        p.sync();
    }
    // Pivots are eliminated KB at a time.  Inside a block of KB pivots only the block's own columns and rows are brought up
    // to date pivot by pivot; everything behind them then receives the KB rank-1 updates in one pass, every lane holding an
    // RI x RJ patch in registers -- each element still sees a(i,j) -= l(i,k) * u(k,j) for k in ascending order with the same
    // operands, so the factors are the same bits as with one full update per pivot, but the patch is read and written once
    // per KB pivots and the multipliers are shared across it (the rank-1 form needed four shared-memory accesses per
    // multiply-add).  Only fronts swept by the whole block (or a warp on a large front) go this way; the small fronts of the
    // lower local levels, a few lanes each, keep one full update per pivot (measured: no faster in patches).
    // (G == 1 with a large front only happens in the scalar walk of the GPU-less tests: they cover the blocked form too.)
    constexpr int KBMAX = 4, KB = 4, RI = 2, RJ = 2;
    const bool blocked = G > 32 || ((G == 32 || G == 1) && d >= 48);
    const int GI = G < 32 ? G : 32, GJ = G / GI; // lanes along a column, lane groups across columns
    const int li = lane % GI, lj = lane / GI;
    for (int32_t k = 0; !blocked && k < s; ++k)
    {
        double *colk = A + (int64_t)k * d;
        const double piv = colk[k];
        if (lane == 0 && (!(fabs(piv) > pivot_tol) || std::isinf(piv))) *fail = 1;
        const double inv = 1.0 / piv;
        // (no barrier here: the scaling below leaves the pivot itself alone, and the last update ended with one)
        for (int32_t i = k + 1 + lane; i < d; i += G) colk[i] *= inv;
        p.sync();
        const int32_t m = d - k - 1;
        for (int32_t e = lane; e < m * m; e += G)
        {
            const int32_t j = k + 1 + e / m, i = k + 1 + e % m;
            A[(int64_t)j * d + i] -= colk[i] * A[(int64_t)j * d + k];
        }
        p.sync();
    }
    for (int32_t k0 = 0; blocked && k0 < s; k0 += KB)
    {
        const int32_t k1 = (k0 + KB < s) ? k0 + KB : s;
        for (int32_t k = k0; k < k1; ++k)
        {
            double *colk = A + (int64_t)k * d;
            const double piv = colk[k];
            if (lane == 0 && (!(fabs(piv) > pivot_tol) || std::isinf(piv))) *fail = 1;
            const double inv = 1.0 / piv;
            // (no barrier here: the scaling leaves the pivot itself alone, and the last update ended with one)
            for (int32_t i = k + 1 + lane; i < d; i += G) colk[i] *= inv;
            p.sync();
            if (k + 1 < k1)
            {
                for (int32_t j = k + 1; j < k1; ++j)
                {   // the block's columns right of the pivot
                    double *colj = A + (int64_t)j * d;
                    const double ukj = colj[k];
                    for (int32_t i = k + 1 + lane; i < d; i += G) colj[i] -= colk[i] * ukj;
                }
                for (int32_t i = k + 1; i < k1; ++i)
                {   // the block's rows below the pivot, right of the block
                    const double lik = colk[i];
                    for (int32_t j = k1 + lane; j < d; j += G) A[(int64_t)j * d + i] -= lik * A[(int64_t)j * d + k];
                }
                p.sync();
            }
        }
        const int32_t m = d - k1, kb = k1 - k0;
        if (m > 0)
        {
            for (int32_t jt = lj * RJ; jt < m; jt += GJ * RJ)
                for (int32_t it = li; it < m; it += GI * RI)
                {
                    double acc[RI][RJ], lv[KBMAX][RI], uv[KBMAX][RJ];
                    DAE_UNROLL
                    for (int b = 0; b < RJ; ++b)
                        DAE_UNROLL
                        for (int a = 0; a < RI; ++a)
                            acc[a][b] = (it + a * GI < m && jt + b < m) ? A[(int64_t)(k1 + jt + b) * d + k1 + it + a * GI] : 0.0;
                    DAE_UNROLL
                    for (int kk = 0; kk < KBMAX; ++kk)
                    {
                        DAE_UNROLL
                        for (int a = 0; a < RI; ++a) lv[kk][a] = (kk < kb && it + a * GI < m) ? A[(int64_t)(k0 + kk) * d + k1 + it + a * GI] : 0.0;
                        DAE_UNROLL
                        for (int b = 0; b < RJ; ++b) uv[kk][b] = (kk < kb && jt + b < m) ? A[(int64_t)(k1 + jt + b) * d + k0 + kk] : 0.0;
                    }
                    DAE_UNROLL
                    for (int kk = 0; kk < KBMAX; ++kk)
                        if (kk < kb)
                        {
                            DAE_UNROLL
                            for (int b = 0; b < RJ; ++b)
                                DAE_UNROLL
                                for (int a = 0; a < RI; ++a) acc[a][b] -= lv[kk][a] * uv[kk][b];
                        }
                    DAE_UNROLL
                    for (int b = 0; b < RJ; ++b)
                        DAE_UNROLL
                        for (int a = 0; a < RI; ++a)
                            if (it + a * GI < m && jt + b < m) A[(int64_t)(k1 + jt + b) * d + k1 + it + a * GI] = acc[a][b];
                }
            p.sync();
        }
    }
    double *Lp = LUunit + (uint32_t)F[F_LUOFF];
    for (int32_t e = lane; e < d * s; e += G) Lp[e] = A[e];
    double *Ut = Lp + (int64_t)d * s;
    for (int32_t e = lane; e < r * s; e += G)
    {
        const int32_t a = e % r, k = e / r;
        Ut[e] = A[(int64_t)(s + a) * d + k];
    }
    if (F[F_EXTCB] != -1)
    {
        double *C = CBg + (uint32_t)F[F_EXTCB];
        for (int32_t e = lane; e < r * r; e += G)
        {
            const int32_t a = e % r, b = e / r;
            C[e] = A[(int64_t)(s + b) * d + s + a];
        }
    }
}

// One-thread fast path for a front of order exactly D (compile-time strides): extend-add in the
// arena, one load of the D x D block into registers, elimination fully unrolled, and the
// factor panels / contribution block written straight from registers.
template <int D>
DAE_HD void factor_front_fixed(const int32_t *B, const int32_t *F, double *W, double *LUunit, double *CBg,
                               double pivot_tol, int *fail)
{
    const int32_t s = F[F_S], r = D - s;
    double *A = W + F[F_WOFF];
    const int32_t *fronts = B + B[H_FRONTS];
    const int32_t nch = F[F_NCHILD];
    for (int32_t k = 0; k < nch; ++k)
    {
        const int32_t v = B[F[F_CHILD] + k];
        if (v >= 0)
        {
            const int32_t *Fc = fronts + v * F_SIZE;
            const int32_t sc = Fc[F_S], rc = Fc[F_R], ldc = sc + rc;
            const double *Cc = W + Fc[F_WOFF] + sc * ldc + sc;
            const int32_t *rel = B + Fc[F_ROWS];
            for (int32_t b = 0; b < rc; ++b)
            {
                double *Ab = A + rel[b] * D;
                for (int32_t a = 0; a < rc; ++a) Ab[rel[a]] += Cc[b * ldc + a];
            }
        }
        else
        {
            const int32_t *X = B - v;
            const int32_t rc = X[X_R];
            const double *Cc = CBg + (uint32_t)X[X_CB];
            const int32_t *rel = B + X[X_REL];
            for (int32_t b = 0; b < rc; ++b)
            {
                double *Ab = A + rel[b] * D;
                for (int32_t a = 0; a < rc; ++a) Ab[rel[a]] += DAE_LDCG(Cc + b * rc + a);
            }
        }
    }
    double a[D][D];
    DAE_UNROLL
    for (int j = 0; j < D; ++j)
        DAE_UNROLL
        for (int i = 0; i < D; ++i) a[j][i] = A[j * D + i];
    bool bad = false;
    DAE_UNROLL
    for (int k = 0; k < D; ++k)
    {
        if (k < s)
        {
            const double piv = a[k][k];
            bad |= !(fabs(piv) > pivot_tol) || std::isinf(piv);
            const double inv = 1.0 / piv;
            DAE_UNROLL
            for (int i = k + 1; i < D; ++i) a[k][i] *= inv;
            DAE_UNROLL
            for (int j = k + 1; j < D; ++j)
                DAE_UNROLL
                for (int i = k + 1; i < D; ++i) a[j][i] -= a[k][i] * a[j][k];
        }
    }
    if (bad) *fail = 1;
    double *Lp = LUunit + (uint32_t)F[F_LUOFF];
    double *Ut = Lp + D * s;
    const bool ext = F[F_EXTCB] != -1;
    double *C = CBg + (uint32_t)F[F_EXTCB];
    DAE_UNROLL
    for (int j = 0; j < D; ++j)
    {
        if (j < s)
        {
            DAE_UNROLL
            for (int i = 0; i < D; ++i) Lp[j * D + i] = a[j][i];
        }
        else
        {
            DAE_UNROLL
            for (int i = 0; i < D; ++i)
            {
                if (i < s) Ut[i * r + (j - s)] = a[j][i];
                else
                {
                    A[j * D + i] = a[j][i]; // contribution block stays in the arena for the parent
                    if (ext) C[(j - s) * r + (i - s)] = a[j][i];
                }
            }
        }
    }
}

template <int DMAX>
DAE_HD void factor_front_small(const int32_t *B, const int32_t *F, double *W, double *LUunit, double *CBg,
                               double pivot_tol, int *fail)
{
    switch (F[F_S] + F[F_R])
    {
    case 1: factor_front_fixed<1>(B, F, W, LUunit, CBg, pivot_tol, fail); break;
    case 2: factor_front_fixed<2>(B, F, W, LUunit, CBg, pivot_tol, fail); break;
    case 3: factor_front_fixed<3>(B, F, W, LUunit, CBg, pivot_tol, fail); break;
    default: factor_front_fixed<4>(B, F, W, LUunit, CBg, pivot_tol, fail); break;
    }
}

// Forward substitution of one front: gather the children's update vectors, solve with the unit
// lower triangle of the pivot block, push the remainder into this front's update vector.
// Y: pivot unknowns of the unit, U: update-vector arena, LUs: the unit's factor storage.
template <class P>
DAE_HD void forward_front(const P &p, const int32_t *B, const int32_t *F, double *Y, double *U, const double *LUs,
                          double *UPDg)
{
    const int lane = p.lane(), G = p.width();
    const int32_t s = F[F_S], r = F[F_R], d = s + r;
    double *y1 = Y + F[F_YOFF];
    double *u = U + F[F_UOFF];
    const int32_t *fronts = B + B[H_FRONTS];
    for (int32_t k = 0; k < F[F_NCHILD]; ++k)
    {
        const int32_t v = B[F[F_CHILD] + k];
        const double *uc; const int32_t *rel; int32_t rc;
        if (v >= 0)
        {
            const int32_t *Fc = fronts + (int64_t)v * F_SIZE;
            rc = Fc[F_R]; uc = U + Fc[F_UOFF]; rel = B + Fc[F_ROWS];
        }
        else
        {
            const int32_t *X = B - v;
            rc = X[X_R]; uc = UPDg + X[X_UPD]; rel = B + X[X_REL];
        }
        for (int32_t a = lane; a < rc; a += G)
        {
            const int32_t ra = rel[a];
            const double uv = (v >= 0) ? uc[a] : DAE_LDCG(uc + a);
            if (ra < s) y1[ra] += uv; else u[ra - s] += uv;
        }
        p.sync();
    }
    const double *Lp = LUs + (uint32_t)F[F_LUOFF];
    for (int32_t k = 0; k < s; ++k)
    {
        const double yk = y1[k];
        const double *colk = Lp + (int64_t)k * d;
        for (int32_t i = k + 1 + lane; i < d; i += G)
        {
            const double v = colk[i] * yk;
            if (i < s) y1[i] -= v; else u[i - s] -= v;
        }
        p.sync();
    }
    if (F[F_EXTUPD] != -1)
        for (int32_t a = lane; a < r; a += G) UPDg[F[F_EXTUPD] + a] = u[a];
}

// Backward substitution of one front.  X = [Y | xr]: the unit's pivot unknowns followed by the
// (already final) unknowns of the rows that lie above the unit.
template <class P>
DAE_HD void backward_front(const P &p, const int32_t *B, const int32_t *F, double *X, const double *LUs)
{
    const int lane = p.lane(), G = p.width();
    const int32_t s = F[F_S], r = F[F_R], d = s + r;
    double *y1 = X + F[F_YOFF];
    const double *Lp = LUs + (uint32_t)F[F_LUOFF];
    const double *Ut = Lp + (int64_t)d * s;
    const int32_t *src = B + F[F_ROWS] + r;
    for (int32_t k = lane; k < s; k += G)
    {
        double acc = y1[k];
        const double *utk = Ut + (int64_t)k * r;
        for (int32_t a = 0; a < r; ++a) acc -= utk[a] * X[src[a]];
        y1[k] = acc;
    }
    p.sync();
    for (int32_t j = s - 1; j >= 0; --j)
    {
        const double *colj = Lp + (int64_t)j * d;
        if (lane == 0) y1[j] = y1[j] / colj[j];
        p.sync();
        const double xj = y1[j];
        for (int32_t i = lane; i < j; i += G) y1[i] -= colj[i] * xj;
        p.sync();
    }
}

} // namespace unit_ops
} // namespace daecpp_b200
