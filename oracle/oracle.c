/* TEST INFRASTRUCTURE ONLY -- not part of the product path.
 *
 * Plain-C restatement of the arithmetic on dae-cpp's Newton linear-solve hot path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load it.
 * Parity status: PINNED -- tests/test_oracle.py checks every function here against
 * oracle/_ref/libdaeref.so (the unmodified reference headers compiled from
 * /root/reference) and against the reference's own test literals
 * (tests/test_sparse-matrix.cpp:183-229, tests/test_time-integrator.cpp:29-65,
 * tests/test_linear-solver.cpp:26-61).
 *
 * Citations are relative to /root/reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* ---- a3: sparse_matrix::convert (dae-cpp/sparse-matrix.hpp:124-152) ------------------
 * Eigen's findOrInsertCoeff (Eigen/src/SparseCore/SparseMatrix.h:231-265) keeps each
 * column sorted by row and does `+=` into an existing slot, so the stored value of a slot
 * is 0.0 + a_k1 + a_k2 + ... over its triplets in INPUT order.  We restate that as a
 * stable counting sort by (column,row) followed by a left-to-right segmented sum. */
static int64_t coo_to_csc(int32_t n, int64_t nT, const double *A, const int32_t *ti, const int32_t *tj,
                          int32_t *outer, int32_t *inner, double *vals)
{
    int64_t *cnt = (int64_t *)calloc((size_t)n + 1, sizeof(int64_t));
    int64_t *byrow = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nT ? nT : 1));
    int64_t *bycol = (int64_t *)malloc(sizeof(int64_t) * (size_t)(nT ? nT : 1));
    int64_t k, p;
    /* stable sort by row ... */
    for (k = 0; k < nT; ++k) cnt[ti[k] + 1]++;
    for (p = 0; p < n; ++p) cnt[p + 1] += cnt[p];
    for (k = 0; k < nT; ++k) byrow[cnt[ti[k]]++] = k;
    /* ... then stable sort by column: result ordered by (col,row,input index) */
    memset(cnt, 0, sizeof(int64_t) * ((size_t)n + 1));
    for (k = 0; k < nT; ++k) cnt[tj[k] + 1]++;
    for (p = 0; p < n; ++p) cnt[p + 1] += cnt[p];
    for (k = 0; k < nT; ++k) { int64_t s = byrow[k]; bycol[cnt[tj[s]]++] = s; }
    int64_t nnz = 0;
    int32_t col = 0;
    outer[0] = 0;
    for (k = 0; k < nT; ++k)
    {
        int64_t s = bycol[k];
        while (col < tj[s]) outer[++col] = (int32_t)nnz;
        if (nnz > outer[col] && inner[nnz - 1] == ti[s])
            vals[nnz - 1] += A[s];
        else
        {
            inner[nnz] = ti[s];
            vals[nnz] = 0.0 + A[s];
            nnz++;
        }
    }
    while (col < n) outer[++col] = (int32_t)nnz;
    free(cnt); free(byrow); free(bycol);
    return nnz;
}

int64_t oracle_convert(int32_t n, int64_t nT, const double *A, const int32_t *ti, const int32_t *tj,
                       int32_t *outer, int32_t *inner, double *vals)
{
    return coo_to_csc(n, nT, A, ti, tj, outer, inner, vals);
}

/* ---- a6: Jb = J - M*alpha on the union pattern (dae-cpp/solver.hpp:523,
 * Eigen/src/SparseCore/SparseCwiseBinaryOp.h:75-105,867-869).  Both present: J - (M*alpha);
 * J only: J - 0; M only: 0 - (M*alpha).  Explicit zeros stay.  Two roundings, no FMA. */
int64_t oracle_assemble_jb(int32_t n, int64_t nJ, const double *JA, const int32_t *Ji, const int32_t *Jj,
                           int64_t nM, const double *MA, const int32_t *Mi, const int32_t *Mj, double alpha,
                           int32_t *outer, int32_t *inner, double *vals)
{
    int32_t *jo = (int32_t *)malloc(sizeof(int32_t) * ((size_t)n + 1));
    int32_t *mo = (int32_t *)malloc(sizeof(int32_t) * ((size_t)n + 1));
    int32_t *ji = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nJ ? nJ : 1));
    int32_t *mi = (int32_t *)malloc(sizeof(int32_t) * (size_t)(nM ? nM : 1));
    double *jv = (double *)malloc(sizeof(double) * (size_t)(nJ ? nJ : 1));
    double *mv = (double *)malloc(sizeof(double) * (size_t)(nM ? nM : 1));
    coo_to_csc(n, nJ, JA, Ji, Jj, jo, ji, jv);
    coo_to_csc(n, nM, MA, Mi, Mj, mo, mi, mv);
    int64_t nnz = 0;
    outer[0] = 0;
    for (int32_t c = 0; c < n; ++c)
    {
        int32_t a = jo[c], ae = jo[c + 1], b = mo[c], be = mo[c + 1];
        while (a < ae || b < be)
        {
            if (b >= be || (a < ae && ji[a] < mi[b]))
            {
                inner[nnz] = ji[a];
                vals[nnz++] = jv[a] - 0.0;
                a++;
            }
            else if (a >= ae || mi[b] < ji[a])
            {
                volatile double m = mv[b] * alpha;
                inner[nnz] = mi[b];
                vals[nnz++] = 0.0 - m;
                b++;
            }
            else
            {
                volatile double m = mv[b] * alpha;
                inner[nnz] = ji[a];
                vals[nnz++] = jv[a] - m;
                a++; b++;
            }
        }
        outer[c + 1] = (int32_t)nnz;
    }
    free(jo); free(mo); free(ji); free(mi); free(jv); free(mv);
    return nnz;
}

/* ---- a4: time_derivative_approx (dae-cpp/solver.hpp:141-217).  Coefficients are formed
 * exactly as the reference's expressions associate (left to right), then applied per
 * element as  alpha*xk - c0*x0 + c1*x1 - c2*x2 + c3*x3.  coef[0]=alpha, coef[1..4]=c0..c3. */
double oracle_bdf_coefficients(int32_t order, const double *dt, double *coef)
{
    const double h0 = dt[0], h1 = dt[1], h2 = dt[2], h3 = dt[3];
    const double h01 = h0 + h1, h12 = h1 + h2, h23 = h2 + h3;
    const double h012 = h0 + h1 + h2, h123 = h1 + h2 + h3, h0123 = h0 + h1 + h2 + h3;
    coef[0] = coef[1] = coef[2] = coef[3] = coef[4] = 0.0;
    switch (order)
    {
    case 1:
        coef[0] = 1.0 / h0;
        coef[1] = coef[0];
        break;
    case 2:
        coef[0] = (2.0 * h0 + h1) / (h0 * h01);
        coef[1] = h01 / (h0 * h1);
        coef[2] = h0 / (h1 * h01);
        break;
    case 3:
        coef[0] = (3.0 * h0 * h0 + h1 * h12 + 2.0 * h0 * (2.0 * h1 + h2)) / (h0 * h01 * h012);
        coef[1] = h01 * h012 / (h0 * h1 * h12);
        coef[2] = h0 * h012 / (h1 * h2 * h01);
        coef[3] = h0 * h01 / (h2 * h12 * h012);
        break;
    case 4:
        coef[0] = (4.0 * h0 * h0 * h0 + h1 * h12 * h123 + 3.0 * h0 * h0 * (3.0 * h1 + 2.0 * h2 + h3) +
                   2.0 * h0 * (3.0 * h1 * h1 + h2 * h23 + 2.0 * h1 * (2.0 * h2 + h3))) /
                  (h0 * h01 * h012 * h0123);
        coef[1] = h01 * h012 * h0123 / (h0 * h1 * h12 * h123);
        coef[2] = h0 * h012 * h0123 / (h1 * h01 * h2 * h23);
        coef[3] = h0 * h01 * h0123 / (h2 * h12 * h012 * h3);
        coef[4] = h0 * h01 * h012 / (h3 * h23 * h123 * h0123);
        break;
    default:
        return NAN;
    }
    return coef[0];
}

double oracle_time_derivative(int32_t n, int32_t order, const double *dt, const double *xk,
                              const double *xh /* 4 x n */, double *dxdt)
{
    double c[5];
    double alpha = oracle_bdf_coefficients(order, dt, c);
    const double *x0 = xh, *x1 = xh + n, *x2 = xh + 2 * (size_t)n, *x3 = xh + 3 * (size_t)n;
    for (int32_t i = 0; i < n; ++i)
    {
        volatile double r;
        if (order == 1)
        {
            r = xk[i] - x0[i];       /* (xk - x0) * alpha : solver.hpp:163 */
            r = r * alpha;
        }
        else
        {
            volatile double t;
            r = alpha * xk[i];
            t = c[1] * x0[i]; r = r - t;
            t = c[2] * x1[i]; r = r + t;
            if (order >= 3) { t = c[3] * x2[i]; r = r - t; }
            if (order >= 4) { t = c[4] * x3[i]; r = r + t; }
        }
        dxdt[i] = r;
    }
    return alpha;
}

/* ---- a5: b = M*dxdt - f (dae-cpp/solver.hpp:516-518): column-major SpMV
 * (Eigen/src/SparseCore/SparseDenseProduct.h:96-112) accumulates, per row, the products in
 * increasing column order starting from 0; then one subtraction of f. */
void oracle_residual(int32_t n, const int32_t *mo, const int32_t *mi, const double *mv,
                     const double *dxdt, const double *f, double *b)
{
    for (int32_t r = 0; r < n; ++r) b[r] = 0.0;
    for (int32_t c = 0; c < n; ++c)
        for (int32_t p = mo[c]; p < mo[c + 1]; ++p)
        {
            volatile double t = mv[p] * dxdt[c];
            b[mi[p]] += t;
        }
    for (int32_t r = 0; r < n; ++r) b[r] -= f[r];
}

/* ---- a11: error check + update (dae-cpp/solver.hpp:590-619).  Returns bit0 = diverged,
 * bit1 = not converged.  On divergence the reference stops updating at the offending
 * element; xk is discarded by the caller's rollback (solver.hpp:643) so only the flags and
 * the elements before it are comparable. */
int oracle_error_check_update(int32_t n, const double *dx, const double *x0, double *xk,
                              double atol, double rtol, double max_err_abs)
{
    int flags = 0;
    for (int32_t i = 0; i < n; ++i)
    {
        double e = fabs(dx[i]);
        if (e > max_err_abs || isnan(dx[i])) { flags |= 1; break; }
        double xa = fabs(x0[i]);
        if (xa > 1e-14 && e / xa > rtol) flags |= 2;
        if (e > atol) flags |= 2;
        xk[i] += dx[i];
    }
    return flags;
}

/* ---- a12: history shift + variability (dae-cpp/solver.hpp:698-721). xh = 4 x n. */
double oracle_accept_step(int32_t n, const double *xk, double *xh, int variability_on, double var_tol)
{
    double *x0 = xh, *x1 = xh + n, *x2 = xh + 2 * (size_t)n, *x3 = xh + 3 * (size_t)n;
    double v = 0.0;
    for (int32_t i = 0; i < n; ++i)
    {
        x3[i] = x2[i]; x2[i] = x1[i]; x1[i] = x0[i]; x0[i] = xk[i];
        if (variability_on && fabs(x1[i]) > var_tol && fabs(x0[i]) > var_tol)
        {
            double rc = fabs((x1[i] - x0[i]) / x1[i]);
            if (rc > v) v = rc;
        }
    }
    return v;
}

/* ---- a8/a9: what Eigen::SparseLU's compute()+solve() deliver (dae-cpp/solver.hpp:537,558):
 * P A Q = L U with row partial pivoting, then two triangular solves.  The oracle does not
 * mimic SuperLU's supernodes or COLAMD (factors are never compared); it is a
 * left-looking column LU with max-abs row pivoting on natural column order, O(n^2) scans,
 * meant for n <= ~2e4.  Returns 0, or 2 when a column has no nonzero pivot. */
int oracle_lu_solve(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals,
                    const double *b, double *x)
{
    /* L stored by columns (unit diagonal implicit) in original row ids; U by columns in pivot
     * positions. */
    size_t cap = (size_t)outer[n] * 4 + 16 * (size_t)n + 64;
    int32_t *Lp = (int32_t *)malloc(sizeof(int32_t) * ((size_t)n + 1)), *Up = (int32_t *)malloc(sizeof(int32_t) * ((size_t)n + 1));
    int32_t *Li = (int32_t *)malloc(sizeof(int32_t) * cap), *Ui = (int32_t *)malloc(sizeof(int32_t) * cap);
    double *Lx = (double *)malloc(sizeof(double) * cap), *Ux = (double *)malloc(sizeof(double) * cap);
    size_t lcap = cap, ucap = cap;
    int32_t *pinv = (int32_t *)malloc(sizeof(int32_t) * (size_t)n); /* row -> pivot position or -1 */
    int32_t *prow = (int32_t *)malloc(sizeof(int32_t) * (size_t)n); /* pivot position -> row */
    double *w = (double *)calloc((size_t)n, sizeof(double));
    double *udiag = (double *)malloc(sizeof(double) * (size_t)n);
    char *mark = (char *)calloc((size_t)n, 1);
    int32_t *list = (int32_t *)malloc(sizeof(int32_t) * (size_t)n);
    int rc = 0;
    for (int32_t r = 0; r < n; ++r) pinv[r] = -1;
    Lp[0] = Up[0] = 0;
    for (int32_t j = 0; j < n && !rc; ++j)
    {
        int32_t nl = 0;
        int32_t kmax = -1;
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
        {
            int32_t r = inner[p];
            w[r] = vals[p];
            if (!mark[r]) { mark[r] = 1; list[nl++] = r; }
            if (pinv[r] > kmax) kmax = pinv[r];
        }
        int32_t un = Up[j];
        /* eliminate with previous pivots in increasing pivot order; kmax grows with fill */
        for (int32_t k = 0; k <= kmax; ++k)
        {
            int32_t r = prow[k];
            if (!mark[r]) continue;
            double ukj = w[r];
            if ((size_t)un + 1 >= ucap) { ucap *= 2; Ui = (int32_t *)realloc(Ui, sizeof(int32_t) * ucap); Ux = (double *)realloc(Ux, sizeof(double) * ucap); }
            Ui[un] = k; Ux[un++] = ukj;
            if (ukj != 0.0)
                for (int32_t p = Lp[k]; p < Lp[k + 1]; ++p)
                {
                    int32_t rr = Li[p];
                    if (!mark[rr]) { mark[rr] = 1; list[nl++] = rr; w[rr] = 0.0; if (pinv[rr] > kmax) kmax = pinv[rr]; }
                    w[rr] -= Lx[p] * ukj;
                }
        }
        Up[j + 1] = un;
        /* pivot: largest magnitude among not-yet-pivotal rows */
        int32_t piv = -1; double best = -1.0;
        for (int32_t q = 0; q < nl; ++q)
        {
            int32_t r = list[q];
            if (pinv[r] < 0 && fabs(w[r]) > best) { best = fabs(w[r]); piv = r; }
        }
        if (piv < 0 || best == 0.0 || isnan(best)) { rc = 2; }
        else
        {
            double d = w[piv];
            udiag[j] = d; pinv[piv] = j; prow[j] = piv;
            int32_t ln = Lp[j];
            for (int32_t q = 0; q < nl; ++q)
            {
                int32_t r = list[q];
                if (pinv[r] < 0)
                {
                    if ((size_t)ln + 1 >= lcap) { lcap *= 2; Li = (int32_t *)realloc(Li, sizeof(int32_t) * lcap); Lx = (double *)realloc(Lx, sizeof(double) * lcap); }
                    Li[ln] = r; Lx[ln++] = w[r] / d;
                }
            }
            Lp[j + 1] = ln;
        }
        for (int32_t q = 0; q < nl; ++q) { mark[list[q]] = 0; w[list[q]] = 0.0; }
    }
    if (!rc)
    {
        /* forward: y (indexed by pivot position) */
        double *y = (double *)malloc(sizeof(double) * (size_t)n);
        double *bb = (double *)malloc(sizeof(double) * (size_t)n);
        memcpy(bb, b, sizeof(double) * (size_t)n);
        for (int32_t k = 0; k < n; ++k)
        {
            double yk = bb[prow[k]];
            y[k] = yk;
            for (int32_t p = Lp[k]; p < Lp[k + 1]; ++p) bb[Li[p]] -= Lx[p] * yk;
        }
        /* backward: U x = y, U column j has entries at pivot positions k<j plus udiag[j] */
        for (int32_t j = n - 1; j >= 0; --j)
        {
            double xj = y[j] / udiag[j];
            x[j] = xj;
            for (int32_t p = Up[j]; p < Up[j + 1]; ++p) y[Ui[p]] -= Ux[p] * xj;
        }
        free(y); free(bb);
    }
    free(Lp); free(Up); free(Li); free(Ui); free(Lx); free(Ux); free(pinv); free(prow);
    free(w); free(udiag); free(mark); free(list);
    return rc;
}

/* CSC mat-vec and the parity metric of SURVEY.md 8(c).2 :
 * eta = ||b - A x||_inf / (||A||_inf ||x||_inf + ||b||_inf);  *res2 = ||b - A x||_2 */
double oracle_backward_error(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals,
                             const double *x, const double *b, double *res2)
{
    double *r = (double *)malloc(sizeof(double) * (size_t)n);
    double *rs = (double *)calloc((size_t)n, sizeof(double));
    for (int32_t i = 0; i < n; ++i) r[i] = b[i];
    for (int32_t c = 0; c < n; ++c)
        for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
        {
            r[inner[p]] -= vals[p] * x[c];
            rs[inner[p]] += fabs(vals[p]);
        }
    double rinf = 0, ainf = 0, xinf = 0, binf = 0, s2 = 0;
    for (int32_t i = 0; i < n; ++i)
    {
        if (fabs(r[i]) > rinf || isnan(r[i])) rinf = fabs(r[i]);
        if (rs[i] > ainf) ainf = rs[i];
        if (fabs(x[i]) > xinf) xinf = fabs(x[i]);
        if (fabs(b[i]) > binf) binf = fabs(b[i]);
        s2 += r[i] * r[i];
    }
    if (res2) *res2 = sqrt(s2);
    free(r); free(rs);
    return rinf / (ainf * xinf + binf);
}
