// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// C-ABI shim around the UNMODIFIED reference headers (dae-cpp 2.2.0 + vendored
// Eigen 3.4.90), compiled from /root/reference where it lies into
// oracle/_ref/libdaeref.so by oracle/Makefile.  It exposes the reference's own
// implementation of every step of the Newton linear-solve hot path so that
//   * oracle/oracle.c (the CPU restatement) can be pinned against it, and
//   * tests/ and bench.py's cpu_baseline leg can use it as the checker / timed
//     CPU baseline ("kind": "reference").
// Nothing under dae-cpp_b200/ may link or load this file.
//
// Reference call sites exercised (paths relative to /root/reference):
//   sparse_matrix::convert              dae-cpp/sparse-matrix.hpp:124-152
//   Jb -= M_*alpha                      dae-cpp/solver.hpp:523
//   b = M_*dxdt; b -= f_                dae-cpp/solver.hpp:516-518
//   time_derivative_approx              dae-cpp/solver.hpp:141-217
//   linsolver.compute / solve           dae-cpp/solver.hpp:537,558
#include <dae-cpp/solver.hpp>

#include <chrono>
#include <cstdint>
#include <cstring>

using namespace daecpp;
using clk = std::chrono::steady_clock;

static double ms_since(clk::time_point t0)
{
    return std::chrono::duration<double, std::milli>(clk::now() - t0).count();
}

static sparse_matrix make_triplets(int64_t nT, const double *A, const int32_t *i, const int32_t *j)
{
    sparse_matrix S;
    S.A.assign(A, A + nT);
    S.i.assign(i, i + nT);
    S.j.assign(j, j + nT);
    return S;
}

static int64_t export_csc(core::eimat &M, int32_t *outer, int32_t *inner, double *vals, int64_t cap)
{
    M.makeCompressed();
    const int64_t nnz = M.nonZeros();
    if (outer)
        std::memcpy(outer, M.outerIndexPtr(), sizeof(int32_t) * (M.cols() + 1));
    if (nnz > cap)
        return -nnz;
    if (inner)
        std::memcpy(inner, M.innerIndexPtr(), sizeof(int32_t) * nnz);
    if (vals)
        std::memcpy(vals, M.valuePtr(), sizeof(double) * nnz);
    return nnz;
}

extern "C" {

// triplets -> compressed columns exactly as sparse_matrix::convert does it.
int64_t daeref_convert(int32_t n, int64_t nT, const double *A, const int32_t *i, const int32_t *j,
                       int32_t *outer, int32_t *inner, double *vals, int64_t cap)
{
    sparse_matrix S = make_triplets(nT, A, i, j);
    core::eimat M = S.convert(n);
    return export_csc(M, outer, inner, vals, cap);
}

// Jb = J.convert(n); Jb -= M.convert(n) * alpha  (solver.hpp:502,523)
int64_t daeref_assemble_jb(int32_t n, int64_t nJ, const double *JA, const int32_t *Ji, const int32_t *Jj,
                           int64_t nM, const double *MA, const int32_t *Mi, const int32_t *Mj, double alpha,
                           int32_t *outer, int32_t *inner, double *vals, int64_t cap)
{
    sparse_matrix J = make_triplets(nJ, JA, Ji, Jj);
    sparse_matrix M = make_triplets(nM, MA, Mi, Mj);
    core::eimat Jb = J.convert(n);
    core::eimat M_ = M.convert(n);
    Jb -= M_ * alpha;
    return export_csc(Jb, outer, inner, vals, cap);
}

// b = M*dxdt - f   (solver.hpp:516-518)
void daeref_residual(int32_t n, int64_t nM, const double *MA, const int32_t *Mi, const int32_t *Mj,
                     const double *dxdt, const double *f, double *b)
{
    sparse_matrix M = make_triplets(nM, MA, Mi, Mj);
    core::eimat M_ = M.convert(n);
    core::eivec d = Eigen::Map<const core::eivec>(dxdt, n);
    core::eivec f_ = Eigen::Map<const core::eivec>(f, n);
    core::eivec r = M_ * d;
    r -= f_;
    std::memcpy(b, r.data(), sizeof(double) * n);
}

// BDF time derivative; x_hist is 4 contiguous vectors of length n (solver.hpp:141-217)
double daeref_time_derivative(int32_t n, int32_t order, const double *dt4, const double *xk,
                              const double *x_hist, double *dxdt)
{
    core::detail::SolverState st(n);
    st.order = order;
    for (int k = 0; k < 4; ++k)
    {
        st.dt[k] = dt4[k];
        st.x[k].assign(x_hist + (size_t)k * n, x_hist + (size_t)(k + 1) * n);
    }
    core::rvec xk_(xk, xk + n);
    core::eivec d(n);
    double alpha = core::detail::time_derivative_approx(d, xk_, st, n);
    std::memcpy(dxdt, d.data(), sizeof(double) * n);
    return alpha;
}

// Eigen::SparseLU on a CSC matrix: compute() (= analyzePattern + factorize) then solve().
// times_ms[0..2] = analyze, factorize, solve (optional). Returns 0 on success, 2/3 as the
// solver's exit codes for decomposition/solve failure (typedefs.hpp:47-54).
int daeref_sparselu(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals,
                    const double *b, double *x, int nrhs_repeat, double *times_ms, int64_t *nnz_lu)
{
    Eigen::Map<const core::eimat> A(n, n, outer[n], outer, inner, vals);
    core::eimat Jb = A;
    Eigen::SparseLU<core::eimat> lin;
    auto t0 = clk::now();
    lin.analyzePattern(Jb);
    double ta = ms_since(t0);
    t0 = clk::now();
    lin.factorize(Jb);
    double tf = ms_since(t0);
    if (lin.info() != Eigen::Success)
        return 2;
    core::eivec rhs = Eigen::Map<const core::eivec>(b, n);
    core::eivec dx;
    t0 = clk::now();
    for (int r = 0; r < (nrhs_repeat > 0 ? nrhs_repeat : 1); ++r)
        dx = lin.solve(rhs);
    double ts = ms_since(t0) / (nrhs_repeat > 0 ? nrhs_repeat : 1);
    if (lin.info() != Eigen::Success)
        return 3;
    std::memcpy(x, dx.data(), sizeof(double) * n);
    if (times_ms)
    {
        times_ms[0] = ta;
        times_ms[1] = tf;
        times_ms[2] = ts;
    }
    if (nnz_lu)
    {
        nnz_lu[0] = lin.nnzL();
        nnz_lu[1] = lin.nnzU();
    }
    return 0;
}

// One full Newton linear step as the reference does it, timed with its own buckets:
// convert(J), convert(M) [optional], Jb -= M*alpha, compute(), solve().
// times_ms: [convert, linear_algebra, compute, solve]
int daeref_newton_step(int32_t n, int64_t nJ, const double *JA, const int32_t *Ji, const int32_t *Jj,
                       int64_t nM, const double *MA, const int32_t *Mi, const int32_t *Mj, double alpha,
                       const double *dxdt, const double *f, double *dx_out, double *times_ms)
{
    sparse_matrix J = make_triplets(nJ, JA, Ji, Jj);
    sparse_matrix M = make_triplets(nM, MA, Mi, Mj);
    auto t0 = clk::now();
    core::eimat M_ = M.convert(n);
    core::eimat Jb = J.convert(n);
    double tc = ms_since(t0);
    t0 = clk::now();
    core::eivec d = Eigen::Map<const core::eivec>(dxdt, n);
    core::eivec f_ = Eigen::Map<const core::eivec>(f, n);
    core::eivec b = M_ * d;
    b -= f_;
    Jb -= M_ * alpha;
    double tl = ms_since(t0);
    Eigen::SparseLU<core::eimat> lin;
    t0 = clk::now();
    lin.compute(Jb);
    double tf = ms_since(t0);
    if (lin.info() != Eigen::Success)
        return 2;
    t0 = clk::now();
    core::eivec dx = lin.solve(b);
    double ts = ms_since(t0);
    if (lin.info() != Eigen::Success)
        return 3;
    std::memcpy(dx_out, dx.data(), sizeof(double) * n);
    if (times_ms)
    {
        times_ms[0] = tc;
        times_ms[1] = tl;
        times_ms[2] = tf;
        times_ms[3] = ts;
    }
    return 0;
}

// The same step as a session, the way the reference's solve() holds its objects (dae-cpp/solver.hpp:325-337): the
// triplet buffers, the converted mass matrix (is_mass_matrix_static: M.convert runs once, solver.hpp:477-481) and the
// Eigen::SparseLU object live across iterations; one iteration is Jb = J.convert(n) (:502), b = M_*dxdt - f
// (:516-518), Jb -= M_*alpha (:523), linsolver.compute(Jb) (:537), dx = linsolver.solve(b) (:558).
struct RefSession
{
    int32_t n;
    sparse_matrix J, M;
    core::eimat M_;
    Eigen::SparseLU<core::eimat> lin;
    core::eivec d, f_, b, dx;
};
void *daeref_newton_open(int32_t n, int64_t nJ, const double *JA, const int32_t *Ji, const int32_t *Jj, int64_t nM, const double *MA,
                         const int32_t *Mi, const int32_t *Mj, const double *dxdt, const double *f)
{
    RefSession *s = new RefSession();
    s->n = n;
    s->J = make_triplets(nJ, JA, Ji, Jj);
    s->M = make_triplets(nM, MA, Mi, Mj);
    s->M_ = s->M.convert(n);
    s->d = Eigen::Map<const core::eivec>(dxdt, n);
    s->f_ = Eigen::Map<const core::eivec>(f, n);
    return s;
}
int daeref_newton_iterate(void *h, double alpha, double *dx_out, double *times_ms)
{
    RefSession *s = static_cast<RefSession *>(h);
    auto t0 = clk::now();
    core::eimat Jb = s->J.convert(s->n);
    double tc = ms_since(t0);
    t0 = clk::now();
    s->b = s->M_ * s->d;
    s->b -= s->f_;
    Jb -= s->M_ * alpha;
    double tl = ms_since(t0);
    t0 = clk::now();
    s->lin.compute(Jb);
    double tf = ms_since(t0);
    if (s->lin.info() != Eigen::Success)
        return 2;
    t0 = clk::now();
    s->dx = s->lin.solve(s->b);
    double ts = ms_since(t0);
    if (s->lin.info() != Eigen::Success)
        return 3;
    if (dx_out)
        std::memcpy(dx_out, s->dx.data(), sizeof(double) * s->n);
    if (times_ms)
    {
        times_ms[0] = tc;
        times_ms[1] = tl;
        times_ms[2] = tf;
        times_ms[3] = ts;
    }
    return 0;
}
void daeref_newton_close(void *h) { delete static_cast<RefSession *>(h); }

const char *daeref_version() { return "dae-cpp 2.2.0 reference headers + Eigen 3.4.90 (unmodified)"; }

} // extern "C"
