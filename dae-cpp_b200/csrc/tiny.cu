// Tiny systems (n <= 32): one whole Newton iteration of dae-cpp/solver.hpp:431-627 in ONE launch of one CTA, with
// every host<->device hand-off through page-locked memory the kernel maps directly: no copy node, no second launch,
// no stream synchronisation -- the host writes f and the triplet values into mapped buffers, launches, and spins on
// a sequence word the kernel stores behind its results.
//
// Why a separate path: BASELINE configs[0] (examples/flame_propagation as shipped) is a SCALAR equation, and most of
// the reference's own test suite integrates systems of order <= 4.  There the general path's handful of launches and
// copies per iteration (50-60 us) is pure latency; the arithmetic is nothing.  This path keeps the arithmetic on the
// device (there is no CPU fallback anywhere in this library) and cuts the latency to one launch + one PCIe round trip.
//
// What the kernel does, in the reference's order and with its roundings where the result is compared bit for bit:
//   assembly  Jb = J - (M*alpha) on a dense n x n image; duplicate triplets summed in input order from +0.0
//             (sparse-matrix.hpp:124-152, solver.hpp:523)
//   residual  b_i = sum_j M_ij * dxdt_j - f_i in increasing column order, dxdt = BDF(xk, x[0..3]) (solver.hpp:141-217, 516-518)
//   factor    dense LU with partial (row) pivoting -- the role of Eigen::SparseLU::compute (solver.hpp:537); the factors
//             stay in device memory for the iterations that reuse them
//   solve     dx = U^-1 L^-1 P b (solver.hpp:558)
//   update    error checks, xk += dx (solver.hpp:590-619) and the variability accept_step would measure (solver.hpp:703-717)
#include <chrono>
#include <cstring>

#include "bdf.hpp"
#include "ctx.hpp"

namespace daecpp_b200
{
namespace
{
constexpr int TN = TINY_MAX_N;
constexpr int TINY_THREADS = 256;
constexpr int TINY_STAGE = 1024; // triplet values staged in shared memory (longer streams are read in place)

struct TinyArgs
{
    int32_t n, do_refactor, var_on, seq, nTJ, nTM;
    Bdf bdf;          // weights, order and the device history pointers
    double atol, rtol, max_err_abs, var_tol, floor;
    // dense slot lists: for entry e = i + j*n the triplets jidx[jptr[e] .. jptr[e+1]) in input order
    const int32_t *jptr, *jidx, *mptr, *midx;
    // mapped page-locked memory (device addresses)
    const double *f, *jv, *mv; // in
    double *xk_out;            // out: the new iterate for the host
    int32_t *res;              // out: [0] flags, [1] pivot failure, [2] sequence word
    double *var_out;           // out: speculative variability
    // device memory
    double *xk, *b, *dx, *lu;  // lu: n*n factors (column major), then the row permutation as n doubles
    const double *x0;
};

__global__ void __launch_bounds__(TINY_THREADS) k_tiny_iteration(TinyArgs a)
{
    __shared__ double A[TN * TN];
    __shared__ double sjv[TINY_STAGE], smv[TINY_STAGE], sf[TN];
    __shared__ double sdxdt[TN], sb[TN], sx[TN];
    __shared__ int sperm[TN];
    __shared__ int sflag[2];
    const int t = threadIdx.x, n = a.n;
    if (t < 2) sflag[t] = 0;
    // every value that lives in the host's mapped block is fetched once, all loads in flight together: one PCIe round trip
    const bool stj = a.nTJ <= TINY_STAGE, stm = a.nTM <= TINY_STAGE;
    if (stj && a.do_refactor) for (int k = t; k < a.nTJ; k += TINY_THREADS) sjv[k] = a.jv[k];
    if (stm) for (int k = t; k < a.nTM; k += TINY_THREADS) smv[k] = a.mv[k];
    if (t < n) sf[t] = a.f[t];
    __syncthreads();
    const double *jv = stj ? sjv : a.jv, *mv = stm ? smv : a.mv;
    // ---- dxdt (one thread per unknown) and, if asked, the dense image of Jb
    if (t < n) sdxdt[t] = bdf_at(a.bdf, t);
    if (a.do_refactor)
    {
        for (int e = t; e < n * n; e += TINY_THREADS)
        {
            const int j0 = a.jptr[e], j1 = a.jptr[e + 1], m0 = a.mptr[e], m1 = a.mptr[e + 1];
            double vj = 0.0, vm = 0.0;
            for (int p = j0; p < j1; ++p) vj = __dadd_rn(vj, jv[a.jidx[p]]);
            for (int p = m0; p < m1; ++p) vm = __dadd_rn(vm, mv[a.midx[p]]);
            double v = 0.0; // structurally absent
            if (j1 > j0 && m1 > m0) v = __dsub_rn(vj, __dmul_rn(vm, a.bdf.c[0]));
            else if (j1 > j0) v = __dsub_rn(vj, 0.0);
            else if (m1 > m0) v = __dsub_rn(0.0, __dmul_rn(vm, a.bdf.c[0]));
            A[e] = v;
        }
    }
    __syncthreads();
    // ---- b = M*dxdt - f, row by row, columns ascending, from +0.0
    if (t < n)
    {
        double acc = 0.0;
        for (int j = 0; j < n; ++j)
        {
            const int e = t + j * n, m0 = a.mptr[e], m1 = a.mptr[e + 1];
            if (m1 > m0)
            {
                double vm = 0.0;
                for (int p = m0; p < m1; ++p) vm = __dadd_rn(vm, mv[a.midx[p]]);
                acc = __dadd_rn(acc, __dmul_rn(vm, sdxdt[j]));
            }
        }
        const double bi = __dsub_rn(acc, sf[t]);
        sb[t] = bi;
        a.b[t] = bi;
    }
    // ---- factorisation: dense LU, partial pivoting (largest magnitude in the column, the diagonal on ties)
    if (a.do_refactor)
    {
        if (t < n) sperm[t] = t;
        __syncthreads();
        for (int k = 0; k < n; ++k)
        {
            if (t == 0)
            {   // n <= 32: a serial pivot search costs less than a reduction with its barriers
                int p = k;
                double best = fabs(A[k + k * n]);
                for (int i = k + 1; i < n; ++i)
                {
                    const double v = fabs(A[i + k * n]);
                    if (v > best) { best = v; p = i; }
                }
                if (!(best > 0.0) || best > 1.7976931348623157e308) sflag[1] = 1;
                sflag[0] = p;
            }
            __syncthreads();
            const int p = sflag[0];
            if (p != k)
            {
                for (int j = t; j < n; j += TINY_THREADS)
                {
                    const double tmp = A[k + j * n];
                    A[k + j * n] = A[p + j * n];
                    A[p + j * n] = tmp;
                }
                if (t == 0) { const int q = sperm[k]; sperm[k] = sperm[p]; sperm[p] = q; }
            }
            __syncthreads();
            const double piv = A[k + k * n];
            for (int i = k + 1 + t; i < n; i += TINY_THREADS) A[i + k * n] = A[i + k * n] / piv;
            __syncthreads();
            const int m = n - k - 1;
            for (int e = t; e < m * m; e += TINY_THREADS)
            {
                const int i = k + 1 + e % m, j = k + 1 + e / m;
                A[i + j * n] = fma(-A[i + k * n], A[k + j * n], A[i + j * n]);
            }
            __syncthreads();
        }
        for (int e = t; e < n * n; e += TINY_THREADS) a.lu[e] = A[e];
        if (t < n) a.lu[n * n + t] = (double)sperm[t];
        if (t == 0) sflag[0] = 0;
    }
    else
    {
        for (int e = t; e < n * n; e += TINY_THREADS) A[e] = a.lu[e];
        if (t < n) sperm[t] = (int)a.lu[n * n + t];
    }
    __syncthreads();
    // ---- triangular solves (thread 0: at most 32 x 32)
    if (t == 0)
    {
        for (int i = 0; i < n; ++i)
        {
            double y = sb[sperm[i]];
            for (int j = 0; j < i; ++j) y = fma(-A[i + j * n], sx[j], y);
            sx[i] = y;
        }
        for (int i = n - 1; i >= 0; --i)
        {
            double y = sx[i];
            for (int j = i + 1; j < n; ++j) y = fma(-A[i + j * n], sx[j], y);
            sx[i] = y / A[i + i * n];
        }
    }
    __syncthreads();
    // ---- error check, update, speculative variability (solver.hpp:590-619, 703-717)
    int flag = 0;
    double v = 0.0;
    if (t < n)
    {
        const double d = sx[t], e = fabs(d), xa0 = a.x0[t], xa = fabs(xa0);
        a.dx[t] = d;
        if (e > a.max_err_abs || isnan(d)) flag |= DAECPP_B200_DIVERGED;
        if ((xa > 1e-14 && __ddiv_rn(e, xa) > a.rtol) || e > a.atol) flag |= DAECPP_B200_NOT_CONVERGED;
        const double xn = __dadd_rn(a.xk[t], d);
        a.xk[t] = xn;
        a.xk_out[t] = xn;
        if (a.var_on && xa > a.var_tol && fabs(xn) > a.var_tol)
        {
            v = fabs(__ddiv_rn(__dsub_rn(xa0, xn), xa0));
            if (!(v == v)) v = 0.0;
        }
    }
    if (t < 32)
    {
        const unsigned any = __reduce_or_sync(0xffffffffu, (unsigned)flag);
        unsigned long long bits = (unsigned long long)__double_as_longlong(v);
        for (int o = 16; o > 0; o >>= 1)
        {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
            bits = other > bits ? other : bits;
        }
        if (t == 0)
        {
            a.res[0] = (int32_t)any;
            a.res[1] = sflag[1];
            *a.var_out = __longlong_as_double((long long)bits);
        }
    }
    __syncthreads();
    if (t == 0)
    {   // results first, then the word the host spins on
        __threadfence_system();
        *reinterpret_cast<volatile int32_t *>(a.res + 2) = a.seq;
    }
}

// dense slot lists of a triplet stream: per entry e = i + j*n the triplet indices in input order
void slot_lists(int32_t n, const std::vector<int32_t> &ti, const std::vector<int32_t> &tj, std::vector<int32_t> &ptr, std::vector<int32_t> &idx)
{
    ptr.assign((size_t)n * n + 1, 0);
    for (size_t k = 0; k < ti.size(); ++k) ptr[(size_t)ti[k] + (size_t)tj[k] * n + 1]++;
    for (size_t e = 0; e < (size_t)n * n; ++e) ptr[e + 1] += ptr[e];
    idx.assign(ti.size(), 0);
    std::vector<int32_t> fill(ptr.begin(), ptr.end() - 1);
    for (size_t k = 0; k < ti.size(); ++k) idx[(size_t)fill[(size_t)ti[k] + (size_t)tj[k] * n]++] = (int32_t)k;
}

template <class T> int upload_small(daecpp_b200_ctx *c, DevBuf<T> &d, const std::vector<T> &h)
{
    DAECPP_CUDA_TRY(c, d.reserve(h.size() ? h.size() : 1));
    if (!h.empty()) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(d.p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, c->stream));
    return DAECPP_B200_OK;
}

int ensure_mapped(daecpp_b200_ctx *c, size_t nTJ, size_t nTM)
{
    TinyState &T = c->tiny;
    const size_t n = (size_t)c->n;
    if (T.h && nTJ <= T.capJ && nTM <= T.capM) return DAECPP_B200_OK;
    // the device may still be reading the old block
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    std::vector<double> keepJ, keepM;
    if (T.h) { keepJ.assign(T.h_jv, T.h_jv + T.capJ); keepM.assign(T.h_mv, T.h_mv + T.capM); cudaFreeHost(T.h); T.h = nullptr; }
    const size_t capJ = std::max<size_t>(nTJ + nTJ / 2 + 8, T.capJ), capM = std::max<size_t>(nTM + nTM / 2 + 8, T.capM);
    const size_t doubles = 2 * n + capJ + capM + 2; // f, xk_out, jv, mv, var + 1 spare
    const size_t bytes = sizeof(double) * doubles + sizeof(int32_t) * 8;
    DAECPP_CUDA_TRY(c, cudaHostAlloc((void **)&T.h, bytes, cudaHostAllocMapped));
    std::memset(T.h, 0, bytes);
    void *dp = nullptr;
    DAECPP_CUDA_TRY(c, cudaHostGetDevicePointer(&dp, T.h, 0));
    T.d = static_cast<char *>(dp);
    double *hd = reinterpret_cast<double *>(T.h);
    T.h_f = hd; T.h_xk = hd + n; T.h_jv = hd + 2 * n; T.h_mv = T.h_jv + capJ; T.h_var = T.h_mv + capM;
    T.h_res = reinterpret_cast<int32_t *>(hd + doubles);
    if (!keepJ.empty()) std::memcpy(T.h_jv, keepJ.data(), sizeof(double) * keepJ.size());
    if (!keepM.empty()) std::memcpy(T.h_mv, keepM.data(), sizeof(double) * keepM.size());
    T.capJ = capJ; T.capM = capM;
    return DAECPP_B200_OK;
}

template <class P> P *dev_of(const TinyState &T, P *host_ptr) { return reinterpret_cast<P *>(T.d + (reinterpret_cast<char *>(host_ptr) - T.h)); }

} // namespace

bool tiny_eligible(const daecpp_b200_ctx *c) { return c->tiny.enabled && c->n <= TINY_MAX_N && !c->lanes; }

// Triplet hand-off of a tiny system: the (i, j) stream is compared with the cached one on the HOST, the values go to
// the mapped block, nothing is launched.  The regular device-side staging is brought up to date only when an entry
// point outside this path asks for it (tiny_flush).  Returns EINDEX for an index out of range.
int tiny_set_triplets(daecpp_b200_ctx *c, bool jac, const double *A, const int32_t *i, const int32_t *j, int64_t nT, bool values_only)
{
    TinyState &T = c->tiny;
    std::vector<int32_t> &ci = jac ? T.Ji : T.Mi, &cj = jac ? T.Jj : T.Mj;
    bool &have = jac ? T.haveJ : T.haveM;
    if (values_only)
    {
        if (!have || (int64_t)ci.size() != nT) { c->err = "set_jacobian_values: no cached pattern of this length"; return DAECPP_B200_EINVAL; }
    }
    else
    {
        bool same = have && (int64_t)ci.size() == nT;
        for (int64_t k = 0; same && k < nT; ++k) same = ci[(size_t)k] == i[k] && cj[(size_t)k] == j[k];
        if (!same)
        {
            for (int64_t k = 0; k < nT; ++k)
                if ((uint32_t)i[k] >= (uint32_t)c->n || (uint32_t)j[k] >= (uint32_t)c->n)
                {
                    have = false;
                    c->err = "sparse matrix index out of boundaries";
                    return DAECPP_B200_EINDEX;
                }
            ci.assign(i, i + nT); cj.assign(j, j + nT);
            std::vector<int32_t> ptr, idx;
            slot_lists(c->n, ci, cj, ptr, idx);
            int rc;
            if ((rc = upload_small(c, jac ? T.jptr : T.mptr, ptr))) return rc;
            if ((rc = upload_small(c, jac ? T.jidx : T.midx, idx))) return rc;
            DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream)); // ptr / idx are locals
            have = true;
            T.have_factor = false;
        }
    }
    int rc = ensure_mapped(c, jac ? (size_t)nT : T.Ji.size(), jac ? T.Mi.size() : (size_t)nT);
    if (rc) return rc;
    if (nT > 0) std::memcpy(jac ? T.h_jv : T.h_mv, A, sizeof(double) * (size_t)nT);
    (jac ? T.staleJ : T.staleM) = true;
    if (jac) c->direct_csc = false;
    return DAECPP_B200_OK;
}

// An entry point outside the tiny path needs the regular device-side triplet staging: replay the cached streams.
int tiny_flush(daecpp_b200_ctx *c)
{
    TinyState &T = c->tiny;
    int rc = DAECPP_B200_OK;
    if (T.staleM && T.haveM)
    {
        T.staleM = false;
        if ((rc = assemble_triplets(c, c->M, T.h_mv, T.Mi.data(), T.Mj.data(), (int64_t)T.Mi.size(), false))) return rc;
        if ((rc = build_mass_csr(c))) return rc;
    }
    if (T.staleJ && T.haveJ)
    {
        T.staleJ = false;
        if ((rc = assemble_triplets(c, c->J, T.h_jv, T.Ji.data(), T.Jj.data(), (int64_t)T.Ji.size(), true))) return rc;
        c->have_factor = false;
    }
    return rc;
}

// One Newton iteration; w: BDF weights.  xk_host may be NULL.
int tiny_iteration(daecpp_b200_ctx *c, const double *f_host, int32_t do_refactor, int32_t order, const double w[5], double atol, double rtol,
                   double max_err_abs, double *xk_host, int32_t *flags)
{
    TinyState &T = c->tiny;
    if (!T.haveJ || !T.haveM) { c->err = "set_mass/set_jacobian must precede"; return DAECPP_B200_EINVAL; }
    if (!do_refactor && !T.have_factor) { c->err = "no valid factorisation"; return DAECPP_B200_EINVAL; }
    int rc = ensure_mapped(c, T.Ji.size(), T.Mi.size());
    if (rc) return rc;
    const size_t n = (size_t)c->n;
    DAECPP_CUDA_TRY(c, T.lu.reserve(n * n + n));
    std::memcpy(T.h_f, f_host, sizeof(double) * n);
    TinyArgs a{};
    a.n = c->n; a.do_refactor = do_refactor ? 1 : 0; a.var_on = 1; a.seq = (int32_t)++T.seq;
    a.nTJ = (int32_t)T.Ji.size(); a.nTM = (int32_t)T.Mi.size();
    a.bdf = make_bdf(c, order, w);
    a.atol = atol; a.rtol = rtol; a.max_err_abs = max_err_abs; a.var_tol = T.var_tol; a.floor = c->refine_floor;
    a.jptr = T.jptr.p; a.jidx = T.jidx.p; a.mptr = T.mptr.p; a.midx = T.midx.p;
    a.f = dev_of(T, T.h_f); a.jv = dev_of(T, T.h_jv); a.mv = dev_of(T, T.h_mv);
    a.xk_out = dev_of(T, T.h_xk); a.res = dev_of(T, T.h_res); a.var_out = dev_of(T, T.h_var);
    a.xk = c->xk; a.b = c->b; a.dx = c->dx; a.lu = T.lu.p; a.x0 = c->xh[0];
    k_tiny_iteration<<<1, TINY_THREADS, 0, c->stream>>>(a);
    c->kernels += 1;
    {   // spin on the sequence word; a long wait falls back to the stream
        volatile int32_t *seq = T.h_res + 2;
        const auto t0 = std::chrono::steady_clock::now();
        bool seen = false;
        for (uint32_t spins = 0;; ++spins)
        {
            if (*seq == a.seq) { seen = true; break; }
            __builtin_ia32_pause();
            if ((spins & 1023u) == 1023u && std::chrono::steady_clock::now() - t0 > std::chrono::milliseconds(2)) break;
        }
        if (!seen)
        {
            DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
            if (*seq != a.seq) { c->err = "tiny iteration did not complete"; return DAECPP_B200_ECUDA; }
        }
    }
    if (do_refactor)
    {
        T.have_factor = T.h_res[1] == 0;
        c->n_refactor += 1;
        c->last_alpha = w[0];
    }
    c->n_solve += 1;
    T.spec_var = *T.h_var;
    T.spec_valid = true;
    c->have_factor = false; // the regular path's factors (if any) are not these
    std::memcpy(c->h_scalars + 2, w, sizeof(double) * 5);
    c->h_flags[6] = order;
    if (do_refactor && !T.have_factor) return DAECPP_B200_EDECOMP;
    *flags = T.h_res[0] & (DAECPP_B200_DIVERGED | DAECPP_B200_NOT_CONVERGED);
    if (xk_host) std::memcpy(xk_host, T.h_xk, sizeof(double) * n);
    return DAECPP_B200_OK;
}

void tiny_release(daecpp_b200_ctx *c)
{
    TinyState &T = c->tiny;
    if (T.h) cudaFreeHost(T.h);
    T.h = nullptr;
    T.jptr.release(); T.jidx.release(); T.mptr.release(); T.midx.release(); T.lu.release();
}

} // namespace daecpp_b200
