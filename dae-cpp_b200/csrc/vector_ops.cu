// Fused, HBM-resident vector kernels of the Newton iteration.
//
//   k_residual : dxdt = BDF(xk, x[0..3]) (dae-cpp/solver.hpp:141-217) fused with
//                b = M*dxdt - f            (dae-cpp/solver.hpp:516-518)
//   k_update   : error checks and xk += dx (dae-cpp/solver.hpp:590-619)
//   k_accept   : history shift + max relative variability (dae-cpp/solver.hpp:698-721)
//
// Arithmetic is written with __dmul_rn/__dadd_rn/__dsub_rn so that nvcc cannot contract to
// FMA: the reference's default build has none, and b / xk are compared bit for bit.
#include "bdf.hpp"
#include "ctx.hpp"

namespace daecpp_b200
{
namespace
{
constexpr int TPB = 256;

__global__ void k_dxdt(Bdf s, int32_t n, double *__restrict__ out)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = bdf_at(s, i);
}

// One thread per row; the row's mass entries are visited in increasing column order.
__global__ void k_residual(Bdf s, int32_t n, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ rowslot,
                           const int32_t *__restrict__ rowcol, const double *__restrict__ mvals,
                           const double *__restrict__ f, double *__restrict__ b, int32_t *__restrict__ clear)
{
    if (clear && blockIdx.x == 0 && threadIdx.x < 3) clear[threadIdx.x] = 0;
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    double acc = 0.0;
    const int32_t e = rowptr[r + 1];
    for (int32_t p = rowptr[r]; p < e; ++p)
        acc = __dadd_rn(acc, __dmul_rn(mvals[rowslot[p]], bdf_at(s, rowcol[p])));
    b[r] = __dsub_rn(acc, f[r]);
}

// M with exactly one entry per row on the diagonal and slots in row order (MassMatrixIdentity, any
// diagonal mass): no index is read, two rows per thread with 16-byte accesses.  The sum still
// starts from +0.0 like the general row loop, so b is bit-identical.
__global__ void k_residual_diag(Bdf s, int32_t n, const double *__restrict__ mvals, const double *__restrict__ f,
                                double *__restrict__ b, int32_t *__restrict__ clear)
{
    // first kernel of a fused iteration: the iteration's flag words are cleared here (every later
    // kernel waits for this grid before it can raise one), which saves a memset node on the stream
    if (clear && blockIdx.x == 0 && threadIdx.x < 3) clear[threadIdx.x] = 0;
    // the factor kernel that follows on the stream is launched programmatically dependent (chain.cu):
    // let it start launching now; it waits for this grid's memory before touching global data
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int64_t i = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 2;
    if (i + 1 < n)
    {
        const double2 xk = *reinterpret_cast<const double2 *>(s.xk + i), x0 = *reinterpret_cast<const double2 *>(s.x0 + i);
        double2 x1 = make_double2(0, 0), x2 = x1, x3 = x1;
        if (s.order >= 2) x1 = *reinterpret_cast<const double2 *>(s.x1 + i);
        if (s.order >= 3) x2 = *reinterpret_cast<const double2 *>(s.x2 + i);
        if (s.order >= 4) x3 = *reinterpret_cast<const double2 *>(s.x3 + i);
        const double2 m = *reinterpret_cast<const double2 *>(mvals + i), ff = *reinterpret_cast<const double2 *>(f + i);
        double2 r;
        r.x = __dsub_rn(__dadd_rn(0.0, __dmul_rn(m.x, bdf_one(s, xk.x, x0.x, x1.x, x2.x, x3.x))), ff.x);
        r.y = __dsub_rn(__dadd_rn(0.0, __dmul_rn(m.y, bdf_one(s, xk.y, x0.y, x1.y, x2.y, x3.y))), ff.y);
        *reinterpret_cast<double2 *>(b + i) = r;
    }
    else if (i < n)
        b[i] = __dsub_rn(__dadd_rn(0.0, __dmul_rn(mvals[i], bdf_at(s, i))), f[i]);
}
__global__ void k_check_diag(int32_t n, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ rowslot,
                             const int32_t *__restrict__ rowcol, int32_t *__restrict__ flag)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    if (rowptr[r] != r || rowptr[r + 1] != r + 1 || rowslot[r] != r || rowcol[r] != r) *flag = 0;
}

__global__ void k_update(int32_t n, const double *__restrict__ dx, const double *__restrict__ x0, double *__restrict__ xk,
                         double atol, double rtol, double max_err_abs, int32_t *__restrict__ flags)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int f = 0;
    if (i < n)
    {
        const double d = dx[i];
        const double e = fabs(d);
        if (e > max_err_abs || isnan(d)) f |= DAECPP_B200_DIVERGED;
        const double xa = fabs(x0[i]);
        if (xa > 1e-14 && __ddiv_rn(e, xa) > rtol) f |= DAECPP_B200_NOT_CONVERGED;
        if (e > atol) f |= DAECPP_B200_NOT_CONVERGED;
        xk[i] = __dadd_rn(xk[i], d);
    }
    // one atomic per warp that has something to report
    const unsigned any = __reduce_or_sync(0xffffffffu, (unsigned)f);
    if (any && (threadIdx.x & 31) == 0) atomicOr(&flags[2], (int32_t)any);
}

__global__ void k_accept(int32_t n, const double *__restrict__ xk, const double *__restrict__ x0_old,
                         double *__restrict__ x0_new, int variability_on, double tol,
                         unsigned long long *__restrict__ vmax_bits)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    double v = 0.0;
    if (i < n)
    {
        const double a = xk[i];
        x0_new[i] = a;
        if (variability_on)
        {
            const double p = x0_old[i];
            if (fabs(p) > tol && fabs(a) > tol) v = fabs(__ddiv_rn(__dsub_rn(p, a), p));
            if (!(v == v)) v = 0.0; // the reference's `rel_change > variability` ignores NaN
        }
    }
    if (!variability_on) return;
    // non-negative doubles order like their bit patterns; NaN (sign clear) compares above inf
    unsigned long long bits = (unsigned long long)__double_as_longlong(v);
    for (int o = 16; o > 0; o >>= 1)
    {
        unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
        bits = other > bits ? other : bits;
    }
    if ((threadIdx.x & 31) == 0 && bits) atomicMax(vmax_bits, bits);
}

} // namespace

Bdf make_bdf(daecpp_b200_ctx *c, int32_t order, const double coef[5])
{
    Bdf s;
    for (int k = 0; k < 5; ++k) s.c[k] = coef[k];
    s.xk = c->xk; s.x0 = c->xh[0]; s.x1 = c->xh[1]; s.x2 = c->xh[2]; s.x3 = c->xh[3];
    s.order = order;
    return s;
}

int launch_dxdt(daecpp_b200_ctx *c, int32_t order, const double coef[5])
{
    k_dxdt<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(make_bdf(c, order, coef), c->n, c->dxdt);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

// after build_mass_csr: is M diagonal with slots in row order?  (one synchronisation per mass pattern)
int classify_mass(daecpp_b200_ctx *c)
{
    c->m_diag = false;
    if (c->M.nnz != c->n) return DAECPP_B200_OK;
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 7, 0xff, sizeof(int32_t), c->stream));
    k_check_diag<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->m_rowptr.p, c->m_rowslot.p, c->m_rowcol.p, c->d_flags + 7);
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 7, c->d_flags + 7, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    c->m_diag = c->h_flags[7] != 0;
    c->kernels += 1;
    return DAECPP_B200_OK;
}

int launch_residual(daecpp_b200_ctx *c, int32_t order, const double coef[5], bool clear_flags)
{
    if (c->m_diag)
    {
        const int pairs = (c->n + 1) / 2;
        k_residual_diag<<<(pairs + TPB - 1) / TPB, TPB, 0, c->stream>>>(make_bdf(c, order, coef), c->n, c->M.vals.p, c->f, c->b, clear_flags ? c->d_flags + 2 : nullptr);
        c->kernels += 1;
        DAECPP_CUDA_TRY(c, cudaGetLastError());
        return DAECPP_B200_OK;
    }
    k_residual<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(make_bdf(c, order, coef), c->n, c->m_rowptr.p, c->m_rowslot.p,
                                                             c->m_rowcol.p, c->M.vals.p, c->f, c->b, clear_flags ? c->d_flags + 2 : nullptr);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

int launch_update(daecpp_b200_ctx *c, double atol, double rtol, double max_err_abs)
{
    k_update<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->dx, c->xh[0], c->xk, atol, rtol, max_err_abs, c->d_flags);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

int launch_accept(daecpp_b200_ctx *c, int32_t variability_on, double tol)
{
    // rotate the history by pointer: the oldest buffer becomes the new x[0]
    double *oldest = c->xh[3];
    const double *x0_old = c->xh[0];
    c->xh[3] = c->xh[2]; c->xh[2] = c->xh[1]; c->xh[1] = c->xh[0]; c->xh[0] = oldest;
    k_accept<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->xk, x0_old, oldest, variability_on, tol,
                                                           (unsigned long long *)c->d_scalars);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

} // namespace daecpp_b200
