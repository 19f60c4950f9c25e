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

// partial diagonal: dslot[r] = slot of M_rr, -1 for a row without elements; *flag = 0 if a row holds anything else
__global__ void k_diag_slots(int32_t n, const int32_t *__restrict__ rowptr, const int32_t *__restrict__ rowslot,
                             const int32_t *__restrict__ rowcol, int32_t *__restrict__ dslot, int32_t *__restrict__ flag)
{
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    const int32_t p = rowptr[r], cnt = rowptr[r + 1] - p;
    int32_t s = -1;
    if (cnt == 1 && rowcol[p] == r) s = rowslot[p];
    else if (cnt != 0) *flag = 0;
    dslot[r] = s;
}

__global__ void k_update(int32_t n, const double *__restrict__ dx, const double *__restrict__ x0, double *__restrict__ xk,
                         double atol, double rtol, double max_err_abs, double floor, int32_t *__restrict__ flags)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int f = 0;
    if (i < n)
    {
        const double d = dx[i];
        const double e = fabs(d);
        if (e > max_err_abs || isnan(d)) f |= DAECPP_B200_DIVERGED;
        const double xa = fabs(x0[i]);
        if ((xa > 1e-14 && __ddiv_rn(e, xa) > rtol) || e > atol)
            f |= e > floor ? (DAECPP_B200_NOT_CONVERGED | DAECPP_B200_ABOVE_FLOOR) : DAECPP_B200_NOT_CONVERGED;
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


// ---- guarded iterative refinement (ctx.hpp: refine_steps) -------------------------------------------------
// One step r = b - Jb*dx, Jb*e = r, dx += e for the Newton update just computed; Jb = J - (M*alpha) is applied
// row by row from the triplet values (deterministic order, fused multiply-adds: r is wanted as accurately as
// FP64 gives it, it is not compared bit for bit with anything).  Works on a single system or on the lanes of a
// batch (blockIdx.y, idle lanes skipped).
struct RefineArgs
{
    int32_t n, lanes;
    int64_t s_vec, s_hist, s_jv, s_mv;
    const int32_t *mode, *lane_list;
    const int32_t *jrowptr, *jrowslot, *jrowcol, *jsrc; // jsrc: slot -> triplet whose value it is (values stay in triplet order), or null
    const double *jv;
    const int32_t *mrowptr, *mrowslot, *mrowcol;
    const double *mv;
    const int32_t *mslot; // partial-diagonal mass: row -> slot of M_rr or -1 (null: slot = row)
    double alpha;
    const double *alpha_arr;
    const double *b; // the right-hand side of the solve, or null: diagonal mass, b_i = M_ii*dxdt(xk_i - dx_i) - f_i is formed again
    Bdf bdf;
    const double *lane_w;
    const int32_t *lane_order, *lane_hbase;
    const double *hist;
    const double *f, *dx_in;
    double *xk, *dx, *r;
    const double *e;
    const double *x0;
    double atol, rtol, max_err_abs, floor;
    int32_t *flags; // single: d_flags (word [2]); lanes: [lanes][4] (word [0])
    unsigned long long *var;
    int32_t var_on;
    double var_tol;
    int32_t update; // 0: dx += e only (linear-solver entry: no iterate, no flags)
};

__device__ __forceinline__ Bdf refine_bdf(const RefineArgs &a, int lane)
{
    if (!a.lane_w) return a.bdf;
    Bdf s;
    const int hb = a.lane_hbase[lane];
    const int64_t off = (int64_t)lane * a.s_vec;
    s.order = a.lane_order[lane];
#pragma unroll
    for (int k = 0; k < 5; ++k) s.c[k] = a.lane_w[(int64_t)k * a.lanes + lane];
    s.xk = a.xk + off;
    s.x0 = a.hist + (int64_t)(hb & 3) * a.s_hist + off; s.x1 = a.hist + (int64_t)((hb + 1) & 3) * a.s_hist + off;
    s.x2 = a.hist + (int64_t)((hb + 2) & 3) * a.s_hist + off; s.x3 = a.hist + (int64_t)((hb + 3) & 3) * a.s_hist + off;
    return s;
}

__global__ void k_refine_residual(RefineArgs a)
{
    const int lane = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    if (a.mode && a.mode[lane] == 0) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const int64_t off = (int64_t)lane * a.s_vec;
    const double *dx = a.dx_in + off;
    const double *mv = a.mv + (int64_t)lane * a.s_mv;
    double acc;
    if (a.b) acc = a.b[off + i];
    else
    {   // xk already holds xk + dx
        const Bdf s = refine_bdf(a, lane);
        const double xo = __dsub_rn(s.xk[i], dx[i]);
        const double x1 = s.order >= 2 ? s.x1[i] : 0.0, x2 = s.order >= 3 ? s.x2[i] : 0.0, x3 = s.order >= 4 ? s.x3[i] : 0.0;
        const int64_t slot = a.mslot ? (int64_t)a.mslot[i] : i;
        acc = slot >= 0 ? __dsub_rn(__dadd_rn(0.0, __dmul_rn(mv[slot], bdf_one(s, xo, s.x0[i], x1, x2, x3))), a.f[off + i]) : __dsub_rn(0.0, a.f[off + i]);
    }
    const double *jv = a.jv + (int64_t)lane * a.s_jv;
    for (int32_t p = a.jrowptr[i], e = a.jrowptr[i + 1]; p < e; ++p)
    {
        const int32_t slot = a.jrowslot[p];
        acc = fma(-jv[a.jsrc ? a.jsrc[slot] : slot], dx[a.jrowcol[p]], acc);
    }
    const double alpha = a.alpha_arr ? a.alpha_arr[lane] : a.alpha;
    for (int32_t p = a.mrowptr[i], e = a.mrowptr[i + 1]; p < e; ++p)
        acc = fma(__dmul_rn(mv[a.mrowslot[p]], alpha), dx[a.mrowcol[p]], acc);
    a.r[off + i] = acc;
}

// dx += e, xk += e and the error checks of dae-cpp/solver.hpp:590-619 on the corrected update
__global__ void k_refine_apply(RefineArgs a)
{
    const int lane = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    if (a.mode && a.mode[lane] == 0) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t off = (int64_t)lane * a.s_vec;
    int f = 0;
    double v = 0.0;
    if (i < a.n)
    {
        const double e = a.e[off + i];
        if (!a.update)
        {   // linear-solver entry: corrections at the level of the last bits are rounding noise of the correction
            // solve itself; a solution that is already that accurate is left exactly as it is
            const double d0 = a.dx[off + i];
            if (fabs(e) > 4.0 * 2.220446049250313e-16 * fabs(d0)) a.dx[off + i] = __dadd_rn(d0, e);
            return;
        }
        const double d = __dadd_rn(a.dx[off + i], e);
        a.dx[off + i] = d;
        const double *x0 = a.lane_hbase ? a.hist + (int64_t)(a.lane_hbase[lane] & 3) * a.s_hist + off : a.x0;
        const double xn = __dadd_rn(a.xk[off + i], e);
        a.xk[off + i] = xn;
        const double ad = fabs(d), xa = fabs(x0[i]);
        if (ad > a.max_err_abs || isnan(d)) f |= DAECPP_B200_DIVERGED;
        if ((xa > 1e-14 && __ddiv_rn(ad, xa) > a.rtol) || ad > a.atol)
            f |= ad > a.floor ? (DAECPP_B200_NOT_CONVERGED | DAECPP_B200_ABOVE_FLOOR) : DAECPP_B200_NOT_CONVERGED;
        if (a.var_on && xa > a.var_tol && fabs(xn) > a.var_tol) v = fabs(__ddiv_rn(__dsub_rn(x0[i], xn), x0[i]));
    }
    const unsigned any = __reduce_or_sync(0xffffffffu, (unsigned)f);
    if (any && (threadIdx.x & 31) == 0) atomicOr(a.lane_hbase ? &a.flags[4 * lane] : &a.flags[2], (int32_t)any);
    if (a.var_on && a.var)
    {
        unsigned long long bits = (unsigned long long)__double_as_longlong(v);
        for (int o = 16; o > 0; o >>= 1)
        {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
            bits = other > bits ? other : bits;
        }
        if ((threadIdx.x & 31) == 0 && bits) atomicMax(a.var + lane, bits);
    }
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
    c->m_pdiag = false;
    if (c->M.nnz > c->n) return DAECPP_B200_OK;
    const bool full = c->M.nnz == c->n;
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 7, 0xff, sizeof(int32_t), c->stream));
    if (full) k_check_diag<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->m_rowptr.p, c->m_rowslot.p, c->m_rowcol.p, c->d_flags + 7);
    else
    {
        DAECPP_CUDA_TRY(c, c->m_dslot.reserve((size_t)c->n));
        k_diag_slots<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->m_rowptr.p, c->m_rowslot.p, c->m_rowcol.p, c->m_dslot.p, c->d_flags + 7);
    }
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 7, c->d_flags + 7, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    (full ? c->m_diag : c->m_pdiag) = c->h_flags[7] != 0;
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
    k_update<<<(c->n + TPB - 1) / TPB, TPB, 0, c->stream>>>(c->n, c->dx, c->xh[0], c->xk, atol, rtol, max_err_abs, c->refine_floor, c->d_flags);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}


int refine_last_solve(daecpp_b200_ctx *c, double alpha, double atol, double rtol, double max_err_abs, bool b_is_stored, bool update)
{
    int rc;
    const LaneSet *L = c->lanes;
    cudaStream_t st = c->stream;
    if ((rc = build_jac_csr(c))) return rc;
    c->h_flags[4] = 0; // (the sort behind build_jac_csr parks a count there)
    if (!c->jvals_from_triplets && !L && (rc = ensure_csc_values(c, c->J))) return rc;
    double *r, *e;
    if (L) { r = L->refine_r; e = L->refine_e; }
    else
    {
        const size_t ns = ((size_t)c->n + 31) / 32 * 32;
        DAECPP_CUDA_TRY(c, c->refine_r.reserve(ns));
        DAECPP_CUDA_TRY(c, c->refine_e.reserve(ns));
        r = c->refine_r.p; e = c->refine_e.p;
    }
    if (!r || !e) { c->err = "refinement buffers missing"; return DAECPP_B200_EINVAL; }
    RefineArgs a{};
    a.n = c->n;
    a.lanes = L ? L->count : 0;
    a.jrowptr = c->j_rowptr.p; a.jrowslot = c->j_rowslot.p; a.jrowcol = c->j_rowcol.p;
    a.jsrc = c->jvals_from_triplets ? c->J.src.p : nullptr;
    a.mrowptr = c->m_rowptr.p; a.mrowslot = c->m_rowslot.p; a.mrowcol = c->m_rowcol.p;
    a.mslot = c->m_pdiag ? c->m_dslot.p : nullptr;
    a.alpha = alpha;
    a.f = c->f; a.dx_in = c->dx; a.dx = c->dx; a.xk = c->xk; a.r = r; a.e = e;
    a.b = b_is_stored ? c->b : nullptr;
    a.atol = atol; a.rtol = rtol; a.max_err_abs = max_err_abs; a.floor = c->refine_floor;
    a.update = update ? 1 : 0;
    if (L)
    {
        a.s_vec = L->s_vec; a.s_hist = L->s_hist; a.s_jv = L->s_jv; a.s_mv = L->s_mv;
        a.mode = L->mode; a.lane_list = L->list_active; a.jv = L->jv; a.mv = L->mv; a.alpha_arr = L->w;
        a.lane_w = L->w; a.lane_order = L->order; a.lane_hbase = L->hbase; a.hist = L->hist;
        a.flags = L->flags; a.var = L->var; a.var_on = L->var_on; a.var_tol = L->var_tol;
    }
    else
    {
        a.jv = c->jvals_from_triplets ? c->J.A.p : c->J.vals.p;
        a.mv = c->M.vals.p;
        a.bdf = c->bdf_pending;
        a.x0 = c->xh[0];
        a.flags = c->d_flags;
    }
    const dim3 grid((unsigned)((c->n + TPB - 1) / TPB), (unsigned)(L ? (L->list_active ? L->n_active : L->count) : 1));
    k_refine_residual<<<grid, TPB, 0, st>>>(a);
    // Jb*e = r with the factors at hand: the solve reads c->b and writes c->dx
    double *keep_b = c->b, *keep_dx = c->dx;
    c->b = r; c->dx = e;
    const bool keep_fuse = c->fuse_residual_now, keep_pub = c->publish_flags;
    c->fuse_residual_now = false; c->publish_flags = false;
    rc = numeric_solve(c, false);
    c->fuse_residual_now = keep_fuse; c->publish_flags = keep_pub;
    c->b = keep_b; c->dx = keep_dx;
    if (rc) return rc;
    if (L) DAECPP_CUDA_TRY(c, cudaMemsetAsync(L->flags, 0, sizeof(int32_t) * 4 * (size_t)L->count + sizeof(unsigned long long) * (size_t)L->count, st));
    else DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 2, 0, sizeof(int32_t), st));
    c->flags_clean = false;
    k_refine_apply<<<grid, TPB, 0, st>>>(a);
    c->kernels += 2;
    c->n_refine += 1;
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
