// Numeric multifrontal LU refactorisation and triangular solves on the device.
//
// Role replaced: Eigen::SparseLU::factorize / _solve_impl as called from
// dae-cpp/solver.hpp:537,558 (Eigen/src/SparseLU/SparseLU.h:611-814, 338-358).  The algorithm
// is different by design: the reference interleaves symbolic and numeric work column by column
// with partial pivoting and is strictly sequential; here the pattern-only analysis
// (symbolic.cpp) fixes a shallow assembly tree once, and every refactorisation is a sweep over
// its levels in which all fronts of a level are processed concurrently:
//
//   scatter   Jb = J - (M*alpha) (bit-exact, solver.hpp:523) written straight into front storage
//   level l   for every front: extend-add the children's contribution blocks, dense LU of the
//             pivot block (static pivoting), panel solves, Schur complement -> own contribution
//   solve     forward sweep up the tree with per-front update vectors (deterministic, no
//             atomics), backward sweep down the tree
//
// A front is worked on by a cooperative group of G threads (4/8/32 lanes of a warp, or a whole
// CTA); G is picked per level from the largest front in it.
#include <algorithm>
#include <chrono>

#include "ctx.hpp"

namespace daecpp_b200
{
namespace
{

struct FrontTables
{
    const int32_t *sn_first, *sn_rows, *sn_rel, *child_ptr, *child_idx, *level_sn;
    const int64_t *sn_rowptr, *sn_lu, *sn_cb;
};

template <int G>
struct Grp
{
    static constexpr int FRONTS_PER_CTA = (G <= 32) ? (128 / G) : 1;
    static constexpr int THREADS = (G <= 32) ? 128 : G;
    __device__ static __forceinline__ int lane() { return (G <= 32) ? (threadIdx.x % G) : threadIdx.x; }
    __device__ static __forceinline__ int slot() { return (G <= 32) ? (threadIdx.x / G) : 0; }
    __device__ static __forceinline__ void sync()
    {
        if (G > 32) __syncthreads();
        else if (G == 32) __syncwarp();
        else
        {
            const unsigned base = ((threadIdx.x & 31) / G) * G;
            __syncwarp((((1u << G) - 1u)) << base);
        }
    }
};

// Jb slot values on the union pattern, bit-exact with `Jb -= M_*alpha` (solver.hpp:523): parity export.
__global__ void k_jb_values(int64_t nnzU, const int32_t *__restrict__ uJ, const int32_t *__restrict__ uM,
                            const double *__restrict__ jv, const double *__restrict__ mv, double alpha,
                            double *__restrict__ uvals)
{
    int64_t u = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (u >= nnzU) return;
    const int32_t a = uJ[u], b = uM[u];
    const double m = (b >= 0) ? __dmul_rn(mv[b], alpha) : 0.0;
    const double j = (a >= 0) ? jv[a] : 0.0;
    uvals[u] = __dsub_rn(j, m);
}

template <int G>
__global__ void __launch_bounds__(Grp<G>::THREADS) k_factor_level(FrontTables T, int32_t begin, int32_t count,
                                                                  double *__restrict__ LU, double *__restrict__ CB,
                                                                  double pivot_tol, int32_t *__restrict__ flags)
{
    const int32_t idx = blockIdx.x * Grp<G>::FRONTS_PER_CTA + Grp<G>::slot();
    if (idx >= count) return;
    const int t = Grp<G>::lane();
    const int32_t f = T.level_sn[begin + idx];
    const int32_t s = T.sn_first[f + 1] - T.sn_first[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - T.sn_rowptr[f]);
    const int32_t ld = s + r;
    double *Lp = LU + T.sn_lu[f];
    double *Ut = Lp + (int64_t)ld * s;
    double *C = CB + T.sn_cb[f];
    const int32_t rr = r * r;

    for (int32_t e = t; e < rr; e += G) C[e] = 0.0;
    Grp<G>::sync();

    for (int32_t ci = T.child_ptr[f]; ci < T.child_ptr[f + 1]; ++ci)
    {
        const int32_t c = T.child_idx[ci];
        const int64_t rp = T.sn_rowptr[c];
        const int32_t rc = (int32_t)(T.sn_rowptr[c + 1] - rp);
        const double *Cc = CB + T.sn_cb[c];
        const int32_t *rel = T.sn_rel + rp;
        for (int32_t e = t; e < rc * rc; e += G)
        {
            const int32_t a = e % rc, bc = e / rc;
            const int32_t ra = rel[a], rb = rel[bc];
            const double v = Cc[e];
            if (rb < s) Lp[(int64_t)rb * ld + ra] += v;
            else if (ra < s) Ut[(int64_t)ra * r + (rb - s)] += v;
            else C[(int64_t)(rb - s) * r + (ra - s)] += v;
        }
        Grp<G>::sync();
    }

    // right-looking elimination; the rank-1 updates walk columns by warp and rows by lane (coalesced,
    // no integer division per element)
    constexpr int LW = (G >= 32) ? 32 : G;          // lanes along a column
    constexpr int NW = (G >= 32) ? G / 32 : 1;      // column groups
    const int lane = t % LW, wid = t / LW;
    for (int32_t k = 0; k < s; ++k)
    {
        double *colk = Lp + (int64_t)k * ld;
        const double d = colk[k];
        if (t == 0 && (!(fabs(d) > pivot_tol) || isinf(d))) atomicOr(&flags[3], 1);
        const double inv = 1.0 / d;
        for (int32_t i = k + 1 + t; i < ld; i += G) colk[i] *= inv;
        Grp<G>::sync();
        for (int32_t j = k + 1 + wid; j < s; j += NW)
        {
            double *colj = Lp + (int64_t)j * ld;
            const double ukj = colj[k];
            for (int32_t i = k + 1 + lane; i < ld; i += LW) colj[i] -= colk[i] * ukj;
        }
        const double *utk = Ut + (int64_t)k * r;
        for (int32_t j = k + 1 + wid; j < s; j += NW)
        {
            double *utj = Ut + (int64_t)j * r;
            const double ljk = colk[j];
            for (int32_t i = lane; i < r; i += LW) utj[i] -= utk[i] * ljk;
        }
        Grp<G>::sync();
    }

    // Schur complement C -= F21 * F12, 4 x 1 register blocking along the rows
    for (int32_t bc = wid; bc < r; bc += NW)
        for (int32_t a0 = lane * 4; a0 < r; a0 += LW * 4)
        {
            double acc[4] = {0.0, 0.0, 0.0, 0.0};
            const int na = min(4, r - a0);
            for (int32_t k = 0; k < s; ++k)
            {
                const double u = Ut[(int64_t)k * r + bc];
                const double *l = Lp + (int64_t)k * ld + s + a0;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (q < na) acc[q] += l[q] * u;
            }
            for (int q = 0; q < na; ++q) C[(int64_t)bc * r + a0 + q] -= acc[q];
        }
}

// ---- blocked LU of big fronts -----------------------------------------------------------------
// The fronts at the top of the assembly tree (orders of several hundred to ~1500 on a 1024^2 grid)
// carry almost all the flops of the factorisation.  They are factored level by level, all fronts of
// a level in the same launches, by a right-looking blocked LU with panels of NB columns:
//     panel   one CTA per front: unblocked elimination of columns [k0, k1), rows [k0, ld)
//     trsm    one thread per trailing column: U12 = L11^-1 * A12 (unit lower L11 from shared memory)
//     gemm    64 x 64 tiles over the whole trailing matrix (F11/F21 remainder, F12 remainder and the
//             contribution block): F -= L[:, k0:k1) * U[k0:k1, :), 4 x 4 register blocking, FP64 FMA
// The frontal matrix lives in three pieces (symbolic.hpp): L panel (ld x s, column major), transposed
// U panel (r x s) and contribution block (r x r); fr_at() maps (i, j) of the front onto them.
constexpr int NB = 32;

// Programmatic dependent launch for the panel / trailing-update pairs of a block column: a kernel lets its
// successor start launching at once (its CTAs read the static front tables and then wait), and waits itself
// for its predecessor's memory before touching factor storage.  No-ops in a plain launch.
__device__ __forceinline__ void bf_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void bf_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
template <class... P, class... A> inline void bf_launch_pdl_smem(cudaStream_t st, void (*kernel)(P...), dim3 grid, int threads, size_t smem, A... args);
template <class... P, class... A> inline void bf_launch_pdl(cudaStream_t st, void (*kernel)(P...), dim3 grid, int threads, A... args)
{
    bf_launch_pdl_smem(st, kernel, grid, threads, 0, args...);
}
template <class... P, class... A> inline void bf_launch_pdl_smem(cudaStream_t st, void (*kernel)(P...), dim3 grid, int threads, size_t smem, A... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(threads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<P>(args)...);
}

struct FrontRef
{
    double *Lp, *Ut, *C;
    int32_t s, r, ld;
};
__device__ __forceinline__ FrontRef front_ref(const FrontTables &T, int32_t f, double *LU, double *CB)
{
    FrontRef F;
    F.s = T.sn_first[f + 1] - T.sn_first[f];
    F.r = (int32_t)(T.sn_rowptr[f + 1] - T.sn_rowptr[f]);
    F.ld = F.s + F.r;
    F.Lp = LU + T.sn_lu[f];
    F.Ut = F.Lp + (int64_t)F.ld * F.s;
    F.C = CB + T.sn_cb[f];
    return F;
}
__device__ __forceinline__ double *fr_at(const FrontRef &F, int32_t i, int32_t j)
{
    if (j < F.s) return F.Lp + (int64_t)j * F.ld + i;
    if (i < F.s) return F.Ut + (int64_t)i * F.r + (j - F.s);
    return F.C + (int64_t)(j - F.s) * F.r + (i - F.s);
}

// zero the contribution blocks of a level (grid.y = front)
__global__ void k_bf_zero(FrontTables T, int32_t begin, double *__restrict__ CB)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int64_t r = T.sn_rowptr[f + 1] - T.sn_rowptr[f];
    double *C = CB + T.sn_cb[f];
    bf_pdl_trigger();
    bf_pdl_wait();
    for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < r * r; e += (int64_t)gridDim.x * blockDim.x) C[e] = 0.0;
}
// extend-add of the `ord`-th child of every front of the level (grid.y = front).  Children are added
// one launch after the other: two children may hit the same entry, and the sum must not depend on timing.
__global__ void __launch_bounds__(256) k_bf_extend(FrontTables T, int32_t begin, int32_t ord, double *__restrict__ LU,
                                                   double *__restrict__ CB)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int32_t ci = T.child_ptr[f] + ord;
    if (ci >= T.child_ptr[f + 1]) return;
    const FrontRef F = front_ref(T, f, LU, CB);
    const int32_t s = F.s, r = F.r, ld = F.ld;
    const int32_t c = T.child_idx[ci];
    const int64_t rp = T.sn_rowptr[c];
    const int32_t rc = (int32_t)(T.sn_rowptr[c + 1] - rp);
    const double *Cc = CB + T.sn_cb[c];
    const int32_t *rel = T.sn_rel + rp;
    const int lane = threadIdx.x % 32, wid = threadIdx.x / 32, NW = blockDim.x / 32;
    bf_pdl_trigger();
    bf_pdl_wait();
    for (int32_t bc = blockIdx.x * NW + wid; bc < rc; bc += gridDim.x * NW)
    {
        const int32_t rb = rel[bc];
        constexpr int U = 4; // independent entries per lane: all loads of a batch first, then the stores
        for (int32_t a0 = blockIdx.z * 32 * U; a0 < rc; a0 += gridDim.z * 32 * U)
        {
            double *dst[U];
            double v[U], old[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
            {
                const int32_t a = a0 + u * 32 + lane;
                dst[u] = nullptr;
                if (a < rc)
                {
                    const int32_t ra = rel[a];
                    v[u] = Cc[(int64_t)bc * rc + a];
                    dst[u] = rb < s ? F.Lp + (int64_t)rb * ld + ra : ra < s ? F.Ut + (int64_t)ra * r + (rb - s) : F.C + (int64_t)(rb - s) * r + (ra - s);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (dst[u]) old[u] = *dst[u];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (dst[u]) *dst[u] = old[u] + v[u]; // distinct (a, bc) of one child hit distinct entries
        }
    }
}

// LU of the NB x NB diagonal block held in shared memory (identity padded), by all nt threads of the CTA.
// The block is a dependent chain of NB pivots, so what matters is the latency of one pivot step, and a step
// is kept to: read the pivot, reciprocal (every thread for itself), one fused update per element straight
// from the UNSCALED pivot column (a_ij -= (a_ik * r) * a_kj), one barrier.  The multipliers are scaled once
// at the end.  (Measured on B200: one warp holding the block in registers and broadcasting the pivot row by
// shuffles needs ~1300 cycles per pivot, a full-precision division plus two barriers per pivot ~700.)
// The operations per element and their order do not depend on nt, so every CTA that repeats the
// factorisation gets the same bits.  Rinv[k] = 1 / U(k,k), also used by the row solves.
// look: the reciprocal of pivot k+1 is formed by every thread for itself while pivot k is applied --
// d(k+1) = a(k+1,k+1) - (a(k+1,k) * r(k)) * a(k,k+1), the very operations the element's owner would perform -- so the
// division runs beside the update instead of between two barriers.  The owner leaves that element alone and one
// thread stores d(k) at the start of step k, when nobody reads it any more.  Same bits either way.
__device__ __forceinline__ void bf_block_lu(double (*A)[NB + 1], double *Rinv, int nb, int t, int nt, double pivot_tol, int32_t *flags,
                                            bool look = false)
{
    bool bad = false;
    const int i = t % NB;
    double d_cur = 1.0, r_cur = 1.0;
    if (look && nb > 0)
    {
        d_cur = A[0][0];
        bad = !(fabs(d_cur) > pivot_tol) || isinf(d_cur);
        r_cur = 1.0 / d_cur;
    }
#pragma unroll 1
    for (int k = 0; k < nb; ++k)
    {
        double r;
        if (look)
        {
            r = r_cur;
            if (t == 0) A[k][k] = d_cur;
        }
        else
        {
            const double d = A[k][k];
            bad = bad || !(fabs(d) > pivot_tol) || isinf(d);
            r = 1.0 / d;
        }
        // each thread keeps the same (i, j) elements for every pivot; the loop is unrolled so that all
        // shared-memory loads of a step are issued together, behind the reciprocal
        const double l = __dmul_rn(A[i][k], r);
        const bool more = look && k + 1 < nb;
        double dn = 1.0;
        if (more) dn = __fma_rn(-__dmul_rn(A[k + 1][k], r), A[k][k + 1], A[k + 1][k + 1]);
#pragma unroll
        for (int m = 0; m < NB / 4; ++m)
        {
            const int j = t / NB + m * (nt / NB);
            if (j < NB && j > k && i > k && !(more && i == k + 1 && j == k + 1)) A[i][j] = __fma_rn(-l, A[k][j], A[i][j]);
        }
        if (more)
        {
            bad = bad || !(fabs(dn) > pivot_tol) || isinf(dn);
            d_cur = dn;
            r_cur = 1.0 / dn;
        }
        __syncthreads();
    }
    if (flags && t == 0 && bad) atomicOr(&flags[3], 1);
    if (t < NB) Rinv[t] = 1.0 / A[t][t];
    __syncthreads();
#pragma unroll 4
    for (int e = t; e < NB * NB; e += nt)
    {
        const int ii = e % NB, j = e / NB;
        if (ii > j && j < nb) A[ii][j] *= Rinv[j];
    }
    __syncthreads();
}
// Panel solves against the factored diagonal block A, one thread per row of L21 = A21 * U11^-1 (rows) or per
// column of U12 = L11^-1 * A12 (!rows).  S[k][j] = U(k, k+1+j) resp. L(k+1+j, k), zero beyond the block: the
// thread's NB entries are shifted left by one register per pivot, the finished entry is stored at once, and
// the zero padding makes every step the same NB - 1 updates -- a rolled loop again.  Rows of S start on 16-byte
// boundaries, so a step's coefficients arrive as NB/2 broadcast 16-byte loads issued together, ahead of the FMAs
// (8-byte loads interleaved with the FMAs made the step wait for shared memory 31 times).
__device__ __forceinline__ void bf_fill_pad(double (*S)[NB], const double (*A)[NB + 1], bool rows, int t, int nt)
{
#pragma unroll 8
    for (int e = t; e < NB * NB; e += nt)
    {
        const int k = e / NB, j = e % NB, m = k + 1 + j;
        S[k][j] = m < NB ? (rows ? A[k][m] : A[m][k]) : 0.0;
    }
    __syncthreads();
}
__device__ __forceinline__ void bf_panel_solve(const FrontRef &F, const double (*S)[NB], const double *Rinv, bool rows,
                                               int32_t k0, int nb, int32_t q)
{
    // the thread's NB entries sit at base[k * stride]: a row of the L panel (stride ld), or a column of U12
    // that lies in the L panel (q < s: contiguous) or in the transposed U panel (stride r)
    double *base;
    int64_t stride;
    if (rows) { base = F.Lp + (int64_t)k0 * F.ld + q; stride = F.ld; }
    else if (q < F.s) { base = F.Lp + (int64_t)q * F.ld + k0; stride = 1; }
    else { base = F.Ut + (int64_t)k0 * F.r + (q - F.s); stride = F.r; }
    double x[NB];
#pragma unroll
    for (int k = 0; k < NB; ++k) x[k] = k < nb ? base[k * stride] : 0.0;
#pragma unroll 1
    for (int k = 0; k < nb; ++k)
    {
        double2 c[NB / 2];
        const double2 *row = reinterpret_cast<const double2 *>(S[k]);
#pragma unroll
        for (int j = 0; j < NB / 2; ++j) c[j] = row[j];
        const double xk = rows ? x[0] * Rinv[k] : x[0];
        base[k * stride] = xk;
#pragma unroll
        for (int j = 0; j < NB - 1; ++j) x[j] = x[j + 1] - xk * ((j & 1) ? c[j / 2].y : c[j / 2].x);
        x[NB - 1] = 0.0;
    }
}
template <int NT>
__device__ __forceinline__ void bf_load_diag(double (*A)[NB + 1], const FrontRef &F, int32_t k0, int nb, int t)
{
    constexpr int IT = NB * NB / NT;
    double v[IT]; // all loads of the thread in flight before the first shared-memory store
#pragma unroll
    for (int it = 0; it < IT; ++it)
    {
        const int e = t + it * NT, i = e % NB, j = e / NB; // column j contiguous in i
        v[it] = (i < nb && j < nb) ? F.Lp[(int64_t)(k0 + j) * F.ld + (k0 + i)] : (i == j ? 1.0 : 0.0);
    }
#pragma unroll
    for (int it = 0; it < IT; ++it)
    {
        const int e = t + it * NT;
        A[e % NB][e / NB] = v[it];
    }
    __syncthreads();
}
__device__ __forceinline__ void bf_store_diag(const double (*A)[NB + 1], const FrontRef &F, int32_t k0, int nb, int t, int nt)
{
    for (int e = t; e < NB * NB; e += nt)
    {
        const int i = e % NB, j = e / NB;
        if (i < nb && j < nb) F.Lp[(int64_t)(k0 + j) * F.ld + (k0 + i)] = A[i][j];
    }
}

// panel, step 1 (separate-launch variant): LU of the NB x NB diagonal block, one CTA per front
__global__ void __launch_bounds__(128) k_bf_diag(FrontTables T, int32_t begin, int32_t k0, double *__restrict__ LU,
                                                 double *__restrict__ CB, double pivot_tol, int32_t *__restrict__ flags, bool look)
{
    const int32_t f = T.level_sn[begin + blockIdx.x];
    const FrontRef F = front_ref(T, f, LU, CB);
    if (k0 >= F.s) return;
    const int32_t nb = min(k0 + NB, F.s) - k0;
    __shared__ double A[NB][NB + 1];
    __shared__ double Rinv[NB];
    const int t = threadIdx.x;
    bf_load_diag<128>(A, F, k0, nb, t);
    bf_block_lu(A, Rinv, nb, t, 128, pivot_tol, flags, look);
    bf_store_diag(A, F, k0, nb, t, 128);
}

// panel, step 2 (rows below the diagonal block): one thread per row, L21 = A21 * U11^-1 in registers
__global__ void __launch_bounds__(128) k_bf_rows(FrontTables T, int32_t begin, int32_t k0, double *__restrict__ LU,
                                                 double *__restrict__ CB)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const FrontRef F = front_ref(T, f, LU, CB);
    if (k0 >= F.s) return;
    const int32_t k1 = min(k0 + NB, F.s), nb = k1 - k0;
    if ((int64_t)blockIdx.x * blockDim.x + k1 >= F.ld) return;
    __shared__ double A[NB][NB + 1];
    __shared__ double Rinv[NB];
    __shared__ __align__(16) double P[NB][NB];
    const int t = threadIdx.x;
    bf_load_diag<128>(A, F, k0, nb, t); // factored already (k_bf_diag)
    if (t < NB) Rinv[t] = 1.0 / A[t][t];
    bf_fill_pad(P, A, true, t, 128);
    const int32_t i = k1 + blockIdx.x * blockDim.x + t;
    if (i < F.ld) bf_panel_solve(F, P, Rinv, true, k0, nb, i);
}

// panel, step 3 (columns right of the diagonal block): one thread per column, U12 = L11^-1 * A12
__global__ void __launch_bounds__(128) k_bf_trsm(FrontTables T, int32_t begin, int32_t k0, double *__restrict__ LU,
                                                 double *__restrict__ CB)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const FrontRef F = front_ref(T, f, LU, CB);
    if (k0 >= F.s) return;
    const int32_t k1 = min(k0 + NB, F.s), nb = k1 - k0;
    if ((int64_t)blockIdx.x * blockDim.x + k1 >= F.ld) return;
    __shared__ double A[NB][NB + 1];
    __shared__ __align__(16) double P[NB][NB];
    const int t = threadIdx.x;
    bf_load_diag<128>(A, F, k0, nb, t);
    bf_fill_pad(P, A, false, t, 128);
    const int32_t j = k1 + blockIdx.x * blockDim.x + t;
    if (j < F.ld) bf_panel_solve(F, P, nullptr, false, k0, nb, j);
}

// panel in ONE launch: every CTA factors the (still unfactored) diagonal block for itself in shared memory
// -- 2 us of redundant work instead of a dependent launch -- and then solves its 128 rows of L21
// (even blockIdx.x) or its 128 columns of U12 (odd blockIdx.x) out of registers.  The factored
// diagonal block is written back by the trailing update's first CTA (k_bf_gemm), when no CTA of this
// launch can still be reading the unfactored one: factored a second time there (diagbuf == nullptr), or -- the
// front's first CTA of this launch leaves its copy in diagbuf and reports the pivots -- just copied into place.
__global__ void __launch_bounds__(128) k_bf_panel(FrontTables T, int32_t begin, int32_t k0, double *__restrict__ LU,
                                                  double *__restrict__ CB, double *__restrict__ diagbuf, double pivot_tol,
                                                  int32_t *__restrict__ flags, bool look)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const FrontRef F = front_ref(T, f, LU, CB);
    if (k0 >= F.s) return;
    const int32_t k1 = min(k0 + NB, F.s), nb = k1 - k0;
    const bool rows = (blockIdx.x & 1) == 0;
    const int32_t blk = blockIdx.x >> 1;
    if ((int64_t)blk * 128 + k1 >= F.ld) return;
    __shared__ double A[NB][NB + 1];
    __shared__ double Rinv[NB];
    __shared__ __align__(16) double P[NB][NB];
    const int t = threadIdx.x;
    const int32_t q = k1 + blk * 128 + t;
    bf_pdl_trigger();
    bf_pdl_wait(); // everything above came from the static tables
    if (q < F.ld)
    {   // pull the thread's panel entries towards L2 while the diagonal block is being factored
        const double *base = rows ? F.Lp + (int64_t)k0 * F.ld + q : q < F.s ? F.Lp + (int64_t)q * F.ld + k0 : F.Ut + (int64_t)k0 * F.r + (q - F.s);
        const int64_t stride = rows ? F.ld : q < F.s ? 1 : F.r;
        if (stride != 1)
        {
#pragma unroll 8
            for (int k = 0; k < nb; ++k) asm volatile("prefetch.global.L2 [%0];" ::"l"(base + k * stride));
        }
        else
        {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base + nb - 1));
        }
    }
    bf_load_diag<128>(A, F, k0, nb, t);
    const bool keeper = diagbuf != nullptr && blockIdx.x == 0;
    bf_block_lu(A, Rinv, nb, t, 128, keeper ? pivot_tol : 0.0, keeper ? flags : nullptr, look);
    if (keeper)
    {
        double *D = diagbuf + (int64_t)blockIdx.y * (NB * NB);
#pragma unroll
        for (int e = t; e < NB * NB; e += 128) D[e] = A[e % NB][e / NB];
    }
    bf_fill_pad(P, A, rows, t, 128);
    if (q < F.ld) bf_panel_solve(F, P, Rinv, rows, k0, nb, q);
}

// FP64 tensor-core step: D(8x8) += A(8x4) * B(4x8).  Lane l holds A(l/4, l%4), B(l%4, l/4), D(l/4, 2*(l%4) + {0,1}).
__device__ __forceinline__ void dmma884(double &d0, double &d1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

// Trailing update F -= L[:, k0:k1) * U[k0:k1, :) in 64 x 64 tiles.  DMMA: the eight warps of the CTA hold
// 16 x 32 sub-tiles as 2 x 4 accumulator fragments of mma.m8n8k4.f64 (DMMA), operands from shared memory
// (row stride 68 doubles: the 4 k x 8 rows of a fragment load fall in 32 different banks per half warp).
// !DMMA: 4 x 4 register blocking on the FP64 FMA pipe (kept for A/B runs, option bf_dmma = 0).
// strip != 0: only the part of the trailing matrix that later block columns of THIS front read -- rows or columns inside the
// pivot block -- is updated per block column; the contribution block gets all its updates in one pass at the end (k_bf_schur).
template <bool DMMA>
__global__ void __launch_bounds__(256) k_bf_gemm(FrontTables T, int32_t begin, int32_t k0, double *__restrict__ LU,
                                                 double *__restrict__ CB, double pivot_tol, int32_t *__restrict__ flags, int32_t strip,
                                                 const double *__restrict__ diagbuf, bool look)
{
    const int32_t f = T.level_sn[begin + blockIdx.z];
    const FrontRef F = front_ref(T, f, LU, CB);
    if (k0 >= F.s) return;
    const int32_t k1 = min(k0 + NB, F.s), nb = k1 - k0;
    const int32_t i0 = k1 + blockIdx.x * 64, j0 = k1 + blockIdx.y * 64;
    const int t = threadIdx.x;
    bf_pdl_trigger();
    bf_pdl_wait();
    if (flags && blockIdx.x == gridDim.x - 1)
    {   // one extra CTA per front (fused panel): the panel launch is complete, nobody reads the unfactored
        // diagonal block any more, so it is factored once more and written back -- beside the tiles, not before one
        if (blockIdx.y != 0) return;
        if (diagbuf && k1 < F.ld)
        {   // the panel launch kept its factored copy (and reported the pivots)
            const double *D = diagbuf + (int64_t)blockIdx.z * (NB * NB);
            double v[NB * NB / 256];
#pragma unroll
            for (int it = 0; it < NB * NB / 256; ++it) v[it] = D[t + it * 256];
#pragma unroll
            for (int it = 0; it < NB * NB / 256; ++it)
            {
                const int e = t + it * 256, i = e % NB, j = e / NB;
                if (i < nb && j < nb) F.Lp[(int64_t)(k0 + j) * F.ld + (k0 + i)] = v[it];
            }
            return;
        }
        __shared__ double Dg[NB][NB + 1];
        __shared__ double Rinv[NB];
        bf_load_diag<256>(Dg, F, k0, nb, t);
        bf_block_lu(Dg, Rinv, nb, t, 256, pivot_tol, flags, look);
        bf_store_diag(Dg, F, k0, nb, t, 256);
        return;
    }
    if (i0 >= F.ld || j0 >= F.ld) return;
    if (strip && i0 >= F.s && j0 >= F.s) return; // the tile lies inside the contribution block
    constexpr int LDS = DMMA ? 68 : 65;
    __shared__ double A[NB][LDS], Bm[NB][LDS];
    // DMMA: the read-modify-write of the tile -- addresses and old values (independent loads) requested first, so that
    // they travel together with the operands instead of after the barrier behind them
    const int w = t >> 5, lane = t & 31, g = lane >> 2, c = lane & 3;
    const int wi = (w & 3) * 16, wj = (w >> 2) * 32;
    double *dst[2][4][2];
    double old[2][4][2];
    if (DMMA)
    {
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nj = 0; nj < 4; ++nj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                {
                    const int32_t i = i0 + wi + mi * 8 + g, j = j0 + wj + nj * 8 + 2 * c + e;
                    dst[mi][nj][e] = (i < F.ld && j < F.ld && !(strip && i >= F.s && j >= F.s)) ? fr_at(F, i, j) : nullptr;
                    old[mi][nj][e] = dst[mi][nj][e] ? *dst[mi][nj][e] : 0.0;
                }
    }
    {   // operands: all 2 x 8 loads of a thread in flight before the first shared-memory store
        constexpr int IT = NB * 64 / 256;
        double va[IT], vb[IT];
        const bool in_l = j0 + 63 < F.s; // U(k0+k, j0+q): column major in the L panel (k fast), row major in the U panel (q fast)
#pragma unroll
        for (int it = 0; it < IT; ++it)
        {
            const int e = t + it * 256;
            const int k = e / 64, q = e % 64;
            // A[k][q] = L(i0+q, k0+k): consecutive q are consecutive addresses of column k0+k
            va[it] = (k < nb && i0 + q < F.ld) ? F.Lp[(int64_t)(k0 + k) * F.ld + (i0 + q)] : 0.0;
            const int kb = in_l ? e % NB : k, qb = in_l ? e / NB : q;
            vb[it] = (kb < nb && j0 + qb < F.ld) ? *fr_at(F, k0 + kb, j0 + qb) : 0.0;
        }
#pragma unroll
        for (int it = 0; it < IT; ++it)
        {
            const int e = t + it * 256;
            A[e / 64][e % 64] = va[it];
            if (in_l) Bm[e % NB][e / NB] = vb[it]; else Bm[e / 64][e % 64] = vb[it];
        }
    }
    __syncthreads();
    if (DMMA)
    {
        double acc[2][4][2];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nj = 0; nj < 4; ++nj) acc[mi][nj][0] = acc[mi][nj][1] = 0.0;
#pragma unroll
        for (int kk = 0; kk < NB; kk += 4)
        {
            double av[2], bv[4];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi) av[mi] = A[kk + c][wi + mi * 8 + g];
#pragma unroll
            for (int nj = 0; nj < 4; ++nj) bv[nj] = Bm[kk + c][wj + nj * 8 + g];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                for (int nj = 0; nj < 4; ++nj) dmma884(acc[mi][nj][0], acc[mi][nj][1], av[mi], bv[nj]);
        }
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nj = 0; nj < 4; ++nj)
#pragma unroll
                for (int e = 0; e < 2; ++e)
                    if (dst[mi][nj][e]) *dst[mi][nj][e] = old[mi][nj][e] - acc[mi][nj][e];
        return;
    }
    const int ti = (t % 16) * 4, tj = (t / 16) * 4;
    double acc[4][4];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.0;
#pragma unroll 8
    for (int k = 0; k < NB; ++k)
    {
        double av[4], bv[4];
#pragma unroll
        for (int a = 0; a < 4; ++a) { av[a] = A[k][ti + a]; bv[a] = Bm[k][tj + a]; }
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] += av[a] * bv[b];
    }
#pragma unroll
    for (int b = 0; b < 4; ++b)
    {
        const int32_t j = j0 + tj + b;
        if (j >= F.ld) continue;
#pragma unroll
        for (int a = 0; a < 4; ++a)
        {
            const int32_t i = i0 + ti + a;
            if (i < F.ld && !(strip && i >= F.s && j >= F.s)) *fr_at(F, i, j) -= acc[a][b];
        }
    }
}

// Contribution block of every front of a level in ONE pass: C -= L21 * U12 over all s pivots, 64 x 64 tiles on the FP64
// tensor cores.  Per tile the block columns are applied in order, each as "old - (sum of its 32 products)" -- the
// arithmetic of k_bf_gemm applied block column by block column, so the result is the same bit for bit -- but the tile
// is read and written once instead of once per block column.
__global__ void __launch_bounds__(256) k_bf_schur(FrontTables T, int32_t begin, double *__restrict__ LU, double *__restrict__ CB)
{
    const int32_t f = T.level_sn[begin + blockIdx.z];
    const FrontRef F = front_ref(T, f, LU, CB);
    const int32_t i0 = F.s + blockIdx.x * 64, j0 = F.s + blockIdx.y * 64;
    const int t = threadIdx.x;
    bf_pdl_trigger();
    bf_pdl_wait();
    if (i0 >= F.ld || j0 >= F.ld) return;
    constexpr int LDS = 68;
    __shared__ double A[NB][LDS], Bm[NB][LDS];
    const int w = t >> 5, lane = t & 31, g = lane >> 2, c = lane & 3;
    const int wi = (w & 3) * 16, wj = (w >> 2) * 32;
    double *dst[2][4][2];
    double v[2][4][2];
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int nj = 0; nj < 4; ++nj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
            {
                const int32_t i = i0 + wi + mi * 8 + g, j = j0 + wj + nj * 8 + 2 * c + e;
                dst[mi][nj][e] = (i < F.ld && j < F.ld) ? F.C + (int64_t)(j - F.s) * F.r + (i - F.s) : nullptr;
                v[mi][nj][e] = dst[mi][nj][e] ? *dst[mi][nj][e] : 0.0;
            }
    constexpr int IT = NB * 64 / 256;
    double va[IT], vb[IT];
    auto fetch = [&](int32_t k0) {
        const int32_t nb = min(NB, F.s - k0);
#pragma unroll
        for (int it = 0; it < IT; ++it)
        {
            const int e = t + it * 256;
            const int k = e / 64, q = e % 64;
            va[it] = (k < nb && i0 + q < F.ld) ? F.Lp[(int64_t)(k0 + k) * F.ld + (i0 + q)] : 0.0;
            vb[it] = (k < nb && j0 + q < F.ld) ? F.Ut[(int64_t)(k0 + k) * F.r + (j0 - F.s + q)] : 0.0;
        }
    };
    auto park = [&]() {
#pragma unroll
        for (int it = 0; it < IT; ++it)
        {
            const int e = t + it * 256;
            A[e / 64][e % 64] = va[it];
            Bm[e / 64][e % 64] = vb[it];
        }
    };
    fetch(0);
    park();
    __syncthreads();
    for (int32_t k0 = 0; k0 < F.s; k0 += NB)
    {
        const bool more = k0 + NB < F.s;
        if (more) fetch(k0 + NB); // the next block column's operands are in flight during this one's MMAs
        double acc[2][4][2];
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nj = 0; nj < 4; ++nj) acc[mi][nj][0] = acc[mi][nj][1] = 0.0;
#pragma unroll
        for (int kk = 0; kk < NB; kk += 4)
        {
            double av[2], bv[4];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi) av[mi] = A[kk + c][wi + mi * 8 + g];
#pragma unroll
            for (int nj = 0; nj < 4; ++nj) bv[nj] = Bm[kk + c][wj + nj * 8 + g];
#pragma unroll
            for (int mi = 0; mi < 2; ++mi)
#pragma unroll
                for (int nj = 0; nj < 4; ++nj) dmma884(acc[mi][nj][0], acc[mi][nj][1], av[mi], bv[nj]);
        }
#pragma unroll
        for (int mi = 0; mi < 2; ++mi)
#pragma unroll
            for (int nj = 0; nj < 4; ++nj)
#pragma unroll
                for (int e = 0; e < 2; ++e) v[mi][nj][e] = v[mi][nj][e] - acc[mi][nj][e];
        __syncthreads(); // everybody is done reading this block column's operands
        if (more) park();
        __syncthreads();
    }
#pragma unroll
    for (int mi = 0; mi < 2; ++mi)
#pragma unroll
        for (int nj = 0; nj < 4; ++nj)
#pragma unroll
            for (int e = 0; e < 2; ++e)
                if (dst[mi][nj][e]) *dst[mi][nj][e] = v[mi][nj][e];
}

__global__ void k_permute_in(int32_t n, const int32_t *__restrict__ perm, const double *__restrict__ b, double *__restrict__ y)
{
    int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) y[k] = b[perm[k]];
}
__global__ void k_permute_out(int32_t n, const int32_t *__restrict__ perm, const double *__restrict__ y, double *__restrict__ dx)
{
    int32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < n) dx[perm[k]] = y[k];
}

template <int G>
__global__ void __launch_bounds__(Grp<G>::THREADS) k_forward_level(FrontTables T, int32_t begin, int32_t count,
                                                                   const double *__restrict__ LU, double *__restrict__ upd,
                                                                   double *__restrict__ y)
{
    const int32_t idx = blockIdx.x * Grp<G>::FRONTS_PER_CTA + Grp<G>::slot();
    if (idx >= count) return;
    const int t = Grp<G>::lane();
    const int32_t f = T.level_sn[begin + idx];
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    double *u = upd + rp0;
    double *y1 = y + first;
    for (int32_t a = t; a < r; a += G) u[a] = 0.0;
    Grp<G>::sync();
    for (int32_t ci = T.child_ptr[f]; ci < T.child_ptr[f + 1]; ++ci)
    {
        const int32_t c = T.child_idx[ci];
        const int64_t rp = T.sn_rowptr[c];
        const int32_t rc = (int32_t)(T.sn_rowptr[c + 1] - rp);
        const double *uc = upd + rp;
        const int32_t *rel = T.sn_rel + rp;
        for (int32_t a = t; a < rc; a += G)
        {
            const int32_t ra = rel[a];
            if (ra < s) y1[ra] += uc[a]; else u[ra - s] += uc[a];
        }
        Grp<G>::sync();
    }
    for (int32_t k = 0; k < s; ++k)
    {
        const double yk = y1[k];
        const double *colk = Lp + (int64_t)k * ld;
        for (int32_t i = k + 1 + t; i < ld; i += G)
        {
            const double v = colk[i] * yk;
            if (i < s) y1[i] -= v; else u[i - s] -= v;
        }
        Grp<G>::sync();
    }
}

template <int G>
__global__ void __launch_bounds__(Grp<G>::THREADS) k_backward_level(FrontTables T, int32_t begin, int32_t count,
                                                                    const double *__restrict__ LU, double *__restrict__ y)
{
    const int32_t idx = blockIdx.x * Grp<G>::FRONTS_PER_CTA + Grp<G>::slot();
    if (idx >= count) return;
    const int t = Grp<G>::lane();
    const int32_t f = T.level_sn[begin + idx];
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    const double *Ut = Lp + (int64_t)ld * s;
    const int32_t *rows = T.sn_rows + rp0;
    double *y1 = y + first;
    for (int32_t k = t; k < s; k += G)
    {
        double acc = y1[k];
        const double *utk = Ut + (int64_t)k * r;
        for (int32_t a = 0; a < r; ++a) acc -= utk[a] * y[rows[a]];
        y1[k] = acc;
    }
    Grp<G>::sync();
    for (int32_t j = s - 1; j >= 0; --j)
    {
        const double *colj = Lp + (int64_t)j * ld;
        if (t == 0) y1[j] = y1[j] / colj[j];
        Grp<G>::sync();
        const double xj = y1[j];
        for (int32_t i = t; i < j; i += G) y1[i] -= colj[i] * xj;
        Grp<G>::sync();
    }
}

// ---- triangular solves of big fronts, blocked like the factorisation --------------------------------
// Per level two launches each way.  The triangular part (pivot rows only) runs one CTA per front: a
// block of NB pivots is solved by one warp out of registers (the NB x NB diagonal block is loaded with
// independent loads first, the dependent chain is shuffles only), then the whole CTA applies the block
// to the remaining pivot rows, one thread per row, factor columns read contiguously.  The rectangular
// part (L21 forward, U12 backward) is a dense matrix-vector product spread over many CTAs per front --
// near the root there are only one or two fronts per level and a single SM cannot stream their panels.
constexpr int GV_ROWS = 64, GV_SEG = 4;

__global__ void __launch_bounds__(512) k_bf_forward_tri(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                        double *__restrict__ upd, double *__restrict__ y)
{
    const int32_t f = T.level_sn[begin + blockIdx.x];
    const int t = threadIdx.x, G = blockDim.x;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    double *u = upd + rp0;
    double *y1 = y + first;
    bf_pdl_trigger();
    bf_pdl_wait();
    __shared__ double yb[NB];
    for (int32_t a = t; a < r; a += G) u[a] = 0.0;
    __syncthreads();
    for (int32_t ci = T.child_ptr[f]; ci < T.child_ptr[f + 1]; ++ci)
    {
        const int32_t c = T.child_idx[ci];
        const int64_t rp = T.sn_rowptr[c];
        const int32_t rc = (int32_t)(T.sn_rowptr[c + 1] - rp);
        const double *uc = upd + rp;
        const int32_t *rel = T.sn_rel + rp;
        for (int32_t a = t; a < rc; a += G)
        {
            const int32_t ra = rel[a];
            if (ra < s) y1[ra] += uc[a]; else u[ra - s] += uc[a];
        }
        __syncthreads();
    }
    for (int32_t k0 = 0; k0 < s; k0 += NB)
    {
        const int32_t k1 = min(k0 + NB, s);
        if (t < 32)
        {   // unit lower triangular block: lane i owns row k0 + i
            const int32_t i = k0 + t;
            double l[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) l[q] = (i < k1 && q < t) ? Lp[(int64_t)(k0 + q) * ld + i] : 0.0;
            double yi = (i < k1) ? y1[i] : 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q)
            {
                const double yk = __shfl_sync(0xffffffffu, yi, q);
                if (q < t) yi -= l[q] * yk;
            }
            if (i < k1) { y1[i] = yi; yb[t] = yi; }
        }
        __syncthreads();
        for (int32_t i = k1 + t; i < s; i += G)
        {
            double lv[NB]; // the NB factor entries of the row in flight together
#pragma unroll
            for (int q = 0; q < NB; ++q) lv[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + i] : 0.0;
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += lv[q] * ((k0 + q < k1) ? yb[q] : 0.0);
            y1[i] -= acc;
        }
        __syncthreads();
    }
}

// u -= L21 * y1.  CTA = GV_ROWS rows x GV_SEG interleaved pivot segments, partial sums folded in a fixed order.
__global__ void __launch_bounds__(GV_ROWS *GV_SEG) k_bf_forward_gemv(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                                     double *__restrict__ upd, const double *__restrict__ y)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t a0 = blockIdx.x * GV_ROWS;
    if (a0 >= r) return;
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    const double *y1 = y + first;
    const int ri = threadIdx.x % GV_ROWS, seg = threadIdx.x / GV_ROWS;
    const int32_t a = a0 + ri;
    bf_pdl_trigger();
    bf_pdl_wait();
    __shared__ double part[GV_SEG][GV_ROWS];
    double acc0 = 0.0, acc1 = 0.0;
    if (a < r)
    {
        const double *row = Lp + s + a;
        int32_t k = seg;
        for (; k + GV_SEG < s; k += 2 * GV_SEG)
        {
            acc0 += row[(int64_t)k * ld] * y1[k];
            acc1 += row[(int64_t)(k + GV_SEG) * ld] * y1[k + GV_SEG];
        }
        if (k < s) acc0 += row[(int64_t)k * ld] * y1[k];
    }
    part[seg][ri] = acc0 + acc1;
    __syncthreads();
    if (seg == 0 && a < r)
    {
        double sum = part[0][ri];
#pragma unroll
        for (int q = 1; q < GV_SEG; ++q) sum += part[q][ri];
        upd[rp0 + a] -= sum;
    }
}

// y1 -= U12 * x_ext: one warp per pivot row, lanes along the (contiguous) row of the U panel
__global__ void __launch_bounds__(256) k_bf_backward_gemv(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                          double *__restrict__ y)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int32_t k = blockIdx.x * 8 + wid;
    if (k >= s) return;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    if (r == 0) return;
    const int32_t ld = s + r;
    bf_pdl_trigger();
    bf_pdl_wait();
    const double *utk = LU + T.sn_lu[f] + (int64_t)ld * s + (int64_t)k * r;
    const int32_t *rows = T.sn_rows + rp0;
    double acc0 = 0.0, acc1 = 0.0;
    int32_t a = lane;
    for (; a + 32 < r; a += 64)
    {
        acc0 += utk[a] * y[rows[a]];
        acc1 += utk[a + 32] * y[rows[a + 32]];
    }
    if (a < r) acc0 += utk[a] * y[rows[a]];
    double acc = acc0 + acc1;
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) y[first + k] -= acc;
}

__global__ void __launch_bounds__(512) k_bf_backward_tri(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                         double *__restrict__ y)
{
    const int32_t f = T.level_sn[begin + blockIdx.x];
    const int t = threadIdx.x, G = blockDim.x;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - T.sn_rowptr[f]);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    double *y1 = y + first;
    __shared__ double xb[NB];
    bf_pdl_trigger();
    bf_pdl_wait();
    for (int32_t k1 = s; k1 > 0; k1 -= NB)
    {
        const int32_t k0 = max(k1 - NB, 0);
        if (t < 32)
        {   // upper triangular block (diagonal of U in place): lane i owns row k0 + i
            const int32_t i = k0 + t, nb = k1 - k0;
            double uu[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) uu[q] = (q < nb && t <= q) ? Lp[(int64_t)(k0 + q) * ld + i] : 1.0;
            // reciprocal of the lane's own pivot up front: a division inside the chain costs more than the
            // other NB - 1 steps of the block together
            const double rinv = 1.0 / ((i < k1) ? Lp[(int64_t)i * ld + i] : 1.0);
            double yi = (i < k1) ? y1[i] : 0.0;
#pragma unroll
            for (int q = NB - 1; q >= 0; --q)
            {
                if (q < nb)
                {
                    if (t == q) yi = yi * rinv;
                    const double xj = __shfl_sync(0xffffffffu, yi, q);
                    if (t < q) yi -= uu[q] * xj;
                }
            }
            if (i < k1) { y1[i] = yi; xb[t] = yi; }
        }
        __syncthreads();
        for (int32_t i = t; i < k0; i += G)
        {
            double uv[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) uv[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + i] : 0.0;
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += uv[q] * ((k0 + q < k1) ? xb[q] : 0.0);
            y1[i] -= acc;
        }
        __syncthreads();
    }
}

// ---- the same sweeps with the memory latency taken out of the dependent chain (option bf_fast_solve) --------------------
// Every operation of the kernels above in the same order, so the results are the same bits.  What changes: the pivot
// unknowns of the front live in shared memory for the whole sweep (the round-1 kernels went through global memory twice
// per block of NB pivots), the factor entries a block needs -- the diagonal block of warp 0, one panel row per thread -- are
// loaded BEFORE the chain they do not depend on (next block's diagonal block while this block's rows are applied), and
// the matrix-vector kernels issue their loads in batches of 16 instead of two per loop trip.


// Loads the compiler must not sink towards their uses (it serialised the batches of the matrix-vector kernels into
// load - multiply-add pairs, one memory round trip each): volatile asm keeps the requests together, in program order.
__device__ __forceinline__ double ld_nc_f64(const double *p)
{
    double v;
    asm volatile("ld.global.nc.f64 %0, [%1];" : "=d"(v) : "l"(p));
    return v;
}
__device__ __forceinline__ double ld_f64(const double *p)
{
    double v;
    asm volatile("ld.global.f64 %0, [%1];" : "=d"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ int32_t ld_nc_s32(const int32_t *p)
{
    int32_t v;
    asm volatile("ld.global.nc.s32 %0, [%1];" : "=r"(v) : "l"(p));
    return v;
}

// Block b of NB pivots belongs to warp b % TRI_W, which holds the block's triangle in registers from the start of the
// kernel (the factors are final long before the sweep: requested before the wait for the level below) -- no memory
// round trip sits between two dependent chains any more.
// (TRI_T threads: 256 for pivot blocks of up to 288, 128 / 64 on levels whose fronts have at most 128 / 64 pivots -- the
// kernel needs ~250 registers per thread, so the smaller CTAs are what lets two or four fronts share an SM there)
template <int TRI_T>
__global__ void __launch_bounds__(TRI_T) k_bf_forward_tri2(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                           double *__restrict__ upd, double *__restrict__ y)
{
    constexpr int TRI_W = TRI_T / 32;
    extern __shared__ double ys[]; // the front's pivot unknowns
    const int32_t f = T.level_sn[begin + blockIdx.x];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    double *u = upd + rp0;
    double *y1 = y + first;
    const int32_t cb = T.child_ptr[f], ce = T.child_ptr[f + 1];
    // the first children's table entries, all requested together
    constexpr int PC = 4;
    int64_t c_rp[PC];
    int32_t c_rc[PC];
#pragma unroll
    for (int q = 0; q < PC; ++q)
    {
        c_rp[q] = 0; c_rc[q] = 0;
        if (cb + q < ce)
        {
            const int32_t c = T.child_idx[cb + q];
            c_rp[q] = T.sn_rowptr[c];
            c_rc[q] = (int32_t)(T.sn_rowptr[c + 1] - c_rp[q]);
        }
    }
    __shared__ double yb[NB];
    double l[NB];  // the warp's block: lane i holds row i of the unit lower triangle
    double lv[NB]; // every thread: its row of the panel below the current block
    auto load_block = [&](int32_t k0) {
        const int32_t k1 = min(k0 + NB, s), i = k0 + lane;
#pragma unroll
        for (int q = 0; q < NB; ++q) l[q] = (i < k1 && q < lane) ? Lp[(int64_t)(k0 + q) * ld + i] : 0.0;
    };
    if (w * NB < s) load_block(w * NB);
    for (int32_t c = t; c + 1 < s; c += TRI_T)
    {   // the rest of the pivot block's lower triangle on its way to L2 (one request per 128-byte line of a column)
        const double *col = Lp + (int64_t)c * ld;
        for (int32_t i = c + 1; i < s; i += 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(col + i));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(col + s - 1));
    }
    bf_pdl_trigger();
    bf_pdl_wait();
    for (int32_t a = t; a < s; a += TRI_T) ys[a] = y1[a];
    for (int32_t a = t; a < r; a += TRI_T) u[a] = 0.0;
    __syncthreads();
    for (int32_t ci = cb; ci < ce; ++ci)
    {
        int64_t rp; int32_t rc;
        if (ci - cb < PC) { rp = c_rp[0]; rc = c_rc[0]; }
        else { const int32_t c = T.child_idx[ci]; rp = T.sn_rowptr[c]; rc = (int32_t)(T.sn_rowptr[c + 1] - rp); }
#pragma unroll
        for (int q = 0; q + 1 < PC; ++q) { c_rp[q] = c_rp[q + 1]; c_rc[q] = c_rc[q + 1]; }
        const double *uc = upd + rp;
        const int32_t *rel = T.sn_rel + rp;
        for (int32_t a = t; a < rc; a += TRI_T)
        {
            const int32_t ra = rel[a];
            if (ra < s) ys[ra] += uc[a]; else u[ra - s] += uc[a];
        }
        __syncthreads();
    }
    int32_t blk = 0;
    for (int32_t k0 = 0; k0 < s; k0 += NB, ++blk)
    {
        const int32_t k1 = min(k0 + NB, s);
        const int32_t i = k1 + t;
        if (i < s)
        {
#pragma unroll
            for (int q = 0; q < NB; ++q) lv[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + i] : 0.0;
        }
        if (w == (blk % TRI_W))
        {
            const int32_t ib = k0 + lane;
            double yi = (ib < k1) ? ys[ib] : 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q)
            {
                const double yk = __shfl_sync(0xffffffffu, yi, q);
                if (q < lane) yi -= l[q] * yk;
            }
            if (ib < k1) { ys[ib] = yi; yb[lane] = yi; }
            if (k0 + TRI_W * NB < s) load_block(k0 + TRI_W * NB); // the warp's next block, TRI_W blocks ahead
        }
        __syncthreads();
        if (i < s)
        {
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += lv[q] * ((k0 + q < k1) ? yb[q] : 0.0);
            ys[i] -= acc;
        }
        for (int32_t i2 = i + TRI_T; i2 < s; i2 += TRI_T)
        {
            double lw[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) lw[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + i2] : 0.0;
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += lw[q] * ((k0 + q < k1) ? yb[q] : 0.0);
            ys[i2] -= acc;
        }
        __syncthreads();
    }
    for (int32_t a = t; a < s; a += TRI_T) y1[a] = ys[a];
}

__global__ void __launch_bounds__(GV_ROWS *GV_SEG) k_bf_forward_gemv2(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                                      double *__restrict__ upd, const double *__restrict__ y)
{
    extern __shared__ double ysh[]; // the front's pivot unknowns
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    const int32_t a0 = blockIdx.x * GV_ROWS;
    if (a0 >= r) return;
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    const double *y1 = y + first;
    const int ri = threadIdx.x % GV_ROWS, seg = threadIdx.x / GV_ROWS;
    const int32_t a = a0 + ri;
    constexpr int UB = 8;
    const double *row = Lp + s + min(a, r - 1);
    // the factors are final: the first batch travels while the level below finishes
    double v0[UB], v1[UB];
    const bool first_batch = seg + (2 * UB - 1) * GV_SEG < s;
    if (first_batch)
    {
#pragma unroll
        for (int q = 0; q < UB; ++q)
        {
            v0[q] = ld_nc_f64(row + (int64_t)(seg + 2 * q * GV_SEG) * ld);
            v1[q] = ld_nc_f64(row + (int64_t)(seg + (2 * q + 1) * GV_SEG) * ld);
        }
    }
    bf_pdl_trigger();
    bf_pdl_wait();
    __shared__ double part[GV_SEG][GV_ROWS];
    for (int32_t k = threadIdx.x; k < s; k += GV_ROWS * GV_SEG) ysh[k] = y1[k];
    __syncthreads();
    double acc0 = 0.0, acc1 = 0.0;
    if (a < r)
    {
        int32_t k = seg;
        for (; k + (2 * UB - 1) * GV_SEG < s; k += 2 * UB * GV_SEG)
        {
            if (k != seg)
            {
#pragma unroll
                for (int q = 0; q < UB; ++q)
                {
                    v0[q] = ld_nc_f64(row + (int64_t)(k + 2 * q * GV_SEG) * ld);
                    v1[q] = ld_nc_f64(row + (int64_t)(k + (2 * q + 1) * GV_SEG) * ld);
                }
            }
#pragma unroll
            for (int q = 0; q < UB; ++q)
            {
                acc0 += v0[q] * ysh[k + 2 * q * GV_SEG];
                acc1 += v1[q] * ysh[k + (2 * q + 1) * GV_SEG];
            }
        }
        for (; k + GV_SEG < s; k += 2 * GV_SEG)
        {
            acc0 += row[(int64_t)k * ld] * ysh[k];
            acc1 += row[(int64_t)(k + GV_SEG) * ld] * ysh[k + GV_SEG];
        }
        if (k < s) acc0 += row[(int64_t)k * ld] * ysh[k];
    }
    part[seg][ri] = acc0 + acc1;
    __syncthreads();
    if (seg == 0 && a < r)
    {
        double sum = part[0][ri];
#pragma unroll
        for (int q = 1; q < GV_SEG; ++q) sum += part[q][ri];
        upd[rp0 + a] -= sum;
    }
}

__global__ void __launch_bounds__(256) k_bf_backward_gemv2(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                           double *__restrict__ y)
{
    const int32_t f = T.level_sn[begin + blockIdx.y];
    const int lane = threadIdx.x % 32, wid = threadIdx.x / 32;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int32_t k = blockIdx.x * 8 + wid;
    if (k >= s) return;
    const int64_t rp0 = T.sn_rowptr[f];
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - rp0);
    if (r == 0) return;
    const int32_t ld = s + r;
    const double *utk = LU + T.sn_lu[f] + (int64_t)ld * s + (int64_t)k * r;
    const int32_t *rows = T.sn_rows + rp0;
    // row indices and factor entries are static: the first batch is requested before the wait for the level above
    constexpr int UB = 4;
    int32_t i0[UB], i1[UB];
    double v0[UB], v1[UB];
    const bool first_batch = lane + 64 * (UB - 1) + 32 < r;
    if (first_batch)
    {
#pragma unroll
        for (int q = 0; q < UB; ++q)
        {
            i0[q] = ld_nc_s32(rows + lane + 64 * q);
            i1[q] = ld_nc_s32(rows + lane + 64 * q + 32);
            v0[q] = ld_nc_f64(utk + lane + 64 * q);
            v1[q] = ld_nc_f64(utk + lane + 64 * q + 32);
        }
    }
    bf_pdl_trigger();
    bf_pdl_wait();
    double acc0 = 0.0, acc1 = 0.0;
    int32_t a = lane;
    for (; a + 64 * (UB - 1) + 32 < r; a += 64 * UB)
    {
        double x0[UB], x1[UB];
        if (a != lane)
        {
#pragma unroll
            for (int q = 0; q < UB; ++q)
            {
                i0[q] = ld_nc_s32(rows + a + 64 * q);
                i1[q] = ld_nc_s32(rows + a + 64 * q + 32);
                v0[q] = ld_nc_f64(utk + a + 64 * q);
                v1[q] = ld_nc_f64(utk + a + 64 * q + 32);
            }
        }
#pragma unroll
        for (int q = 0; q < UB; ++q) { x0[q] = ld_f64(y + i0[q]); x1[q] = ld_f64(y + i1[q]); }
#pragma unroll
        for (int q = 0; q < UB; ++q) { acc0 += v0[q] * x0[q]; acc1 += v1[q] * x1[q]; }
    }
    for (; a + 32 < r; a += 64)
    {
        acc0 += utk[a] * y[rows[a]];
        acc1 += utk[a + 32] * y[rows[a + 32]];
    }
    if (a < r) acc0 += utk[a] * y[rows[a]];
    double acc = acc0 + acc1;
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) y[first + k] -= acc;
}

template <int TRI_T>
__global__ void __launch_bounds__(TRI_T) k_bf_backward_tri2(FrontTables T, int32_t begin, const double *__restrict__ LU,
                                                            double *__restrict__ y)
{
    constexpr int TRI_W = TRI_T / 32;
    extern __shared__ double ys[];
    const int32_t f = T.level_sn[begin + blockIdx.x];
    const int t = threadIdx.x, lane = t & 31, w = t >> 5;
    const int32_t first = T.sn_first[f];
    const int32_t s = T.sn_first[f + 1] - first;
    const int32_t r = (int32_t)(T.sn_rowptr[f + 1] - T.sn_rowptr[f]);
    const int32_t ld = s + r;
    const double *Lp = LU + T.sn_lu[f];
    double *y1 = y + first;
    __shared__ double xb[NB];
    double uu[NB], uv[NB];
    double rinv = 1.0;
    // blocks are counted from the end of the pivot range; block b belongs to warp b % TRI_W, which holds its triangle
    // and the reciprocals of its pivots from the start of the kernel
    auto load_block = [&](int32_t k1) {
        const int32_t k0 = max(k1 - NB, 0), i = k0 + lane, nb = k1 - k0;
#pragma unroll
        for (int q = 0; q < NB; ++q) uu[q] = (q < nb && lane <= q) ? Lp[(int64_t)(k0 + q) * ld + i] : 1.0;
        rinv = 1.0 / ((i < k1) ? Lp[(int64_t)i * ld + i] : 1.0);
    };
    if (s - w * NB > 0) load_block(s - w * NB);
    for (int32_t c = t + 1; c < s; c += TRI_T)
    {   // the pivot block's upper triangle on its way to L2
        const double *col = Lp + (int64_t)c * ld;
        for (int32_t i = 0; i < c; i += 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(col + i));
        asm volatile("prefetch.global.L2 [%0];" ::"l"(col + c - 1));
    }
    bf_pdl_trigger();
    bf_pdl_wait();
    for (int32_t a = t; a < s; a += TRI_T) ys[a] = y1[a];
    __syncthreads();
    int32_t blk = 0;
    for (int32_t k1 = s; k1 > 0; k1 -= NB, ++blk)
    {
        const int32_t k0 = max(k1 - NB, 0);
        if (t < k0)
        {
#pragma unroll
            for (int q = 0; q < NB; ++q) uv[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + t] : 0.0;
        }
        if (w == (blk % TRI_W))
        {
            const int32_t i = k0 + lane, nb = k1 - k0;
            double yi = (i < k1) ? ys[i] : 0.0;
#pragma unroll
            for (int q = NB - 1; q >= 0; --q)
            {
                if (q < nb)
                {
                    if (lane == q) yi = yi * rinv;
                    const double xj = __shfl_sync(0xffffffffu, yi, q);
                    if (lane < q) yi -= uu[q] * xj;
                }
            }
            if (i < k1) { ys[i] = yi; xb[lane] = yi; }
            if (k1 - TRI_W * NB > 0) load_block(k1 - TRI_W * NB);
        }
        __syncthreads();
        if (t < k0)
        {
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += uv[q] * ((k0 + q < k1) ? xb[q] : 0.0);
            ys[t] -= acc;
        }
        for (int32_t i2 = t + TRI_T; i2 < k0; i2 += TRI_T)
        {
            double uw[NB];
#pragma unroll
            for (int q = 0; q < NB; ++q) uw[q] = (k0 + q < k1) ? Lp[(int64_t)(k0 + q) * ld + i2] : 0.0;
            double acc = 0.0;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc += uw[q] * ((k0 + q < k1) ? xb[q] : 0.0);
            ys[i2] -= acc;
        }
        __syncthreads();
    }
    for (int32_t a = t; a < s; a += TRI_T) y1[a] = ys[a];
}

FrontTables tables(daecpp_b200_ctx *c)
{
    FrontTables T;
    T.sn_first = c->dS.sn_first.p; T.sn_rows = c->dS.sn_rows.p; T.sn_rel = c->dS.sn_rel.p;
    T.child_ptr = c->dS.child_ptr.p; T.child_idx = c->dS.child_idx.p; T.level_sn = c->dS.big_sn.p;
    T.sn_rowptr = c->dS.sn_rowptr.p; T.sn_lu = c->dS.sn_lu.p; T.sn_cb = c->dS.sn_cb.p;
    return T;
}

template <class T>
int upload(daecpp_b200_ctx *c, DevBuf<T> &d, const std::vector<T> &h)
{
    DAECPP_CUDA_TRY(c, d.reserve(h.size() ? h.size() : 1));
    if (!h.empty())
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(d.p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, c->stream));
    return DAECPP_B200_OK;
}

int pick_group(int32_t max_dim)
{
    if (max_dim <= 4) return 4;
    if (max_dim <= 8) return 8;
    if (max_dim <= 48) return 32;
    if (max_dim <= 384) return 256;
    return 1024; // the few fronts at the top of the tree: one full-size CTA each
}

#define DISPATCH_GROUP(G, KERNEL, count, ...)                                                                     \
    switch (G)                                                                                                    \
    {                                                                                                             \
    case 4: KERNEL<4><<<((count) + Grp<4>::FRONTS_PER_CTA - 1) / Grp<4>::FRONTS_PER_CTA, Grp<4>::THREADS, 0, st>>>(__VA_ARGS__); break;       \
    case 8: KERNEL<8><<<((count) + Grp<8>::FRONTS_PER_CTA - 1) / Grp<8>::FRONTS_PER_CTA, Grp<8>::THREADS, 0, st>>>(__VA_ARGS__); break;       \
    case 32: KERNEL<32><<<((count) + Grp<32>::FRONTS_PER_CTA - 1) / Grp<32>::FRONTS_PER_CTA, Grp<32>::THREADS, 0, st>>>(__VA_ARGS__); break;  \
    case 1024: KERNEL<1024><<<(count), Grp<1024>::THREADS, 0, st>>>(__VA_ARGS__); break;                         \
    default: KERNEL<256><<<(count), Grp<256>::THREADS, 0, st>>>(__VA_ARGS__); break;                             \
    }

} // namespace

// Download the pattern of Jb, analyse it on the host, ship the schedule back.
int upload_symbolic(daecpp_b200_ctx *c)
{
    const int32_t n = c->n;
    cudaStream_t st = c->stream;
    auto t0 = std::chrono::steady_clock::now();
    std::vector<int32_t> outer((size_t)n + 1), inner((size_t)(c->nnzU ? c->nnzU : 1));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(outer.data(), c->u_outer.p, sizeof(int32_t) * ((size_t)n + 1), cudaMemcpyDeviceToHost, st));
    if (c->nnzU)
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(inner.data(), c->u_inner.p, sizeof(int32_t) * c->nnzU, cudaMemcpyDeviceToHost, st));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
    {   // banded patterns (1-D problems) take the block-chain path: no per-front tables at all
        int rc = chain_try_setup(c, n, outer.data(), inner.data());
        if (rc) return rc;
        if (c->CH.active)
        {
            DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
            c->have_symbolic = true;
            c->have_factor = false;
            c->n_symbolic += 1;
            c->ms_symbolic = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
            return DAECPP_B200_OK;
        }
    }
    const char *err = nullptr;
    // a diagonal with holes: the row matching of the analysis prefers big entries, so hand it the magnitudes of the
    // matrix about to be factored (Jb = J - alpha*M with the alpha the caller announced; static afterwards)
    std::vector<double> weights;
    {
        int32_t missing = 0;
        for (int32_t j = 0; j < n; ++j)
            if (!std::binary_search(inner.begin() + outer[j], inner.begin() + outer[j + 1], j)) ++missing;
        if (missing && c->sopt.transversal && !c->direct_csc && c->nnzU)
        {
            int rcw = jb_values(c, c->alpha_hint);
            if (rcw) return rcw;
            weights.resize((size_t)c->nnzU);
            DAECPP_CUDA_TRY(c, cudaMemcpyAsync(weights.data(), c->u_vals.p, sizeof(double) * c->nnzU, cudaMemcpyDeviceToHost, st));
            DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
            for (double &v : weights) v = std::fabs(v);
        }
    }
    c->sopt.weights = weights.empty() ? nullptr : weights.data();
    const bool analysed = analyse_pattern(n, outer.data(), inner.data(), c->sopt, c->S, &err);
    c->sopt.weights = nullptr;
    if (!analysed)
    {
        c->err = err ? err : "symbolic analysis failed";
        return DAECPP_B200_ENOMEM;
    }
    Symbolic &S = c->S;
    int rc;
    if ((rc = upload(c, c->dS.perm, S.perm))) return rc;
    if ((rc = upload(c, c->dS.iperm, S.iperm))) return rc;
    {
        std::vector<int32_t> pr(S.perm);
        if (!S.row_src.empty())
            for (int32_t &v : pr) v = S.row_src[(size_t)v];
        if ((rc = upload(c, c->dS.perm_rhs, pr))) return rc;
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st)); // pr is a local
    }
    if (S.nbig)
    { // the global front tables are only read by the per-level kernels of big fronts
        if ((rc = upload(c, c->dS.sn_first, S.sn_first))) return rc;
        if ((rc = upload(c, c->dS.sn_rows, S.sn_rows))) return rc;
        if ((rc = upload(c, c->dS.sn_rel, S.sn_rel))) return rc;
        if ((rc = upload(c, c->dS.child_ptr, S.child_ptr))) return rc;
        if ((rc = upload(c, c->dS.child_idx, S.child_idx))) return rc;
        if ((rc = upload(c, c->dS.big_sn, S.big_sn))) return rc;
        if ((rc = upload(c, c->dS.sn_rowptr, S.sn_rowptr))) return rc;
        if ((rc = upload(c, c->dS.sn_lu, S.sn_lu))) return rc;
        if ((rc = upload(c, c->dS.sn_cb, S.sn_cb))) return rc;
        if ((rc = upload(c, c->dS.big_u, S.big_u))) return rc;
        if ((rc = upload(c, c->dS.big_dst, S.big_dst))) return rc;
        DAECPP_CUDA_TRY(c, c->dS.big_sj.reserve(S.big_u.size() + 1));
        DAECPP_CUDA_TRY(c, c->dS.big_sm.reserve(S.big_u.size() + 1));
    }
    if ((rc = upload(c, c->dS.blob, S.blob))) return rc;
    if ((rc = upload(c, c->dS.unit_blob, S.unit_blob))) return rc;
    if ((rc = upload(c, c->dS.ent_u, S.ent_u))) return rc;
    if ((rc = upload(c, c->dS.ent_dst, S.ent_dst))) return rc;
    DAECPP_CUDA_TRY(c, c->dS.ent_sj.reserve(S.ent_u.size() + 1));
    DAECPP_CUDA_TRY(c, c->dS.ent_sm.reserve(S.ent_u.size() + 1));
    const size_t nsync = 1 + 2 * (size_t)S.nunit + 1;
    DAECPP_CUDA_TRY(c, c->dS.sync_words.reserve(nsync));
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->dS.sync_words.p, 0, sizeof(int32_t) * nsync, st));
    c->ticket_base = 0;
    c->solve_epoch = 0;
    DAECPP_CUDA_TRY(c, c->LU.reserve((size_t)S.lu_size + 1));
    DAECPP_CUDA_TRY(c, c->CB.reserve((size_t)S.cb_size + 1));
    DAECPP_CUDA_TRY(c, c->upd.reserve((size_t)S.upd_size + 1));
    c->groups.clear();
    int32_t max_level_fronts = 0;
    for (const LaunchGroup &g : S.groups)
    {
        GroupLaunch L;
        L.kind = g.kind; L.begin = g.begin; L.count = g.end - g.begin; L.max_dim = g.max_dim; L.overlap = g.same_stage;
        L.group = pick_group(g.max_dim);
        if (L.kind != 0)
            for (int32_t q = 0; q < L.count; ++q)
            {
                const int32_t f = S.big_sn[L.begin + q];
                L.max_s = std::max(L.max_s, S.s_of(f));
                L.max_r = std::max(L.max_r, S.r_of(f));
                L.max_children = std::max(L.max_children, S.child_ptr[f + 1] - S.child_ptr[f]);
                max_level_fronts = std::max(max_level_fronts, L.count);
            }
        c->groups.push_back(L);
    }
    if (max_level_fronts) DAECPP_CUDA_TRY(c, c->bf_diag.reserve((size_t)max_level_fronts * NB * NB));
    // the fused assembly reads the staged triplet values when every slot has exactly one triplet
    c->jvals_from_triplets = !c->direct_csc && c->J.no_duplicates && c->J.nnz > 0;
    if ((rc = unit_setup(c))) return rc;
    if ((rc = compose_entries(c))) return rc;
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
    c->have_symbolic = true;
    c->have_factor = false;
    c->n_symbolic += 1;
    c->ms_symbolic = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return DAECPP_B200_OK;
}

// vals[] of a triplet set (compressed-column values) -- only the parity exports and the
// duplicate-summing path need them; the fused assembly reads the triplet values directly.
int ensure_csc_values(daecpp_b200_ctx *c, TripletSet &T)
{
    if (T.vals_valid || !T.have_pattern) return DAECPP_B200_OK;
    int rc = regather_values(c, T);
    if (rc) return rc;
    T.vals_valid = true;
    return DAECPP_B200_OK;
}

int jb_values(daecpp_b200_ctx *c, double alpha)
{
    if (c->direct_csc || !c->nnzU) return DAECPP_B200_OK;
    int rc = ensure_csc_values(c, c->J);
    if (rc) return rc;
    if ((rc = ensure_csc_values(c, c->M))) return rc;
    k_jb_values<<<(int)((c->nnzU + 255) / 256), 256, 0, c->stream>>>(c->nnzU, c->u_J.p, c->u_M.p, c->J.vals.p, c->M.vals.p,
                                                                  alpha, c->u_vals.p);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

// Only the values of Jb = J - (M*alpha) on the union pattern (parity export, no factorisation).
int assemble_values_only(daecpp_b200_ctx *c, double alpha)
{
    int rc = jb_values(c, alpha);
    if (rc) return rc;
    c->n_refactor += 1;
    c->last_alpha = alpha;
    c->have_factor = false;
    return DAECPP_B200_OK;
}

// The context's side stream and its fork / join events (created on first use).
static int ensure_side_stream(daecpp_b200_ctx *c)
{
    if (c->stream_zero) return DAECPP_B200_OK;
    DAECPP_CUDA_TRY(c, cudaStreamCreateWithFlags(&c->stream_zero, cudaStreamNonBlocking));
    DAECPP_CUDA_TRY(c, cudaEventCreateWithFlags(&c->ev_z0, cudaEventDisableTiming));
    DAECPP_CUDA_TRY(c, cudaEventCreateWithFlags(&c->ev_z1, cudaEventDisableTiming));
    DAECPP_CUDA_TRY(c, cudaEventCreateWithFlags(&c->ev_u0, cudaEventDisableTiming));
    DAECPP_CUDA_TRY(c, cudaEventCreateWithFlags(&c->ev_u1, cudaEventDisableTiming));
    return DAECPP_B200_OK;
}
// A launch of subtree units in the same stage of the tree as the level of big fronts behind it: neither depends on the
// other, and the units are a chain of dependent latencies on a few SMs' worth of lanes -- they run beside the level.
static bool overlaps_next(const daecpp_b200_ctx *c, size_t l)
{
    return c->bf_overlap_units && c->blocked_lu && c->groups[l].kind == 0 && c->groups[l].overlap && l + 1 < c->groups.size() &&
           c->groups[l + 1].kind != 0;
}
#define FORK_SIDE(c, st) do { DAECPP_CUDA_TRY(c, cudaEventRecord((c)->ev_u0, st)); DAECPP_CUDA_TRY(c, cudaStreamWaitEvent((c)->stream_zero, (c)->ev_u0, 0)); } while (0)
#define JOIN_SIDE(c, st) do { DAECPP_CUDA_TRY(c, cudaEventRecord((c)->ev_u1, (c)->stream_zero)); DAECPP_CUDA_TRY(c, cudaStreamWaitEvent(st, (c)->ev_u1, 0)); } while (0)

int numeric_refactor_async(daecpp_b200_ctx *c, double alpha, bool clear_flags)
{
    cudaStream_t st = c->stream;
    Symbolic &S = c->S;
    int rc;
    if (!c->jvals_from_triplets && !c->direct_csc && (rc = ensure_csc_values(c, c->J))) return rc;
    c->flags_clean = false;
    if (clear_flags) DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 3, 0, 2 * sizeof(int32_t), st));
    if (c->CH.active)
    {
        if ((rc = chain_refactor_async(c, alpha))) return rc;
        c->n_refactor += 1;
        c->last_alpha = alpha;
        return DAECPP_B200_OK;
    }
    const bool side = S.nbig && c->blocked_lu && c->bf_zero_ahead;
    if (S.nbig && !side)
    {
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->LU.p + S.lu_big_begin, 0, sizeof(double) * (size_t)(S.lu_size - S.lu_big_begin), st));
        if ((rc = launch_scatter_big(c, alpha, st))) return rc;
    }
    FrontTables T = tables(c);
    bool zero_ahead = false, zero_waited = false;
    if (side)
    {   // the big fronts' factor storage (zeroed, matrix entries scattered) and the contribution blocks of ALL big fronts
        // (zeroed) are prepared on a side stream while the unit launches (a chain of dependent latencies that leaves the
        // memory system idle) run on the main one; the first level of big fronts waits for it
        if ((rc = ensure_side_stream(c))) return rc;
        DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev_z0, st)); // behind everything that still reads the blocks
        DAECPP_CUDA_TRY(c, cudaStreamWaitEvent(c->stream_zero, c->ev_z0, 0));
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->LU.p + S.lu_big_begin, 0, sizeof(double) * (size_t)(S.lu_size - S.lu_big_begin), c->stream_zero));
        if ((rc = launch_scatter_big(c, alpha, c->stream_zero))) return rc;
        for (int32_t b0 = 0; b0 < S.nbig; b0 += 32768)
            k_bf_zero<<<dim3(64, std::min(32768, S.nbig - b0)), 256, 0, c->stream_zero>>>(T, b0, c->CB.p);
        DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev_z1, c->stream_zero));
        c->kernels += 1;
        zero_ahead = true;
    }
    bool join_pending = false;
    for (size_t gl = 0; gl < c->groups.size(); ++gl)
    {
        const GroupLaunch &L = c->groups[gl];
        if (L.kind != 0 && zero_ahead && !zero_waited) { DAECPP_CUDA_TRY(c, cudaStreamWaitEvent(st, c->ev_z1, 0)); zero_waited = true; }
        if (L.kind == 0)
        {
            if (gl > 0 && overlaps_next(c, gl))
            {   // beside the level of big fronts that follows; joined behind that level
                if ((rc = ensure_side_stream(c))) return rc;
                FORK_SIDE(c, st);
                if ((rc = launch_unit_factor(c, L, alpha, c->stream_zero))) return rc;
                join_pending = true;
                continue;
            }
            if ((rc = launch_unit_factor(c, L, alpha))) return rc;
        }
        else if (c->blocked_lu && L.max_dim >= c->blocked_lu_min_dim)
        {   // blocked right-looking LU, all fronts of the level in the same launches
            const int32_t max_s = L.max_s, max_ld = L.max_dim, max_children = L.max_children, max_r = L.max_r;
            const bool look = c->bf_lu_lookahead;
            double *dbuf = c->bf_diag_copy ? c->bf_diag.p : nullptr;
            // contribution blocks updated once, behind the last block column (k_bf_schur), not once per block column
            const bool schur_level = c->bf_schur || (c->bf_schur_fronts > 0 && L.count >= c->bf_schur_fronts && max_s > NB);
            const int32_t strip = (schur_level && c->bf_dmma && max_r > 0) ? 1 : 0;
            if (!zero_ahead) bf_launch_pdl(st, k_bf_zero, dim3(64, L.count), 256, T, L.begin, c->CB.p);
            // few fronts in the level: the rows of a child's column are spread over grid.z as well, so that a large
            // contribution block is not added by a handful of warps one batch after the other
            const int ez = L.count > 4 ? 1 : std::max(1, std::min(4, (max_ld + 255) / 256));
            for (int32_t ord = 0; ord < max_children; ++ord)
                bf_launch_pdl(st, k_bf_extend, dim3(std::max(1, std::min(256, (max_ld + 7) / 8)), L.count, ez), 256, T, L.begin, ord, c->LU.p, c->CB.p);
            c->kernels += (zero_ahead ? 0 : 1) + max_children;
            for (int32_t k0 = 0; k0 < max_s; k0 += NB)
            {
                // largest trailing order among the fronts still being factored: a front with fewer pivots than
                // the level's maximum ends its last block column early and has MORE rows and columns behind it
                int32_t rest = 0;
                for (int32_t q = 0; q < L.count; ++q)
                {
                    const int32_t f = S.big_sn[L.begin + q], sf = S.s_of(f);
                    if (sf > k0) rest = std::max(rest, sf + S.r_of(f) - std::min(k0 + NB, sf));
                }
                if (rest <= 0 || !c->bf_fused_panel)
                {
                    k_bf_diag<<<L.count, 128, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags, look);
                    c->kernels += 1;
                    if (rest <= 0) continue;
                    k_bf_rows<<<dim3((rest + 127) / 128, L.count), 128, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p);
                    k_bf_trsm<<<dim3((rest + 127) / 128, L.count), 128, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p);
                    const int tiles = (rest + 63) / 64;
                    if (c->bf_dmma) k_bf_gemm<true><<<dim3(tiles, tiles, L.count), 256, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, 0.0, nullptr, strip, nullptr, look);
                    else k_bf_gemm<false><<<dim3(tiles, tiles, L.count), 256, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, 0.0, nullptr, strip, nullptr, look);
                    c->kernels += 3;
                    continue;
                }
                // two launches per block column: panel (diagonal block factored redundantly by every CTA), trailing update
                const int tiles = (rest + 63) / 64;
                const dim3 gp(2 * ((rest + 127) / 128), L.count), gg(tiles + 1, tiles, L.count);
                if (c->bf_pdl)
                {
                    bf_launch_pdl(st, k_bf_panel, gp, 128, T, L.begin, k0, c->LU.p, c->CB.p, dbuf, c->pivot_tol, c->d_flags, look);
                    if (c->bf_dmma) bf_launch_pdl(st, k_bf_gemm<true>, gg, 256, T, L.begin, k0, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags, strip, dbuf, look);
                    else bf_launch_pdl(st, k_bf_gemm<false>, gg, 256, T, L.begin, k0, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags, strip, dbuf, look);
                }
                else
                {
                    k_bf_panel<<<gp, 128, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, dbuf, c->pivot_tol, c->d_flags, look);
                    if (c->bf_dmma) k_bf_gemm<true><<<gg, 256, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags, strip, dbuf, look);
                    else k_bf_gemm<false><<<gg, 256, 0, st>>>(T, L.begin, k0, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags, strip, dbuf, look);
                }
                c->kernels += 2;
            }
            if (strip)
            {
                const int32_t ct = (max_r + 63) / 64;
                bf_launch_pdl(st, k_bf_schur, dim3(ct, ct, L.count), 256, T, L.begin, c->LU.p, c->CB.p);
                c->kernels += 1;
            }
        }
        else
        {
            DISPATCH_GROUP(L.group, k_factor_level, L.count, T, L.begin, L.count, c->LU.p, c->CB.p, c->pivot_tol, c->d_flags);
            c->kernels += 1;
        }
        if (join_pending && L.kind != 0) { JOIN_SIDE(c, st); join_pending = false; }
    }
    if (join_pending) JOIN_SIDE(c, st);
    if (zero_ahead && !zero_waited) DAECPP_CUDA_TRY(c, cudaStreamWaitEvent(st, c->ev_z1, 0));
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    c->n_refactor += 1;
    c->last_alpha = alpha;
    return DAECPP_B200_OK;
}

int numeric_refactor(daecpp_b200_ctx *c, double alpha)
{
    cudaStream_t st = c->stream;
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev0, st));
    int rc = numeric_refactor_async(c, alpha, true);
    if (rc) return rc;
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev1, st));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 3, c->d_flags + 3, 2 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->ms_last_refactor = ms;
    if (c->h_flags[4]) { c->err = "unit dependency wait timed out"; c->have_factor = false; return DAECPP_B200_ECUDA; }
    if (c->h_flags[3])
    {
        c->have_factor = false;
        return DAECPP_B200_EDECOMP;
    }
    c->have_factor = true;
    return DAECPP_B200_OK;
}

// dx = Jb^-1 b, all on the stream, no synchronisation.  With no big fronts the permutations (and,
// if asked, the error checks and xk += dx) are fused into the unit kernels.
int numeric_solve(daecpp_b200_ctx *c, bool fuse_update, double atol, double rtol, double max_err_abs)
{
    cudaStream_t st = c->stream;
    const int32_t n = c->n;
    int rc;
    if (c->CH.active)
    {
        if ((rc = chain_solve(c, fuse_update, atol, rtol, max_err_abs))) return rc;
        c->n_solve += 1;
        return DAECPP_B200_OK;
    }
    FrontTables T = tables(c);
    const bool direct = c->S.nbig == 0;
    c->solve_epoch += 1;
    if (!direct)
    {
        k_permute_in<<<(n + 255) / 256, 256, 0, st>>>(n, c->dS.perm_rhs.p, c->b, c->work);
        c->kernels += 1;
    }
    bool join_pending = false;
    for (size_t l = 0; l < c->groups.size(); ++l)
    {
        const GroupLaunch &L = c->groups[l];
        if (L.kind == 0)
        {
            if (l > 0 && overlaps_next(c, l))
            {   // same stage as the level of big fronts behind it: beside that level, joined behind it
                if ((rc = ensure_side_stream(c))) return rc;
                FORK_SIDE(c, st);
                if ((rc = launch_unit_forward(c, L, direct, c->stream_zero))) return rc;
                join_pending = true;
                continue;
            }
            if ((rc = launch_unit_forward(c, L, direct))) return rc;
        }
        else if (c->blocked_lu && L.max_dim >= c->blocked_lu_min_dim)
        {
            const size_t ysm = sizeof(double) * (size_t)L.max_s;
            if (c->bf_fast_solve && ysm <= 40 * 1024)
            {
                if (L.max_s <= 64) bf_launch_pdl_smem(st, k_bf_forward_tri2<64>, dim3(L.count), 64, ysm, T, L.begin, c->LU.p, c->upd.p, c->work);
                else if (L.max_s <= 128) bf_launch_pdl_smem(st, k_bf_forward_tri2<128>, dim3(L.count), 128, ysm, T, L.begin, c->LU.p, c->upd.p, c->work);
                else bf_launch_pdl_smem(st, k_bf_forward_tri2<256>, dim3(L.count), 256, ysm, T, L.begin, c->LU.p, c->upd.p, c->work);
                if (L.max_r > 0)
                    bf_launch_pdl_smem(st, k_bf_forward_gemv2, dim3((L.max_r + GV_ROWS - 1) / GV_ROWS, L.count), GV_ROWS * GV_SEG, ysm, T, L.begin, c->LU.p, c->upd.p, c->work);
                c->kernels += L.max_r > 0 ? 2 : 1;
            }
            else
            {
                bf_launch_pdl(st, k_bf_forward_tri, dim3(L.count), 512, T, L.begin, c->LU.p, c->upd.p, c->work);
                bf_launch_pdl(st, k_bf_forward_gemv, dim3((L.max_dim + GV_ROWS - 1) / GV_ROWS, L.count), GV_ROWS * GV_SEG, T, L.begin, c->LU.p, c->upd.p, c->work);
                c->kernels += 2;
            }
        }
        else
        {
            DISPATCH_GROUP(L.group, k_forward_level, L.count, T, L.begin, L.count, c->LU.p, c->upd.p, c->work);
            c->kernels += 1;
        }
        if (join_pending && L.kind != 0) { JOIN_SIDE(c, st); join_pending = false; }
    }
    if (join_pending) { JOIN_SIDE(c, st); join_pending = false; }
    for (size_t l = c->groups.size(); l-- > 0;)
    {
        const GroupLaunch &L = c->groups[l];
        if (L.kind != 0 && l > 1 && overlaps_next(c, l - 1))
        {   // the launch of units in front of this level (next in the backward order) is independent of it: started beside it
            if ((rc = ensure_side_stream(c))) return rc;
            FORK_SIDE(c, st);
            if ((rc = launch_unit_backward(c, c->groups[l - 1], direct, fuse_update, atol, rtol, max_err_abs, c->stream_zero))) return rc;
            join_pending = true;
        }
        if (L.kind == 0)
        {
            if (join_pending) { JOIN_SIDE(c, st); join_pending = false; continue; } // ran beside the level behind it
            if ((rc = launch_unit_backward(c, L, direct, fuse_update, atol, rtol, max_err_abs))) return rc;
        }
        else if (c->blocked_lu && L.max_dim >= c->blocked_lu_min_dim)
        {
            const size_t ysm = sizeof(double) * (size_t)L.max_s;
            if (c->bf_fast_solve && ysm <= 40 * 1024)
            {
                if (L.max_r > 0) bf_launch_pdl(st, k_bf_backward_gemv2, dim3((L.max_s + 7) / 8, L.count), 256, T, L.begin, c->LU.p, c->work);
                if (L.max_s <= 64) bf_launch_pdl_smem(st, k_bf_backward_tri2<64>, dim3(L.count), 64, ysm, T, L.begin, c->LU.p, c->work);
                else if (L.max_s <= 128) bf_launch_pdl_smem(st, k_bf_backward_tri2<128>, dim3(L.count), 128, ysm, T, L.begin, c->LU.p, c->work);
                else bf_launch_pdl_smem(st, k_bf_backward_tri2<256>, dim3(L.count), 256, ysm, T, L.begin, c->LU.p, c->work);
                c->kernels += L.max_r > 0 ? 2 : 1;
                continue;
            }
            bf_launch_pdl(st, k_bf_backward_gemv, dim3((L.max_dim + 7) / 8, L.count), 256, T, L.begin, c->LU.p, c->work);
            bf_launch_pdl(st, k_bf_backward_tri, dim3(L.count), 512, T, L.begin, c->LU.p, c->work);
            c->kernels += 2;
        }
        else
        {
            DISPATCH_GROUP(L.group, k_backward_level, L.count, T, L.begin, L.count, c->LU.p, c->work);
            c->kernels += 1;
        }
    }
    if (!direct)
    {
        k_permute_out<<<(n + 255) / 256, 256, 0, st>>>(n, c->dS.perm.p, c->work, c->dx);
        c->kernels += 1;
        if (fuse_update && (rc = launch_update(c, atol, rtol, max_err_abs))) return rc;
    }
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    c->n_solve += 1;
    return DAECPP_B200_OK;
}

} // namespace daecpp_b200
