// Block-chain numeric LU refactorisation and triangular solves (chain.hpp, chain_ops.hpp).
//
// Role replaced: Eigen::SparseLU::factorize / _solve_impl behind linsolver.compute(Jb) and
// linsolver.solve(b) (dae-cpp/solver.hpp:537,558) for block-tridiagonal patterns, with the
// assembly Jb = J - M*alpha (solver.hpp:523), the permutation and the error check / update loop
// (solver.hpp:590-619) fused into the first and the last kernel.
//
// One CTA owns a tile of T*Q0 consecutive block rows.  The matrix values of the tile are gathered
// through the slot map into shared memory with coalesced loads, each thread eliminates its own
// chunk of Q0 rows out of shared memory and streams the factors to HBM in a [step][field][thread]
// layout (every store and every later load of the solves is a full 256-byte warp transaction),
// two more levels are swept inside the CTA, and one row of the grid-level reduced system leaves
// the CTA.  That reduced system is factored by the same kernel run on it (one CTA for up to
// T*Q0 tiles).  No index table other than the slot map is read: the hierarchy is implicit.
// HBM traffic per refactorisation: slot map + matrix values in, 5*B*B doubles per block row out.
#include "bdf.hpp"
#include "chain.hpp"
#include "chain_ops.hpp"
#include "ctx.hpp"

namespace daecpp_b200
{
namespace
{
using namespace chain;

// one level of the hierarchy as a kernel sees it
struct LevelView
{
    int32_t rows, ntiles, is_top;
    double *F, *Y;
    const double *red_in;
    double *red_out;
    const double *rv_in;
    double *rv_out;
    double *xlev;      // solution of this level's rows (levels >= 1)
    const double *xup; // solution of the level above = this level's separators
};

struct ChainArgs
{
    LevelView lv;  // the level this launch sweeps, one CTA per tile
    LevelView top; // fuse_top: the single-tile top level, swept by the CTA that finishes last
    int32_t fuse_top, n, trace_tiles, packed, prefetch;
    int32_t top_pcr;      // the top level is solved by parallel cyclic reduction straight from the reduced system (no top factors)
    int32_t extra_factor; // forward launch: CTA number lv.ntiles factors the top level (deferred from the factor launch)
    uint32_t *counter; // [2 * members] tiles finished (factor, forward)
    // factor
    const int32_t *sj, *sm, *flags, *desc;
    const double *jv, *mv;
    double alpha, tol, rel;
    // solve
    const double *rhs;
    const int32_t *perm;
    double *out, *xk;
    const double *x0;
    double atol, rtol, max_err_abs, floor;
    int32_t update;
    int32_t *dflags;
    // fused residual (diagonal mass): rhs = M*dxdt - f formed while staging (vector_ops.cu: k_residual_diag)
    int32_t fuse_res;
    int32_t *cancel;   // pipelined resident iterations: non-zero = the previous iteration ended in a way the host must look at first; every CTA of this one leaves at once
    int32_t top_cta;     // forward: one extra CTA does nothing but the grid-level solve (cyclic reduction); the tiles only count themselves off
    int32_t pf_rhs;      // factor: pull the sources of the coming forward sweep's right-hand side (a.bdf, a.fvec, a.mdiag) into L2
    int32_t stage_early; // factor: the matrix values are not produced by the kernel in front: stage them before the dependency wait
    int32_t rhs_early;   // forward: the launch in front is this iteration's factor kernel, which does not touch the right-hand side's sources
    int32_t refine_on; // (the publisher raises `cancel` where the host would refine)
    const int32_t *mslot; // partial-diagonal mass: row -> slot of M_rr or -1 (null: slot = row)
    Bdf bdf;
    const double *mdiag, *fvec;
    double *b_out; // optional copy of b to HBM
    int32_t seq;
    double *xk_host;     // mapped pinned copy of xk for the host (small systems), written with the fused update
    int32_t *host_flags; // mapped host mirror of dflags[2..4], written by the CTA that finishes the iteration last
    // ensemble: lane = blockIdx.y, strides in elements; per-lane control words (ctx.hpp: LaneSet), null for a
    // single system
    const double *alpha_arr;
    int64_t s_jv, s_mv, s_F, s_Y, s_red, s_rv, s_x, s_vec, s_hist;
    int32_t lanes;      // lanes of the batch (stride of the per-lane control arrays)
    int32_t grid_lanes; // lanes this launch covers = gridDim.y
    const int32_t *lane_list; // [grid_lanes] lane of every grid row, null: the row index itself
    const int32_t *lane_mode, *lane_hbase, *lane_order;
    const double *lane_w; // [5][lanes]
    double *hist;
    int32_t *lane_flags;
    unsigned long long *lane_var;
    int32_t var_on;
    double var_tol;
    long long *trace; // debug ("trace_units"): [kernel][tile][8] %globaltimer stamps, member 0
};

// flag words of a lane addressed like the context's d_flags: [2] update flags, [3] pivot failure, [4] wait time-out
__device__ __forceinline__ int32_t *flag_words(const ChainArgs &a, int64_t mb) { return a.lane_flags ? a.lane_flags + 4 * mb - 2 : a.dflags; }
__device__ __forceinline__ int lane_mode(const ChainArgs &a, int64_t mb) { return a.lane_mode ? a.lane_mode[mb] : 2; }

constexpr int TRACE_SLOTS = 8, TRACE_KERNELS = 5; // factor tiles, factor top, forward tiles, top solve, backward tiles
__device__ __forceinline__ void stamp(const ChainArgs &a, int kernel, int tile, int slot)
{
    if (a.trace && threadIdx.x == 0 && blockIdx.y == 0 && !a.lane_list)
    {
        long long ns;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns));
        a.trace[((size_t)kernel * a.trace_tiles + tile) * TRACE_SLOTS + slot] = ns;
    }
}

// Programmatic dependent launch: the chain kernels of one Newton iteration follow each other on the
// stream with nothing in between, so each one lets its successor start launching as soon as its own
// loads are issued, and waits for its predecessor's memory before touching global data.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// true in exactly one CTA per member: the one whose tile finished last (its view of the other
// tiles' results is complete after the fence)
__device__ __forceinline__ bool last_tile_done(uint32_t *counter, int ntiles, int t, bool sys = false)
{
    __shared__ int s_last;
    if (sys) __threadfence_system(); // this CTA stored to mapped host memory
    else __threadfence();
    __syncthreads();
    if (t == 0)
    {
        const uint32_t ticket = atomicAdd(counter, 1u);
        s_last = (ticket == (uint32_t)ntiles - 1u);
        if (s_last) *counter = 0u;
    }
    __syncthreads();
    if (s_last) __threadfence();
    return s_last != 0;
}

// Affine tiles whose only mass entries sit on the diagonal of the diagonal blocks (every identity-like
// mass matrix): the J values of the whole tile are put in flight at once with LDGSTS (cp.async, 8 bytes,
// global -> shared, no register staging), the few M values ride in registers meanwhile, and M*alpha is
// subtracted in place by the thread that issued the copy.  One DRAM round trip for the whole tile.
// Entries without a mass term are J's bits unchanged, which is what `J - 0.0` gives (solver.hpp:523).
template <int B>
__device__ __forceinline__ bool stage_matrix_affine_async(int t, int tile, double *S, const int32_t *desc, const double *jv,
                                                          const double *mvals, double alpha, bool allow_packed)
{
    using D = Dim<B>;
    constexpr int PER = D::ROWS + 1, NFE = 3 * D::BB, NT = D::T, IT = (PER + NT - 1) / NT;
    __shared__ int32_t sd[NFE * DESC];
    for (int k = t; k < NFE * DESC; k += NT) sd[k] = desc[(size_t)tile * NFE * DESC + k];
    __syncthreads();
    bool simple = true, packed = allow_packed && sd[DJLO] <= sd[DJHI];
#pragma unroll
    for (int fe = 0; fe < NFE; ++fe)
    {
        const bool diag = fe >= D::BB && fe < 2 * D::BB && ((fe - D::BB) / B == (fe - D::BB) % B);
        if (!diag && sd[fe * DESC + DMLO] <= sd[fe * DESC + DMHI]) simple = false;
        // one contiguous run: field fe of row rr sits at j0 + NFE*rr + fe
        packed = packed && sd[fe * DESC + DJLO] <= sd[fe * DESC + DJHI] && sd[fe * DESC + DJ0] == sd[DJ0] + fe && sd[fe * DESC + DJS] == NFE;
    }
    if (!simple) return false; // uniform over the CTA: the caller takes the register-staged path
    const uint32_t s0 = (uint32_t)__cvta_generic_to_shared(S);
    double m[IT][B];
    if (packed)
    {   // row-major triplet streams (flame-1D: (i,i-1),(i,i),(i,i+1) per row): consecutive lanes copy
        // consecutive doubles (every sector requested once, fully used), the scatter is on the
        // shared-memory side
        const double *src = jv + sd[DJ0];
        for (int idx = t; idx < NFE * PER; idx += NT)
        {
            const int rr = idx / NFE, fe = idx - rr * NFE;
            const int32_t *e = sd + fe * DESC;
            const int q = fe * D::RP + sidx<B>(rr - 1);
            if (rr >= e[DJLO] && rr <= e[DJHI])
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s0 + 8u * (uint32_t)q), "l"(src + idx) : "memory");
            else
                S[q] = rr >= e[DONE] ? 1.0 : 0.0;
        }
    }
    else
    {
#pragma unroll
        for (int it = 0; it < IT; ++it)
        {
            const int rr = t + it * NT;
            if (rr < PER)
            {
                const int p = sidx<B>(rr - 1);
#pragma unroll
                for (int fe = 0; fe < NFE; ++fe)
                {
                    const int32_t *e = sd + fe * DESC;
                    if (rr >= e[DJLO] && rr <= e[DJHI])
                        asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(s0 + 8u * (uint32_t)(fe * D::RP + p)),
                                     "l"(jv + (e[DJ0] + e[DJS] * rr))
                                     : "memory");
                    else
                        S[(size_t)fe * D::RP + p] = rr >= e[DONE] ? 1.0 : 0.0;
                }
            }
        }
    }
    // the few M values (diagonal of the diagonal blocks) ride in registers meanwhile
#pragma unroll
    for (int it = 0; it < IT; ++it)
    {
        const int rr = t + it * NT;
#pragma unroll
        for (int r = 0; r < B; ++r)
        {
            const int32_t *e = sd + (D::BB + r * B + r) * DESC;
            m[it][r] = (rr >= e[DMLO] && rr <= e[DMHI]) ? mvals[e[DM0] + e[DMS] * rr] : 0.0;
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_all;" ::: "memory");
    if (packed) __syncthreads(); // there the copy of an element and its mass correction belong to different threads
#pragma unroll
    for (int it = 0; it < IT; ++it)
    {
        const int rr = t + it * NT;
        if (rr < PER)
        {
            const int p = sidx<B>(rr - 1);
#pragma unroll
            for (int r = 0; r < B; ++r)
            {
                const int32_t *e = sd + (D::BB + r * B + r) * DESC;
                if (rr >= e[DMLO] && rr <= e[DMHI])
                {
                    double *q = S + (size_t)(D::BB + r * B + r) * D::RP + p;
                    *q = __dsub_rn(*q, __dmul_rn(m[it][r], alpha));
                }
            }
        }
    }
    return true;
}

// ---- bulk-copy staging (B = 1, interior tiles of a row-major triplet stream) ------------------------------
// When every row of the tile has its three entries and they sit in one contiguous run of the triplet values
// (JacobianMatrixShape emits (i,i-1),(i,i),(i,i+1) row by row: flame-1D), a thread's chunk of Q0 rows is 48
// consecutive doubles.  Each thread then fetches its own chunk with ONE bulk asynchronous copy
// (cp.async.bulk global -> shared, completion counted in bytes on an mbarrier; UBLKCP in the SASS) instead of 48
// eight-byte LDGSTS, and eliminates straight out of that array-of-rows image: chunk t sits at S + 50 t doubles
// (400 bytes: a multiple of 16 as the bulk copy wants; consecutive chunks start two 8-byte banks apart, a 2-way
// conflict on three loads per elimination step).  Source windows are aligned down to 16 bytes; `delta` is the
// 0- or 8-byte offset of the data inside the window, the same for every chunk of the tile.
struct PackedRows1
{
    const double *Cb;   // first interior row of the chunk: row k, field f at Cb[3 k + f]
    const double *prev; // the row before the chunk (previous chunk's separator, or the halo row of the tile)
    __device__ __forceinline__ Mat<1> get(int f, int k) const
    {
        Mat<1> X;
        X.a[0] = k >= 0 ? Cb[3 * k + f] : prev[f];
        return X;
    }
    __device__ __forceinline__ Mat<1> A(int k) const { return get(0, k); }
    __device__ __forceinline__ Mat<1> D(int k) const { return get(1, k); }
    __device__ __forceinline__ Mat<1> C(int k) const { return get(2, k); }
    __device__ __forceinline__ Mat<1> Hinit() const { return get(2, -1); }
};
constexpr int BULK_STRIDE = 50;                       // doubles between consecutive chunks of the image
constexpr int BULK_HALO = BULK_STRIDE * Cfg<1>::T;    // the tile's halo row sits behind the chunks
static_assert(BULK_HALO + 3 <= 3 * Dim<1>::RP, "the bulk image must fit the staged-tile buffer");

// Returns the chunk's row accessor through `rows`; false (uniform over the CTA) when the tile does not qualify.
__device__ __forceinline__ bool stage_matrix_bulk1(int t, int tile, double *S, const int32_t *desc, const double *jv, const double *mvals,
                                                   double alpha, PackedRows1 &rows)
{
    using D = Dim<1>;
    constexpr int PER = D::ROWS + 1, NFE = 3, NT = D::T, Q0 = D::Q0;
    __shared__ int32_t sd[NFE * DESC];
    __shared__ __align__(8) unsigned long long mbar;
    for (int k = t; k < NFE * DESC; k += NT) sd[k] = desc[(size_t)tile * NFE * DESC + k];
    if (t == 0)
    {
        const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mb) : "memory");
    }
    __syncthreads();
    bool ok = true;
#pragma unroll
    for (int fe = 0; fe < NFE; ++fe)
    {
        const int32_t *e = sd + fe * DESC;
        ok = ok && e[DJLO] == 0 && e[DJHI] == PER - 1 && e[DONE] >= PER && e[DJ0] == sd[DJ0] + fe && e[DJS] == NFE;
        const bool diag = fe == 1;
        ok = ok && (diag ? (e[DMLO] <= 1 && e[DMHI] == PER - 1) : (e[DMLO] > e[DMHI])); // a mass entry on every diagonal, none elsewhere
    }
    if (!ok) return false;
    const uint32_t mb = (uint32_t)__cvta_generic_to_shared(&mbar);
    constexpr uint32_t BYTES = BULK_STRIDE * 8;
    if (t == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(BYTES * NT) : "memory");
    const double *src = jv + sd[DJ0] + NFE * (1 + Q0 * t); // first interior row of this thread's chunk
    const uintptr_t sa = reinterpret_cast<uintptr_t>(src);
    const uint32_t delta = (uint32_t)(sa & 15u);
    const uint32_t dst = (uint32_t)__cvta_generic_to_shared(S + BULK_STRIDE * t);
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(sa - delta), "r"(BYTES), "r"(mb)
                 : "memory");
    // meanwhile: the halo row and this chunk's mass entries (diagonal field) through ordinary loads
    if (t < NFE) S[BULK_HALO + t] = jv[sd[DJ0] + t];
    const int32_t *em = sd + 1 * DESC;
    double m[Q0];
#pragma unroll
    for (int k = 0; k < Q0; ++k) m[k] = mvals[em[DM0] + em[DMS] * (1 + Q0 * t + k)];
    {   // wait for every chunk of the tile (the previous chunk's separator row is read as row -1)
        uint32_t done = 0;
        while (!done)
            asm volatile("{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(done) : "r"(mb) : "memory");
    }
    double *cb = S + BULK_STRIDE * t + (delta >> 3);
#pragma unroll
    for (int k = 0; k < Q0; ++k) cb[3 * k + 1] = __dsub_rn(cb[3 * k + 1], __dmul_rn(m[k], alpha)); // Jb = J - (M*alpha), solver.hpp:523
    rows.Cb = cb;
    rows.prev = t > 0 ? cb - BULK_STRIDE + 3 * (Q0 - 1) : S + BULK_HALO;
    return true;
}

// (wait_here: the caller left the dependency wait to this function -- it comes after the staging, which reads only the
// resident matrix values and writes shared memory; returns false if the iteration turned out to be cancelled)
template <int B, bool FIRST>
__device__ __forceinline__ bool factor_tile(const ChainArgs &a, const LevelView &lv, FactorShared<B> &sh, int t, int tile,
                                            int64_t mb, bool trigger, bool wait_here = false)
{
    using D = Dim<B>;
    double *Ft = lv.F + mb * a.s_F + (size_t)tile * D::TILE_F;
    const int nvalid = min(D::ROWS, lv.rows - tile * D::ROWS);
    const int tk = FIRST ? 0 : 1;
    stamp(a, tk, tile, 0);
    bool bulk = false; // the tile was staged by per-chunk bulk copies (B = 1): level 0 reads the array-of-rows image
    PackedRows1 prow{nullptr, nullptr};
    if (FIRST)
    {
        const double alpha = a.alpha_arr ? a.alpha_arr[mb] : a.alpha;
        bool affine = true;
#pragma unroll
        for (int fe = 0; fe < 3 * D::BB; ++fe) affine = affine && (a.flags[tile * 3 * D::BB + fe] & TF_AFF);
        if (affine)
        {
            if constexpr (B == 1)
                if (a.packed >= 2) bulk = stage_matrix_bulk1(t, tile, sh.S, a.desc, a.jv + mb * a.s_jv, a.mv + mb * a.s_mv, alpha, prow);
            if (!bulk && !stage_matrix_affine_async<B>(t, tile, sh.S, a.desc, a.jv + mb * a.s_jv, a.mv + mb * a.s_mv, alpha, a.packed != 0))
                stage_matrix_affine<B>(t, D::T, tile, sh.S, a.desc, a.jv + mb * a.s_jv, a.mv + mb * a.s_mv, alpha);
        }
        else stage_matrix_first<B>(t, D::T, tile, sh.S, a.sj, a.sm, a.flags, a.jv + mb * a.s_jv, a.mv + mb * a.s_mv, alpha);
    }
    else
        stage_matrix_reduced<B>(t, D::T, tile, lv.rows, sh.S, lv.red_in + mb * a.s_red);
    __syncthreads();
    if (wait_here)
    {
        pdl_wait();
        if (a.cancel && *reinterpret_cast<volatile int32_t *>(a.cancel)) return false;
    }
    stamp(a, tk, tile, 1);
    if (trigger) pdl_trigger();
    int fail = 0;
    {
        Mat<B> G, Dk, SL, H;
        if constexpr (B == 1)
        {
            if (bulk) chunk_factor<1, D::Q0 - 1>(prow, FacRef<1, D::T>{Ft + t}, a.tol, a.rel, fail, G, Dk, SL, H); // interior tile: every chunk is live
            else factor_phase0<B>(t, nvalid, sh, Ft, a.tol, a.rel, fail, G, Dk, SL, H);
        }
        else
            factor_phase0<B>(t, nvalid, sh, Ft, a.tol, a.rel, fail, G, Dk, SL, H);
        __syncthreads();
        factor_store1<B>(t, sh, G, Dk, SL, H);
    }
    __syncthreads();
    stamp(a, tk, tile, 2);
    factor_phase1<B>(t, nvalid, sh, Ft, a.tol, a.rel, fail);
    __syncthreads();
    stamp(a, tk, tile, 3);
    if (t == 0) factor_phase2<B>(tile, lv.ntiles, lv.is_top != 0, sh, Ft, lv.red_out + mb * a.s_red, a.tol, a.rel, fail);
    if (fail) atomicOr(&flag_words(a, mb)[3], 1);
    stamp(a, tk, tile, 4);
    return true;
}

template <int B, bool FIRST>
__global__ void __launch_bounds__(Cfg<B>::T) k_chain_factor(ChainArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    FactorShared<B> &sh = *reinterpret_cast<FactorShared<B> *>(smem);
    const int t = threadIdx.x, tile = blockIdx.x;
    const int64_t mb = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    if (FIRST && a.trace && t == 0 && tile == 0 && mb == 0)
    {   // keep the previous iteration's first and last stamp of tile 0: the idle time between iterations
        a.trace[6] = a.trace[5];
        a.trace[7] = a.trace[((size_t)4 * a.trace_tiles) * TRACE_SLOTS + 3];
    }
    stamp(a, FIRST ? 0 : 1, tile, 5);
    if (FIRST && a.pf_rhs)
    {   // the forward sweep that follows forms b = M*dxdt - f for these rows: by then its CTAs find the operands in L2
        // instead of waiting for DRAM with nothing else to do (prefetches neither read nor order anything)
        using D = Dim<B>;
        const int64_t k0 = (int64_t)tile * D::ROWS * B;
        const int bytes = (int)(min((int64_t)D::ROWS * B, (int64_t)a.n - k0) * sizeof(double));
        const char *src[6] = {reinterpret_cast<const char *>(a.fvec + k0), reinterpret_cast<const char *>(a.bdf.x0 + k0),
                              reinterpret_cast<const char *>(a.mdiag + k0), a.bdf.order >= 2 ? reinterpret_cast<const char *>(a.bdf.x1 + k0) : nullptr,
                              a.bdf.order >= 3 ? reinterpret_cast<const char *>(a.bdf.x2 + k0) : nullptr,
                              a.bdf.order >= 4 ? reinterpret_cast<const char *>(a.bdf.x3 + k0) : nullptr};
#pragma unroll
        for (int v = 0; v < 6; ++v)
            if (src[v])
                for (int off = t * 128; off < bytes; off += Cfg<B>::T * 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(src[v] + off));
    }
    const bool early = FIRST && a.stage_early; // single system, values resident: the wait moves behind the staging
    if (!early)
    {
        pdl_wait();
        if (a.cancel && *reinterpret_cast<volatile int32_t *>(a.cancel)) return;
        if (lane_mode(a, mb) != 2) return; // this lane keeps its factors (uniform over the lane's CTAs)
    }
    if (!factor_tile<B, FIRST>(a, a.lv, sh, t, tile, mb, true, early)) return;
    if (a.fuse_top && last_tile_done(a.counter + 2 * mb, a.lv.ntiles, t))
        factor_tile<B, false>(a, a.top, sh, t, 0, mb, false); // the grid-level reduced system, same CTA
}

// backward sweep of one tile and the write-out of its unknowns; sh.xR / sh.xL, sh.F12 and sh.Y12 are set
template <int B, bool FIRST>
__device__ __forceinline__ void backward_tail(const ChainArgs &a, const LevelView &lv, SolveShared<B> &sh, int t, int tile,
                                              int64_t mb, int nvalid, double *Ft, double *Yt)
{
    using D = Dim<B>;
    if (t == 0) backward_phase2<B>(sh);
    __syncthreads();
    backward_phase1<B>(t, nvalid, sh);
    __syncthreads();
    stamp(a, FIRST ? 4 : 3, tile, 6);
    backward_phase0<B>(t, nvalid, sh, Ft, Yt);
    __syncthreads();
    stamp(a, FIRST ? 4 : 3, tile, 7);
    constexpr int U = 8;
    const int total = B * nvalid;
    if (FIRST)
    {
        if (!a.out) return;
        double *__restrict__ out = a.out + mb * a.s_vec;
        double *__restrict__ xk = a.xk + mb * a.s_vec;
        const double *__restrict__ x0 = (a.lane_hbase ? a.hist + (int64_t)a.lane_hbase[mb] * a.s_hist : a.x0) + mb * a.s_vec;
        int flag = 0;
        double vmax = 0.0;
        for (int i0 = t; i0 < total; i0 += D::T * U)
        {
            int64_t o[U];
            double v[U], xa[U], xo[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
            {
                const int idx = i0 + u * D::T;
                const int q = idx / nvalid, row = idx - q * nvalid;
                const int64_t k = ((int64_t)tile * D::ROWS + row) * B + q;
                const bool ok = idx < total && k < a.n;
                o[u] = ok ? (a.perm ? (int64_t)a.perm[k] : k) : -1;
                v[u] = ok ? sh.R[(size_t)q * D::RP + sidx<B>(row)] : 0.0;
            }
            if (a.update)
            {
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (o[u] >= 0) { xa[u] = x0[o[u]]; xo[u] = xk[o[u]]; }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
            {
                if (o[u] < 0) continue;
                out[o[u]] = v[u];
                if (a.update)
                {   // dae-cpp/solver.hpp:590-619
                    const double e = fabs(v[u]), ax = fabs(xa[u]);
                    if (e > a.max_err_abs || isnan(v[u])) flag |= DAECPP_B200_DIVERGED;
                    if ((ax > 1e-14 && __ddiv_rn(e, ax) > a.rtol) || e > a.atol)
                        flag |= e > a.floor ? (DAECPP_B200_NOT_CONVERGED | DAECPP_B200_ABOVE_FLOOR) : DAECPP_B200_NOT_CONVERGED;
                    const double xn = __dadd_rn(xo[u], v[u]);
                    xk[o[u]] = xn;
                    if (a.xk_host) a.xk_host[o[u]] = xn;
                    if (a.var_on && fabs(xa[u]) > a.var_tol && fabs(xn) > a.var_tol)
                    {   // what accept_step would measure if this iterate is accepted (solver.hpp:703-717, vector_ops.cu: k_accept)
                        const double rel = fabs(__ddiv_rn(__dsub_rn(xa[u], xn), xa[u]));
                        if (rel > vmax) vmax = rel;
                    }
                }
            }
        }
        if (flag) atomicOr(&flag_words(a, mb)[2], flag);
        if (a.var_on && a.lane_var)
        {   // non-negative doubles order like their bit patterns
            unsigned long long bits = (unsigned long long)__double_as_longlong(vmax);
            for (int o2 = 16; o2 > 0; o2 >>= 1)
            {
                const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o2);
                bits = other > bits ? other : bits;
            }
            if ((t & 31) == 0 && bits) atomicMax(a.lane_var + mb, bits);
        }
        if (a.host_flags && last_tile_done(a.counter + 4, lv.ntiles, t, a.xk_host != nullptr))
        {
            if (t == 0)
            {   // published, reset for the next iteration, then the sequence number the host may spin on
                int32_t fw[3];
                for (int k = 0; k < 3; ++k) a.host_flags[k] = fw[k] = atomicExch(&a.dflags[2 + k], 0);
                if (a.cancel)
                {   // the iteration queued behind this one must not run if the host has something to do first:
                    // a failed pivot, a dependency time-out, or an update the host would refine (api.cu: wants_refinement)
                    const bool refine = a.refine_on && (fw[0] & DAECPP_B200_NOT_CONVERGED) && !(fw[0] & (DAECPP_B200_ABOVE_FLOOR | DAECPP_B200_DIVERGED));
                    if (fw[1] || fw[2] || refine) *a.cancel = 1;
                }
                __threadfence_system();
                *reinterpret_cast<volatile int32_t *>(a.host_flags + 3) = a.seq;
            }
        }
    }
    else
    {
        double *xl = lv.xlev + mb * a.s_x;
        for (int idx = t; idx < total; idx += D::T)
        {
            const int q = idx / nvalid, row = idx - q * nvalid;
            xl[((size_t)tile * D::ROWS + row) * B + q] = sh.R[(size_t)q * D::RP + sidx<B>(row)];
        }
    }
}

// stage_rhs_first with the residual formed on the fly: b_o = (0 + M_oo * dxdt_o) - f_o, the roundings of
// k_residual_diag (dae-cpp/solver.hpp:516-518 for a diagonal mass matrix).  All loads of a batch first.
template <int B>
__device__ __forceinline__ void stage_rhs_residual(int t, int tile, int n, double *R, const ChainArgs &a, int64_t mb)
{
    using D = Dim<B>;
    constexpr int TOTAL = B * D::ROWS, U = 4, NT = D::T;
    Bdf s = a.bdf;
    if (a.lane_w)
    {   // the lane's own BDF order, weights and history ring
        const int hb = a.lane_hbase[mb];
        s.order = a.lane_order[mb];
#pragma unroll
        for (int k = 0; k < 5; ++k) s.c[k] = a.lane_w[(int64_t)k * a.lanes + mb];
        s.xk = a.xk;
        s.x0 = a.hist + (int64_t)(hb & 3) * a.s_hist; s.x1 = a.hist + (int64_t)((hb + 1) & 3) * a.s_hist;
        s.x2 = a.hist + (int64_t)((hb + 2) & 3) * a.s_hist; s.x3 = a.hist + (int64_t)((hb + 3) & 3) * a.s_hist;
    }
    const int64_t sv = mb * a.s_vec;
    for (int i0 = t; i0 < TOTAL; i0 += NT * U)
    {
        int64_t o[U];
        double xk[U], x0[U], x1[U], x2[U], x3[U], m[U], f[U];
        int64_t ms[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * NT;
            const int q = idx / D::ROWS, row = idx - q * D::ROWS;
            const int64_t k = ((int64_t)tile * D::ROWS + row) * B + q;
            o[u] = (idx < TOTAL && k < n) ? (a.perm ? (int64_t)a.perm[k] : k) : -1;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) ms[u] = o[u] < 0 ? -1 : (a.mslot ? (int64_t)a.mslot[o[u]] : o[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const bool ok = o[u] >= 0;
            const int64_t j = ok ? o[u] + sv : 0;
            xk[u] = ok ? s.xk[j] : 0.0; x0[u] = ok ? s.x0[j] : 0.0;
            x1[u] = (ok && s.order >= 2) ? s.x1[j] : 0.0;
            x2[u] = (ok && s.order >= 3) ? s.x2[j] : 0.0;
            x3[u] = (ok && s.order >= 4) ? s.x3[j] : 0.0;
            m[u] = ms[u] >= 0 ? a.mdiag[ms[u] + mb * a.s_mv] : 0.0;
            f[u] = ok ? a.fvec[j] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * NT;
            if (idx < TOTAL)
            {
                const int q = idx / D::ROWS, row = idx - q * D::ROWS;
                // (a row without a mass element: the product M*dxdt holds +0.0 there)
                const double md = ms[u] >= 0 ? __dadd_rn(0.0, __dmul_rn(m[u], bdf_one(s, xk[u], x0[u], x1[u], x2[u], x3[u]))) : 0.0;
                const double v = o[u] >= 0 ? __dsub_rn(md, f[u]) : 0.0;
                R[(size_t)q * D::RP + sidx<B>(row)] = v;
                if (a.b_out && o[u] >= 0) a.b_out[o[u] + sv] = v;
            }
        }
    }
}

// ---- top level by parallel cyclic reduction ----------------------------------------------------------
// The grid-level system (one row per tile: C3 489 rows) is too small to occupy the GPU and too serial
// for one CTA: the chunked LU sweeps it in three dependent levels, forward and backward, ~35 dependent
// steps each way with a handful of threads busy.  Cyclic reduction takes log2(rows) steps with every
// thread busy: at stride s row i eliminates its couplings to rows i-s and i+s
//     al = -A_i D_{i-s}^-1,  ga = -C_i D_{i+s}^-1
//     A_i' = al A_{i-s},  C_i' = ga C_{i+s},  D_i' = D_i + al C_{i-s} + ga A_{i+s},  r_i' = r_i + al r_{i-s} + ga r_{i+s}
// and after the last step x_i = D_i^-1 r_i.  It works on the reduced rows the tiles left in HBM
// (A = G, D = Df - SL(next), C = Hf(next), chain_ops.hpp), so no top-level factors exist in this mode;
// pivot blocks are inverted with the same in-block partial pivoting and raise the same failure flag.
template <int B> struct Pcr
{
    static constexpr int ROWS = (B == 1) ? 512 : (B == 2 ? 128 : 64);
    static constexpr int NR = (ROWS + Cfg<B>::T - 1) / Cfg<B>::T;
    static_assert(sizeof(double) * ROWS * (3 * B * B + B) <= sizeof(SolveShared<B>), "cyclic reduction works inside the solve kernel's shared memory");
};
template <int B> __device__ __forceinline__ Mat<B> pcr_ldm(const double *p, int i)
{
    Mat<B> X;
#pragma unroll
    for (int e = 0; e < B * B; ++e) X.a[e] = p[e * Pcr<B>::ROWS + i];
    return X;
}
template <int B> __device__ __forceinline__ Vec<B> pcr_ldv(const double *p, int i)
{
    Vec<B> v;
#pragma unroll
    for (int q = 0; q < B; ++q) v.a[q] = p[q * Pcr<B>::ROWS + i];
    return v;
}
// Pivot-block inverse of the cyclic reduction.  B = 1: hardware reciprocal seed + two Newton steps, no
// special-case branch (the library reciprocal's slow-path call would serialise the independent chains);
// a zero or non-finite pivot is caught by the explicit test.  B > 1: Gauss-Jordan with in-block pivoting.
template <int B> __device__ __forceinline__ Mat<B> pcr_inv(const Mat<B> &X, double tol, int &fail)
{
    if constexpr (B == 1)
    {
        const double d = X.a[0];
        fail |= bad_pivot(d, tol) ? 1 : 0;
        double y;
        asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
        y = fma(y, fma(-d, y, 1.0), y);
        y = fma(y, fma(-d, y, 1.0), y);
        Mat<B> r;
        r.a[0] = y;
        return r;
    }
    else
        return minv<B>(X, tol, fail);
}

template <int B>
__device__ __noinline__ void pcr_top(int R, const double *__restrict__ red, const double *__restrict__ rv, double *__restrict__ xl,
                                     double tol, int32_t *fail_flag, double *__restrict__ W, int t, long long *trace, bool dry = false)
{
    auto mark = [&](int slot) {
        if (trace && t == 0) { long long ns; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(ns)); trace[slot] = ns; }
    };
    using D = Dim<B>;
    constexpr int BB = D::BB, PR = Pcr<B>::ROWS, NR = Pcr<B>::NR, NT = D::T;
    double *sA = W, *sD = W + BB * PR, *sC = W + 2 * BB * PR, *sr = W + 3 * BB * PR;
    {   // all global loads of the thread first (one L2 round trip), then the stores
        double va[NR][BB], vd[NR][BB], vd2[NR][BB], vc[NR][BB], vy[NR][B], vt[NR][B];
#pragma unroll
        for (int k = 0; k < NR; ++k)
        {
            const int i = t + k * NT;
            const bool in = i < R, nx = i + 1 < R;
#pragma unroll
            for (int e = 0; e < BB; ++e)
            {
                va[k][e] = in ? red[((size_t)0 * R + i) * BB + e] : 0.0;
                vd[k][e] = in ? red[((size_t)1 * R + i) * BB + e] : 0.0;
                vd2[k][e] = nx ? red[((size_t)2 * R + i + 1) * BB + e] : 0.0;
                vc[k][e] = nx ? red[((size_t)3 * R + i + 1) * BB + e] : 0.0;
            }
#pragma unroll
            for (int q = 0; q < B; ++q)
            {
                vy[k][q] = in ? rv[((size_t)0 * R + i) * B + q] : 0.0;
                vt[k][q] = nx ? rv[((size_t)1 * R + i + 1) * B + q] : 0.0;
            }
        }
#pragma unroll
        for (int k = 0; k < NR; ++k)
        {
            const int i = t + k * NT;
            if (i >= R) continue;
#pragma unroll
            for (int e = 0; e < BB; ++e) { sA[e * PR + i] = va[k][e]; sD[e * PR + i] = vd[k][e] - vd2[k][e]; sC[e * PR + i] = vc[k][e]; }
#pragma unroll
            for (int q = 0; q < B; ++q) sr[q * PR + i] = vy[k][q] - vt[k][q];
        }
    }
    __syncthreads();
    mark(1);
    int fail = 0;
#pragma unroll 1 // one copy of the step: the code runs once per launch in one CTA, every copy is an instruction-cache miss
    for (int s = 1; s < R; s <<= 1)
    {
        if (s == 8) mark(2);
        Mat<B> nA[NR], nD[NR], nC[NR];
        Vec<B> nr[NR];
        // Branch-free: a missing neighbour is the row itself with a zero multiplier, so the NR x 2
        // eliminations of a thread are independent instruction streams the scheduler can interleave
        // (one warp per scheduler here: there is no other latency hiding).
#pragma unroll
        for (int k = 0; k < NR; ++k)
        {
            const int i = min(t + k * NT, R - 1);
            const int im = max(i - s, 0), ip = min(i + s, R - 1);
            const double ml = (i - s >= 0) ? -1.0 : 0.0, mr = (i + s < R) ? -1.0 : 0.0;
            Mat<B> al = mmul<B>(pcr_ldm<B>(sA, i), pcr_inv<B>(pcr_ldm<B>(sD, im), tol, fail));
            Mat<B> ga = mmul<B>(pcr_ldm<B>(sC, i), pcr_inv<B>(pcr_ldm<B>(sD, ip), tol, fail));
#pragma unroll
            for (int e = 0; e < BB; ++e) { al.a[e] *= ml; ga.a[e] *= mr; }
            Mat<B> Di = pcr_ldm<B>(sD, i);
            Vec<B> ri = pcr_ldv<B>(sr, i);
            madd<B>(Di, mmul<B>(al, pcr_ldm<B>(sC, im)));
            madd<B>(Di, mmul<B>(ga, pcr_ldm<B>(sA, ip)));
            vadd<B>(ri, mv<B>(al, pcr_ldv<B>(sr, im)));
            vadd<B>(ri, mv<B>(ga, pcr_ldv<B>(sr, ip)));
            nA[k] = mmul<B>(al, pcr_ldm<B>(sA, im));
            nC[k] = mmul<B>(ga, pcr_ldm<B>(sC, ip));
            nD[k] = Di; nr[k] = ri;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < NR; ++k)
        {
            const int i = t + k * NT;
            if (i >= R) continue;
#pragma unroll
            for (int e = 0; e < BB; ++e) { sA[e * PR + i] = nA[k].a[e]; sD[e * PR + i] = nD[k].a[e]; sC[e * PR + i] = nC[k].a[e]; }
#pragma unroll
            for (int q = 0; q < B; ++q) sr[q * PR + i] = nr[k].a[q];
        }
        __syncthreads();
    }
    mark(3);
    for (int i = t; i < R; i += NT)
    {
        const Vec<B> x = mv<B>(pcr_inv<B>(pcr_ldm<B>(sD, i), tol, fail), pcr_ldv<B>(sr, i));
        if (dry) continue; // (a rehearsal on whatever the arrays hold: nothing leaves the CTA)
#pragma unroll
        for (int q = 0; q < B; ++q) xl[(size_t)i * B + q] = x.a[q];
    }
    if (fail && !dry) atomicOr(fail_flag, 1);
}

// forward sweep of one tile; the single tile of the top level carries on into the backward sweep
template <int B, bool FIRST>
__device__ __forceinline__ void forward_tile(const ChainArgs &a, const LevelView &lv, SolveShared<B> &sh, int t, int tile,
                                             int64_t mb, bool trigger, bool rhs_staged = false)
{
    using D = Dim<B>;
    double *Ft = lv.F + mb * a.s_F + (size_t)tile * D::TILE_F;
    double *Yt = lv.Y + mb * a.s_Y + (size_t)tile * D::TILE_Y;
    const int nvalid = min(D::ROWS, lv.rows - tile * D::ROWS);
    const int tk = FIRST ? 2 : 3;
    stamp(a, tk, tile, 0);
    if (rhs_staged) {}
    else if (FIRST && a.fuse_res) stage_rhs_residual<B>(t, tile, a.n, sh.R, a, mb);
    else if (FIRST) stage_rhs_first<B>(t, D::T, tile, a.n, sh.R, a.rhs + mb * a.s_vec, a.perm);
    else stage_rhs_reduced<B>(t, D::T, tile, lv.rows, sh.R, lv.rv_in + mb * a.s_rv);
    load_small_factors<B>(t, D::T, sh, Ft);
    __syncthreads();
    stamp(a, tk, tile, 1);
    if (trigger) pdl_trigger();
    forward_phase0<B>(t, nvalid, sh, Ft, Yt);
    __syncthreads();
    stamp(a, tk, tile, 2);
    forward_phase1<B>(t, nvalid, sh);
    __syncthreads();
    if (t == 0) forward_phase2<B>(tile, lv.ntiles, lv.is_top != 0, sh, lv.rv_out + mb * a.s_rv);
    __syncthreads();
    stamp(a, tk, tile, 3);
    if (!lv.is_top)
    {   // the level-1/2 forward results wait in HBM for the backward kernel
        copy_small(t, D::T, Yt + D::Y0, sh.Y12, D::Y1 + D::Y2);
        stamp(a, tk, tile, 4);
        return;
    }
    backward_tail<B, FIRST>(a, lv, sh, t, tile, mb, nvalid, Ft, Yt);
    stamp(a, tk, tile, 4);
}

template <int B, bool FIRST>
__global__ void __launch_bounds__(Cfg<B>::T, (B <= 2 ? 4 : 1)) k_chain_forward(ChainArgs a)
{
    extern __shared__ __align__(16) unsigned char smem[];
    SolveShared<B> &sh = *reinterpret_cast<SolveShared<B> *>(smem);
    const int t = threadIdx.x, tile = blockIdx.x;
    const int64_t mb = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    stamp(a, FIRST ? 2 : 3, tile, 5);
    // The launch in front is this iteration's factor kernel (it passed its own dependency wait before it let this grid start):
    // the right-hand side's sources -- iterate, history, f, mass values -- are final, only the factors are still being written
    if (FIRST && a.top_cta && tile == a.lv.ntiles)
    {   // The CTA of the grid-level system.  Which tile finishes last is not known in advance, so the cyclic reduction used
        // to start cold -- instruction cache, address arithmetic, the reduced matrix rows -- in whatever CTA that was, with
        // every other SM waiting for it.  This CTA rehearses it while the tiles sweep (results discarded), then waits for
        // their count and runs it for real.
        pdl_wait();
        if (a.cancel && *reinterpret_cast<volatile int32_t *>(a.cancel)) return;
        pdl_trigger();
        double *W = reinterpret_cast<double *>(smem);
        pcr_top<B>(a.top.rows, a.top.red_in + mb * a.s_red, a.top.rv_in + mb * a.s_rv, a.top.xlev + mb * a.s_x, a.tol, nullptr, W, t, nullptr, true);
        __syncthreads();
        if (t == 0)
        {
            volatile uint32_t *cnt = a.counter + 2 * mb + 1;
            long long t0, t1;
            asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
            while (*cnt < (uint32_t)a.lv.ntiles)
            {
                asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t1));
                if (t1 - t0 > 2000000000ll) { atomicOr(&flag_words(a, mb)[4], 1); break; } // (2 s: something is badly wrong; do not hang the GPU)
            }
            *cnt = 0u;
        }
        __syncthreads();
        __threadfence();
        stamp(a, 3, 0, 0);
        pcr_top<B>(a.top.rows, a.top.red_in + mb * a.s_red, a.top.rv_in + mb * a.s_rv, a.top.xlev + mb * a.s_x, a.tol, &flag_words(a, mb)[3], W, t,
                   a.trace ? a.trace + (size_t)3 * a.trace_tiles * TRACE_SLOTS : nullptr);
        stamp(a, 3, 0, 4);
        return;
    }
    const bool early = FIRST && a.rhs_early && !(a.extra_factor && tile == a.lv.ntiles);
    if (early)
    {
        using D = Dim<B>;
        if (a.fuse_res) stage_rhs_residual<B>(t, tile, a.n, sh.R, a, mb);
        else stage_rhs_first<B>(t, D::T, tile, a.n, sh.R, a.rhs + mb * a.s_vec, a.perm);
    }
    pdl_wait();
    if (a.cancel && *reinterpret_cast<volatile int32_t *>(a.cancel)) return;
    if (lane_mode(a, mb) == 0) return; // idle lane
    if (a.extra_factor && tile == a.lv.ntiles)
    {   // uniform over the CTA: the grid-level reduced system is factored here while the other CTAs sweep forward
        factor_tile<B, false>(a, a.top, *reinterpret_cast<FactorShared<B> *>(smem), t, 0, mb, true);
    }
    else
        forward_tile<B, FIRST>(a, a.lv, sh, t, tile, mb, true, early);
    if (FIRST && a.top_cta)
    {   // count this tile off: its reduced right-hand side is in global memory
        __threadfence();
        __syncthreads();
        if (t == 0) atomicAdd(a.counter + 2 * mb + 1, 1u);
        return;
    }
    if (a.fuse_top && last_tile_done(a.counter + 2 * mb + 1, a.lv.ntiles + a.extra_factor, t))
    {
        if (a.top_pcr)
        {
            stamp(a, 3, 0, 0);
            pcr_top<B>(a.top.rows, a.top.red_in + mb * a.s_red, a.top.rv_in + mb * a.s_rv, a.top.xlev + mb * a.s_x, a.tol, &flag_words(a, mb)[3],
                       reinterpret_cast<double *>(smem), t, a.trace ? a.trace + (size_t)3 * a.trace_tiles * TRACE_SLOTS : nullptr);
            stamp(a, 3, 0, 4);
        }
        else
            forward_tile<B, false>(a, a.top, sh, t, 0, mb, false); // top level: forward, then its backward sweep
    }
}

template <int B, bool FIRST>
__global__ void __launch_bounds__(Cfg<B>::T) k_chain_backward(ChainArgs a)
{
    using D = Dim<B>;
    extern __shared__ __align__(16) unsigned char smem[];
    SolveShared<B> &sh = *reinterpret_cast<SolveShared<B> *>(smem);
    const int t = threadIdx.x, tile = blockIdx.x;
    const int64_t mb = a.lane_list ? a.lane_list[blockIdx.y] : blockIdx.y;
    const LevelView &lv = a.lv;
    double *Ft = lv.F + mb * a.s_F + (size_t)tile * D::TILE_F;
    double *Yt = lv.Y + mb * a.s_Y + (size_t)tile * D::TILE_Y;
    const int nvalid = min(D::ROWS, lv.rows - tile * D::ROWS);
    if (lane_mode(a, mb) == 0) return; // idle lane
    stamp(a, 4, tile, 0);
    // The factors are two launches old (the forward kernel waited for the factor kernel before it let
    // this one start), so they may be touched before the dependency wait: while the top level is being
    // solved by the last CTA of the forward launch, pull this tile's factors into L2 and the small
    // levels into shared memory.
    if (a.prefetch)
    {
        const char *f0 = reinterpret_cast<const char *>(Ft);
        for (int off = t * 128; off < (int)(D::TILE_F * sizeof(double)); off += D::T * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(f0 + off));
        if (FIRST && a.update && !a.perm && a.prefetch > 1)
        {   // operands of the fused update (solver.hpp:590-619): x[0] and xk of this tile's rows
            const int64_t k0 = (int64_t)tile * D::ROWS * B;
            const int bytes = (int)(min((int64_t)D::ROWS * B, (int64_t)a.n - k0) * sizeof(double));
            const char *p0 = reinterpret_cast<const char *>(a.x0 + mb * a.s_vec + k0);
            const char *p1 = reinterpret_cast<const char *>(a.xk + mb * a.s_vec + k0);
            for (int off = t * 128; off < bytes; off += D::T * 128)
            {
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p0 + off));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p1 + off));
            }
        }
    }
    load_small_factors<B>(t, D::T, sh, Ft);
    pdl_wait();
    if (a.cancel && *reinterpret_cast<volatile int32_t *>(a.cancel)) return;
    stamp(a, 4, tile, 1);
    if (t < B)
    {
        const double *xup = lv.xup + mb * a.s_x;
        sh.xR[t] = xup[(size_t)tile * B + t];
        sh.xL[t] = tile > 0 ? xup[(size_t)(tile - 1) * B + t] : 0.0;
    }
    copy_small(t, D::T, sh.Y12, Yt + D::Y0, D::Y1 + D::Y2);
    __syncthreads();
    stamp(a, 4, tile, 2);
    pdl_trigger();
    backward_tail<B, FIRST>(a, lv, sh, t, tile, mb, nvalid, Ft, Yt);
    stamp(a, 4, tile, 3);
}

// slot map -> sources of the values: sj = index into the staged triplet values (or J's
// compressed values), sm = slot of M; negative codes (-1 absent, -2 one) pass through
__global__ void k_chain_compose(int64_t ne, int32_t per_fe, const int32_t *__restrict__ ent_u, const int32_t *__restrict__ uJ,
                                const int32_t *__restrict__ uM, const int32_t *__restrict__ srcJ, int32_t *__restrict__ sj,
                                int32_t *__restrict__ sm, int32_t *__restrict__ flags)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ne) return;
    const int32_t u = ent_u[e];
    int32_t j = u, m = -1;
    if (u >= 0 && uJ)
    {
        const int32_t a = uJ[u];
        j = a < 0 ? -1 : (srcJ ? srcJ[a] : a);
        m = uM[u];
    }
    sj[e] = j;
    sm[e] = m;
    if (m >= 0) atomicOr(&flags[e / per_fe], TF_MASS);
}

// one block per (tile, field element): is the value source an arithmetic progression over a range of
// rows, structural zeros before it and zeros / identity padding behind it?  (chain_ops.hpp, DESC)
__global__ void k_chain_affine(int32_t per_fe, const int32_t *__restrict__ sj, const int32_t *__restrict__ sm,
                               int32_t *__restrict__ flags, int32_t *__restrict__ desc)
{
    const size_t base = (size_t)blockIdx.x * per_fe;
    const int32_t fl = flags[blockIdx.x];
    __shared__ int32_t r[5]; // jlo, jhi, mlo, mhi, one
    if (threadIdx.x == 0) { r[0] = per_fe; r[1] = -1; r[2] = per_fe; r[3] = -1; r[4] = per_fe; }
    __syncthreads();
    for (int rr = threadIdx.x; rr < per_fe; rr += blockDim.x)
    {
        const int32_t s = sj[base + rr], m = sm[base + rr];
        if (s >= 0) { atomicMin(&r[0], rr); atomicMax(&r[1], rr); }
        if (m >= 0) { atomicMin(&r[2], rr); atomicMax(&r[3], rr); }
        if (s == -2) atomicMin(&r[4], rr);
    }
    __syncthreads();
    const int32_t jlo = r[0], jhi = r[1], mlo = r[2], mhi = r[3], one = r[4];
    const int32_t sb = jhi > jlo ? sj[base + jlo + 1] - sj[base + jlo] : 0;
    const int32_t s0 = jhi >= jlo ? sj[base + jlo] - sb * jlo : 0;
    const int32_t mb = mhi > mlo ? sm[base + mlo + 1] - sm[base + mlo] : 0;
    const int32_t m0 = mhi >= mlo ? sm[base + mlo] - mb * mlo : 0;
    int ok = 1;
    for (int rr = threadIdx.x; rr < per_fe; rr += blockDim.x)
    {
        const int32_t s = sj[base + rr], m = sm[base + rr];
        if (rr >= jlo && rr <= jhi) ok &= (s == s0 + sb * rr);
        else ok &= (s == (rr >= one ? -2 : -1));
        if (rr >= mlo && rr <= mhi) ok &= (m == m0 + mb * rr);
        else ok &= (m < 0);
    }
    ok = __syncthreads_and(ok);
    if (threadIdx.x == 0)
    {
        int32_t *d = desc + (size_t)blockIdx.x * DESC;
        d[DJ0] = s0; d[DJS] = sb; d[DM0] = m0; d[DMS] = mb;
        d[DJLO] = jlo; d[DJHI] = jhi; d[DMLO] = mlo; d[DMHI] = mhi; d[DONE] = one; d[9] = 0;
        flags[blockIdx.x] = ok ? (fl | TF_AFF) : (fl & ~TF_AFF);
    }
}

} // namespace

namespace
{

template <int B> constexpr size_t factor_smem() { return sizeof(FactorShared<B>); }
template <int B> constexpr size_t solve_smem() { return sizeof(SolveShared<B>); }

template <int B> int set_attributes(daecpp_b200_ctx *c)
{
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_factor<B, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)factor_smem<B>()));
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_factor<B, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)factor_smem<B>()));
    // the level-0 forward launch may carry the deferred top-level factorisation in one extra CTA
    const size_t fwd = factor_smem<B>() > solve_smem<B>() ? factor_smem<B>() : solve_smem<B>();
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_forward<B, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)fwd));
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_forward<B, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)solve_smem<B>()));
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_backward<B, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)solve_smem<B>()));
    DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_chain_backward<B, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)solve_smem<B>()));
    return DAECPP_B200_OK;
}

template <class T> int upload_vec(daecpp_b200_ctx *c, DevBuf<T> &d, const std::vector<T> &h)
{
    DAECPP_CUDA_TRY(c, d.reserve(h.size() ? h.size() : 1));
    if (!h.empty())
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(d.p, h.data(), sizeof(T) * h.size(), cudaMemcpyHostToDevice, c->stream));
    return DAECPP_B200_OK;
}

LevelView level_view(daecpp_b200_ctx *c, int g)
{
    const Plan &P = c->CH;
    const Level &L = P.levels[g];
    LevelView v{};
    v.rows = L.rows; v.ntiles = L.ntiles;
    v.is_top = (g + 1 == (int)P.levels.size());
    v.F = c->dC.F.p + L.fac0;
    v.Y = c->dC.Y.p + L.y0;
    v.red_out = c->dC.red.p + L.red0;
    v.rv_out = c->dC.rv.p + L.rv0;
    if (g > 0)
    {
        v.red_in = c->dC.red.p + P.levels[g - 1].red0;
        v.rv_in = c->dC.rv.p + P.levels[g - 1].rv0;
        v.xlev = c->dC.x.p + L.x0;
    }
    if (!v.is_top) v.xup = c->dC.x.p + P.levels[g + 1].x0;
    return v;
}

// fuse: the level above g is the single-tile top level and is swept by the last CTA of this launch
ChainArgs level_args(daecpp_b200_ctx *c, int g, bool fuse)
{
    ChainArgs a{};
    a.lv = level_view(c, g);
    a.n = c->CH.n;
    a.fuse_top = fuse ? 1 : 0;
    if (fuse) a.top = level_view(c, g + 1);
    a.counter = reinterpret_cast<uint32_t *>(c->dC.counter.p);
    a.dflags = c->d_flags;
    a.cancel = c->cancel_armed ? c->d_flags + 5 : nullptr;
    a.refine_on = (c->refine_steps > 0 && !c->direct_csc && !c->refine_futile) ? 1 : 0;
    if (const LaneSet *L = c->lanes)
    {
        const Plan &P = c->CH;
        a.lanes = L->count;
        a.grid_lanes = L->list_active ? L->n_active : L->count;
        a.lane_list = L->list_active;
        a.alpha_arr = L->w; // w[0][lane]
        a.s_jv = L->s_jv; a.s_mv = L->s_mv; a.s_vec = L->s_vec; a.s_hist = L->s_hist;
        a.s_F = P.fac_doubles; a.s_Y = P.y_doubles; a.s_red = P.red_doubles; a.s_rv = P.rv_doubles; a.s_x = P.x_doubles;
        a.lane_mode = L->mode; a.lane_hbase = L->hbase; a.lane_order = L->order; a.lane_w = L->w;
        a.hist = L->hist; a.lane_flags = L->flags; a.lane_var = L->var; a.var_on = L->var_on; a.var_tol = L->var_tol;
    }
    a.tol = c->pivot_tol;
    a.rel = c->pivot_rel;
    a.packed = c->chain_packed;
    a.prefetch = c->chain_prefetch;
    if (c->trace_units && c->CH.levels.size() == 2)
    {   // debug stamps, two-level hierarchies only (kernel ids would collide otherwise)
        const size_t m = (size_t)TRACE_KERNELS * c->CH.levels[0].ntiles * TRACE_SLOTS;
        if (c->trace.cap < m)
        {
            if (c->trace.reserve(m) == cudaSuccess) cudaMemsetAsync(c->trace.p, 0, sizeof(long long) * m, c->stream);
        }
        a.trace = c->trace.p;
        a.trace_tiles = c->CH.levels[0].ntiles;
    }
    return a;
}

template <class K> void launch_pdl(daecpp_b200_ctx *c, K kernel, const ChainArgs &a, int threads, size_t smem)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(a.lv.ntiles + a.extra_factor + a.top_cta, a.grid_lanes > 0 ? a.grid_lanes : 1, 1);
    cfg.blockDim = dim3(threads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = c->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, kernel, a);
}
template <int B> void launch_factor(daecpp_b200_ctx *c, const ChainArgs &a, bool first)
{
    if (first) launch_pdl(c, k_chain_factor<B, true>, a, Cfg<B>::T, factor_smem<B>());
    else launch_pdl(c, k_chain_factor<B, false>, a, Cfg<B>::T, factor_smem<B>());
}
template <int B> void launch_forward(daecpp_b200_ctx *c, const ChainArgs &a, bool first)
{
    const size_t smem = a.extra_factor ? (factor_smem<B>() > solve_smem<B>() ? factor_smem<B>() : solve_smem<B>()) : solve_smem<B>();
    if (first) launch_pdl(c, k_chain_forward<B, true>, a, Cfg<B>::T, smem);
    else launch_pdl(c, k_chain_forward<B, false>, a, Cfg<B>::T, smem);
}
template <int B> void launch_backward(daecpp_b200_ctx *c, const ChainArgs &a, bool first)
{
    if (first) launch_pdl(c, k_chain_backward<B, true>, a, Cfg<B>::T, solve_smem<B>());
    else launch_pdl(c, k_chain_backward<B, false>, a, Cfg<B>::T, solve_smem<B>());
}

#define CHAIN_DISPATCH(B, CALL)              \
    switch (B)                               \
    {                                        \
    case 1: CALL(1); break;                  \
    case 2: CALL(2); break;                  \
    default: CALL(3); break;                 \
    }

} // namespace

// Recognise the pattern as a block chain; on success ship the plan and compose the slot map.
// c->CH.active tells the caller whether the chain path took the pattern.
int chain_try_setup(daecpp_b200_ctx *c, int32_t n, const int32_t *outer, const int32_t *inner)
{
    Plan &P = c->CH;
    P.active = false;
    if (!c->chain_enabled || n < c->chain_min_n) return DAECPP_B200_OK;
    if (!detect(n, outer, inner, c->chain_max_b, P)) return DAECPP_B200_OK;
    int rc;
    if ((rc = upload_vec(c, c->dC.ent_u, P.ent_u))) return rc;
    if ((rc = upload_vec(c, c->dC.flags, P.tile_flags))) return rc;
    if (!P.identity_perm && (rc = upload_vec(c, c->dC.perm, P.perm))) return rc;
    DAECPP_CUDA_TRY(c, c->dC.sj.reserve(P.ent_u.size() + 1));
    DAECPP_CUDA_TRY(c, c->dC.sm.reserve(P.ent_u.size() + 1));
    DAECPP_CUDA_TRY(c, c->dC.F.reserve((size_t)P.fac_doubles + 1));
    DAECPP_CUDA_TRY(c, c->dC.Y.reserve((size_t)P.y_doubles + 1));
    DAECPP_CUDA_TRY(c, c->dC.red.reserve((size_t)P.red_doubles + 1));
    DAECPP_CUDA_TRY(c, c->dC.rv.reserve((size_t)P.rv_doubles + 1));
    DAECPP_CUDA_TRY(c, c->dC.x.reserve((size_t)P.x_doubles + 1));
    DAECPP_CUDA_TRY(c, c->dC.counter.reserve(8));
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->dC.counter.p, 0, 8 * sizeof(int32_t), c->stream));
#define SETATTR(B) rc = set_attributes<B>(c)
    CHAIN_DISPATCH(P.B, SETATTR)
#undef SETATTR
    if (rc) return rc;
    // sources of the matrix values (same rule as the general path: the staged triplet values when
    // every slot has exactly one triplet, else J's duplicate-summed compressed values)
    c->jvals_from_triplets = !c->direct_csc && c->J.no_duplicates && c->J.nnz > 0;
    const int32_t *uJ = c->direct_csc ? nullptr : c->u_J.p;
    const int32_t *srcJ = (!c->direct_csc && c->jvals_from_triplets) ? c->J.src.p : nullptr;
    const int64_t ne = (int64_t)P.ent_u.size();
    const int32_t per_fe = rows_per_tile(P.B) + 1;
    k_chain_compose<<<(int)((ne + 255) / 256), 256, 0, c->stream>>>(ne, per_fe, c->dC.ent_u.p, uJ, c->u_M.p, srcJ, c->dC.sj.p,
                                                                  c->dC.sm.p, c->dC.flags.p);
    DAECPP_CUDA_TRY(c, c->dC.desc.reserve(P.tile_flags.size() * DESC + 4));
    if (c->chain_affine)
        k_chain_affine<<<(int)P.tile_flags.size(), 256, 0, c->stream>>>(per_fe, c->dC.sj.p, c->dC.sm.p, c->dC.flags.p, c->dC.desc.p);
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    c->kernels += 2;
    return DAECPP_B200_OK;
}

// With exactly two levels (n up to ROWS^2 block rows) the top level rides in the last CTA of the level-0
// launches: one launch per refactorisation, two per solve.
// chain_fuse_top bit 0: refactorisation, bit 1: solve.  Measured on B200 (C3): fusing the solve's top level
// saves 3.6 us, fusing the factorisation's costs 5.6 us (profiles/r01_v6_*), hence the default 2.
static bool fuse_two_levels(const daecpp_b200_ctx *c, int bit) { return (c->chain_fuse_top & bit) && c->CH.levels.size() == 2; }

int chain_refactor_async(daecpp_b200_ctx *c, double alpha)
{
    const Plan &P = c->CH;
    const double *jv = c->direct_csc ? c->u_vals.p : (c->jvals_from_triplets ? c->J.A.p : c->J.vals.p);
    const bool fuse = fuse_two_levels(c, 1);
    if (c->sm_count == 0) cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, c->device);
    // at most 4 CTAs per SM with the factor kernel's shared memory: the forward launch must stay one wave
    const bool defer = !fuse && c->chain_may_defer && c->chain_defer_top && fuse_two_levels(c, 2) && P.levels[0].ntiles + 1 <= 4 * c->sm_count;
    // ... or no top-level factorisation at all: the solve reduces the top system cyclically (pcr_top)
    const int pcr_rows = P.B == 1 ? Pcr<1>::ROWS : (P.B == 2 ? Pcr<2>::ROWS : Pcr<3>::ROWS);
    const bool pcr = !fuse && c->chain_may_defer && c->chain_top_pcr && fuse_two_levels(c, 2) && P.levels[1].rows <= pcr_rows;
    c->chain_top_mode = pcr ? 1 : 0;
    c->chain_top_pending = defer && !pcr;
    const int nlaunch = (fuse || defer || pcr) ? 1 : (int)P.levels.size();
    for (int g = 0; g < nlaunch; ++g)
    {
        ChainArgs a = level_args(c, g, fuse);
        if (c->lanes && c->lanes->list_refactor) { a.lane_list = c->lanes->list_refactor; a.grid_lanes = c->lanes->n_refactor; }
        if (g == 0)
        {
            a.sj = c->dC.sj.p; a.sm = c->dC.sm.p; a.flags = c->dC.flags.p; a.desc = c->dC.desc.p;
            a.jv = c->lanes ? c->lanes->jv : jv; a.mv = c->lanes ? c->lanes->mv : c->M.vals.p; a.alpha = alpha;
            a.stage_early = (!c->lanes && c->values_quiescent && c->jvals_from_triplets) ? 1 : 0;
            if (!c->lanes && c->rhs_prefetch_hint && P.identity_perm && c->m_diag)
            {   // (single system, natural order, full diagonal mass: the rows of a tile are one contiguous range of every source)
                a.pf_rhs = 1; a.bdf = c->bdf_pending; a.fvec = c->f; a.mdiag = c->M.vals.p;
            }
        }
#define LF(B) launch_factor<B>(c, a, g == 0)
        CHAIN_DISPATCH(P.B, LF)
#undef LF
        c->kernels += 1;
    }
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

// b -> dx (and, fused into the last kernel, the error checks and xk += dx)
int chain_solve(daecpp_b200_ctx *c, bool fuse_update, double atol, double rtol, double max_err_abs)
{
    const Plan &P = c->CH;
    const int nlev = (int)P.levels.size();
    const bool fuse = fuse_two_levels(c, 2);
    auto first_args = [&](ChainArgs &a) {
        a.rhs = c->b;
        a.perm = P.identity_perm ? nullptr : c->dC.perm.p;
        a.out = c->dx; a.xk = c->xk; a.x0 = c->xh[0];
        a.atol = atol; a.rtol = rtol; a.max_err_abs = max_err_abs; a.floor = c->refine_floor;
        a.update = fuse_update ? 1 : 0;
        a.rhs_early = c->forward_follows_factor ? 1 : 0;
        if (c->fuse_residual_now)
        {
            a.fuse_res = 1; a.bdf = c->bdf_pending; a.mdiag = c->lanes ? c->lanes->mv : c->M.vals.p; a.fvec = c->f;
            a.mslot = c->m_pdiag ? c->m_dslot.p : nullptr;
            a.b_out = c->keep_b ? c->b : nullptr;
        }
        a.host_flags = c->publish_flags ? c->h_flags_dev + 2 + 8 * c->publish_slot : nullptr;
        if (a.host_flags) { c->flags_published = true; a.seq = (int32_t)++c->flag_seq; }
        if (a.host_flags && fuse_update && c->xk_host_target) { a.xk_host = c->xk_host_target; c->xk_direct_used = true; }
    };
    if ((c->chain_top_pending || c->chain_top_mode == 1) && !fuse)
    {   // (options changed in between) factor the top level on its own
        c->chain_top_mode = 0;
        ChainArgs a = level_args(c, 1, false);
#define LF(B) launch_factor<B>(c, a, false)
        CHAIN_DISPATCH(P.B, LF)
#undef LF
        c->kernels += 1;
        c->chain_top_pending = false;
    }
    for (int g = 0; g < (fuse ? 1 : nlev); ++g)
    {
        ChainArgs a = level_args(c, g, fuse);
        if (g == 0) first_args(a);
        if (g == 0 && c->chain_top_pending) { a.extra_factor = 1; c->chain_top_pending = false; }
        if (g == 0 && c->chain_top_mode == 1) a.top_pcr = 1;
        // a CTA of its own for the grid-level solve: single system (with lanes the spinning CTAs of later lanes could keep
        // tiles of earlier ones off the SMs), every CTA of the launch resident at once
        if (g == 0 && a.top_pcr && fuse && !a.extra_factor && !c->lanes && c->chain_top_cta && c->sm_count > 0 &&
            a.lv.ntiles + 1 <= 4 * c->sm_count)
            a.top_cta = 1;
#define LW(B) launch_forward<B>(c, a, g == 0)
        CHAIN_DISPATCH(P.B, LW)
#undef LW
        c->kernels += 1;
    }
    for (int g = nlev - 2; g >= 0; --g)
    {
        ChainArgs a = level_args(c, g, false);
        if (g == 0) first_args(a);
#define LB(B) launch_backward<B>(c, a, g == 0)
        CHAIN_DISPATCH(P.B, LB)
#undef LB
        c->kernels += 1;
    }
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

} // namespace daecpp_b200
