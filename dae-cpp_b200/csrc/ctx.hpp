// Internal context layout shared by the .cu translation units of libdaecpp_b200.so.
#pragma once
#include "bdf.hpp"
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>

#include "../../include/daecpp_b200.h"
#include "chain.hpp"
#include "symbolic.hpp"

namespace daecpp_b200
{

// Device memory of the library.  With pooling on (daecpp_b200_set_pooling) a freed block of 4 MB or more is kept, per
// device and size, for the next request of that size: cudaFree of gigabytes costs tens to hundreds of milliseconds
// (more with peer mappings in the process), and an ensemble driver builds the same buffers sweep after sweep.
cudaError_t dev_alloc(void **p, size_t bytes);
void dev_free(void *p);
void dev_cache_release(); // hand every cached block back to the driver

// Grow-only device buffer.
template <class T>
struct DevBuf
{
    T *p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t count)
    {
        if (count <= cap) return cudaSuccess;
        if (p) dev_free(p);
        p = nullptr; cap = 0;
        size_t want = count + count / 8 + 16;
        cudaError_t e = dev_alloc((void **)&p, want * sizeof(T));
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) dev_free(p); p = nullptr; cap = 0; }
};

// One triplet stream (mass matrix or Jacobian) and its compressed-column form on the device.
struct TripletSet
{
    // staging of the latest upload
    DevBuf<double> A;
    DevBuf<int32_t> ti, tj;
    int64_t nT = 0;
    // cached pattern (valid while the (i,j) stream repeats)
    bool have_pattern = false;
    int64_t nT_cached = -1;
    DevBuf<int32_t> ti_cached, tj_cached;
    DevBuf<int32_t> src;     // [nT]  triplet indices ordered by (col,row,input order)
    DevBuf<int32_t> src_ptr; // [nnz+1] first position in src of every slot
    DevBuf<int32_t> outer;   // [n+1]
    DevBuf<int32_t> inner;   // [nnz]
    DevBuf<double> vals;     // [nnz]
    int64_t nnz = 0;
    bool no_duplicates = false;
    bool vals_valid = false; // vals[] reflects the latest upload
    uint64_t pattern_version = 0;
    void release()
    {
        A.release(); ti.release(); tj.release(); ti_cached.release(); tj_cached.release();
        src.release(); src_ptr.release(); outer.release(); inner.release(); vals.release();
    }
};

struct DeviceSymbolic
{
    // global front tables (used by the per-level kernels that handle big fronts)
    DevBuf<int32_t> perm_rhs; // row of the right-hand side that feeds elimination position k (= perm unless rows were moved to fill the diagonal)
    DevBuf<int32_t> perm, iperm, sn_first, sn_rows, sn_rel, child_ptr, child_idx, big_sn;
    DevBuf<int64_t> sn_rowptr, sn_lu, sn_cb;
    // subtree-unit schedule (schedule.hpp)
    DevBuf<int32_t> blob, unit_blob, ent_u, ent_dst, ent_sj, ent_sm, big_u, big_sj, big_sm;
    DevBuf<uint32_t> big_dst;
    DevBuf<int32_t> sync_words; // [0] ticket, [1..nunit] dep counters, [1+nunit..] backward done epochs
    void release()
    {
        perm.release(); perm_rhs.release(); iperm.release(); sn_first.release(); sn_rows.release(); sn_rel.release();
        child_ptr.release(); child_idx.release(); big_sn.release(); sn_rowptr.release(); sn_lu.release();
        sn_cb.release(); blob.release(); unit_blob.release(); ent_u.release(); ent_dst.release();
        ent_sj.release(); ent_sm.release(); big_u.release(); big_sj.release(); big_sm.release(); big_dst.release();
        sync_words.release();
    }
};

// device side of the block-chain plan (chain.hpp)
struct DeviceChain
{
    DevBuf<int32_t> ent_u, sj, sm, flags, perm, desc, counter;
    DevBuf<double> F, Y, red, rv, x;
    void release()
    {
        ent_u.release(); sj.release(); sm.release(); flags.release(); perm.release(); desc.release(); counter.release();
        F.release(); Y.release(); red.release(); rv.release(); x.release();
    }
};

// Lock-step batched ensemble (api: daecpp_b200_batch_*, batch.cu): `count` systems of the same pattern share
// one plan; every per-system array of the chain path becomes [count][...] and the chain kernels take the
// lane as blockIdx.y.  Per-lane control words are device arrays refreshed once per Newton round.
struct LaneSet
{
    int32_t count = 0;
    int64_t s_vec = 0, s_hist = 0, s_jv = 0, s_mv = 0;
    double *hist = nullptr;                    // [4][count][s_vec] ring of accepted states; x[k] of a lane = slot (hbase + k) & 3
    const int32_t *mode = nullptr;             // [count] DAECPP_B200_LANE_*
    const int32_t *hbase = nullptr;            // [count]
    const int32_t *order = nullptr;            // [count]
    // compacted lane lists of the round: the kernels are launched over these (grid.y = list length), so a round with
    // few lanes in flight costs few CTAs
    const int32_t *list_active = nullptr, *list_refactor = nullptr;
    int32_t n_active = 0, n_refactor = 0;
    const double *w = nullptr;                 // [5][count] BDF weights, w[0] = alpha
    int32_t *flags = nullptr;                  // [count][4]: update flags, pivot failure, wait time-out, spare
    unsigned long long *var = nullptr;         // [count] bit pattern of the lane's maximum variability
    const double *jv = nullptr, *mv = nullptr; // triplet values of J [count][s_jv], compressed values of M [count][s_mv]
    int32_t var_on = 0;
    double var_tol = 0.0;
    double *f = nullptr, *b = nullptr, *dx = nullptr, *xk = nullptr; // lane slabs [count][s_vec]
    double *refine_r = nullptr, *refine_e = nullptr;                 // [count][s_vec], allocated on first use
};

// Tiny systems (tiny.cu): one-launch Newton iteration with mapped page-locked hand-off.
constexpr int TINY_MAX_N = 32;
struct TinyState
{
    bool enabled = true;
    std::vector<int32_t> Ji, Jj, Mi, Mj; // the cached (i, j) streams, compared on the host
    bool haveJ = false, haveM = false;
    bool staleJ = false, staleM = false; // the regular device-side triplet staging lags behind the mapped values
    DevBuf<int32_t> jptr, jidx, mptr, midx; // dense slot lists
    DevBuf<double> lu;                      // n*n factors + the row permutation
    bool have_factor = false;
    // one mapped page-locked block: f[n] | xk[n] | jv[capJ] | mv[capM] | var | ... | res[8]
    char *h = nullptr, *d = nullptr;
    size_t capJ = 0, capM = 0;
    double *h_f = nullptr, *h_xk = nullptr, *h_jv = nullptr, *h_mv = nullptr, *h_var = nullptr;
    int32_t *h_res = nullptr;
    uint32_t seq = 0;
    double var_tol = 1e-4; // tolerance of the speculative variability (SolverOptions::variability_tolerance default)
    double spec_var = 0.0;
    bool spec_valid = false;
};

struct GroupLaunch
{
    int32_t kind = 0, begin = 0, count = 0;
    int32_t group = 32, max_dim = 0; // kind 1: threads cooperating on one front
    int32_t grid = 0;                // kind 0: persistent blocks
    int32_t overlap = 0;             // kind 0: independent of the big-front group behind it (symbolic.hpp: same_stage)
    int32_t max_s = 0, max_r = 0, max_children = 0; // kind 1: over the fronts of the level (pivots, rows below, children)
};

} // namespace daecpp_b200

// Everything daecpp_b200_set_option can change, with its default.  A pooled context handed to a new owner gets a fresh
// copy (api.cu: recycle), so an option added here cannot leak from one owner to the next.
struct daecpp_b200_options
{
    // plan-affecting (a change invalidates the symbolic analysis)
    daecpp_b200::SymbolicOptions sopt;
    int unit_threads = 0;  // threads per subtree unit; 0: 256, or 128 where the automatic unit budget hit its ceiling (symbolic.cpp)
    bool chain_enabled = true, chain_affine = true;
    int chain_min_n = 256, chain_max_b = 3;
    // chain path
    bool lean_flags = true;            // flag words published to a mapped host mirror by the CTA that finishes the iteration
    bool spin_flags = true;            // resident iteration: spin on that word instead of cudaStreamSynchronize
    bool pipeline_iterations = true;   // daecpp_b200_iterations_resident: iteration k+1 enqueued while k runs
    bool host_xk_direct = true;        // small systems: the fused update stores xk straight into the caller's pinned buffer
    int64_t host_xk_direct_max_n = 131072;
    bool chain_fuse_residual = true;   // (partial) diagonal mass: b = M*dxdt - f is formed while the forward kernel stages its right-hand side
    bool keep_b = false;               // fused residual: also write b to HBM (get_vector(b) after a fused iteration)
    int chain_prefetch = 0;            // backward kernel: L2 prefetch of the tile's factors before the dependency wait
    int chain_packed = 2;              // staging of affine tiles (chain.cu): 2 = one bulk copy per thread chunk where the tile qualifies,
                                       // 1 = contiguous-run LDGSTS, 0 = per-field strided copies
    bool chain_defer_top = true;       // two-level chains inside a fused iteration: the top level is factored by one
                                       // extra CTA of the forward launch, concurrently with the tiles' forward sweeps
    bool chain_top_pcr = true;         // ... or solved by parallel cyclic reduction with no top-level factors at all (chain.cu: pcr_top)
    int chain_fuse_top = 2;            // bit 0: factor, bit 1: solve (top level swept by the last CTA of the level-0 launch)
    bool chain_top_cta = true;         // fused iteration, cyclic-reduction top level: a dedicated CTA of the forward launch rehearses and runs it
    // general multifrontal path
    bool blocked_lu = true;            // big fronts: blocked multi-CTA LU (numeric.cu) instead of one CTA per front
    int blocked_lu_min_dim = 96;       // level groups whose largest front is at least this order
    bool bf_pdl = true;                // big fronts: panel / trailing-update launches chained by programmatic dependent launch
    bool bf_dmma = true;               // big fronts: trailing update on the FP64 tensor cores (mma.m8n8k4.f64), 0 = FP64 FMA tiles
    bool bf_fused_panel = true;        // big fronts: panel in one launch (numeric.cu: k_bf_panel), 0 = diag / rows / trsm launches
    bool bf_fast_solve = true;         // big fronts: triangular sweeps with the pivot unknowns in shared memory and the factor loads of a
                                       // block issued ahead of the dependent chain, matrix-vector kernels with batched loads (same
                                       // operations in the same order: bit-identical to 0 = the round-1 kernels)
    bool bf_diag_copy = true;          // big fronts, fused panel: the factored diagonal block is handed from the panel launch to the trailing
                                       // update through a side buffer and copied into place instead of being factored a second time
    bool bf_lu_lookahead = true;       // big fronts: the reciprocal of pivot k+1 is formed by every thread during the update of pivot k
                                       // (off the barrier path); same bits
    bool bf_schur = false;             // big fronts: contribution block updated in one pass after the last block column (k_bf_schur).
                                       // Bit-identical and it reads / writes the block once instead of once per block column, but measured
                                       // slower on C4 (11.8 vs 11.1 ms): a block column costs ~40 us of launch + dependent latency whatever
                                       // its trailing update touches, so the extra pass per level is not paid back
    bool bf_zero_ahead = true;         // big fronts: all contribution blocks zeroed by one launch on a side stream at the start of the
                                       // refactorisation, under the latency-bound unit launches, instead of one launch per level
    bool bf_overlap_units = true;      // a launch of subtree units that sits in the same stage of the tree as the level of big fronts behind it
                                       // runs beside that level on the side stream (they do not depend on each other)
    int bf_schur_fronts = 0;           // ... only on levels with at least this many fronts and more than one block column (0: never;
                                       // measured slower at 30, 100 and 400 as well)
    // numerics
    // Guarded iterative refinement (one step of r = b - Jb*dx, Jb*e = r, dx += e): taken only when the Newton update
    // failed the convergence test on elements whose |dx| is below refine_floor, i.e. where the outcome of the
    // reference's relative test (solver.hpp:603-609) hangs on the componentwise accuracy of the linear solve
    int refine_steps = 1;
    double refine_floor = 1e-11;
    double pivot_tol = 0.0;
    // chain path: a pivot below pivot_rel x (largest entry of its own matrix row) has lost every significant digit
    // to cancellation and is reported like a zero pivot (Eigen would have pivoted away from it)
    double pivot_rel = 1e-14;
    // hand-off / debugging
    bool assemble_only = false;
    bool pinned_io = false;            // host vectors handed to the ABI are page-locked: no staging copies
    bool phase_timing = false;         // per-phase device times of iteration_resident
    bool trace_units = false;          // debug: per-unit globaltimer stamps of the factor kernel
    int trace_unit_index = 0;          // debug: unit (of the first launch group) whose local levels are stamped one by one
};

struct daecpp_b200_ctx : daecpp_b200_options
{
    int32_t n = 0;
    int device = 0;
    cudaStream_t stream = nullptr;
    std::string err;

    // state vectors (solver.hpp:301-337 equivalents), all length n, resident in HBM
    double *xk = nullptr, *xh[4] = {nullptr, nullptr, nullptr, nullptr};
    double *f = nullptr, *b = nullptr, *dx = nullptr, *work = nullptr, *dxdt = nullptr;
    int32_t *d_flags = nullptr; // [8] device flag words
    int32_t *h_flags = nullptr; // pinned mirror (mapped: the last kernel of a fused chain iteration writes it directly)
    int32_t *h_flags_dev = nullptr;
    bool publish_flags = false, flags_published = false;
    // pipelined resident iterations (api.cu: daecpp_b200_iterations_resident): iteration k+1 is enqueued before the host has
    // seen k's flags; the mirror has two slots ([2..5] and [10..13]) and the device cancels k+1 if k needs the host
    bool values_quiescent = false;       // resident iterations: nothing on the stream in front of the factor kernel writes the matrix values
    bool rhs_prefetch_hint = false;      // the fused residual of this iteration will be formed by the forward launch: bdf_pending is set
    bool forward_follows_factor = false; // the forward launch being built comes right behind this iteration's factor launch
    int publish_slot = 0;
    bool cancel_armed = false;
    // small systems: the fused update also stores xk straight into the caller's pinned buffer (mapped), so a host
    // iteration needs neither a device-to-host copy node nor a stream synchronisation
    const double *xk_host_cached = nullptr;
    double *xk_host_dev_cached = nullptr, *xk_host_target = nullptr;
    bool xk_direct_used = false;
    uint32_t flag_seq = 0;             // sequence number written behind the published flags (h_flags[5])
    bool flags_clean = false;          // d_flags[2..4] are zero (the publishing CTA of a fused chain iteration resets them)
    bool residual_was_fused = false;   // the last fused iteration formed b inside the forward kernel (b is not in HBM)
    bool fuse_residual_now = false;    // set by the fused iteration for the solve that follows
    daecpp_b200::Bdf bdf_pending{};
    double *d_scalars = nullptr, *h_scalars = nullptr; // [8]
    double *h_stage = nullptr;  // pinned staging for n doubles
    size_t h_stage_cap = 0;

    daecpp_b200::TripletSet M, J;
    // CSR view of M for the residual (row-wise accumulation in column order)
    daecpp_b200::DevBuf<int32_t> m_rowptr, m_rowslot, m_rowcol;
    uint64_t m_csr_version = ~0ull;
    // CSR view of J (slots of the compressed-column form by row): only the guarded refinement step reads it
    daecpp_b200::DevBuf<int32_t> j_rowptr, j_rowslot, j_rowcol;
    uint64_t j_csr_version = ~0ull;
    daecpp_b200::DevBuf<double> refine_r, refine_e; // residual and correction of the refinement step
    bool m_diag = false; // M is diagonal with slots in row order: index-free residual kernel
    bool m_pdiag = false; // every stored element of M lies on the diagonal, some rows have none (a DAE's algebraic rows, or M = 0):
                          // m_dslot[row] = slot of M_rr or -1, and the residual is still formed inside the forward kernel
    daecpp_b200::DevBuf<int32_t> m_dslot;

    // union pattern of J and M = pattern of Jb
    uint64_t union_J_version = ~0ull, union_M_version = ~0ull;
    int64_t nnzU = 0;
    daecpp_b200::DevBuf<int32_t> u_outer, u_inner, u_J, u_M;
    daecpp_b200::DevBuf<double> u_vals;
    bool direct_csc = false; // matrix given through the linear-solver-only entry

    // symbolic + numeric factor
    bool have_symbolic = false, have_factor = false;
    daecpp_b200::Symbolic S;
    daecpp_b200::DeviceSymbolic dS;
    std::vector<daecpp_b200::GroupLaunch> groups;
    // block-chain path (banded patterns): taken instead of the multifrontal schedule when the
    // pattern is recognised as a chain of blocks of order <= chain_max_b
    daecpp_b200::chain::Plan CH;
    daecpp_b200::DeviceChain dC;
    const daecpp_b200::LaneSet *lanes = nullptr; // set by a batch context around its launches (batch.cu)
    daecpp_b200::TinyState tiny;
    int sm_count = 0;
    bool chain_may_defer = false;      // set by the fused iteration around its refactor + solve pair
    int chain_top_mode = 0;            // 0: top-level LU factors are current, 1: there are none (cyclic reduction at solve time)
    bool chain_top_pending = false;    // level-0 factors are current, the top level's are not yet
    uint32_t ticket_base = 0; // value of the ticket word before the next unit launch
    int32_t solve_epoch = 0;
    int unit_threads_eff = 256; // what unit_threads resolved to for the current plan
    int unit_grid_factor = 0, unit_grid_solve = 0; // resident blocks (SMs x occupancy)
    daecpp_b200::DevBuf<long long> trace;
    bool jvals_from_triplets = false; // fused assembly reads the staged triplet values directly
    daecpp_b200::DevBuf<double> LU, CB, upd;
    daecpp_b200::DevBuf<double> bf_diag; // big fronts: factored diagonal block of the current block column, per front of the level (bf_diag_copy)
    double last_alpha = 0.0;
    double alpha_hint = 0.0; // alpha of the factorisation about to follow a symbolic analysis (row matching weights)

    // scratch for the cold (sorting) path
    daecpp_b200::DevBuf<uint64_t> keys_a, keys_b;
    daecpp_b200::DevBuf<int32_t> pay_a, pay_b, scan;
    daecpp_b200::DevBuf<char> cub_tmp;

    // guarded refinement (options: refine_steps, refine_floor)
    double n_refine = 0;
    bool refine_futile = false; // the last refinement of this Newton loop left the outcome as it was: not repeated until the step changes

    // statistics
    double n_symbolic = 0, n_refactor = 0, n_solve = 0, n_cold = 0, kernels = 0;
    double ms_symbolic = 0, ms_last_refactor = 0, ms_last_solve = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_mark = nullptr;
    cudaStream_t stream_zero = nullptr;            // big fronts: contribution blocks zeroed beside the unit launches (bf_zero_ahead)
    cudaEvent_t ev_z0 = nullptr, ev_z1 = nullptr, ev_u0 = nullptr, ev_u1 = nullptr;
    // per-phase device times of iteration_resident (option "phase_timing"):
    // 0 assembly, 1 residual, 2 refactor, 3 solve, 4 update
    cudaEvent_t pev[6] = {nullptr, nullptr, nullptr, nullptr, nullptr, nullptr};
    double ms_phase[5] = {0, 0, 0, 0, 0};
    double n_phase_samples = 0;
};

// internal bit of the update flags: an element failed a convergence test with |dx| above the refinement floor
// (never returned through the ABI)
#define DAECPP_B200_ABOVE_FLOOR 4

namespace daecpp_b200
{
inline const char *base_name(const char *path)
{
    const char *b = path;
    for (const char *q = path; *q; ++q)
        if (*q == '/') b = q + 1;
    return b;
}
} // namespace daecpp_b200

#define DAECPP_CUDA_TRY(ctx, call)                                                                        \
    do                                                                                                    \
    {                                                                                                     \
        cudaError_t e__ = (call);                                                                         \
        if (e__ != cudaSuccess)                                                                           \
        {                                                                                                 \
            (ctx)->err = std::string(#call) + ": " + cudaGetErrorString(e__) + " (" + daecpp_b200::base_name(__FILE__) + ":" + std::to_string(__LINE__) + ")"; \
            return (e__ == cudaErrorMemoryAllocation) ? DAECPP_B200_ENOMEM : DAECPP_B200_ECUDA;           \
        }                                                                                                 \
    } while (0)

namespace daecpp_b200
{
// assembly.cu
int assemble_triplets(daecpp_b200_ctx *c, TripletSet &T, const double *A, const int32_t *i, const int32_t *j,
                      int64_t nT, bool lazy_vals);
int build_union(daecpp_b200_ctx *c);
int build_mass_csr(daecpp_b200_ctx *c);
int build_jac_csr(daecpp_b200_ctx *c);
int regather_values(daecpp_b200_ctx *c, TripletSet &T); // hot path on the resident staging values, async
// numeric.cu
int upload_symbolic(daecpp_b200_ctx *c);
int numeric_refactor(daecpp_b200_ctx *c, double alpha);
int numeric_refactor_async(daecpp_b200_ctx *c, double alpha, bool clear_flags = true); // enqueue only; pivot flag left in d_flags[3]
int assemble_values_only(daecpp_b200_ctx *c, double alpha);
int numeric_solve(daecpp_b200_ctx *c, bool fuse_update = false, double atol = 0, double rtol = 0,
                  double max_err_abs = 0); // b -> dx (and, fused, the error checks and xk += dx)
int ensure_csc_values(daecpp_b200_ctx *c, TripletSet &T); // J.vals / M.vals up to date (parity exports)
int jb_values(daecpp_b200_ctx *c, double alpha);          // u_vals = J - alpha*M on the union pattern
// subtree.cu
int unit_setup(daecpp_b200_ctx *c);
int compose_entries(daecpp_b200_ctx *c);
int launch_scatter_big(daecpp_b200_ctx *c, double alpha, cudaStream_t st);
int launch_unit_factor(daecpp_b200_ctx *c, const GroupLaunch &g, double alpha, cudaStream_t st = nullptr);
int launch_unit_forward(daecpp_b200_ctx *c, const GroupLaunch &g, bool direct, cudaStream_t st = nullptr);
int launch_unit_backward(daecpp_b200_ctx *c, const GroupLaunch &g, bool direct, bool update, double atol, double rtol,
                         double max_err_abs, cudaStream_t st = nullptr);
// chain.cu
int chain_try_setup(daecpp_b200_ctx *c, int32_t n, const int32_t *outer, const int32_t *inner);
int chain_refactor_async(daecpp_b200_ctx *c, double alpha);
int chain_solve(daecpp_b200_ctx *c, bool fuse_update, double atol, double rtol, double max_err_abs);
// tiny.cu
bool tiny_eligible(const daecpp_b200_ctx *c);
int tiny_set_triplets(daecpp_b200_ctx *c, bool jac, const double *A, const int32_t *i, const int32_t *j, int64_t nT, bool values_only);
int tiny_flush(daecpp_b200_ctx *c);
int tiny_iteration(daecpp_b200_ctx *c, const double *f_host, int32_t do_refactor, int32_t order, const double w[5], double atol, double rtol,
                   double max_err_abs, double *xk_host, int32_t *flags);
void tiny_release(daecpp_b200_ctx *c);
// vector_ops.cu
int launch_residual(daecpp_b200_ctx *c, int32_t order, const double coef[5], bool clear_flags = false);
Bdf make_bdf(daecpp_b200_ctx *c, int32_t order, const double coef[5]);
int classify_mass(daecpp_b200_ctx *c);
int launch_dxdt(daecpp_b200_ctx *c, int32_t order, const double coef[5]);
int launch_update(daecpp_b200_ctx *c, double atol, double rtol, double max_err_abs);
int launch_accept(daecpp_b200_ctx *c, int32_t variability_on, double tol);
// guarded refinement of the last solve (single system, or the lanes of c->lanes whose mode word is not idle):
// r = b - Jb*dx, Jb*e = r with the current factors, dx += e, xk += e, flags (and variability) recomputed
int refine_last_solve(daecpp_b200_ctx *c, double alpha, double atol, double rtol, double max_err_abs, bool b_is_stored, bool update = true);
} // namespace daecpp_b200
