/*
 * daecpp_b200 -- C ABI of the B200-native Newton linear-solve hot path for dae-cpp.
 *
 * This is the drop-in boundary.  The reference (dae-cpp 2.2.0) has no FFI: its integrator
 * `core::detail::solve` owns an `Eigen::SparseLU<eimat> linsolver`, `eimat M_, Jb` and
 * `eivec f_, b, dx, dxdt` as locals (dae-cpp/solver.hpp:325-337) and works on them inline.
 * Each entry point below replaces one of those inline regions; the citation names it
 * (paths relative to the reference tree).  The overlay header dae-cpp_b200/dae-cpp/solver.hpp
 * is the only caller a user needs; INTEGRATION.md shows the binding.
 *
 * Conventions
 *   - plain pointers and sizes; host arrays are owned by the caller and are only read/written
 *     during the call (all calls return after the stream work they need has completed, unless
 *     noted "async");
 *   - one context = one DAE system on one GPU, used by one host thread at a time; no globals,
 *     so an ensemble may own many contexts;
 *   - return value: DAECPP_B200_OK (0) or a DAECPP_B200_E* code; numerical failures use the
 *     codes that map onto the reference's exit_code values 2 and 3 (dae-cpp/typedefs.hpp:47-54);
 *   - all values FP64, all indices int32 (the reference's default int_type / float_type,
 *     dae-cpp/typedefs.hpp:57-69).
 *   - there is NO CPU fallback: without a CUDA device every call fails with
 *     DAECPP_B200_ECUDA.
 */
#ifndef DAECPP_B200_H
#define DAECPP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct daecpp_b200_ctx daecpp_b200_ctx;

enum
{
    DAECPP_B200_OK = 0,
    DAECPP_B200_EDECOMP = 2,  /* factorisation failed (zero / non-finite pivot): solver.hpp:541-551 */
    DAECPP_B200_ESOLVE = 3,   /* solution not finite:                            solver.hpp:562-572 */
    DAECPP_B200_EINDEX = 20,  /* triplet index out of range: sparse-matrix.hpp:135-136 ASSERTs      */
    DAECPP_B200_EINVAL = 21,  /* bad argument / call order                                            */
    DAECPP_B200_ECUDA = 22,   /* CUDA runtime error or no device                                      */
    DAECPP_B200_ENOMEM = 23
};

/* flags written by daecpp_b200_solve_update */
enum
{
    DAECPP_B200_DIVERGED = 1,     /* |dx_i| > max_err_abs or NaN   (solver.hpp:596-601) */
    DAECPP_B200_NOT_CONVERGED = 2 /* rel/abs tests failed somewhere (solver.hpp:603-614) */
};

/* ---- life cycle ---------------------------------------------------------------------- */
/* Replaces the locals at solver.hpp:301-337 (SolverState history, xk, M_, Jb, f_, b, dx). */
int daecpp_b200_create(int32_t n, int device, daecpp_b200_ctx **out);
void daecpp_b200_destroy(daecpp_b200_ctx *ctx);
const char *daecpp_b200_last_error(const daecpp_b200_ctx *ctx); /* NULL ctx: last create() error */
const char *daecpp_b200_version(void);

/* Pinned host memory for sparse_matrix::A/i/j (sparse-matrix.hpp:36-40) so that the triplet
 * upload is a true async DMA.  Falls back to nothing: returns NULL on failure. */
void *daecpp_b200_host_alloc(size_t bytes);
void daecpp_b200_host_free(void *p);

/* Page-lock an existing host range (the storage behind sparse_matrix::A/i/j keeps its
 * std::vector type, sparse-matrix.hpp:36-40; the overlay solver registers it once it is large
 * and stable).  Failure is not fatal: the copies then go through pageable memory. */
int daecpp_b200_host_register(void *p, size_t bytes);
int daecpp_b200_host_unregister(void *p);

/* Process-wide, per device: make host threads SLEEP while they wait for the device instead of
 * spinning (cudaDeviceScheduleBlockingSync).  Meant for ensembles (dae-cpp/ensemble.hpp): with many
 * contexts in flight the host cores are needed by the user functors of other members. */
int daecpp_b200_set_blocking_sync(int device, int on);
/* Process-wide: keep destroyed contexts and freed page-locked buffers for the next create / host_alloc
 * of the same size (ensembles create and destroy one context per member; cudaFree would synchronise
 * the device under the other members).  A recycled context keeps its pattern caches and symbolic plan.
 * on = 0 switches pooling off and releases everything pooled. */
int daecpp_b200_set_pooling(int on);

/* ---- state --------------------------------------------------------------------------- */
/* xk = x[0] = x0, order/dt history cleared       (solver.hpp:310-313) */
int daecpp_b200_set_state(daecpp_b200_ctx *ctx, const double *x0);
/* download the Newton iterate xk for the host functors / Solution Manager (solver.hpp:455,500,659) */
int daecpp_b200_get_xk(daecpp_b200_ctx *ctx, double *xk_host);
/* xk = x[0]  (roll back after a failed step, solver.hpp:643,671) */
int daecpp_b200_rollback(daecpp_b200_ctx *ctx);

/* ---- triplet hand-off (replaces sparse_matrix::convert, sparse-matrix.hpp:124-152) ---- */
/* M_ = M.convert(n)   (solver.hpp:481).  nT may be 0 (MassMatrixZero). */
int daecpp_b200_set_mass(daecpp_b200_ctx *ctx, const double *A, const int32_t *i, const int32_t *j, int64_t nT);
/* Jb = J.convert(n)   (solver.hpp:502).  A pattern equal to the previous call's takes the
 * cached path (no sort); a new pattern triggers sort + symbolic analysis at the next refactor. */
int daecpp_b200_set_jacobian(daecpp_b200_ctx *ctx, const double *A, const int32_t *i, const int32_t *j, int64_t nT);

/* ---- per-iteration path -------------------------------------------------------------- */
/* dxdt = BDF(order, dt[0..3]) of (xk, x[0..3]);  b = M*dxdt - f   (solver.hpp:439,516-518).
 * f_host: the RHS values (f__[k].val(), solver.hpp:461-464).  *alpha_out receives the BDF
 * derivative d(dxdt)/d(xk) that solver.hpp:439 returns. */
int daecpp_b200_residual(daecpp_b200_ctx *ctx, const double *f_host, int32_t order, const double dt[4],
                         double *alpha_out);
/* Jb -= M_*alpha; linsolver.compute(Jb)   (solver.hpp:523,537).  Symbolic analysis runs only
 * when the union pattern changed.  Returns OK or DAECPP_B200_EDECOMP. */
int daecpp_b200_refactor(daecpp_b200_ctx *ctx, double alpha);
/* dx = linsolver.solve(b); error checks; xk += dx   (solver.hpp:558,590-619).
 * *flags = DIVERGED | NOT_CONVERGED bits.  Returns OK or DAECPP_B200_ESOLVE. */
int daecpp_b200_solve_update(daecpp_b200_ctx *ctx, double atol, double rtol, double max_err_abs, int32_t *flags);
/* x[3]<-x[2]<-x[1]<-x[0]<-xk and max relative variability   (solver.hpp:698-721) */
int daecpp_b200_accept_step(daecpp_b200_ctx *ctx, int32_t variability_on, double variability_tolerance,
                            double *variability_out);

/* ---- one whole Newton iteration on data already resident in HBM -------------------------
 * The device part of solver.hpp:431-627 in one submission, with no host<->device copy except
 * the flag word: re-assemble J from the triplet VALUES of the last set_jacobian upload (cached
 * pattern), b = M*dxdt - f from the f of the last residual upload, then (do_refactor != 0)
 * Jb = J - alpha*M and the numeric refactorisation, the triangular solves, error checks and
 * xk += dx.  This is the quantity SURVEY.md 8(d) times ("triplet values resident in HBM" ->
 * "dx applied, flags ready"); bench.py's `value` is measured through it. */
int daecpp_b200_iteration_resident(daecpp_b200_ctx *ctx, int32_t do_refactor, int32_t order, const double dt[4],
                                   double atol, double rtol, double max_err_abs, int32_t *flags);
/* `count` such iterations back to back; *flags receives the last iteration's.  On the block-chain path the
 * iterations are PIPELINED: iteration k+1 is enqueued while k runs, every iteration still publishes its own
 * flag words and the host still reads each of them in order; where the host has to act between two
 * iterations (a failed pivot -> DAECPP_B200_EDECOMP is returned, a dependency time-out, an update it would
 * refine) the device cancels the queued iteration before it touches anything, so the state on return is
 * exactly that of the one-at-a-time loop (option "pipeline_iterations" = 0 selects that loop).  This is what
 * bench.py times as `value`. */
int daecpp_b200_iterations_resident(daecpp_b200_ctx *ctx, int32_t count, int32_t do_refactor, int32_t order,
                                    const double dt[4], double atol, double rtol, double max_err_abs, int32_t *flags);
/* Values-only Jacobian hand-off: the (i, j) stream is the one of the last
 * set_jacobian call -- JacobianMatrixShape fixes it at construction (jacobian-matrix.hpp:66-92) -- so
 * only the nT values travel.  Asynchronous: A must stay valid until the next synchronising call
 * (iteration_host / refactor).  EINVAL when no pattern of this length is cached. */
int daecpp_b200_set_jacobian_values(daecpp_b200_ctx *ctx, const double *A, int64_t nT);
/* One whole Newton iteration from HOST buffers in one submission and one synchronisation:
 * f upload, b = M*dxdt - f, (do_refactor) Jb = J - alpha*M + refactorisation, solves, error checks,
 * xk += dx, download of the new xk into xk_host (may be NULL).  Same results and status codes as
 * residual + refactor + solve_update + get_xk (solver.hpp:439-619); on DAECPP_B200_EDECOMP the
 * iterate is not meaningful and the caller rolls back exactly as the reference does (:541-551). */
int daecpp_b200_iteration_host(daecpp_b200_ctx *ctx, const double *f_host, int32_t do_refactor, int32_t order,
                               const double dt[4], double atol, double rtol, double max_err_abs, double *xk_host,
                               double *alpha_out, int32_t *flags);
/* CUDA-event stopwatch on the context's stream (torch.cuda.Event cannot see this stream). */
int daecpp_b200_event_mark(daecpp_b200_ctx *ctx);
int daecpp_b200_event_elapsed_ms(daecpp_b200_ctx *ctx, double *ms); /* synchronises the stream */

/* ---- linear-solver-only entry points (what Eigen::SparseLU::compute/solve expose,
 *      Eigen/src/SparseLU/SparseLU.h:210-215,338-358; used by tests/test_linear-solver.cpp) ---- */
int daecpp_b200_set_rhs(daecpp_b200_ctx *ctx, const double *b_host); /* b (device) = b_host */
int daecpp_b200_solve(daecpp_b200_ctx *ctx, double *dx_host);        /* dx = Jb^-1 b; dx_host may be NULL */

/* ---- parity / debug exports ---------------------------------------------------------- */
/* Compressed-column copy of the mass matrix (which=0), the Jacobian (1) or Jb = J - alpha*M as
 * last refactored (2).  outer has n+1 entries; inner/vals are filled up to cap entries.
 * *nnz_out receives the true count. */
int daecpp_b200_get_csc(daecpp_b200_ctx *ctx, int32_t which, int32_t *outer, int32_t *inner, double *vals,
                        int64_t cap, int64_t *nnz_out);
/* which: 0 = b, 1 = dx, 2 = dxdt, 3 = xk, 4..7 = x[0..3] */
int daecpp_b200_get_vector(daecpp_b200_ctx *ctx, int32_t which, double *host);
/* Overwrite a device vector (same selectors except dxdt); lets tests plant a BDF history
 * (SolverState::x, solver.hpp:45) without integrating. */
int daecpp_b200_set_vector(daecpp_b200_ctx *ctx, int32_t which, const double *host);
/* Counters and timings.  Keys: "n_symbolic", "n_refactor", "n_solve", "n_fronts", "n_levels",
 * "nnz_factor", "factor_flops", "max_front", "ms_symbolic", "ms_last_refactor", "ms_last_solve",
 * "kernels_launched", "n_cold_assembly" ... returns EINVAL for an unknown key. */
int daecpp_b200_get_stat(daecpp_b200_ctx *ctx, const char *key, double *value);
/* Debug: with option "trace_units" = 1 the factor kernel stamps %globaltimer (ns) four times per
 * subtree unit (start, staged, dependencies met, done), in ticket order, for the factor kernel and
 * then for the forward and the backward solve kernels (3 x n_units x 4 values).  On the block-chain
 * path (two-level hierarchies) the layout is 5 kernels x tiles x 8 stamps instead
 * (profiles/tools/chain_trace.py reads it). */
int daecpp_b200_debug_unit_trace(daecpp_b200_ctx *ctx, int64_t *out, int64_t cap, int64_t *count);
/* Tunables, by name ("refine_steps", "pivot_tol", "leaf_size", "use_graph" ...).  Block-chain path
 * (banded patterns; every variant computes the same factors, the switches exist for A/B measurements):
 * "chain" (0: always the general multifrontal path), "chain_min_n", "chain_max_block", "chain_affine"
 * (descriptor instead of slot map for tiles whose value sources are arithmetic progressions),
 * "chain_packed" (contiguous-run LDGSTS staging), "chain_fuse_top" (bit 0: factor, bit 1: solve -- top
 * level swept by the CTA that finishes last), "chain_defer_top" (top level factored by an extra CTA of the
 * forward launch inside daecpp_b200_iteration_*), "chain_fuse_residual" (diagonal mass: b formed inside the
 * forward kernel; "keep_b" = 1 also writes it to HBM for daecpp_b200_get_vector), "lean_flags" (flag words
 * delivered through a mapped host word instead of memset + copy nodes).  General path: "blocked_lu",
 * "blocked_lu_min_dim", "bf_fused_panel" (panel of a block column in one launch), "bf_dmma" (trailing update of big
 * fronts on the FP64 tensor cores, default 1; 0 = FP64 FMA tiles), "bf_pdl" (programmatic dependent launch between them),
 * "bf_schur" (contribution block of a big front updated in one pass behind its last block column; bit-identical, default 0:
 * measured slower; "bf_schur_fronts" = N > 0: only on levels with at least N fronts, also measured slower),
 * "bf_fast_solve" (default 1: triangular sweeps of big fronts with the pivot unknowns in shared memory, diagonal blocks held in
 * registers from the start of the kernel and batched loads; 0 = the round-1 kernels; bit-identical), "bf_diag_copy" (default 1:
 * the factored diagonal block of a block column travels from the panel launch to the trailing update through a side buffer
 * instead of being factored a second time; bit-identical), "bf_lu_lookahead" (default 1: reciprocal of the next pivot formed
 * beside the update; bit-identical), "bf_zero_ahead" (default 1: the big fronts' factor storage and contribution blocks are
 * prepared on a side stream of the context while the subtree-unit launches run; 0 = in line, one zeroing launch per level),
 * "bf_overlap_units" (default 1: a launch of subtree units in the same stage of the tree as the level of big fronts behind it
 * runs beside that level on the side stream; scheduling only), "group8_dim" / "warp_dim" / "warp_dim_many" / "warp_min_fronts" (local levels of a subtree unit
 * swept eight lanes / one warp per front up to this front order, larger fronts by the whole block; defaults 8 / 16 / 16 / 4,
 * round 1: 12 / 40), "unit_threads" (0 = automatic: 256, or 128 with 40 KB units where the automatic unit budget hits its ceiling;
 * else a multiple of 32, at most 256).  Numerics: "refine_steps" / "refine_floor" (guarded refinement step), "pivot_rel" (row-relative pivot
 * test of the chain path), "transversal" (row matching before the ordering when the diagonal has structural holes),
 * "tiny" (one-launch path for n <= 32).  "chain_packed": 2 = bulk copies (cp.async.bulk) where the tile qualifies,
 * 1 = contiguous-run LDGSTS, 0 = per-field copies.  "pipeline_iterations": see daecpp_b200_iterations_resident.
 * "chain_top_cta" (default 1): fused iteration of a single system, the grid-level cyclic reduction runs in a CTA of its own
 * that rehearses it while the tiles sweep.
 * Debug: "trace_units" (with "trace_unit_index": the unit whose local levels are stamped one by one), "phase_timing", "pinned_io". */
int daecpp_b200_set_option(daecpp_b200_ctx *ctx, const char *key, double value);

/* ==== lock-step batched ensemble (an ADDITION: the reference has no ensemble API) =========
 * `lanes` independent DAE systems of the same size and the same sparsity pattern (parameter sweeps) live
 * in one batch context on one GPU.  One call of daecpp_b200_batch_round performs ONE Newton iteration of
 * dae-cpp/solver.hpp:431-627 for every lane that asks for it: the kernels of the single-system path run
 * with the lane as the outer grid dimension, each lane with its own BDF order / step history / alpha, its
 * own matrix values, flags and factors, so a round costs one launch sequence and one flag read however
 * many lanes are in flight.  The caller keeps one integrator state per lane (dae-cpp/solver.hpp:38-69,
 * 402-797: dae-cpp/ensemble.hpp does) and tells every round what each lane does next.  Results of a lane are
 * bit-identical to the same iteration run through a single-system context on the same inputs.
 * Patterns the block-chain path does not take (2-D grids, arbitrary sparsity) and Jacobian triplet streams
 * with duplicates are refused with DAECPP_B200_EINVAL: such ensembles run one context per member. */
typedef struct daecpp_b200_batch daecpp_b200_batch;

enum /* daecpp_b200_lane_cmd.mode */
{
    DAECPP_B200_LANE_IDLE = 0,     /* the lane sits this round out (finished, or waiting)                  */
    DAECPP_B200_LANE_SOLVE = 1,    /* Newton iteration that reuses the lane's factors  (solver.hpp:533)    */
    DAECPP_B200_LANE_REFACTOR = 2  /* Newton iteration with Jacobian update + refactorisation              */
};
enum /* daecpp_b200_lane_cmd.pre: applied to the lane before anything else of the round */
{
    DAECPP_B200_LANE_KEEP = 0,
    DAECPP_B200_LANE_ACCEPT = 1,   /* x[3]<-x[2]<-x[1]<-x[0]<-xk   (solver.hpp:698-721)                    */
    DAECPP_B200_LANE_ROLLBACK = 2  /* xk = x[0]                    (solver.hpp:643,671)                    */
};
#define DAECPP_B200_MAX_ROUND_ITERATIONS 8
typedef struct
{
    int32_t mode, pre, order;
    /* Several Newton iterations of one time step in one round (fewer host round trips when few lanes are in
     * flight): after its first iteration the lane goes on, up to max_iterations in all (<= 1: just the one, at most
     * DAECPP_B200_MAX_ROUND_ITERATIONS), for as long as the iteration before ended like the reference's loop
     * continues (solver.hpp:431-627): not converged, not diverged, factorisation fine, no refinement called for.
     * Iteration j refactorises when period <= 1 or (phase + j) % period == 0 -- phase = Newton index of the first
     * iteration, period = Newton_scheme + 1 (solver.hpp:533) -- and reuses the factors otherwise. */
    int32_t period;
    double t;     /* time handed to the lane's functors (solver.hpp:432: state.t after the increment)    */
    double dt[4]; /* SolverState::dt (solver.hpp:41)                                                      */
    int32_t phase, max_iterations;
} daecpp_b200_lane_cmd;
typedef struct
{
    int32_t flags;      /* DAECPP_B200_DIVERGED | DAECPP_B200_NOT_CONVERGED of the lane's LAST iteration of the round */
    int32_t status;     /* DAECPP_B200_OK, or DAECPP_B200_EDECOMP / DAECPP_B200_ESOLVE for this lane      */
    double variability; /* max |(x[0] - xk_new) / x[0]| over the elements the reference looks at
                           (solver.hpp:703-717), computed with the update: what accept_step would return */
    int32_t iterations; /* Newton iterations the lane ran this round (all but the last: not converged)     */
    int32_t reserved;
} daecpp_b200_lane_result;

/* Device functors (opt-in).  The callbacks are given DEVICE pointers and the CUDA stream
 * to launch on; they must only enqueue work on that stream and return 0.  dae-cpp/device-functor.cuh builds
 * them from one templated equations(x, t, i) functor (the device twin of VectorFunctionElements +
 * JacobianMatrixShape, vector-function.hpp:46-72, jacobian-matrix.hpp:66-92).  The lanes to evaluate are
 * lane_list[0 .. n_list); when lane_list is NULL they are the lanes whose mode[lane] is at least min_mode. */
typedef struct
{
    int32_t n, lanes, min_mode, reserved;
    int64_t x_stride, out_stride; /* elements between consecutive lanes                                     */
    const double *x;              /* [lanes][x_stride]  Newton iterates xk                                   */
    const double *t;              /* [lanes]                                                                 */
    const int32_t *mode;          /* [lanes]                                                                 */
    const int32_t *lane_list;     /* [n_list] or NULL                                                        */
    int32_t n_list, reserved2;
    double *out;                  /* rhs: f [lanes][out_stride];  jacobian: triplet values [lanes][out_stride] */
    void *stream;                 /* cudaStream_t                                                            */
} daecpp_b200_functor_args;
typedef int (*daecpp_b200_device_fn)(void *user, const daecpp_b200_functor_args *args);
typedef struct
{
    void *user;
    daecpp_b200_device_fn rhs;      /* f = rhs(xk, t)                          (solver.hpp:450-464)          */
    daecpp_b200_device_fn jacobian; /* values of the Jacobian triplets, in the order of the (i, j) stream
                                       given to daecpp_b200_batch_set_pattern (solver.hpp:499)              */
} daecpp_b200_device_functors;

int daecpp_b200_batch_create(int32_t n, int32_t lanes, int device, daecpp_b200_batch **out);
void daecpp_b200_batch_destroy(daecpp_b200_batch *b);
const char *daecpp_b200_batch_last_error(const daecpp_b200_batch *b);
/* The (i, j) streams every lane's Jacobian / mass functor emits (sparse-matrix.hpp:36-56).  Sorts, builds the
 * union pattern and runs the symbolic analysis once for all lanes. */
int daecpp_b200_batch_set_pattern(daecpp_b200_batch *b, const int32_t *Ji, const int32_t *Jj, int64_t nTJ, const int32_t *Mi,
                                  const int32_t *Mj, int64_t nTM);
/* Triplet VALUES, lane-major host slabs [count][nT] for lanes [first, first + count).  count = 0 (mass only):
 * one set of nTM values shared by all lanes. */
int daecpp_b200_batch_set_mass_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count);
int daecpp_b200_batch_set_jacobian_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count);
int daecpp_b200_batch_set_rhs_values(daecpp_b200_batch *b, const double *f, int32_t first, int32_t count); /* [count][n] */
/* xk = x[0] = x0, history cleared (solver.hpp:310-313);  x0: [count][n].  count = 0: ONE state of n values shared by
 * all lanes (a sweep that starts every member from the same initial condition uploads it once). */
int daecpp_b200_batch_set_state(daecpp_b200_batch *b, const double *x0, int32_t first, int32_t count);
/* which: same selectors as daecpp_b200_get_vector (0 b, 1 dx, 3 xk, 4..7 x[0..3]); out: [count][n] */
int daecpp_b200_batch_get_vector(daecpp_b200_batch *b, int32_t which, double *out, int32_t first, int32_t count);
/* NULL or all-NULL members: functor values come from set_rhs_values / set_jacobian_values */
int daecpp_b200_batch_set_device_functors(daecpp_b200_batch *b, const daecpp_b200_device_functors *fn);
/* One Newton round.  cmd, res: [lanes].  variability_on / variability_tolerance: SolverOptions
 * solution_variability_control / variability_tolerance (solver.hpp:700-717). */
int daecpp_b200_batch_round(daecpp_b200_batch *b, const daecpp_b200_lane_cmd *cmd, double atol, double rtol, double max_err_abs,
                            int32_t variability_on, double variability_tolerance, daecpp_b200_lane_result *res);
/* "rounds", "kernels_launched", "chain_block", "chain_tiles", "ms_device" (CUDA-event time of all rounds so far),
 * "bytes_per_lane" ... */
int daecpp_b200_batch_get_stat(daecpp_b200_batch *b, const char *key, double *value);

/* ==== an entry point that allocates nothing ================================================================
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Maximum relative variability of solver.hpp:703-717 -- max_i |(x0_i - xk_i) / x0_i| over the elements with
 * |x0_i| > tolerance and |xk_i| > tolerance, NaN counted as 0 -- i.e. what daecpp_b200_accept_step reports, computed
 * by a variant of its kernel whose scratch (per-block partial maxima + a ticket word) lives in a file-scope
 * `static __device__` array of fixed size (33 KB per GPU that loads the library, held until exit) instead of
 * memory the context allocated.  xk_dev, x0_dev (n doubles each) and result_dev (1 double) are DEVICE pointers owned
 * by the caller.  The call only enqueues a memset node and one kernel on `stream` (a cudaStream_t; NULL = the
 * default stream) and returns; *result_dev is valid once that stream work has completed.  It calls no cudaMalloc,
 * cudaFree or allocating library routine.  The first call on a device also looks up the scratch array's device
 * address (cudaGetSymbolAddress): make one call per device outside any capture.
 * n > DAECPP_B200_NOALLOC_MAX_N is refused with DAECPP_B200_EINVAL and a message in errbuf (may be NULL).
 * Every call on a device shares the one scratch array: calls that overlap on the GPU (different streams, different
 * host threads) race on it -- serialise them. */
#define DAECPP_B200_NOALLOC_MAX_N 16777216
int daecpp_b200_variability_noalloc(const double *xk_dev, const double *x0_dev, int64_t n, double tolerance, double *result_dev,
                                    void *stream, char *errbuf, size_t errbuf_len);

#ifdef __cplusplus
}
#endif
#endif /* DAECPP_B200_H */
