// Lock-step batched ensemble: `lanes` independent DAE systems of one sparsity pattern in one context
// (include/daecpp_b200.h: daecpp_b200_batch_*).
//
// Every lane is one instance of the per-iteration path of dae-cpp/solver.hpp:431-627; nothing is exchanged
// between lanes.  The symbolic work (triplet sort, union pattern, block-chain plan) is done once on a
// prototype single-system context; the numeric kernels of chain.cu then run with the lane as blockIdx.y
// over lane-major slabs [lanes][...] of triplet values, factors, forward results and state vectors.  A
// round = one small control upload, the functor launches (device functors) or the slabs the caller
// uploaded (host functors), residual, refactorisation of the lanes that asked for one, the triangular
// solves with the fused error check / update / variability, one small result download.
#include <algorithm>
#include <chrono>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "bdf.hpp"
#include "ctx.hpp"

using namespace daecpp_b200;

struct daecpp_b200_batch
{
    daecpp_b200_ctx *c = nullptr; // prototype context: pattern caches, plan, stream
    int32_t n = 0, lanes = 0;
    int device = 0;
    size_t ns = 0;                // lane stride of the state vectors (multiple of 32 doubles)
    std::string err;

    double *slab = nullptr;       // xk, f, b, dx [lanes][ns] each, then the history ring [4][lanes][ns]
    double *xk = nullptr, *f = nullptr, *b = nullptr, *dx = nullptr, *hist = nullptr;
    double *proto_xk = nullptr, *proto_f = nullptr, *proto_b = nullptr, *proto_dx = nullptr;
    DevBuf<double> refine_buf;    // residual and correction of the guarded refinement step, [2][lanes][ns], on first use
    double n_refine_rounds = 0, n_refined_lanes = 0;
    DevBuf<double> jv, mv;        // triplet values of J [lanes][nTJ]; compressed-column values of M [lanes][nnzM]
    int64_t nTJ = 0, nTM = 0, nnzM = 0;
    std::vector<int32_t> m_src, m_src_ptr; // mass: triplets of every slot in input order (host copy)
    bool have_pattern = false, have_mass = false;

    // control words of a round: pinned structure-of-arrays block and its device mirror
    //   mode | pre | order | hbase | active lanes | refactoring lanes | lanes with a pre action | no-refinement marks |
    //   planned modes of iterations 0..7  (i32[L] each, padded to 8 bytes)  | t f64[L] | w f64[5][L]
    char *h_ctl = nullptr, *d_ctl = nullptr;
    size_t ctl_bytes = 0;
    // results: flags i32[4][L] (lane-major [L][4]) | var u64[L]
    char *h_res = nullptr, *d_res = nullptr;
    size_t res_bytes = 0;
    std::vector<int32_t> hbase;
    std::vector<char> refine_futile; // the lane's last refinement in this Newton loop did not change the outcome: no second one
    LaneSet L;

    daecpp_b200_device_functors fn{};
    bool have_fn = false;
    std::vector<char> rhs_loaded, jac_loaded; // host-functor mode: the lane has values to iterate on

    static constexpr int KMAX = DAECPP_B200_MAX_ROUND_ITERATIONS;
    static constexpr size_t BOUNCE_BYTES = (size_t)8 << 20;
    char *h_bounce = nullptr;     // two page-locked bounce buffers for the lane slabs of pageable callers
    cudaEvent_t ev_bounce[2] = {nullptr, nullptr};
    double rounds = 0, ms_device = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    bool ev_pending = false;

    int32_t *ctl_mode() const { return reinterpret_cast<int32_t *>(h_ctl); }
    int32_t *ctl_pre() const { return ctl_mode() + lanes; }
    int32_t *ctl_order() const { return ctl_mode() + 2 * lanes; }
    int32_t *ctl_hbase() const { return ctl_mode() + 3 * lanes; }
    int32_t *ctl_list_active() const { return ctl_mode() + 4 * lanes; }
    int32_t *ctl_list_refactor() const { return ctl_mode() + 5 * lanes; }
    int32_t *ctl_list_pre() const { return ctl_mode() + 6 * lanes; }
    int32_t *ctl_norefine() const { return ctl_mode() + 7 * lanes; }
    int32_t *ctl_plan(int j) const { return ctl_mode() + (8 + j) * (size_t)lanes; }
    size_t ctl_int_bytes() const { return (sizeof(int32_t) * (8 + KMAX) * (size_t)lanes + 7) / 8 * 8; }
    // results: per iteration j of the round  flags i32[L][4] | var u64[L];  behind them  iterations run i32[L] | effective mode i32[L]
    size_t res_stride() const { return sizeof(int32_t) * 4 * (size_t)lanes + sizeof(unsigned long long) * (size_t)lanes; }
    int32_t *res_iters(char *base) const { return reinterpret_cast<int32_t *>(base + KMAX * res_stride()); }
    double *ctl_t() const { return reinterpret_cast<double *>(h_ctl + ctl_int_bytes()); }
    template <class T> T *dev(const T *host_ptr) const { return reinterpret_cast<T *>(d_ctl + (reinterpret_cast<const char *>(host_ptr) - h_ctl)); }
    double *ctl_w() const { return ctl_t() + lanes; }
};

namespace
{
constexpr int TPB = 256;
std::mutex g_bounce_mu;
std::vector<char *> g_bounce_free; // page-locked bounce buffers of destroyed batches (process-wide, a few 16 MB blocks)

#define BATCH_TRY(b, call)                                                                          \
    do                                                                                              \
    {                                                                                               \
        cudaError_t e__ = (call);                                                                   \
        if (e__ != cudaSuccess)                                                                     \
        {                                                                                           \
            (b)->err = std::string(#call) + ": " + cudaGetErrorString(e__);                         \
            return (e__ == cudaErrorMemoryAllocation) ? DAECPP_B200_ENOMEM : DAECPP_B200_ECUDA;     \
        }                                                                                           \
    } while (0)

int bfail(daecpp_b200_batch *b, int code, const std::string &msg)
{
    b->err = msg;
    return code;
}

struct PreArgs
{
    int32_t n, lanes;
    int64_t s_vec, s_hist;
    const int32_t *pre, *hbase, *list;
    double *xk, *hist;
};

// accept: x[0] <- xk into the ring slot the host rotated to (dae-cpp/solver.hpp:698-702, the shift itself is
// the rotation of hbase); rollback: xk <- x[0] (solver.hpp:643,671).  One grid row per lane with such an action.
__global__ void k_batch_pre(PreArgs a)
{
    const int lane = a.list[blockIdx.y];
    const int pre = a.pre[lane];
    if (pre == DAECPP_B200_LANE_KEEP) return;
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    double *x0 = a.hist + (int64_t)a.hbase[lane] * a.s_hist + (int64_t)lane * a.s_vec;
    double *xk = a.xk + (int64_t)lane * a.s_vec;
    if (pre == DAECPP_B200_LANE_ACCEPT) x0[i] = xk[i];
    else xk[i] = x0[i];
}

struct PlanArgs
{
    int32_t lanes, j, n_list;
    const int32_t *list, *plan_j, *norefine;
    const int32_t *prev_flags; // [lanes][4] of iteration j-1
    int32_t *eff, *iters;
};
// Which lanes run iteration j of the round: all planned ones for j = 0; afterwards those whose previous iteration ended
// the way the reference's Newton loop continues (solver.hpp:431-627).  One thread per lane of the round's list.
__global__ void k_batch_plan(PlanArgs a)
{
    // launched with programmatic stream serialisation like the kernels around it (every thread passes the wait)
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= a.n_list) return;
    const int lane = a.list[q];
    int m = a.plan_j[lane];
    if (a.j > 0 && m != DAECPP_B200_LANE_IDLE)
    {
        const int32_t *pf = a.prev_flags + 4 * lane;
        const bool go_on = a.eff[lane] != DAECPP_B200_LANE_IDLE && (pf[0] & DAECPP_B200_NOT_CONVERGED) && !(pf[0] & DAECPP_B200_DIVERGED) &&
                           pf[1] == 0 && ((pf[0] & DAECPP_B200_ABOVE_FLOOR) || a.norefine[lane]);
        if (!go_on) m = DAECPP_B200_LANE_IDLE;
    }
    a.eff[lane] = m;
    if (m != DAECPP_B200_LANE_IDLE) a.iters[lane] += 1;
}

struct BroadcastArgs
{
    int32_t n;
    int64_t s_vec;
    double *x; // lane 0 holds the state, lanes 1.. receive it
};
__global__ void k_batch_broadcast(BroadcastArgs a)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < a.n) a.x[(int64_t)(blockIdx.y + 1) * a.s_vec + i] = a.x[i];
}

struct ResArgs
{
    int32_t n, lanes;
    int64_t s_vec, s_hist, s_mv;
    const int32_t *mode, *hbase, *order, *list;
    const double *w;
    const double *xk, *hist, *mv, *f;
    double *b;
    const int32_t *rowptr, *rowslot, *rowcol;
};

__device__ __forceinline__ Bdf lane_bdf(const ResArgs &a, int lane)
{
    Bdf s;
    const int hb = a.hbase[lane];
    const int64_t off = (int64_t)lane * a.s_vec;
    s.order = a.order[lane];
#pragma unroll
    for (int k = 0; k < 5; ++k) s.c[k] = a.w[(int64_t)k * a.lanes + lane];
    s.xk = a.xk + off;
    s.x0 = a.hist + (int64_t)(hb & 3) * a.s_hist + off; s.x1 = a.hist + (int64_t)((hb + 1) & 3) * a.s_hist + off;
    s.x2 = a.hist + (int64_t)((hb + 2) & 3) * a.s_hist + off; s.x3 = a.hist + (int64_t)((hb + 3) & 3) * a.s_hist + off;
    return s;
}

// b = M*dxdt - f per lane: one thread per row, the row's mass entries in increasing column order, sum from +0.0
// (the arithmetic of vector_ops.cu: k_residual; dae-cpp/solver.hpp:141-217, 516-518)
__global__ void k_batch_residual(ResArgs a)
{
    const int lane = a.list[blockIdx.y];
    if (a.mode[lane] == DAECPP_B200_LANE_IDLE) return; // halted earlier in this round
    const int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= a.n) return;
    const Bdf s = lane_bdf(a, lane);
    const double *mv = a.mv + (int64_t)lane * a.s_mv;
    double acc = 0.0;
    const int32_t e = a.rowptr[r + 1];
    for (int32_t p = a.rowptr[r]; p < e; ++p) acc = __dadd_rn(acc, __dmul_rn(mv[a.rowslot[p]], bdf_at(s, a.rowcol[p])));
    const int64_t off = (int64_t)lane * a.s_vec;
    a.b[off + r] = __dsub_rn(acc, a.f[off + r]);
}

void release(daecpp_b200_batch *b)
{
    if (!b) return;
    cudaSetDevice(b->device);
    if (b->c)
    {
        if (b->c->stream) cudaStreamSynchronize(b->c->stream);
        // hand the prototype its own vectors back before it frees them
        if (b->proto_xk) { b->c->xk = b->proto_xk; b->c->f = b->proto_f; b->c->b = b->proto_b; b->c->dx = b->proto_dx; }
        b->c->lanes = nullptr;
        daecpp_b200_destroy(b->c);
    }
    if (b->slab) dev_free(b->slab);
    b->jv.release(); b->mv.release(); b->refine_buf.release();
    if (b->h_ctl) cudaFreeHost(b->h_ctl);
    if (b->d_ctl) cudaFree(b->d_ctl);
    if (b->h_res) cudaFreeHost(b->h_res);
    if (b->h_bounce)
    {
        std::lock_guard<std::mutex> lk(g_bounce_mu);
        if (g_bounce_free.size() < 16) g_bounce_free.push_back(b->h_bounce);
        else cudaFreeHost(b->h_bounce);
    }
    for (cudaEvent_t e : b->ev_bounce) if (e) cudaEventDestroy(e);
    if (b->d_res) cudaFree(b->d_res);
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    delete b;
}

bool lane_range_ok(const daecpp_b200_batch *b, int32_t first, int32_t count)
{
    return first >= 0 && count >= 0 && (int64_t)first + count <= b->lanes;
}

// Variable-step BDF weights in the reference's association (api.cu: bdf_weights is file-local there)
bool weights(int order, const double *dt, double *w)
{
    const double h0 = dt[0], h1 = dt[1], h2 = dt[2], h3 = dt[3];
    const double s01 = h0 + h1, s12 = h1 + h2, s23 = h2 + h3;
    const double s012 = h0 + h1 + h2, s123 = h1 + h2 + h3, s0123 = h0 + h1 + h2 + h3;
    w[0] = w[1] = w[2] = w[3] = w[4] = 0.0;
    switch (order)
    {
    case 1: w[0] = 1.0 / h0; w[1] = w[0]; return true;
    case 2: w[0] = (2.0 * h0 + h1) / (h0 * s01); w[1] = s01 / (h0 * h1); w[2] = h0 / (h1 * s01); return true;
    case 3:
        w[0] = (3.0 * h0 * h0 + h1 * s12 + 2.0 * h0 * (2.0 * h1 + h2)) / (h0 * s01 * s012);
        w[1] = s01 * s012 / (h0 * h1 * s12); w[2] = h0 * s012 / (h1 * h2 * s01); w[3] = h0 * s01 / (h2 * s12 * s012);
        return true;
    case 4:
        w[0] = (4.0 * h0 * h0 * h0 + h1 * s12 * s123 + 3.0 * h0 * h0 * (3.0 * h1 + 2.0 * h2 + h3) +
                2.0 * h0 * (3.0 * h1 * h1 + h2 * s23 + 2.0 * h1 * (2.0 * h2 + h3))) / (h0 * s01 * s012 * s0123);
        w[1] = s01 * s012 * s0123 / (h0 * h1 * s12 * s123); w[2] = h0 * s012 * s0123 / (h1 * s01 * h2 * s23);
        w[3] = h0 * s01 * s0123 / (h2 * s12 * s012 * h3); w[4] = h0 * s01 * s012 / (h3 * s23 * s123 * s0123);
        return true;
    default: return false;
    }
}

} // namespace

extern "C" {

int daecpp_b200_batch_create(int32_t n, int32_t lanes, int device, daecpp_b200_batch **out)
{
    if (!out) return DAECPP_B200_EINVAL;
    *out = nullptr;
    if (lanes <= 0) return DAECPP_B200_EINVAL;
    daecpp_b200_batch *b = new (std::nothrow) daecpp_b200_batch();
    if (!b) return DAECPP_B200_ENOMEM;
    b->n = n; b->lanes = lanes; b->device = device;
    int rc = daecpp_b200_create(n, device, &b->c);
    if (rc) { delete b; return rc; }
    auto bail = [&](cudaError_t e, const char *what) {
        const int code = e == cudaErrorMemoryAllocation ? DAECPP_B200_ENOMEM : DAECPP_B200_ECUDA;
        (void)what;
        release(b);
        return code;
    };
    cudaError_t e;
    b->ns = ((size_t)n + 31) / 32 * 32;
    const size_t per = (size_t)lanes * b->ns;
    if ((e = dev_alloc((void **)&b->slab, sizeof(double) * 8 * per)) != cudaSuccess) return bail(e, "cudaMalloc(state)");
    b->xk = b->slab; b->f = b->slab + per; b->b = b->slab + 2 * per; b->dx = b->slab + 3 * per; b->hist = b->slab + 4 * per;
    if ((e = cudaMemsetAsync(b->slab, 0, sizeof(double) * 8 * per, b->c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
    b->ctl_bytes = b->ctl_int_bytes() + sizeof(double) * 6 * (size_t)lanes;
    b->res_bytes = daecpp_b200_batch::KMAX * b->res_stride() + sizeof(int32_t) * 2 * (size_t)lanes;
    if ((e = cudaMallocHost((void **)&b->h_ctl, b->ctl_bytes)) != cudaSuccess) return bail(e, "cudaMallocHost");
    if ((e = cudaMalloc((void **)&b->d_ctl, b->ctl_bytes)) != cudaSuccess) return bail(e, "cudaMalloc(ctl)");
    if ((e = cudaMallocHost((void **)&b->h_res, b->res_bytes)) != cudaSuccess) return bail(e, "cudaMallocHost");
    if ((e = cudaMalloc((void **)&b->d_res, b->res_bytes)) != cudaSuccess) return bail(e, "cudaMalloc(res)");
    if ((e = cudaMemsetAsync(b->d_res, 0, b->res_bytes, b->c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
    if ((e = cudaEventCreate(&b->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&b->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    std::memset(b->h_ctl, 0, b->ctl_bytes);
    b->hbase.assign((size_t)lanes, 0);
    b->refine_futile.assign((size_t)lanes, 0);
    b->rhs_loaded.assign((size_t)lanes, 0);
    b->jac_loaded.assign((size_t)lanes, 0);
    // the chain kernels of the prototype read / write the lane slabs
    daecpp_b200_ctx *c = b->c;
    b->proto_xk = c->xk; b->proto_f = c->f; b->proto_b = c->b; b->proto_dx = c->dx;
    LaneSet &L = b->L;
    L.count = lanes;
    L.s_vec = (int64_t)b->ns; L.s_hist = (int64_t)per;
    L.hist = b->hist;
    L.mode = b->res_iters(b->d_res) + lanes; /* the EFFECTIVE mode of the iteration in flight (k_batch_plan) */ L.hbase = b->dev(b->ctl_hbase()); L.order = b->dev(b->ctl_order());
    L.w = b->dev(b->ctl_w());
    L.list_active = b->dev(b->ctl_list_active()); L.list_refactor = b->dev(b->ctl_list_refactor());
    L.flags = reinterpret_cast<int32_t *>(b->d_res);
    L.var = reinterpret_cast<unsigned long long *>(b->d_res + sizeof(int32_t) * 4 * (size_t)lanes);
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
    *out = b;
    return DAECPP_B200_OK;
}

void daecpp_b200_batch_destroy(daecpp_b200_batch *b) { release(b); }

const char *daecpp_b200_batch_last_error(const daecpp_b200_batch *b)
{
    if (!b) return daecpp_b200_last_error(nullptr);
    return b->err.empty() ? daecpp_b200_last_error(b->c) : b->err.c_str();
}

int daecpp_b200_batch_set_pattern(daecpp_b200_batch *b, const int32_t *Ji, const int32_t *Jj, int64_t nTJ, const int32_t *Mi,
                                  const int32_t *Mj, int64_t nTM)
{
    if (!b || nTJ <= 0 || !Ji || !Jj || nTM < 0 || (nTM > 0 && (!Mi || !Mj))) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    daecpp_b200_ctx *c = b->c;
    b->err.clear();
    b->have_pattern = false;
    b->have_mass = false;
    // the prototype analyses the pattern exactly like a single system would (values are irrelevant here)
    std::vector<double> ones((size_t)(nTJ > nTM ? nTJ : nTM), 1.0);
    int rc;
    if ((rc = daecpp_b200_set_mass(c, ones.data(), Mi, Mj, nTM))) return bfail(b, rc, c->err);
    if ((rc = daecpp_b200_set_jacobian(c, ones.data(), Ji, Jj, nTJ))) return bfail(b, rc, c->err);
    if ((rc = build_union(c))) return bfail(b, rc, c->err);
    c->have_symbolic = false;
    if ((rc = upload_symbolic(c))) return bfail(b, rc, c->err);
    if (!c->CH.active)
        return bfail(b, DAECPP_B200_EINVAL, "batched ensembles need a block-chain (banded) pattern; run one context per member instead");
    if (!c->jvals_from_triplets)
        return bfail(b, DAECPP_B200_EINVAL, "batched ensembles need a Jacobian triplet stream without duplicates");
    b->nTJ = nTJ; b->nTM = nTM; b->nnzM = c->M.nnz;
    const chain::Plan &P = c->CH;
    const size_t L = (size_t)b->lanes;
    BATCH_TRY(b, b->jv.reserve(L * (size_t)nTJ));
    BATCH_TRY(b, b->mv.reserve(L * (size_t)(b->nnzM ? b->nnzM : 1)));
    BATCH_TRY(b, cudaMemsetAsync(b->jv.p, 0, sizeof(double) * L * (size_t)nTJ, c->stream));
    BATCH_TRY(b, cudaMemsetAsync(b->mv.p, 0, sizeof(double) * L * (size_t)(b->nnzM ? b->nnzM : 1), c->stream));
    BATCH_TRY(b, c->dC.F.reserve(L * (size_t)P.fac_doubles + 1));
    BATCH_TRY(b, c->dC.Y.reserve(L * (size_t)P.y_doubles + 1));
    BATCH_TRY(b, c->dC.red.reserve(L * (size_t)P.red_doubles + 1));
    BATCH_TRY(b, c->dC.rv.reserve(L * (size_t)P.rv_doubles + 1));
    BATCH_TRY(b, c->dC.x.reserve(L * (size_t)P.x_doubles + 1));
    BATCH_TRY(b, c->dC.counter.reserve(2 * L + 8));
    BATCH_TRY(b, cudaMemsetAsync(c->dC.counter.p, 0, sizeof(int32_t) * (2 * L + 8), c->stream));
    // mass: slot -> triplets in input order, for the host-side duplicate sum of set_mass_values
    b->m_src.assign((size_t)nTM, 0);
    b->m_src_ptr.assign((size_t)b->nnzM + 1, 0);
    if (nTM > 0)
    {
        BATCH_TRY(b, cudaMemcpyAsync(b->m_src.data(), c->M.src.p, sizeof(int32_t) * nTM, cudaMemcpyDeviceToHost, c->stream));
        BATCH_TRY(b, cudaMemcpyAsync(b->m_src_ptr.data(), c->M.src_ptr.p, sizeof(int32_t) * ((size_t)b->nnzM + 1), cudaMemcpyDeviceToHost, c->stream));
    }
    BATCH_TRY(b, cudaStreamSynchronize(c->stream));
    b->L.s_jv = nTJ; b->L.s_mv = b->nnzM;
    b->L.jv = b->jv.p; b->L.mv = b->mv.p;
    b->have_pattern = true;
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_set_mass_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count)
{
    if (!b || !b->have_pattern || (b->nTM > 0 && !A) || !lane_range_ok(b, first, count)) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    daecpp_b200_ctx *c = b->c;
    const bool shared = count == 0;
    if (shared) { first = 0; count = b->lanes; }
    if (b->nnzM == 0) { b->have_mass = true; return DAECPP_B200_OK; }
    // compressed-column values: per slot 0.0 + a_k1 + a_k2 + ... in input order (sparse-matrix.hpp:124-152)
    const int32_t host_lanes = shared ? 1 : count; // shared values travel once and are replicated on the device
    std::vector<double> csc((size_t)host_lanes * (size_t)b->nnzM);
    for (int32_t l = 0; l < host_lanes; ++l)
    {
        const double *src = A + (size_t)l * (size_t)b->nTM;
        double *dst = csc.data() + (size_t)l * (size_t)b->nnzM;
        for (int64_t s = 0; s < b->nnzM; ++s)
        {
            volatile double v = 0.0;
            for (int32_t p = b->m_src_ptr[s]; p < b->m_src_ptr[s + 1]; ++p) v = v + src[b->m_src[p]];
            dst[s] = v;
        }
    }
    BATCH_TRY(b, cudaMemcpyAsync(b->mv.p + (size_t)first * (size_t)b->nnzM, csc.data(), sizeof(double) * csc.size(), cudaMemcpyHostToDevice, c->stream));
    if (shared && count > 1)
    {
        BroadcastArgs a{(int32_t)b->nnzM, (int64_t)b->nnzM, b->mv.p};
        k_batch_broadcast<<<dim3((unsigned)((b->nnzM + TPB - 1) / TPB), (unsigned)(count - 1)), TPB, 0, c->stream>>>(a);
        c->kernels += 1;
    }
    BATCH_TRY(b, cudaStreamSynchronize(c->stream));
    b->have_mass = true;
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_set_jacobian_values(daecpp_b200_batch *b, const double *A, int32_t first, int32_t count)
{
    if (!b || !b->have_pattern || !A || !lane_range_ok(b, first, count)) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    BATCH_TRY(b, cudaMemcpyAsync(b->jv.p + (size_t)first * (size_t)b->nTJ, A, sizeof(double) * (size_t)count * (size_t)b->nTJ, cudaMemcpyHostToDevice, b->c->stream));
    BATCH_TRY(b, cudaStreamSynchronize(b->c->stream));
    for (int32_t l = 0; l < count; ++l) b->jac_loaded[(size_t)first + l] = 1;
    return DAECPP_B200_OK;
}

static int copy_lanes(daecpp_b200_batch *b, double *dev, double *host_out, const double *host_in, int32_t first, int32_t count)
{
    // lane-major host slab [count][n] <-> device [lanes][ns], through two page-locked bounce buffers so that the
    // DMA of one chunk of lanes overlaps the host copy of the previous one (the caller's memory is pageable)
    const size_t n = (size_t)b->n;
    cudaStream_t st = b->c->stream;
    if (!b->h_bounce)
    {
        {   // page-locking 16 MB costs milliseconds: the buffers are kept for the next batch of the process
            std::lock_guard<std::mutex> lk(g_bounce_mu);
            if (!g_bounce_free.empty()) { b->h_bounce = g_bounce_free.back(); g_bounce_free.pop_back(); }
        }
        if (!b->h_bounce) BATCH_TRY(b, cudaMallocHost((void **)&b->h_bounce, 2 * daecpp_b200_batch::BOUNCE_BYTES));
        BATCH_TRY(b, cudaEventCreateWithFlags(&b->ev_bounce[0], cudaEventDisableTiming));
        BATCH_TRY(b, cudaEventCreateWithFlags(&b->ev_bounce[1], cudaEventDisableTiming));
    }
    size_t rows = daecpp_b200_batch::BOUNCE_BYTES / (sizeof(double) * n);
    if (rows == 0)
    {   // one lane does not fit a bounce buffer: plain copies
        for (int32_t l = 0; l < count; ++l)
        {
            if (host_in) BATCH_TRY(b, cudaMemcpyAsync(dev + ((size_t)first + l) * b->ns, host_in + (size_t)l * n, sizeof(double) * n, cudaMemcpyHostToDevice, st));
            else BATCH_TRY(b, cudaMemcpyAsync(host_out + (size_t)l * n, dev + ((size_t)first + l) * b->ns, sizeof(double) * n, cudaMemcpyDeviceToHost, st));
        }
        BATCH_TRY(b, cudaStreamSynchronize(st));
        return DAECPP_B200_OK;
    }
    const int32_t nchunks = (int32_t)(((size_t)count + rows - 1) / rows);
    auto chunk_rows = [&](int32_t k) { return (size_t)std::min<size_t>(rows, (size_t)count - (size_t)k * rows); };
    auto bounce = [&](int32_t k) { return reinterpret_cast<double *>(b->h_bounce + (size_t)(k & 1) * daecpp_b200_batch::BOUNCE_BYTES); };
    for (int32_t k = 0; k <= nchunks; ++k)
    {
        if (host_in)
        {
            if (k < nchunks)
            {
                if (k >= 2) BATCH_TRY(b, cudaEventSynchronize(b->ev_bounce[k & 1])); // the DMA that last read this buffer
                std::memcpy(bounce(k), host_in + (size_t)k * rows * n, sizeof(double) * chunk_rows(k) * n);
                BATCH_TRY(b, cudaMemcpy2DAsync(dev + ((size_t)first + (size_t)k * rows) * b->ns, sizeof(double) * b->ns, bounce(k), sizeof(double) * n,
                                               sizeof(double) * n, chunk_rows(k), cudaMemcpyHostToDevice, st));
                BATCH_TRY(b, cudaEventRecord(b->ev_bounce[k & 1], st));
            }
        }
        else
        {
            if (k < nchunks)
            {
                BATCH_TRY(b, cudaMemcpy2DAsync(bounce(k), sizeof(double) * n, dev + ((size_t)first + (size_t)k * rows) * b->ns, sizeof(double) * b->ns,
                                               sizeof(double) * n, chunk_rows(k), cudaMemcpyDeviceToHost, st));
                BATCH_TRY(b, cudaEventRecord(b->ev_bounce[k & 1], st));
            }
            if (k >= 1)
            {
                BATCH_TRY(b, cudaEventSynchronize(b->ev_bounce[(k - 1) & 1]));
                std::memcpy(host_out + (size_t)(k - 1) * rows * n, bounce(k - 1), sizeof(double) * chunk_rows(k - 1) * n);
            }
        }
    }
    BATCH_TRY(b, cudaStreamSynchronize(st));
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_set_rhs_values(daecpp_b200_batch *b, const double *f, int32_t first, int32_t count)
{
    if (!b || !f || !lane_range_ok(b, first, count)) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    if (count == 0) return DAECPP_B200_OK;
    int rc = copy_lanes(b, b->f, nullptr, f, first, count);
    if (rc) return rc;
    for (int32_t l = 0; l < count; ++l) b->rhs_loaded[(size_t)first + l] = 1;
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_set_state(daecpp_b200_batch *b, const double *x0, int32_t first, int32_t count)
{
    if (!b || !x0 || !lane_range_ok(b, first, count)) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    cudaStream_t st = b->c->stream;
    const size_t per = (size_t)b->lanes * b->ns;
    if (count == 0)
    {   // one state shared by every lane: n doubles travel, the lanes are filled on the device
        first = 0; count = b->lanes;
        int rc = copy_lanes(b, b->xk, nullptr, x0, 0, 1);
        if (rc) return rc;
        BroadcastArgs a{b->n, (int64_t)b->ns, b->xk};
        if (b->lanes > 1) k_batch_broadcast<<<dim3((unsigned)((b->n + TPB - 1) / TPB), (unsigned)(b->lanes - 1)), TPB, 0, st>>>(a);
        b->c->kernels += 1;
    }
    else
    {
        int rc = copy_lanes(b, b->xk, nullptr, x0, first, count);
        if (rc) return rc;
    }
    const size_t off = (size_t)first * b->ns, bytes = sizeof(double) * (size_t)count * b->ns;
    BATCH_TRY(b, cudaMemcpyAsync(b->hist + off, b->xk + off, bytes, cudaMemcpyDeviceToDevice, st));
    for (int k = 1; k < 4; ++k) BATCH_TRY(b, cudaMemsetAsync(b->hist + (size_t)k * per + off, 0, bytes, st));
    BATCH_TRY(b, cudaStreamSynchronize(st));
    for (int32_t l = 0; l < count; ++l) b->hbase[(size_t)first + l] = 0;
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_get_vector(daecpp_b200_batch *b, int32_t which, double *out, int32_t first, int32_t count)
{
    if (!b || !out || !lane_range_ok(b, first, count)) return DAECPP_B200_EINVAL;
    cudaSetDevice(b->device);
    if (count == 0) return DAECPP_B200_OK;
    const size_t per = (size_t)b->lanes * b->ns;
    switch (which)
    {
    case 0: return copy_lanes(b, b->b, out, nullptr, first, count);
    case 1: return copy_lanes(b, b->dx, out, nullptr, first, count);
    case 3: return copy_lanes(b, b->xk, out, nullptr, first, count);
    case 4: case 5: case 6: case 7:
        for (int32_t l = 0; l < count; ++l)
        {   // the ring slot differs from lane to lane
            const int slot = (b->hbase[(size_t)first + l] + (which - 4)) & 3;
            BATCH_TRY(b, cudaMemcpyAsync(out + (size_t)l * (size_t)b->n, b->hist + (size_t)slot * per + ((size_t)first + l) * b->ns,
                                         sizeof(double) * (size_t)b->n, cudaMemcpyDeviceToHost, b->c->stream));
        }
        BATCH_TRY(b, cudaStreamSynchronize(b->c->stream));
        return DAECPP_B200_OK;
    default: return bfail(b, DAECPP_B200_EINVAL, "unknown vector selector");
    }
}

int daecpp_b200_batch_set_device_functors(daecpp_b200_batch *b, const daecpp_b200_device_functors *fn)
{
    if (!b) return DAECPP_B200_EINVAL;
    b->have_fn = fn && fn->rhs && fn->jacobian;
    if (fn && (fn->rhs || fn->jacobian) && !b->have_fn) return bfail(b, DAECPP_B200_EINVAL, "device functors: rhs and jacobian must both be given");
    if (b->have_fn) b->fn = *fn;
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_round(daecpp_b200_batch *b, const daecpp_b200_lane_cmd *cmd, double atol, double rtol, double max_err_abs,
                            int32_t variability_on, double variability_tolerance, daecpp_b200_lane_result *res)
{
    if (!b || !cmd || !res) return DAECPP_B200_EINVAL;
    if (!b->have_pattern || !b->have_mass) return bfail(b, DAECPP_B200_EINVAL, "set_pattern and set_mass_values must precede a round");
    cudaSetDevice(b->device);
    daecpp_b200_ctx *c = b->c;
    cudaStream_t st = c->stream;
    const int32_t Ln = b->lanes;
    constexpr int KMAX = daecpp_b200_batch::KMAX;
    int n_active = 0, n_refactor0 = 0, n_pre = 0, K = 1;
    for (int32_t l = 0; l < Ln; ++l)
    {
        const daecpp_b200_lane_cmd &q = cmd[l];
        if (q.mode < 0 || q.mode > 2 || q.pre < 0 || q.pre > 2) return bfail(b, DAECPP_B200_EINVAL, "bad lane command");
        b->ctl_mode()[l] = q.mode;
        b->ctl_pre()[l] = q.pre;
        if (q.pre == DAECPP_B200_LANE_ACCEPT) b->hbase[l] = (b->hbase[l] + 3) & 3; // the oldest slot receives xk and becomes x[0]
        if (q.pre != DAECPP_B200_LANE_KEEP) { b->refine_futile[l] = 0; b->ctl_list_pre()[n_pre++] = l; } // a new step, or a new attempt at it
        b->ctl_hbase()[l] = b->hbase[l];
        b->ctl_order()[l] = q.order;
        b->ctl_t()[l] = q.t;
        b->ctl_norefine()[l] = (c->refine_steps <= 0 || b->refine_futile[l]) ? 1 : 0;
        double w[5] = {0, 0, 0, 0, 0};
        int kl = 0;
        if (q.mode != DAECPP_B200_LANE_IDLE)
        {
            if (!weights(q.order, q.dt, w)) return bfail(b, DAECPP_B200_EINVAL, "unsupported time integration order");
            kl = q.max_iterations > 1 ? (q.max_iterations < KMAX ? q.max_iterations : KMAX) : 1;
            if (!b->have_fn)
            {
                kl = 1; // host functors: every iteration needs fresh values from the caller
                if (!b->rhs_loaded[l] || (q.mode == DAECPP_B200_LANE_REFACTOR && !b->jac_loaded[l]))
                    return bfail(b, DAECPP_B200_EINVAL, "host-functor lane without uploaded rhs / Jacobian values");
            }
            b->ctl_list_active()[n_active++] = l;
            if (q.mode == DAECPP_B200_LANE_REFACTOR) b->ctl_list_refactor()[n_refactor0++] = l;
            if (kl > K) K = kl;
        }
        for (int j = 0; j < KMAX; ++j)
        {
            int m = DAECPP_B200_LANE_IDLE;
            if (j == 0) m = q.mode;
            else if (j < kl) m = (q.period <= 1 || (q.phase + j) % q.period == 0) ? DAECPP_B200_LANE_REFACTOR : DAECPP_B200_LANE_SOLVE;
            b->ctl_plan(j)[l] = m;
        }
        for (int k = 0; k < 5; ++k) b->ctl_w()[(size_t)k * Ln + l] = w[k];
    }
    BATCH_TRY(b, cudaEventRecord(b->ev0, st));
    BATCH_TRY(b, cudaMemcpyAsync(b->d_ctl, b->h_ctl, b->ctl_bytes, cudaMemcpyHostToDevice, st));
    BATCH_TRY(b, cudaMemsetAsync(b->d_res, 0, b->res_bytes, st));
    const unsigned gx = (unsigned)((b->n + TPB - 1) / TPB);
    if (n_pre)
    {
        PreArgs a{b->n, Ln, (int64_t)b->ns, b->L.s_hist, b->dev(b->ctl_pre()), b->L.hbase, b->dev(b->ctl_list_pre()), b->xk, b->hist};
        k_batch_pre<<<dim3(gx, (unsigned)n_pre), TPB, 0, st>>>(a);
        c->kernels += 1;
    }
    const size_t rstride = b->res_stride();
    auto flags_of = [&](char *base, int j) { return reinterpret_cast<int32_t *>(base + (size_t)j * rstride); };
    auto var_of = [&](char *base, int j) { return reinterpret_cast<unsigned long long *>(base + (size_t)j * rstride + sizeof(int32_t) * 4 * (size_t)Ln); };
    if (n_active > 0)
    {
        const dim3 gv(gx, (unsigned)n_active);
        const double *d_t = b->dev(b->ctl_t());
        b->L.var_on = variability_on ? 1 : 0;
        b->L.var_tol = variability_tolerance;
        b->L.n_active = n_active;
        b->L.list_active = b->dev(b->ctl_list_active());
        const bool fuse_res = (c->m_diag || c->m_pdiag) && c->chain_fuse_residual;
        int rc = DAECPP_B200_OK;
        for (int j = 0; j < K && !rc; ++j)
        {
            // the round's lists stay those of iteration 0 (a superset): lanes that halted on the device leave through
            // their effective mode word.  Iterations after the first refactorise whoever planned to.
            int n_refac = n_refactor0;
            if (j > 0)
            {
                n_refac = 0;
                for (int32_t q = 0; q < n_active; ++q)
                    if (b->ctl_plan(j)[b->ctl_list_active()[q]] == DAECPP_B200_LANE_REFACTOR) ++n_refac;
            }
            b->L.flags = flags_of(b->d_res, j);
            b->L.var = var_of(b->d_res, j);
            {
                PlanArgs pa{Ln, j, n_active, b->L.list_active, b->dev(b->ctl_plan(j)), b->dev(b->ctl_norefine()),
                            j > 0 ? flags_of(b->d_res, j - 1) : nullptr, b->res_iters(b->d_res) + Ln, b->res_iters(b->d_res)};
                {
                    cudaLaunchConfig_t cfg = {};
                    cfg.gridDim = dim3((unsigned)((n_active + 127) / 128), 1, 1);
                    cfg.blockDim = dim3(128, 1, 1);
                    cfg.stream = st;
                    cudaLaunchAttribute at[1];
                    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
                    at[0].val.programmaticStreamSerializationAllowed = 1;
                    cfg.attrs = at;
                    cfg.numAttrs = 1;
                    BATCH_TRY(b, cudaLaunchKernelEx(&cfg, k_batch_plan, pa));
                }
                c->kernels += 1;
            }
            // j == 0: the compacted list of refactoring lanes is exact; later iterations run the factor kernels over the
            // whole active list (the effective mode sorts the lanes out)
            b->L.list_refactor = j == 0 ? b->dev(b->ctl_list_refactor()) : b->L.list_active;
            b->L.n_refactor = j == 0 ? n_refactor0 : n_active;
            if (b->have_fn)
            {
                daecpp_b200_functor_args fa{};
                fa.n = b->n; fa.lanes = Ln; fa.x_stride = (int64_t)b->ns; fa.x = b->xk; fa.t = d_t; fa.mode = b->L.mode; fa.stream = st;
                fa.min_mode = DAECPP_B200_LANE_SOLVE; fa.out = b->f; fa.out_stride = (int64_t)b->ns;
                fa.lane_list = b->L.list_active; fa.n_list = n_active;
                if (b->fn.rhs(b->fn.user, &fa)) return bfail(b, DAECPP_B200_EINVAL, "device rhs functor failed");
                c->kernels += 1;
                if (n_refac)
                {
                    fa.min_mode = DAECPP_B200_LANE_REFACTOR; fa.out = b->jv.p; fa.out_stride = b->nTJ;
                    fa.lane_list = b->L.list_refactor; fa.n_list = b->L.n_refactor;
                    if (b->fn.jacobian(b->fn.user, &fa)) return bfail(b, DAECPP_B200_EINVAL, "device Jacobian functor failed");
                    c->kernels += 1;
                }
            }
            // the prototype's launch code runs over the lane slabs
            c->lanes = &b->L;
            c->xk = b->xk; c->f = b->f; c->b = b->b; c->dx = b->dx;
            if (!fuse_res)
            {
                ResArgs a{b->n, Ln, (int64_t)b->ns, b->L.s_hist, b->L.s_mv, b->L.mode, b->L.hbase, b->L.order, b->L.list_active, b->L.w,
                          b->xk, b->hist, b->mv.p, b->f, b->b, c->m_rowptr.p, c->m_rowslot.p, c->m_rowcol.p};
                k_batch_residual<<<gv, TPB, 0, st>>>(a);
                c->kernels += 1;
            }
            c->chain_may_defer = true; // the solve follows on the stream: the top level may ride in the forward launch
            if (n_refac) rc = chain_refactor_async(c, 0.0);
            c->chain_may_defer = false;
            if (!rc)
            {
                c->fuse_residual_now = fuse_res;
                c->publish_flags = false;
                c->forward_follows_factor = n_refac > 0;
                rc = chain_solve(c, true, atol, rtol, max_err_abs);
                c->fuse_residual_now = false;
                c->forward_follows_factor = false;
            }
            c->lanes = nullptr;
            c->xk = b->proto_xk; c->f = b->proto_f; c->b = b->proto_b; c->dx = b->proto_dx;
        }
        if (rc) return bfail(b, rc, c->err);
    }
    BATCH_TRY(b, cudaMemcpyAsync(b->h_res, b->d_res, b->res_bytes, cudaMemcpyDeviceToHost, st));
    BATCH_TRY(b, cudaStreamSynchronize(st));
    const int32_t *iters = b->res_iters(b->h_res);
    // the results of a lane are those of the last iteration it ran
    std::vector<int32_t> last((size_t)Ln, 0);
    for (int32_t l = 0; l < Ln; ++l) last[l] = iters[l] > 0 ? iters[l] - 1 : 0;
    if (n_active > 0 && c->refine_steps > 0)
    {   // guarded refinement (api.cu: wants_refinement) for the lanes whose update failed the convergence test only on
        // elements below the refinement floor: a second, small pass over those lanes
        int n_ref = 0;
        for (int32_t l = 0; l < Ln; ++l)
        {
            const int32_t *fl0 = flags_of(b->h_res, last[l]) + 4 * l;
            const int32_t f = fl0[0];
            const bool want = cmd[l].mode != DAECPP_B200_LANE_IDLE && fl0[1] == 0 && (f & DAECPP_B200_NOT_CONVERGED) &&
                              !(f & (DAECPP_B200_ABOVE_FLOOR | DAECPP_B200_DIVERGED)) && !b->refine_futile[l];
            b->ctl_mode()[l] = want ? DAECPP_B200_LANE_SOLVE : DAECPP_B200_LANE_IDLE; // (host-side mark; the device reads the plan words)
            b->ctl_plan(0)[l] = b->ctl_mode()[l];
            if (want) b->ctl_list_active()[n_ref++] = l;
        }
        if (n_ref)
        {
            const size_t per = (size_t)Ln * b->ns;
            BATCH_TRY(b, b->refine_buf.reserve(2 * per));
            b->L.refine_r = b->refine_buf.p; b->L.refine_e = b->refine_buf.p + per;
            BATCH_TRY(b, cudaMemcpyAsync(b->d_ctl, b->h_ctl, b->ctl_bytes, cudaMemcpyHostToDevice, st));
            // results of the pass go to a scratch slot (iteration slot KMAX-1 is free unless the round ran KMAX
            // iterations; then slot K-1 is borrowed and what the other lanes left there is put back)
            const int slot = K < KMAX ? KMAX - 1 : K - 1;
            b->L.flags = flags_of(b->d_res, slot);
            b->L.var = var_of(b->d_res, slot);
            std::vector<int32_t> keep_fl;
            std::vector<unsigned long long> keep_var;
            if (slot == K - 1)
            {
                keep_fl.assign(flags_of(b->h_res, slot), flags_of(b->h_res, slot) + 4 * (size_t)Ln);
                keep_var.assign(var_of(b->h_res, slot), var_of(b->h_res, slot) + Ln);
            }
            {
                PlanArgs pa{Ln, 0, n_ref, b->L.list_active, b->dev(b->ctl_plan(0)), b->dev(b->ctl_norefine()), nullptr,
                            b->res_iters(b->d_res) + Ln, b->res_iters(b->d_res)};
                k_batch_plan<<<(n_ref + 127) / 128, 128, 0, st>>>(pa); // sets the effective mode words of the refined lanes
            }
            b->L.n_active = n_ref; b->L.n_refactor = 0;
            c->lanes = &b->L;
            c->xk = b->xk; c->f = b->f; c->b = b->b; c->dx = b->dx;
            const int rc = refine_last_solve(c, 0.0, atol, rtol, max_err_abs, !((c->m_diag || c->m_pdiag) && c->chain_fuse_residual));
            c->lanes = nullptr;
            c->xk = b->proto_xk; c->f = b->proto_f; c->b = b->proto_b; c->dx = b->proto_dx;
            if (rc) return bfail(b, rc, c->err);
            BATCH_TRY(b, cudaMemcpyAsync(b->h_res + (size_t)slot * rstride, b->d_res + (size_t)slot * rstride, rstride, cudaMemcpyDeviceToHost, st));
            BATCH_TRY(b, cudaStreamSynchronize(st));
            int32_t *fls = flags_of(b->h_res, slot);
            unsigned long long *vas = var_of(b->h_res, slot);
            for (int32_t q = 0; q < n_ref; ++q)
            {   // move the refined lanes' results to their own last slot
                const int32_t l = b->ctl_list_active()[q];
                int32_t *dstf = flags_of(b->h_res, last[l]) + 4 * l;
                const int32_t f0 = fls[4 * l];
                if ((f0 & DAECPP_B200_NOT_CONVERGED) && !(f0 & (DAECPP_B200_ABOVE_FLOOR | DAECPP_B200_DIVERGED)))
                    b->refine_futile[l] = 1; // the corrected update fails the same way: the noise is not the solve's
                if (last[l] != slot)
                {
                    for (int k = 0; k < 4; ++k) dstf[k] = fls[4 * l + k];
                    var_of(b->h_res, last[l])[l] = vas[l];
                }
            }
            if (slot == K - 1)
                for (int32_t l = 0; l < Ln; ++l)
                    if (b->ctl_mode()[l] == DAECPP_B200_LANE_IDLE)
                    {   // lanes that were not refined keep what the round left in this slot
                        for (int k = 0; k < 4; ++k) fls[4 * l + k] = keep_fl[4 * (size_t)l + k];
                        vas[l] = keep_var[l];
                    }
            b->n_refine_rounds += 1;
            b->n_refined_lanes += n_ref;
        }
    }
    BATCH_TRY(b, cudaEventRecord(b->ev1, st));
    BATCH_TRY(b, cudaEventSynchronize(b->ev1));
    {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, b->ev0, b->ev1) == cudaSuccess) b->ms_device += ms;
    }
    b->rounds += 1;
    for (int32_t l = 0; l < Ln; ++l)
    {
        daecpp_b200_lane_result &r = res[l];
        const int32_t *fl = flags_of(b->h_res, last[l]) + 4 * l;
        r.flags = fl[0] & (DAECPP_B200_DIVERGED | DAECPP_B200_NOT_CONVERGED);
        r.status = DAECPP_B200_OK;
        r.iterations = cmd[l].mode != DAECPP_B200_LANE_IDLE ? iters[l] : 0;
        r.reserved = 0;
        if (cmd[l].mode != DAECPP_B200_LANE_IDLE && fl[1]) r.status = DAECPP_B200_EDECOMP;
        double v;
        std::memcpy(&v, &var_of(b->h_res, last[l])[l], sizeof(double));
        r.variability = v;
        if (!b->have_fn && cmd[l].mode != DAECPP_B200_LANE_IDLE) { b->rhs_loaded[l] = 0; b->jac_loaded[l] = b->jac_loaded[l] && cmd[l].mode != DAECPP_B200_LANE_REFACTOR; }
    }
    return DAECPP_B200_OK;
}

int daecpp_b200_batch_get_stat(daecpp_b200_batch *b, const char *key, double *value)
{
    if (!b || !key || !value) return DAECPP_B200_EINVAL;
    const std::string k(key);
    if (k == "rounds") *value = b->rounds;
    else if (k == "ms_device") *value = b->ms_device;
    else if (k == "lanes") *value = b->lanes;
    else if (k == "refine_rounds") *value = b->n_refine_rounds;
    else if (k == "refined_lanes") *value = b->n_refined_lanes;
    else if (k == "bytes_per_lane")
        *value = 8.0 * (8.0 * (double)b->ns + (double)b->nTJ + (double)b->nnzM + (double)(b->c->CH.fac_doubles + b->c->CH.y_doubles + b->c->CH.red_doubles + b->c->CH.rv_doubles + b->c->CH.x_doubles));
    else return daecpp_b200_get_stat(b->c, key, value);
    return DAECPP_B200_OK;
}

} // extern "C"
