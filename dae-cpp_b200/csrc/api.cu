// C ABI of libdaecpp_b200.so (see include/daecpp_b200.h for the contract and the reference
// regions each entry point replaces).
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <chrono>
#include <mutex>
#include <unordered_map>
#include <map>
#include <new>
#include <utility>

#include "ctx.hpp"

using namespace daecpp_b200;

static thread_local std::string g_create_error;

// Process-wide pools (ensembles): released contexts and page-locked buffers are kept for the next
// member instead of going through cudaFree / cudaFreeHost, which synchronise the whole device and
// serialise the host threads of the other members in flight.  A recycled context keeps its cached
// sparsity patterns and its symbolic plan: a member with the same pattern starts on the hot path.
static std::mutex g_pool_mu;
static bool g_pool_on = false;
static std::vector<daecpp_b200_ctx *> g_ctx_pool;
static std::vector<std::pair<size_t, void *>> g_host_pool;

// Large device blocks (ctx.hpp: dev_alloc / dev_free).  live: every block of kCacheMin bytes or more handed out;
// idle: freed ones waiting for a request of the same device and size.
namespace
{
constexpr size_t kCacheMin = size_t(4) << 20;
constexpr size_t kCacheMaxIdle = size_t(64) << 30; // per process; beyond it a freed block really is freed
struct Block { int device; size_t bytes; };
std::mutex g_dev_mu;
// (never destroyed: a context may be released after the static destructors of this library have run)
std::unordered_map<void *, Block> &g_dev_live = *new std::unordered_map<void *, Block>();
std::multimap<std::pair<int, size_t>, void *> &g_dev_idle = *new std::multimap<std::pair<int, size_t>, void *>();
size_t g_dev_idle_bytes = 0;
} // namespace

namespace daecpp_b200
{
cudaError_t dev_alloc(void **p, size_t bytes)
{
    if (bytes < kCacheMin) return cudaMalloc(p, bytes);
    int dev = 0;
    cudaGetDevice(&dev);
    {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        auto it = g_dev_idle.find({dev, bytes});
        if (it != g_dev_idle.end())
        {
            *p = it->second;
            g_dev_idle.erase(it);
            g_dev_idle_bytes -= bytes;
            g_dev_live[*p] = Block{dev, bytes};
            return cudaSuccess;
        }
    }
    cudaError_t e = cudaMalloc(p, bytes);
    if (e != cudaSuccess && g_dev_idle_bytes)
    {   // out of memory with blocks parked: give them back and try once more
        cudaGetLastError();
        dev_cache_release();
        e = cudaMalloc(p, bytes);
    }
    if (e == cudaSuccess)
    {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        g_dev_live[*p] = Block{dev, bytes};
    }
    return e;
}

void dev_free(void *p)
{
    if (!p) return;
    {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        auto it = g_dev_live.find(p);
        if (it != g_dev_live.end())
        {
            const Block blk = it->second;
            g_dev_live.erase(it);
            bool keep;
            { std::lock_guard<std::mutex> lp(g_pool_mu); keep = g_pool_on; }
            if (keep && g_dev_idle_bytes + blk.bytes <= kCacheMaxIdle)
            {   // cudaFree waits for the device; a parked block may be handed to another context at once, so wait here too
                cudaDeviceSynchronize();
                g_dev_idle.insert({{blk.device, blk.bytes}, p});
                g_dev_idle_bytes += blk.bytes;
                return;
            }
        }
    }
    cudaFree(p);
}

void dev_cache_release()
{
    std::multimap<std::pair<int, size_t>, void *> idle;
    {
        std::lock_guard<std::mutex> lk(g_dev_mu);
        idle.swap(g_dev_idle);
        g_dev_idle_bytes = 0;
    }
    int cur = 0;
    cudaGetDevice(&cur);
    for (auto &kv : idle) { cudaSetDevice(kv.first.first); cudaFree(kv.second); }
    cudaSetDevice(cur);
}
} // namespace daecpp_b200

namespace
{

int fail(daecpp_b200_ctx *c, int code, const char *msg)
{
    c->err = msg;
    return code;
}

// Variable-step BDF-1..4 weights.  The association of every product/quotient follows the
// expressions at dae-cpp/solver.hpp:160,168-173,178-184,190-199 so that alpha and the weights
// are bit-identical to the reference's (C++ evaluates * and / left to right).
bool bdf_weights(int order, const double *dt, double *w)
{
    const double h0 = dt[0], h1 = dt[1], h2 = dt[2], h3 = dt[3];
    const double s01 = h0 + h1, s12 = h1 + h2, s23 = h2 + h3;
    const double s012 = h0 + h1 + h2, s123 = h1 + h2 + h3, s0123 = h0 + h1 + h2 + h3;
    w[0] = w[1] = w[2] = w[3] = w[4] = 0.0;
    if (order == 1)
    {
        w[0] = 1.0 / h0;
        w[1] = w[0];
    }
    else if (order == 2)
    {
        w[0] = (2.0 * h0 + h1) / (h0 * s01);
        w[1] = s01 / (h0 * h1);
        w[2] = h0 / (h1 * s01);
    }
    else if (order == 3)
    {
        w[0] = (3.0 * h0 * h0 + h1 * s12 + 2.0 * h0 * (2.0 * h1 + h2)) / (h0 * s01 * s012);
        w[1] = s01 * s012 / (h0 * h1 * s12);
        w[2] = h0 * s012 / (h1 * h2 * s01);
        w[3] = h0 * s01 / (h2 * s12 * s012);
    }
    else if (order == 4)
    {
        w[0] = (4.0 * h0 * h0 * h0 + h1 * s12 * s123 + 3.0 * h0 * h0 * (3.0 * h1 + 2.0 * h2 + h3) +
                2.0 * h0 * (3.0 * h1 * h1 + h2 * s23 + 2.0 * h1 * (2.0 * h2 + h3))) /
               (h0 * s01 * s012 * s0123);
        w[1] = s01 * s012 * s0123 / (h0 * h1 * s12 * s123);
        w[2] = h0 * s012 * s0123 / (h1 * s01 * h2 * s23);
        w[3] = h0 * s01 * s0123 / (h2 * s12 * s012 * h3);
        w[4] = h0 * s01 * s012 / (h3 * s23 * s123 * s0123);
    }
    else
        return false;
    return true;
}

int ensure_stage(daecpp_b200_ctx *c, size_t count)
{
    if (count <= c->h_stage_cap) return DAECPP_B200_OK;
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = nullptr; c->h_stage_cap = 0;
    DAECPP_CUDA_TRY(c, cudaMallocHost((void **)&c->h_stage, sizeof(double) * count));
    c->h_stage_cap = count;
    return DAECPP_B200_OK;
}

int upload_vector(daecpp_b200_ctx *c, double *dst, const double *src_host)
{
    if (c->pinned_io)
    { // the caller guarantees page-locked buffers (overlay solver.hpp): one DMA, no staging copy
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(dst, src_host, sizeof(double) * c->n, cudaMemcpyHostToDevice, c->stream));
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        return DAECPP_B200_OK;
    }
    // through the pinned stage so the copy is one DMA regardless of how the caller allocated
    int rc = ensure_stage(c, c->n);
    if (rc) return rc;
    std::memcpy(c->h_stage, src_host, sizeof(double) * c->n);
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(dst, c->h_stage, sizeof(double) * c->n, cudaMemcpyHostToDevice, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return DAECPP_B200_OK;
}

int download_vector(daecpp_b200_ctx *c, double *dst_host, const double *src)
{
    if (c->pinned_io)
    {
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(dst_host, src, sizeof(double) * c->n, cudaMemcpyDeviceToHost, c->stream));
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        return DAECPP_B200_OK;
    }
    int rc = ensure_stage(c, c->n);
    if (rc) return rc;
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_stage, src, sizeof(double) * c->n, cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    std::memcpy(dst_host, c->h_stage, sizeof(double) * c->n);
    return DAECPP_B200_OK;
}

} // namespace

// A pooled context handed to a new owner: numerical state and options back to their defaults, the
// cached patterns (triplet streams, union, symbolic plan) stay -- they are revalidated against the new
// owner's uploads exactly like between two calls of one owner.
static void recycle(daecpp_b200_ctx *c)
{
    const daecpp_b200_options fresh;
    daecpp_b200_options &o = *c;
    // a plan built under other plan-affecting options is not this owner's plan
    if (o.chain_enabled != fresh.chain_enabled || o.chain_min_n != fresh.chain_min_n || o.chain_max_b != fresh.chain_max_b ||
        o.chain_affine != fresh.chain_affine || o.unit_threads != fresh.unit_threads || o.sopt.leaf_size != fresh.sopt.leaf_size ||
        o.sopt.unit_smem_bytes != fresh.sopt.unit_smem_bytes || o.sopt.serial_dim != fresh.sopt.serial_dim ||
        o.sopt.max_super != fresh.sopt.max_super || o.sopt.transversal != fresh.sopt.transversal ||
        o.sopt.warp_dim != fresh.sopt.warp_dim || o.sopt.group8_dim != fresh.sopt.group8_dim || o.sopt.warp_dim_many != fresh.sopt.warp_dim_many ||
        o.sopt.warp_min_fronts != fresh.sopt.warp_min_fronts)
        c->have_symbolic = false;
    o = fresh; // every option back to its default, whatever is added to daecpp_b200_options later
    const daecpp_b200::TinyState tiny_fresh;
    c->tiny.enabled = tiny_fresh.enabled; c->tiny.have_factor = false; c->tiny.spec_valid = false; c->tiny.var_tol = tiny_fresh.var_tol;
    c->err.clear();
    c->have_factor = false;
    c->direct_csc = false;
    c->M.vals_valid = false; c->J.vals_valid = false;
    // transient state of the iteration protocol
    c->xk_host_target = nullptr; c->xk_direct_used = false; c->publish_flags = false; c->flags_published = false; c->flags_clean = false;
    c->fuse_residual_now = false; c->residual_was_fused = false; c->forward_follows_factor = false; c->rhs_prefetch_hint = false;
    c->values_quiescent = false; c->publish_slot = 0; c->cancel_armed = false;
    c->chain_top_mode = 0; c->chain_may_defer = false; c->chain_top_pending = false;
    c->refine_futile = false; c->n_refine = 0; c->alpha_hint = 0.0; c->last_alpha = 0.0;
    // statistics
    c->n_symbolic = c->n_refactor = c->n_solve = c->n_cold = c->kernels = 0;
    c->ms_symbolic = c->ms_last_refactor = c->ms_last_solve = 0;
    for (int k = 0; k < 5; ++k) c->ms_phase[k] = 0;
    c->n_phase_samples = 0;
    for (int k = 0; k < 8; ++k) { c->h_flags[k] = 0; c->h_flags[8 + k] = 0; c->h_scalars[k] = 0.0; }
}

extern "C" {
static bool wants_refinement(const daecpp_b200_ctx *c, int32_t fl);
static int refine_sync(daecpp_b200_ctx *c, double atol, double rtol, double max_err_abs, bool b_is_stored);


const char *daecpp_b200_version(void) { return "daecpp_b200 0.1 (sm_100a)"; }

int daecpp_b200_create(int32_t n, int device, daecpp_b200_ctx **out)
{
    if (!out) return DAECPP_B200_EINVAL;
    *out = nullptr;
    if (n <= 0) { g_create_error = "system size must be positive"; return DAECPP_B200_EINVAL; }
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
    {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e) + " (this library has no CPU fallback)";
        return DAECPP_B200_ECUDA;
    }
    if (device < 0 || device >= ndev) { g_create_error = "device index out of range"; return DAECPP_B200_EINVAL; }
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (size_t k = 0; g_pool_on && k < g_ctx_pool.size(); ++k)
            if (g_ctx_pool[k]->n == n && g_ctx_pool[k]->device == device)
            {
                daecpp_b200_ctx *r = g_ctx_pool[k];
                g_ctx_pool.erase(g_ctx_pool.begin() + k);
                recycle(r);
                *out = r;
                return DAECPP_B200_OK;
            }
    }
    daecpp_b200_ctx *c = new (std::nothrow) daecpp_b200_ctx();
    if (!c) return DAECPP_B200_ENOMEM;
    c->n = n;
    c->device = device;
    auto bail = [&](cudaError_t err, const char *what) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        daecpp_b200_destroy(c);
        return err == cudaErrorMemoryAllocation ? DAECPP_B200_ENOMEM : DAECPP_B200_ECUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return bail(e, "cudaSetDevice");
    if ((e = cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "cudaStreamCreate");
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail(e, "cudaEventCreate");
    if ((e = cudaEventCreate(&c->ev_mark)) != cudaSuccess) return bail(e, "cudaEventCreate");
    for (int k = 0; k < 6; ++k)
        if ((e = cudaEventCreate(&c->pev[k])) != cudaSuccess) return bail(e, "cudaEventCreate");
    double *slab = nullptr;
    const size_t nv = 10; // xk, x[0..3], f, b, dx, work, dxdt
    const size_t ns = ((size_t)n + 31) / 32 * 32; // every vector starts on a 256-byte boundary (16-byte vector accesses, odd n)
    if ((e = cudaMalloc((void **)&slab, sizeof(double) * nv * ns)) != cudaSuccess) return bail(e, "cudaMalloc(state)");
    c->xk = slab;
    for (int k = 0; k < 4; ++k) c->xh[k] = slab + (size_t)(k + 1) * ns;
    c->f = slab + 5 * ns; c->b = slab + 6 * ns; c->dx = slab + 7 * ns;
    c->work = slab + 8 * ns; c->dxdt = slab + 9 * ns;
    if ((e = cudaMemsetAsync(slab, 0, sizeof(double) * nv * ns, c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
    if ((e = cudaMalloc((void **)&c->d_flags, sizeof(int32_t) * 8)) != cudaSuccess) return bail(e, "cudaMalloc(flags)");
    if ((e = cudaMalloc((void **)&c->d_scalars, sizeof(double) * 8)) != cudaSuccess) return bail(e, "cudaMalloc(scalars)");
    if ((e = cudaHostAlloc((void **)&c->h_flags, sizeof(int32_t) * 16, cudaHostAllocMapped)) != cudaSuccess) return bail(e, "cudaHostAlloc");
    if ((e = cudaHostGetDevicePointer((void **)&c->h_flags_dev, c->h_flags, 0)) != cudaSuccess) return bail(e, "cudaHostGetDevicePointer");
    if ((e = cudaMallocHost((void **)&c->h_scalars, sizeof(double) * 8)) != cudaSuccess) return bail(e, "cudaMallocHost");
    if ((e = cudaMemsetAsync(c->d_flags, 0, sizeof(int32_t) * 8, c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
    if ((e = cudaMemsetAsync(c->d_scalars, 0, sizeof(double) * 8, c->stream)) != cudaSuccess) return bail(e, "cudaMemset");
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return bail(e, "cudaStreamSynchronize");
    *out = c;
    return DAECPP_B200_OK;
}

static void destroy_now(daecpp_b200_ctx *c);

void daecpp_b200_destroy(daecpp_b200_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        if (g_pool_on && c->stream && c->xk && c->h_scalars && g_ctx_pool.size() < 1024)
        {
            g_ctx_pool.push_back(c);
            return;
        }
    }
    destroy_now(c);
}

int daecpp_b200_set_pooling(int on)
{
    std::vector<daecpp_b200_ctx *> ctxs;
    std::vector<std::pair<size_t, void *>> bufs;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        g_pool_on = on != 0;
        if (!g_pool_on) { ctxs.swap(g_ctx_pool); bufs.swap(g_host_pool); }
    }
    for (daecpp_b200_ctx *c : ctxs) destroy_now(c);
    for (auto &b : bufs) cudaFreeHost(b.second);
    if (!on) daecpp_b200::dev_cache_release();
    return DAECPP_B200_OK;
}

static void destroy_now(daecpp_b200_ctx *c)
{
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    tiny_release(c);
    c->M.release(); c->J.release(); c->dS.release(); c->dC.release();
    c->m_rowptr.release(); c->m_rowslot.release(); c->m_rowcol.release(); c->m_dslot.release();
    c->u_outer.release(); c->u_inner.release(); c->u_J.release(); c->u_M.release(); c->u_vals.release();
    c->LU.release(); c->CB.release(); c->bf_diag.release(); c->upd.release(); c->trace.release();
    c->keys_a.release(); c->keys_b.release(); c->pay_a.release(); c->pay_b.release(); c->scan.release(); c->cub_tmp.release();
    if (c->xk) cudaFree(c->xk);
    if (c->d_flags) cudaFree(c->d_flags);
    if (c->d_scalars) cudaFree(c->d_scalars);
    if (c->h_flags) cudaFreeHost(c->h_flags);
    if (c->h_scalars) cudaFreeHost(c->h_scalars);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->ev0) cudaEventDestroy(c->ev0);
    if (c->ev1) cudaEventDestroy(c->ev1);
    if (c->ev_mark) cudaEventDestroy(c->ev_mark);
    if (c->ev_z0) cudaEventDestroy(c->ev_z0);
    if (c->ev_z1) cudaEventDestroy(c->ev_z1);
    if (c->ev_u0) cudaEventDestroy(c->ev_u0);
    if (c->ev_u1) cudaEventDestroy(c->ev_u1);
    if (c->stream_zero) { cudaStreamSynchronize(c->stream_zero); cudaStreamDestroy(c->stream_zero); }
    for (int k = 0; k < 6; ++k) if (c->pev[k]) cudaEventDestroy(c->pev[k]);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

const char *daecpp_b200_last_error(const daecpp_b200_ctx *c) { return c ? c->err.c_str() : g_create_error.c_str(); }

// page-locked blocks carry their size in a 64-byte header so that a freed block can be reused
void *daecpp_b200_host_alloc(size_t bytes)
{
    if (!bytes) bytes = 1;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        for (size_t k = 0; g_pool_on && k < g_host_pool.size(); ++k)
            if (g_host_pool[k].first == bytes)
            {
                void *p = g_host_pool[k].second;
                g_host_pool.erase(g_host_pool.begin() + k);
                return static_cast<char *>(p) + 64;
            }
    }
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes + 64) != cudaSuccess) return nullptr;
    *static_cast<size_t *>(p) = bytes;
    return static_cast<char *>(p) + 64;
}
void daecpp_b200_host_free(void *q)
{
    if (!q) return;
    void *p = static_cast<char *>(q) - 64;
    {
        std::lock_guard<std::mutex> lk(g_pool_mu);
        if (g_pool_on && g_host_pool.size() < 4096)
        {
            g_host_pool.emplace_back(*static_cast<size_t *>(p), p);
            return;
        }
    }
    cudaFreeHost(p);
}

int daecpp_b200_set_blocking_sync(int device, int on)
{
    if (cudaSetDevice(device) != cudaSuccess) return DAECPP_B200_ECUDA;
    const cudaError_t e = cudaSetDeviceFlags(on ? cudaDeviceScheduleBlockingSync : cudaDeviceScheduleAuto);
    return e == cudaSuccess ? DAECPP_B200_OK : DAECPP_B200_ECUDA;
}

int daecpp_b200_host_register(void *p, size_t bytes)
{
    if (!p || !bytes) return DAECPP_B200_EINVAL;
    const cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) cudaGetLastError();
    if (std::getenv("DAECPP_B200_TRACE_PIN")) std::fprintf(stderr, "daecpp_b200: page lock of %zu bytes %s\n", bytes, e == cudaSuccess ? "taken" : "refused");
    return e == cudaSuccess ? DAECPP_B200_OK : DAECPP_B200_ECUDA;
}
int daecpp_b200_host_unregister(void *p)
{
    if (!p) return DAECPP_B200_EINVAL;
    const cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) cudaGetLastError();
    if (std::getenv("DAECPP_B200_TRACE_PIN")) std::fprintf(stderr, "daecpp_b200: page lock %s\n", e == cudaSuccess ? "released" : "release failed");
    return e == cudaSuccess ? DAECPP_B200_OK : DAECPP_B200_ECUDA;
}

int daecpp_b200_set_state(daecpp_b200_ctx *c, const double *x0)
{
    if (!c || !x0) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    c->refine_futile = false;
    c->tiny.spec_valid = false;
    int rc = upload_vector(c, c->xk, x0);
    if (rc) return rc;
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->xh[0], c->xk, sizeof(double) * c->n, cudaMemcpyDeviceToDevice, c->stream));
    for (int k = 1; k < 4; ++k) DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->xh[k], 0, sizeof(double) * c->n, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return DAECPP_B200_OK;
}

int daecpp_b200_get_xk(daecpp_b200_ctx *c, double *xk_host)
{
    if (!c || !xk_host) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    return download_vector(c, xk_host, c->xk);
}

int daecpp_b200_rollback(daecpp_b200_ctx *c)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    c->tiny.spec_valid = false;
    c->refine_futile = false;
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->xk, c->xh[0], sizeof(double) * c->n, cudaMemcpyDeviceToDevice, c->stream));
    return DAECPP_B200_OK;
}

int daecpp_b200_set_mass(daecpp_b200_ctx *c, const double *A, const int32_t *i, const int32_t *j, int64_t nT)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (tiny_eligible(c)) return tiny_set_triplets(c, false, A, i, j, nT, false);
    int rc = assemble_triplets(c, c->M, A, i, j, nT, false);
    if (rc) return rc;
    return build_mass_csr(c);
}

int daecpp_b200_set_jacobian(daecpp_b200_ctx *c, const double *A, const int32_t *i, const int32_t *j, int64_t nT)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    c->direct_csc = false;
    if (tiny_eligible(c)) return tiny_set_triplets(c, true, A, i, j, nT, false);
    return assemble_triplets(c, c->J, A, i, j, nT, true);
}

int daecpp_b200_residual(daecpp_b200_ctx *c, const double *f_host, int32_t order, const double dt[4], double *alpha_out)
{
    if (!c || !f_host || !dt) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (int rcf = tiny_flush(c)) return rcf;
    if (!c->M.have_pattern) return fail(c, DAECPP_B200_EINVAL, "set_mass must precede residual");
    double w[5];
    if (!bdf_weights(order, dt, w)) return fail(c, DAECPP_B200_EINVAL, "unsupported time integration order");
    if (alpha_out) *alpha_out = w[0];
    int rc = DAECPP_B200_OK;
    if (c->pinned_io)
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->f, f_host, sizeof(double) * c->n, cudaMemcpyHostToDevice, c->stream));
    else
    {
        if ((rc = ensure_stage(c, c->n))) return rc;
        std::memcpy(c->h_stage, f_host, sizeof(double) * c->n);
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->f, c->h_stage, sizeof(double) * c->n, cudaMemcpyHostToDevice, c->stream));
    }
    rc = launch_residual(c, order, w);
    if (rc) return rc;
    // the stage buffer is reused by the next call
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    std::memcpy(c->h_scalars + 2, w, sizeof(w)); // remembered for get_vector(dxdt)
    c->h_flags[6] = order;
    return DAECPP_B200_OK;
}

int daecpp_b200_refactor(daecpp_b200_ctx *c, double alpha)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    int rc = tiny_flush(c);
    if (rc) return rc;
    if ((rc = build_union(c))) return rc;
    if (c->assemble_only) return assemble_values_only(c, alpha);
    c->alpha_hint = alpha;
    if (!c->have_symbolic && (rc = upload_symbolic(c))) return rc;
    return numeric_refactor(c, alpha);
}

int daecpp_b200_solve_update(daecpp_b200_ctx *c, double atol, double rtol, double max_err_abs, int32_t *flags)
{
    if (!c || !flags) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_factor) return fail(c, DAECPP_B200_EINVAL, "no valid factorisation");
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev0, c->stream));
    c->flags_clean = false;
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 2, 0, sizeof(int32_t), c->stream));
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 4, 0, sizeof(int32_t), c->stream));
    int rc = numeric_solve(c, true, atol, rtol, max_err_abs);
    if (rc) return rc;
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev1, c->stream));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 2, c->d_flags + 2, 3 * sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->ms_last_solve = ms;
    if (c->h_flags[4]) return fail(c, DAECPP_B200_ECUDA, "unit dependency wait timed out");
    if (wants_refinement(c, c->h_flags[2]) && (rc = refine_sync(c, atol, rtol, max_err_abs, true))) return rc;
    *flags = c->h_flags[2] & (DAECPP_B200_DIVERGED | DAECPP_B200_NOT_CONVERGED);
    return DAECPP_B200_OK;
}

int daecpp_b200_accept_step(daecpp_b200_ctx *c, int32_t variability_on, double tol, double *variability_out)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    c->refine_futile = false;
    if (tiny_eligible(c) && c->tiny.spec_valid && (!variability_on || tol == c->tiny.var_tol))
    {   // the tiny iteration measured the variability with its update: shift the history, no read-back
        c->tiny.spec_valid = false;
        int rct = launch_accept(c, 0, tol);
        if (rct) return rct;
        if (variability_out) *variability_out = variability_on ? c->tiny.spec_var : 0.0;
        return DAECPP_B200_OK;
    }
    if (tiny_eligible(c)) { c->tiny.var_tol = tol; c->tiny.spec_valid = false; }
    if (variability_on) DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_scalars, 0, sizeof(double), c->stream));
    int rc = launch_accept(c, variability_on, tol);
    if (rc) return rc;
    double v = 0.0;
    if (variability_on)
    {
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_scalars, c->d_scalars, sizeof(double), cudaMemcpyDeviceToHost, c->stream));
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
        v = c->h_scalars[0];
    }
    if (variability_out) *variability_out = v;
    return DAECPP_B200_OK;
}

// Device part of one Newton iteration, enqueued on the stream (no synchronisation).
static int enqueue_iteration(daecpp_b200_ctx *c, int32_t do_refactor, int32_t order, const double w[5], double atol,
                             double rtol, double max_err_abs)
{
    int rc;
    cudaStream_t st = c->stream;
    const bool pt = c->phase_timing;
    // all flag words of the iteration are cleared up front so that its kernels follow each other
    // on the stream with no other node in between (programmatic dependent launch, chain.cu)
    // chain path: the CTA that finishes the iteration writes the flag words to the mapped host mirror and
    // resets them -- no memset and no copy node around the kernels
    const bool lean = c->lean_flags && c->CH.active && c->have_symbolic;
    if (!lean || !c->flags_clean) DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 2, 0, 3 * sizeof(int32_t), st));
    c->flags_clean = false;
    // the triplet values are resident: the Jacobian's compressed-column values (if anything
    // still needs them) are stale from here on
    if (do_refactor) c->J.vals_valid = false;
    if (pt) cudaEventRecord(c->pev[0], st);
    if (do_refactor && !c->jvals_from_triplets && (rc = ensure_csc_values(c, c->J))) return rc; // duplicate-summing path
    if (pt) cudaEventRecord(c->pev[1], st);
    // diagonal mass on the chain path: b = M*dxdt - f is formed by the forward kernel while it stages its
    // right-hand side (same roundings), the residual kernel and the round trip of b through HBM go away
    const bool fuse_res = lean && (c->m_diag || c->m_pdiag) && c->chain_fuse_residual;
    c->residual_was_fused = fuse_res;
    if (fuse_res) c->bdf_pending = make_bdf(c, order, w);
    else if ((rc = launch_residual(c, order, w, false))) return rc;
    if (pt) cudaEventRecord(c->pev[2], st);
    // the solve follows on the same stream and the flags are read once at the end: a two-level chain
    // may leave its (tiny) top-level factorisation to an extra CTA of the forward launch
    c->chain_may_defer = true;
    c->rhs_prefetch_hint = fuse_res && !pt;
    if (do_refactor) rc = numeric_refactor_async(c, w[0], false);
    c->rhs_prefetch_hint = false;
    c->chain_may_defer = false;
    if (do_refactor && rc) return rc;
    if (pt) cudaEventRecord(c->pev[3], st);
    c->forward_follows_factor = do_refactor && lean && !pt;
    c->fuse_residual_now = fuse_res;
    c->publish_flags = lean; // (a pattern change inside the refactorisation may have left the chain path: then nothing is published)
    c->flags_published = false;
    rc = numeric_solve(c, true, atol, rtol, max_err_abs);
    c->publish_flags = false;
    c->fuse_residual_now = false;
    c->forward_follows_factor = false;
    if (rc) return rc;
    c->flags_clean = c->flags_published;
    if (pt) cudaEventRecord(c->pev[4], st);
    if (pt) cudaEventRecord(c->pev[5], st);
    if (!c->flags_published) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 2, c->d_flags + 2, 3 * sizeof(int32_t), cudaMemcpyDeviceToHost, st));
    return DAECPP_B200_OK;
}

// After the stream has been synchronised: phase timers, flags, status.
// The update failed the convergence test only on elements whose |dx| is below the refinement floor: there the
// reference's relative test (solver.hpp:603-609) is decided by the componentwise accuracy of the linear solve
// (Eigen::SparseLU with partial pivoting delivers it; the static-pivot block elimination is normwise accurate).
// One refinement step restores it; flags are recomputed from the corrected update.
static bool wants_refinement(const daecpp_b200_ctx *c, int32_t fl)
{
    return c->refine_steps > 0 && !c->direct_csc && !c->refine_futile && (fl & DAECPP_B200_NOT_CONVERGED) &&
           !(fl & (DAECPP_B200_ABOVE_FLOOR | DAECPP_B200_DIVERGED));
}
static int refine_sync(daecpp_b200_ctx *c, double atol, double rtol, double max_err_abs, bool b_is_stored)
{
    int rc = refine_last_solve(c, c->last_alpha, atol, rtol, max_err_abs, b_is_stored);
    if (rc) return rc;
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags + 2, c->d_flags + 2, sizeof(int32_t), cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    // the corrected update fails the same way: the noise is not the solve's, no second attempt within this Newton loop
    const int32_t fl = c->h_flags[2];
    c->refine_futile = (fl & DAECPP_B200_NOT_CONVERGED) && !(fl & (DAECPP_B200_ABOVE_FLOOR | DAECPP_B200_DIVERGED));
    return DAECPP_B200_OK;
}

static int finish_iteration(daecpp_b200_ctx *c, int32_t do_refactor, int32_t *flags, double atol, double rtol, double max_err_abs, int slot = 0)
{
    const volatile int32_t *hf = c->h_flags + 2 + 8 * slot; // {update flags, pivot failure, wait time-out} of this iteration
    int32_t f2 = hf[0];
    const int32_t f3 = hf[1], f4 = hf[2];
    if (f4) return fail(c, DAECPP_B200_ECUDA, "unit dependency wait timed out");
    if (f3 == 0 && wants_refinement(c, f2))
    {
        int rc = refine_sync(c, atol, rtol, max_err_abs, !c->residual_was_fused);
        if (rc) return rc;
        f2 = c->h_flags[2];
    }
    if (c->phase_timing)
    {
        for (int k = 0; k < 5; ++k)
        {
            float ms = 0;
            cudaEventElapsedTime(&ms, c->pev[k], c->pev[k + 1]);
            c->ms_phase[k] += ms;
        }
        c->n_phase_samples += 1;
    }
    *flags = f2 & (DAECPP_B200_DIVERGED | DAECPP_B200_NOT_CONVERGED);
    if (do_refactor)
    {
        c->have_factor = f3 == 0;
        if (!c->have_factor) return DAECPP_B200_EDECOMP;
    }
    return DAECPP_B200_OK;
}

// The host side of "flags ready".  When the iteration's last CTA published the flags to the mapped mirror
// (chain path) the host spins on the sequence word it wrote behind them -- a few hundred nanoseconds after
// the write instead of a driver round trip; everything else (and a long wait) synchronises the stream.
static int wait_iteration(daecpp_b200_ctx *c, int slot = 0, uint32_t want_seq = 0)
{
    if (c->flags_published && c->spin_flags && !c->phase_timing)
    {
        volatile int32_t *seq = c->h_flags + 5 + 8 * slot;
        const int32_t want = (int32_t)(want_seq ? want_seq : c->flag_seq);
        const auto t0 = std::chrono::steady_clock::now();
        for (uint32_t spins = 0; *seq != want; ++spins)
        {
            __builtin_ia32_pause();
            if ((spins & 1023u) == 1023u && std::chrono::steady_clock::now() - t0 > std::chrono::microseconds(500)) break;
        }
        if (*seq == want) return DAECPP_B200_OK;
    }
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return DAECPP_B200_OK;
}

static int prepare_iteration(daecpp_b200_ctx *c, int32_t do_refactor, int32_t order, const double dt[4], double w[5])
{
    if (int rcf = tiny_flush(c)) return rcf;
    if (!c->M.have_pattern || !c->J.have_pattern) return fail(c, DAECPP_B200_EINVAL, "set_mass/set_jacobian must precede");
    if (!bdf_weights(order, dt, w)) return fail(c, DAECPP_B200_EINVAL, "unsupported time integration order");
    int rc = build_union(c);
    if (rc) return rc;
    c->alpha_hint = w[0];
    if (!c->have_symbolic && (rc = upload_symbolic(c))) return rc;
    if (!do_refactor && !c->have_factor) return fail(c, DAECPP_B200_EINVAL, "no valid factorisation");
    return DAECPP_B200_OK;
}

int daecpp_b200_iteration_resident(daecpp_b200_ctx *c, int32_t do_refactor, int32_t order, const double dt[4],
                                   double atol, double rtol, double max_err_abs, int32_t *flags)
{
    if (!c || !dt || !flags) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    double w[5];
    int rc = prepare_iteration(c, do_refactor, order, dt, w);
    if (rc) return rc;
    // every entry point returns with its stream work done: nothing in front of this iteration's kernels writes the matrix values
    struct Quiescent { daecpp_b200_ctx *c; ~Quiescent() { c->values_quiescent = false; } } q{c};
    c->values_quiescent = true;
    if ((rc = enqueue_iteration(c, do_refactor, order, w, atol, rtol, max_err_abs))) return rc;
    if ((rc = wait_iteration(c))) return rc;
    return finish_iteration(c, do_refactor, flags, atol, rtol, max_err_abs);
}

int daecpp_b200_iterations_resident(daecpp_b200_ctx *c, int32_t count, int32_t do_refactor, int32_t order, const double dt[4],
                                    double atol, double rtol, double max_err_abs, int32_t *flags)
{
    if (!c || !dt || !flags || count < 1) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    double w[5];
    int rc = prepare_iteration(c, do_refactor, order, dt, w);
    if (rc) return rc;
    // Chain path: iteration k+1 is enqueued while k runs, so the device goes from one to the next without waiting for the
    // host's spin-and-launch turn-around (its kernels start behind k's by programmatic dependent launch).  Every iteration
    // still publishes its own flag words (two mirror slots, alternating) and the host still looks at each; where the host
    // has to act between two iterations -- a failed pivot, a time-out, an update it would refine -- the device raises the
    // cancel word and the queued iteration leaves without touching anything, to be enqueued again afterwards.
    struct Quiescent { daecpp_b200_ctx *c; ~Quiescent() { c->values_quiescent = false; } } q{c};
    c->values_quiescent = true; // (as in daecpp_b200_iteration_resident; between iterations only the iterations' own kernels run)
    const bool pipelined = count > 1 && c->pipeline_iterations && c->lean_flags && c->CH.active && c->have_symbolic && c->spin_flags &&
                           !c->phase_timing && !tiny_eligible(c);
    if (!pipelined)
    {
        for (int32_t k = 0; k < count; ++k)
        {
            if (k && (rc = prepare_iteration(c, do_refactor, order, dt, w))) return rc;
            if ((rc = enqueue_iteration(c, do_refactor, order, w, atol, rtol, max_err_abs))) return rc;
            if ((rc = wait_iteration(c))) return rc;
            if ((rc = finish_iteration(c, do_refactor, flags, atol, rtol, max_err_abs))) return rc;
        }
        return DAECPP_B200_OK;
    }
    cudaStream_t st = c->stream;
    auto disarm = [&](int code) { c->cancel_armed = false; c->publish_slot = 0; return code; };
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 5, 0, sizeof(int32_t), st));
    c->cancel_armed = true;
    uint32_t seq_of[2] = {0, 0};
    bool queued = false; // iteration k is already on the stream
    for (int32_t k = 0; k < count; ++k)
    {
        const int slot = k & 1;
        if (!queued)
        {
            c->publish_slot = slot;
            if ((rc = enqueue_iteration(c, do_refactor, order, w, atol, rtol, max_err_abs))) return disarm(rc);
            if (!c->flags_published)
            {   // (left the chain path: nothing was published, finish this one the plain way and go on unpipelined)
                c->cancel_armed = false; c->publish_slot = 0;
                if ((rc = wait_iteration(c))) return rc;
                if ((rc = finish_iteration(c, do_refactor, flags, atol, rtol, max_err_abs))) return rc;
                return k + 1 < count ? daecpp_b200_iterations_resident(c, count - k - 1, do_refactor, order, dt, atol, rtol, max_err_abs, flags) : DAECPP_B200_OK;
            }
            seq_of[slot] = c->flag_seq;
        }
        queued = false;
        const double keep_counts[3] = {c->n_refactor, c->n_solve, c->kernels};
        if (k + 1 < count)
        {   // the next one goes behind it now
            c->publish_slot = slot ^ 1;
            if ((rc = enqueue_iteration(c, do_refactor, order, w, atol, rtol, max_err_abs))) return disarm(rc);
            queued = c->flags_published;
            seq_of[slot ^ 1] = c->flag_seq;
            if (!queued) { DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st)); return disarm(fail(c, DAECPP_B200_EINVAL, "pattern changed inside a resident iteration block")); }
        }
        if ((rc = wait_iteration(c, slot, seq_of[slot]))) return disarm(rc);
        const volatile int32_t *hf = c->h_flags + 2 + 8 * slot;
        const bool host_turn = hf[1] || hf[2] || wants_refinement(c, hf[0]);
        if (host_turn && queued)
        {   // the queued iteration was cancelled on the device: let it drain, re-arm, and do what the host has to do
            DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
            DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags + 5, 0, sizeof(int32_t), st));
            c->flags_clean = true; // (k's publisher reset the words; the cancelled iteration raised none)
            c->n_refactor = keep_counts[0]; c->n_solve = keep_counts[1]; c->kernels = keep_counts[2]; // it did not happen
            queued = false;
        }
        if ((rc = finish_iteration(c, do_refactor, flags, atol, rtol, max_err_abs, slot)))
        {
            if (queued) cudaStreamSynchronize(st);
            return disarm(rc);
        }
    }
    return disarm(DAECPP_B200_OK);
}

int daecpp_b200_set_jacobian_values(daecpp_b200_ctx *c, const double *A, int64_t nT)
{
    if (!c || (nT > 0 && !A)) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (tiny_eligible(c)) return tiny_set_triplets(c, true, A, nullptr, nullptr, nT, true);
    TripletSet &J = c->J;
    if (!J.have_pattern || J.nT_cached != nT) return fail(c, DAECPP_B200_EINVAL, "set_jacobian_values: no cached pattern of this length");
    if (nT > 0) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(J.A.p, A, sizeof(double) * nT, cudaMemcpyHostToDevice, c->stream));
    J.nT = nT;
    J.vals_valid = false;
    c->have_factor = false;
    return DAECPP_B200_OK;
}

int daecpp_b200_iteration_host(daecpp_b200_ctx *c, const double *f_host, int32_t do_refactor, int32_t order,
                               const double dt[4], double atol, double rtol, double max_err_abs, double *xk_host,
                               double *alpha_out, int32_t *flags)
{
    if (!c || !dt || !flags || !f_host) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    double w[5];
    if (tiny_eligible(c) && c->tiny.haveJ && c->tiny.haveM && !c->phase_timing)
    {   // one launch, mapped hand-off, no synchronisation (tiny.cu)
        if (!bdf_weights(order, dt, w)) return fail(c, DAECPP_B200_EINVAL, "unsupported time integration order");
        if (alpha_out) *alpha_out = w[0];
        return tiny_iteration(c, f_host, do_refactor, order, w, atol, rtol, max_err_abs, xk_host, flags);
    }
    int rc = prepare_iteration(c, do_refactor, order, dt, w);
    if (rc) return rc;
    if (alpha_out) *alpha_out = w[0];
    cudaStream_t st = c->stream;
    const size_t bytes = sizeof(double) * (size_t)c->n;
    if (c->pinned_io)
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->f, f_host, bytes, cudaMemcpyHostToDevice, st));
    else
    {
        if ((rc = ensure_stage(c, c->n))) return rc;
        std::memcpy(c->h_stage, f_host, bytes);
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->f, c->h_stage, bytes, cudaMemcpyHostToDevice, st));
    }
    // small systems handed over in pinned memory: the backward kernel stores xk into the caller's buffer itself
    c->xk_host_target = nullptr;
    c->xk_direct_used = false;
    if (xk_host && c->pinned_io && c->host_xk_direct && c->n <= c->host_xk_direct_max_n && !c->phase_timing)
    {
        if (c->xk_host_cached != xk_host)
        {
            void *dp = nullptr;
            if (cudaHostGetDevicePointer(&dp, xk_host, 0) == cudaSuccess) { c->xk_host_cached = xk_host; c->xk_host_dev_cached = static_cast<double *>(dp); }
            else { cudaGetLastError(); c->xk_host_cached = nullptr; c->xk_host_dev_cached = nullptr; }
        }
        if (c->xk_host_cached == xk_host) c->xk_host_target = c->xk_host_dev_cached;
    }
    rc = enqueue_iteration(c, do_refactor, order, w, atol, rtol, max_err_abs);
    c->xk_host_target = nullptr;
    if (rc) return rc;
    const bool direct = c->xk_direct_used && c->flags_published;
    if (xk_host && !direct)
    {
        if (c->pinned_io)
            DAECPP_CUDA_TRY(c, cudaMemcpyAsync(xk_host, c->xk, bytes, cudaMemcpyDeviceToHost, st));
        else
            DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_stage, c->xk, bytes, cudaMemcpyDeviceToHost, st));
    }
    if (direct) { if ((rc = wait_iteration(c))) return rc; }
    else DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
    std::memcpy(c->h_scalars + 2, w, sizeof(w));
    c->h_flags[6] = order;
    const double refined_before = c->n_refine;
    rc = finish_iteration(c, do_refactor, flags, atol, rtol, max_err_abs);
    if (xk_host && c->n_refine != refined_before && (rc == DAECPP_B200_OK))
    {   // the refinement step moved the iterate after it had been handed over
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->pinned_io ? xk_host : c->h_stage, c->xk, bytes, cudaMemcpyDeviceToHost, st));
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
    }
    if (xk_host && !c->pinned_io) std::memcpy(xk_host, c->h_stage, bytes);
    return rc;
}

int daecpp_b200_event_mark(daecpp_b200_ctx *c)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev_mark, c->stream));
    return DAECPP_B200_OK;
}

int daecpp_b200_event_elapsed_ms(daecpp_b200_ctx *c, double *ms_out)
{
    if (!c || !ms_out) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev1, c->stream));
    DAECPP_CUDA_TRY(c, cudaEventSynchronize(c->ev1));
    float ms = 0;
    DAECPP_CUDA_TRY(c, cudaEventElapsedTime(&ms, c->ev_mark, c->ev1));
    *ms_out = ms;
    return DAECPP_B200_OK;
}

int daecpp_b200_set_rhs(daecpp_b200_ctx *c, const double *b_host)
{
    if (!c || !b_host) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    return upload_vector(c, c->b, b_host);
}

int daecpp_b200_solve(daecpp_b200_ctx *c, double *dx_host)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (!c->have_factor) return fail(c, DAECPP_B200_EINVAL, "no valid factorisation");
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev0, c->stream));
    int rc = numeric_solve(c);
    if (rc) return rc;
    // the linear-solver entry has no Newton loop behind it to absorb an inaccurate solve: one refinement step
    // (r = b - Jb*dx in FP64 with fused multiply-adds, Jb*e = r, dx += e) always runs here.  Eigen::SparseLU reaches
    // the same accuracy by pivoting (Eigen/src/SparseLU/SparseLU_pivotL.h:62-131).
    if (c->refine_steps > 0 && !c->direct_csc && c->M.have_pattern && c->J.have_pattern &&
        (rc = refine_last_solve(c, c->last_alpha, 0, 0, 0, true, false)))
        return rc;
    DAECPP_CUDA_TRY(c, cudaEventRecord(c->ev1, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->ev0, c->ev1);
    c->ms_last_solve = ms;
    // a factorisation that went through without a flagged pivot can still deliver garbage (growth): a solution that
    // is not finite is reported the way the reference reports a failed solve (solver.hpp:562-572)
    if (dx_host)
    {
        rc = download_vector(c, dx_host, c->dx);
        if (rc) return rc;
        for (int32_t k = 0; k < c->n; ++k)
            if (!std::isfinite(dx_host[k])) return fail(c, DAECPP_B200_ESOLVE, "the solution is not finite");
        return DAECPP_B200_OK;
    }
    return DAECPP_B200_OK;
}

int daecpp_b200_get_csc(daecpp_b200_ctx *c, int32_t which, int32_t *outer, int32_t *inner, double *vals, int64_t cap,
                        int64_t *nnz_out)
{
    if (!c) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    if (int rcf = tiny_flush(c)) return rcf;
    const int32_t *d_outer, *d_inner; const double *d_vals; int64_t nnz;
    if (which == 0 || which == 1)
    {
        TripletSet &T = which == 0 ? c->M : c->J;
        if (!T.have_pattern) return fail(c, DAECPP_B200_EINVAL, "matrix not set");
        int rc = ensure_csc_values(c, T);
        if (rc) return rc;
        d_outer = T.outer.p; d_inner = T.inner.p; d_vals = T.vals.p; nnz = T.nnz;
    }
    else if (which == 2)
    {
        if (c->n_refactor == 0) return fail(c, DAECPP_B200_EINVAL, "refactor has not run");
        int rc = build_union(c); // (tiny systems iterate without ever building it)
        if (rc) return rc;
        rc = jb_values(c, c->last_alpha); // computed on demand: the factorisation assembles straight into the fronts
        if (rc) return rc;
        d_outer = c->u_outer.p; d_inner = c->u_inner.p; d_vals = c->u_vals.p; nnz = c->nnzU;
    }
    else
        return fail(c, DAECPP_B200_EINVAL, "unknown matrix selector");
    if (nnz_out) *nnz_out = nnz;
    if (outer) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(outer, d_outer, sizeof(int32_t) * ((size_t)c->n + 1), cudaMemcpyDeviceToHost, c->stream));
    const int64_t m = nnz < cap ? nnz : cap;
    if (inner && m > 0) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(inner, d_inner, sizeof(int32_t) * m, cudaMemcpyDeviceToHost, c->stream));
    if (vals && m > 0) DAECPP_CUDA_TRY(c, cudaMemcpyAsync(vals, d_vals, sizeof(double) * m, cudaMemcpyDeviceToHost, c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    return DAECPP_B200_OK;
}

int daecpp_b200_get_vector(daecpp_b200_ctx *c, int32_t which, double *host)
{
    if (!c || !host) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    const double *src = nullptr;
    switch (which)
    {
    case 0: src = c->b; break;
    case 1: src = c->dx; break;
    case 2:
    {
        int rc = launch_dxdt(c, c->h_flags[6] ? c->h_flags[6] : 1, c->h_scalars + 2);
        if (rc) return rc;
        src = c->dxdt;
        break;
    }
    case 3: src = c->xk; break;
    case 4: case 5: case 6: case 7: src = c->xh[which - 4]; break;
    default: return fail(c, DAECPP_B200_EINVAL, "unknown vector selector");
    }
    return download_vector(c, host, src);
}

int daecpp_b200_set_vector(daecpp_b200_ctx *c, int32_t which, const double *host)
{
    if (!c || !host) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    double *dst = nullptr;
    switch (which)
    {
    case 0: dst = c->b; break;
    case 1: dst = c->dx; break;
    case 3: dst = c->xk; break;
    case 4: case 5: case 6: case 7: dst = c->xh[which - 4]; break;
    default: return fail(c, DAECPP_B200_EINVAL, "unknown vector selector");
    }
    return upload_vector(c, dst, host);
}

int daecpp_b200_debug_unit_trace(daecpp_b200_ctx *c, int64_t *out, int64_t cap, int64_t *count)
{
    if (!c || !count) return DAECPP_B200_EINVAL;
    cudaSetDevice(c->device);
    // subtree units: factor, forward, backward x 4 stamps; chain path: 5 kernels x tiles x 8 stamps (chain.cu)
    const int64_t m = c->CH.active ? (int64_t)5 * c->CH.levels[0].ntiles * 8 : (int64_t)c->S.nunit * 12 + 4 + 256; // (+ the per-level stamps of one unit)
    *count = m;
    if (!c->trace.p || !out) return DAECPP_B200_OK;
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    DAECPP_CUDA_TRY(c, cudaMemcpy(out, c->trace.p, sizeof(int64_t) * (m < cap ? m : cap), cudaMemcpyDeviceToHost));
    return DAECPP_B200_OK;
}

int daecpp_b200_get_stat(daecpp_b200_ctx *c, const char *key, double *value)
{
    if (!c || !key || !value) return DAECPP_B200_EINVAL;
    const std::string k(key);
    if (k == "n_symbolic") *value = c->n_symbolic;
    else if (k == "n_refactor") *value = c->n_refactor;
    else if (k == "n_solve") *value = c->n_solve;
    else if (k == "n_refine") *value = c->n_refine;
    else if (k == "n_cold_assembly") *value = c->n_cold;
    else if (k == "kernels_launched") *value = c->kernels;
    else if (k == "ms_symbolic") *value = c->ms_symbolic;
    else if (k == "ms_last_refactor") *value = c->ms_last_refactor;
    else if (k == "ms_last_solve") *value = c->ms_last_solve;
    else if (k == "chain_block") *value = c->CH.active ? c->CH.B : 0;
    else if (k == "chain_levels") *value = c->CH.active ? (double)c->CH.levels.size() : 0;
    else if (k == "chain_tiles") *value = c->CH.active ? c->CH.levels[0].ntiles : 0;
    else if (k == "chain_identity_perm") *value = c->CH.active && c->CH.identity_perm ? 1 : 0;
    else if (k == "n_fronts") *value = c->CH.active ? c->CH.m : c->S.nsn;
    else if (k == "n_levels") *value = c->CH.active ? 3.0 * (double)c->CH.levels.size() : c->S.nlevels;
    else if (k == "nnz_factor") *value = c->CH.active ? (double)c->CH.nnz_factor : (double)c->S.nnz_factor;
    else if (k == "lu_doubles" && c->CH.active) *value = (double)c->CH.fac_doubles;
    else if (k == "schedule_bytes" && c->CH.active) *value = 12.0 * (double)c->CH.ent_u.size();
    else if (k == "factor_flops") *value = c->S.flops;
    else if (k == "max_front") *value = c->S.max_front;
    else if (k == "lu_doubles") *value = (double)c->S.lu_size;
    else if (k == "cb_doubles") *value = (double)c->S.cb_size;
    else if (k == "nnz_jb") *value = (double)c->nnzU;
    else if (k == "n_missing_diag") *value = c->S.n_missing_diag;
    else if (k == "n_rows_moved") *value = c->S.n_rows_moved;
    else if (k == "ms_phase_assembly") *value = c->n_phase_samples ? c->ms_phase[0] / c->n_phase_samples : 0;
    else if (k == "ms_phase_residual") *value = c->n_phase_samples ? c->ms_phase[1] / c->n_phase_samples : 0;
    else if (k == "ms_phase_refactor") *value = c->n_phase_samples ? c->ms_phase[2] / c->n_phase_samples : 0;
    else if (k == "ms_phase_solve") *value = c->n_phase_samples ? c->ms_phase[3] / c->n_phase_samples : 0;
    else if (k == "ms_phase_update") *value = c->n_phase_samples ? c->ms_phase[4] / c->n_phase_samples : 0;
    else if (k == "n_phase_samples") *value = c->n_phase_samples;
    else if (k == "residual_fused") *value = (c->lean_flags && c->CH.active && (c->m_diag || c->m_pdiag) && c->chain_fuse_residual) ? 1 : 0;
    else if (k == "jvals_from_triplets") *value = c->jvals_from_triplets ? 1 : 0;
    else if (k == "upd_doubles") *value = (double)c->S.upd_size;
    else if (k == "n_units") *value = c->S.nunit;
    else if (k == "n_stages") *value = c->S.nstages;
    else if (k == "n_big_fronts") *value = c->S.nbig;
    else if (k == "n_launch_groups") *value = (double)c->S.groups.size();
    else if (k == "factor_smem_bytes") *value = c->S.factor_smem_bytes;
    else if (k == "solve_smem_bytes") *value = c->S.solve_smem_bytes;
    else if (k == "unit_grid_factor") *value = c->unit_grid_factor;
    else if (k == "unit_grid_solve") *value = c->unit_grid_solve;
    else if (k == "schedule_bytes") *value = 4.0 * (double)(c->S.blob.size() + 3 * c->S.ent_u.size());
    else return fail(c, DAECPP_B200_EINVAL, "unknown stat key");
    return DAECPP_B200_OK;
}

int daecpp_b200_set_option(daecpp_b200_ctx *c, const char *key, double value)
{
    if (!c || !key) return DAECPP_B200_EINVAL;
    const std::string k(key);
    if (k == "refine_steps") c->refine_steps = (int)value;
    else if (k == "refine_floor") c->refine_floor = value;
    else if (k == "pivot_tol") c->pivot_tol = value;
    else if (k == "assemble_only") c->assemble_only = value != 0.0;
    else if (k == "phase_timing")
    {
        c->phase_timing = value != 0.0;
        for (int q = 0; q < 5; ++q) c->ms_phase[q] = 0;
        c->n_phase_samples = 0;
    }
    else if (k == "leaf_size") { c->sopt.leaf_size = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "unit_smem_bytes") { c->sopt.unit_smem_bytes = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "group8_dim") { c->sopt.group8_dim = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "warp_dim") { c->sopt.warp_dim = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "warp_dim_many") { c->sopt.warp_dim_many = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "warp_min_fronts") { c->sopt.warp_min_fronts = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "serial_dim") { c->sopt.serial_dim = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "unit_threads")
    {
        if (value != 0 && (value < 32 || value > 256 || ((int)value & 31))) return fail(c, DAECPP_B200_EINVAL, "unit_threads: 0 (automatic) or a multiple of 32 up to 256");
        c->unit_threads = (int)value; c->have_symbolic = false; c->have_factor = false;
    }
    else if (k == "pinned_io") c->pinned_io = value != 0.0;
    else if (k == "chain") { c->chain_enabled = value != 0.0; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "lean_flags") c->lean_flags = value != 0.0;
    else if (k == "spin_flags") c->spin_flags = value != 0.0;
    else if (k == "host_xk_direct") c->host_xk_direct = value != 0.0;
    else if (k == "host_xk_direct_max_n") c->host_xk_direct_max_n = (int64_t)value;
    else if (k == "chain_fuse_residual") c->chain_fuse_residual = value != 0.0;
    else if (k == "keep_b") c->keep_b = value != 0.0;
    else if (k == "chain_prefetch") c->chain_prefetch = (int)value;
    else if (k == "chain_packed") c->chain_packed = (int)value;
    else if (k == "chain_affine") { c->chain_affine = value != 0.0; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "blocked_lu") c->blocked_lu = value != 0.0;
    else if (k == "blocked_lu_min_dim") c->blocked_lu_min_dim = (int)value;
    else if (k == "chain_fuse_top") c->chain_fuse_top = (int)value;
    else if (k == "chain_top_cta") c->chain_top_cta = value != 0.0;
    else if (k == "bf_fused_panel") c->bf_fused_panel = value != 0.0;
    else if (k == "bf_dmma") c->bf_dmma = value != 0.0;
    else if (k == "bf_schur") c->bf_schur = value != 0.0;
    else if (k == "bf_overlap_units") c->bf_overlap_units = value != 0.0;
    else if (k == "bf_zero_ahead") c->bf_zero_ahead = value != 0.0;
    else if (k == "bf_schur_fronts") c->bf_schur_fronts = (int)value;
    else if (k == "bf_fast_solve") c->bf_fast_solve = value != 0.0;
    else if (k == "bf_diag_copy") c->bf_diag_copy = value != 0.0;
    else if (k == "bf_lu_lookahead") c->bf_lu_lookahead = value != 0.0;
    else if (k == "bf_pdl") c->bf_pdl = value != 0.0;
    else if (k == "chain_defer_top") c->chain_defer_top = value != 0.0;
    else if (k == "chain_top_pcr") c->chain_top_pcr = value != 0.0;
    else if (k == "chain_min_n") { c->chain_min_n = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "chain_max_block") { c->chain_max_b = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "trace_units") c->trace_units = value != 0.0;
    else if (k == "trace_unit_index") c->trace_unit_index = (int)value;
    else if (k == "transversal") { c->sopt.transversal = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else if (k == "pivot_rel") c->pivot_rel = value;
    else if (k == "pipeline_iterations") c->pipeline_iterations = value != 0.0;
    else if (k == "tiny") c->tiny.enabled = value != 0.0;
    else if (k == "max_super") { c->sopt.max_super = (int)value; c->have_symbolic = false; c->have_factor = false; }
    else return fail(c, DAECPP_B200_EINVAL, "unknown option key");
    return DAECPP_B200_OK;
}

} // extern "C"
