// Subtree-unit kernels: numeric LU refactorisation and triangular solves of whole subtrees of
// the assembly tree out of shared memory (schedule.hpp, unit_ops.hpp).
//
// Role replaced: Eigen::SparseLU::factorize / _solve_impl behind linsolver.compute(Jb) and
// linsolver.solve(b) (dae-cpp/solver.hpp:537,558), plus -- fused into the same kernels -- the
// assembly `Jb = J - M*alpha` (solver.hpp:523), the permutations and the error check / update
// loop (solver.hpp:590-619).
//
// One persistent thread block takes units by ticket (children first).  Per unit:
//   1. the unit's integer blob is copied to shared memory (coalesced 16-byte loads);
//   2. factor: the frontal arena is zeroed and the matrix entries J - (M*alpha) are scattered
//      into it straight from the triplet values (bit-exact with the reference's assembly);
//      solve: the unit's factor panels and right-hand-side entries are staged;
//   3. the unit waits until the units below it have published their contribution blocks
//      (dependency counter, bounded spin);
//   4. local levels are swept with __syncthreads between them; a front is handled by 1, 8, 32
//      or all threads depending on the level's largest front;
//   5. factor panels are written once to HBM; only root fronts export a contribution block.
#include "ctx.hpp"
#include <mutex>
#include "unit_ops.hpp"

namespace daecpp_b200
{
namespace
{
using namespace sched;

struct PSerial
{
    __device__ __forceinline__ int lane() const { return 0; }
    __device__ __forceinline__ int width() const { return 1; }
    __device__ __forceinline__ void sync() const {}
};
struct PGroup
{
    int l, w;
    unsigned mask;
    __device__ __forceinline__ int lane() const { return l; }
    __device__ __forceinline__ int width() const { return w; }
    __device__ __forceinline__ void sync() const { __syncwarp(mask); }
};
struct PBlock
{
    __device__ __forceinline__ int lane() const { return threadIdx.x; }
    __device__ __forceinline__ int width() const { return blockDim.x; }
    __device__ __forceinline__ void sync() const { __syncthreads(); }
};

struct UnitArgs
{
    const int32_t *blob, *unit_blob;
    int32_t begin, count;
    uint32_t *ticket;
    uint32_t ticket_base;
    int32_t *dep;   // children-finished counters, indexed by unit
    int32_t *done;  // backward sweep: epoch at which a unit's unknowns became final
    int32_t epoch;
    int32_t *flags; // [3] pivot failure, [4] dependency wait timed out
    long long *trace; // optional: 4 globaltimer stamps per unit (debug)
    long long *trace_levels; // optional: (time, cooperation width, fronts, largest order) per local level of unit trace_unit (debug)
    int32_t trace_unit;
    // factor
    const int32_t *ent_sj, *ent_sm, *ent_dst;
    const double *jvals, *mvals;
    double alpha, pivot_tol;
    int32_t serial_dim;
    double *LU, *CB;
    // solve
    const int32_t *perm; // nullptr: rhs / solution already in elimination order
    const int32_t *perm_rhs; // forward sweep: row of the right-hand side per elimination position (ctx.hpp: DeviceSymbolic)
    const double *rhs;
    double *y, *upd, *out;
    // fused update (solver.hpp:590-619)
    double *xk;
    const double *x0;
    double atol, rtol, max_err_abs, floor;
};

constexpr int SPIN_LIMIT = 1 << 22;

__device__ __forceinline__ void stamp(const UnitArgs &A, int32_t unit, int k)
{
    if (A.trace && threadIdx.x == 0)
    {
        long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        A.trace[(int64_t)unit * 4 + k] = t;
    }
}

__device__ __forceinline__ int32_t ld_volatile(const int32_t *p) { return *(const volatile int32_t *)p; }

// Take the next unit.  Returns false when the launch's units are exhausted.
__device__ __forceinline__ bool next_ticket(const UnitArgs &A, uint32_t *s_t, uint32_t &t)
{
    __syncthreads(); // the previous unit is done with shared memory and with *s_t
    if (threadIdx.x == 0) *s_t = atomicAdd(A.ticket, 1u) - A.ticket_base;
    __syncthreads();
    t = *s_t;
    return t < (uint32_t)A.count;
}

// Asynchronous global -> shared copy of the unit's blob (LDGSTS, 16 bytes per request, L2 only):
// the issuing warps go on to the next phase's loads instead of waiting on registers.
__device__ __forceinline__ void load_blob_async(const int32_t *Bg, int32_t *B, int nints)
{
    const uint32_t dst0 = (uint32_t)__cvta_generic_to_shared(B);
    for (int i = threadIdx.x; i < nints / 4; i += blockDim.x)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst0 + 16u * i), "l"(Bg + 4 * i) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
}
__device__ __forceinline__ void async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__device__ __forceinline__ void wait_counter(const UnitArgs &A, const int32_t *word, int32_t target)
{
    if (threadIdx.x == 0)
    {
        int spins = 0;
        while (ld_volatile(word) != target)
        {
            if (++spins > SPIN_LIMIT) { atomicOr(&A.flags[4], 1); break; }
            __nanosleep(64);
        }
        __threadfence();
    }
    __syncthreads();
}

// Sweep one local level: dispatch on the level's cooperation width.
#define FOR_LEVEL_FRONTS(G, fb, fe, CALL)                                                                       \
    if ((G) == 1)                                                                                               \
    {                                                                                                           \
        PSerial P;                                                                                              \
        for (int32_t fi = (fb) + (int32_t)threadIdx.x; fi < (fe); fi += (int32_t)blockDim.x) { CALL; }          \
    }                                                                                                           \
    else if ((G) == 0)                                                                                          \
    {                                                                                                           \
        PBlock P;                                                                                               \
        for (int32_t fi = (fb); fi < (fe); ++fi) { CALL; }                                                      \
    }                                                                                                           \
    else                                                                                                        \
    {                                                                                                           \
        PGroup P;                                                                                               \
        P.w = (G); P.l = (int)threadIdx.x % (G);                                                                \
        P.mask = (G) == 32 ? 0xffffffffu : (((1u << (G)) - 1u) << ((((int)threadIdx.x & 31) / (G)) * (G)));     \
        const int32_t ng = (int32_t)blockDim.x / (G);                                                           \
        for (int32_t fi = (fb) + (int32_t)threadIdx.x / (G); fi < (fe); fi += ng) { CALL; }                     \
    }

// (register budget: three resident CTAs of 256 threads per SM, what the units' shared memory allows on C4)
__global__ void __launch_bounds__(256, 3) k_unit_factor(UnitArgs A)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t s_t;
    uint32_t t;
    while (next_ticket(A, &s_t, t))
    {
        const int32_t unit = A.begin + (int32_t)t;
        const int32_t *Bg = A.blob + A.unit_blob[unit];
        const int nints = __ldg(Bg + H_NINTS);
        const int nw = __ldg(Bg + H_WDOUBLES);
        int32_t *B = reinterpret_cast<int32_t *>(smem);
        double *W = reinterpret_cast<double *>(smem + (size_t)nints * 4);
        stamp(A, unit, 0);
        load_blob_async(Bg, B, nints);
        for (int i = threadIdx.x; i < nw; i += blockDim.x) W[i] = 0.0;
        __syncthreads();
        {   // Jb = J - (M*alpha) scattered into the frontal arena (two roundings, no FMA: solver.hpp:523)
            const int32_t e0 = __ldg(Bg + H_ENT0), ne = __ldg(Bg + H_ENTCOUNT);
            for (int32_t e = threadIdx.x; e < ne; e += blockDim.x)
            {
                const int32_t sj = __ldg(A.ent_sj + e0 + e), sm = __ldg(A.ent_sm + e0 + e);
                const double j = sj >= 0 ? A.jvals[sj] : 0.0;
                const double m = sm >= 0 ? __dmul_rn(A.mvals[sm], A.alpha) : 0.0;
                W[__ldg(A.ent_dst + e0 + e)] = __dsub_rn(j, m);
            }
        }
        async_wait_all();
        __syncthreads();
        stamp(A, unit, 1);
        if (B[H_NDEP] > 0)
        {
            wait_counter(A, A.dep + unit, B[H_NDEP]);
            if (threadIdx.x == 0) A.dep[unit] = 0;
        }
        else
            __syncthreads();
        stamp(A, unit, 2);
        const int32_t nlev = B[H_NLEV];
        const int32_t *lp = B + B[H_LEVELS], *lg = lp + nlev + 1;
        const int32_t *fronts = B + B[H_FRONTS];
        double *LUu = A.LU + (uint32_t)B[H_LU0];
        int fail = 0;
        for (int32_t l = 0; l < nlev; ++l)
        {
            const int32_t G = lg[l], fb = lp[l], fe = lp[l + 1];
            if (G == 1 && A.serial_dim <= 4)
            {
                for (int32_t fi = fb + (int32_t)threadIdx.x; fi < fe; fi += (int32_t)blockDim.x)
                    unit_ops::factor_front_small<4>(B, fronts + fi * F_SIZE, W, LUu, A.CB, A.pivot_tol, &fail);
            }
            else
            {
                FOR_LEVEL_FRONTS(G, fb, fe,
                                 unit_ops::factor_front(P, B, fronts + (int64_t)fi * F_SIZE, W, LUu, A.CB, A.pivot_tol, &fail));
            }
            __syncthreads();
            if (A.trace_levels && unit == A.trace_unit && threadIdx.x == 0 && l < 60)
            {
                long long now;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
                int32_t md = 0;
                for (int32_t fi = fb; fi < fe; ++fi) md = max(md, fronts[fi * F_SIZE + F_S] + fronts[fi * F_SIZE + F_R]);
                A.trace_levels[4 * l + 0] = now; A.trace_levels[4 * l + 1] = G; A.trace_levels[4 * l + 2] = fe - fb; A.trace_levels[4 * l + 3] = md;
            }
        }
        if (fail) atomicOr(&A.flags[3], 1);
        stamp(A, unit, 3);
        const int32_t parent = B[H_PARENT];
        if (parent >= 0)
        {
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) atomicAdd(A.dep + parent, 1);
        }
    }
}

__global__ void k_unit_forward(UnitArgs A)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t s_t;
    uint32_t t;
    while (next_ticket(A, &s_t, t))
    {
        const int32_t unit = A.begin + (int32_t)t;
        const int32_t *Bg = A.blob + A.unit_blob[unit];
        const int nints = __ldg(Bg + H_NINTS);
        const int ny = __ldg(Bg + H_YDOUBLES), nu = __ldg(Bg + H_UDOUBLES), nlu = __ldg(Bg + H_LUCOUNT);
        int32_t *B = reinterpret_cast<int32_t *>(smem);
        double *Y = reinterpret_cast<double *>(smem + (size_t)nints * 4);
        double *U = Y + ny;
        double *LUs = U + nu;
        stamp(A, unit, 0);
        load_blob_async(Bg, B, nints);
        for (int i = threadIdx.x; i < nu; i += blockDim.x) U[i] = 0.0;
        {
            const double *src = A.LU + (uint32_t)__ldg(Bg + H_LU0);
            for (int i = threadIdx.x; i < nlu; i += blockDim.x) LUs[i] = src[i];
        }
        async_wait_all();
        __syncthreads();
        const int32_t nf = B[H_NF];
        const int32_t *fronts = B + B[H_FRONTS];
        for (int32_t fi = threadIdx.x; fi < nf; fi += blockDim.x)
        {
            const int32_t *F = fronts + (int64_t)fi * F_SIZE;
            for (int32_t k = 0; k < F[F_S]; ++k)
            {
                const int32_t col = F[F_COL0] + k;
                Y[F[F_YOFF] + k] = A.rhs[A.perm_rhs ? A.perm_rhs[col] : col];
            }
        }
        stamp(A, unit, 1);
        if (B[H_NDEP] > 0)
        {
            wait_counter(A, A.dep + unit, B[H_NDEP]);
            if (threadIdx.x == 0) A.dep[unit] = 0;
        }
        else
            __syncthreads();
        stamp(A, unit, 2);
        const int32_t nlev = B[H_NLEV];
        const int32_t *lp = B + B[H_LEVELS], *lg = lp + nlev + 1;
        for (int32_t l = 0; l < nlev; ++l)
        {
            const int32_t G = lg[l], fb = lp[l], fe = lp[l + 1];
            FOR_LEVEL_FRONTS(G, fb, fe, unit_ops::forward_front(P, B, fronts + (int64_t)fi * F_SIZE, Y, U, LUs, A.upd));
            __syncthreads();
        }
        for (int32_t fi = threadIdx.x; fi < nf; fi += blockDim.x)
        {
            const int32_t *F = fronts + (int64_t)fi * F_SIZE;
            for (int32_t k = 0; k < F[F_S]; ++k) A.y[F[F_COL0] + k] = Y[F[F_YOFF] + k];
        }
        stamp(A, unit, 3);
        const int32_t parent = B[H_PARENT];
        if (parent >= 0)
        {
            __threadfence();
            __syncthreads();
            if (threadIdx.x == 0) atomicAdd(A.dep + parent, 1);
        }
    }
}

// UPDATE: also run the reference's error checks and xk += dx on every finished unknown.
template <bool UPDATE>
__global__ void k_unit_backward(UnitArgs A)
{
    extern __shared__ __align__(16) unsigned char smem[];
    __shared__ uint32_t s_t;
    uint32_t t;
    int flag = 0;
    while (next_ticket(A, &s_t, t))
    {
        const int32_t unit = A.begin + A.count - 1 - (int32_t)t; // parents first
        const int32_t *Bg = A.blob + A.unit_blob[unit];
        const int nints = __ldg(Bg + H_NINTS);
        const int ny = __ldg(Bg + H_YDOUBLES), nxr = __ldg(Bg + H_NXR), nlu = __ldg(Bg + H_LUCOUNT);
        int32_t *B = reinterpret_cast<int32_t *>(smem);
        double *X = reinterpret_cast<double *>(smem + (size_t)nints * 4);
        double *LUs = X + ny + nxr;
        stamp(A, unit, 0);
        load_blob_async(Bg, B, nints);
        {
            const double *src = A.LU + (uint32_t)__ldg(Bg + H_LU0);
            for (int i = threadIdx.x; i < nlu; i += blockDim.x) LUs[i] = src[i];
        }
        async_wait_all();
        __syncthreads();
        const int32_t nf = B[H_NF];
        const int32_t *fronts = B + B[H_FRONTS];
        for (int32_t fi = threadIdx.x; fi < nf; fi += blockDim.x)
        {
            const int32_t *F = fronts + (int64_t)fi * F_SIZE;
            for (int32_t k = 0; k < F[F_S]; ++k) X[F[F_YOFF] + k] = A.y[F[F_COL0] + k];
        }
        const int32_t parent = B[H_PARENT];
        stamp(A, unit, 1);
        if (parent >= 0) wait_counter(A, A.done + parent, A.epoch);
        else __syncthreads();
        stamp(A, unit, 2);
        {
            const int32_t *xr = B + B[H_XR];
            for (int32_t a = threadIdx.x; a < nxr; a += blockDim.x) X[ny + a] = __ldcg(A.y + xr[a]);
        }
        __syncthreads();
        const int32_t nlev = B[H_NLEV];
        const int32_t *lp = B + B[H_LEVELS], *lg = lp + nlev + 1;
        for (int32_t l = nlev - 1; l >= 0; --l)
        {
            const int32_t G = lg[l], fb = lp[l], fe = lp[l + 1];
            FOR_LEVEL_FRONTS(G, fb, fe, unit_ops::backward_front(P, B, fronts + (int64_t)fi * F_SIZE, X, LUs));
            __syncthreads();
        }
        for (int32_t fi = threadIdx.x; fi < nf; fi += blockDim.x)
        {
            const int32_t *F = fronts + (int64_t)fi * F_SIZE;
            for (int32_t k = 0; k < F[F_S]; ++k)
            {
                const int32_t col = F[F_COL0] + k;
                const double v = X[F[F_YOFF] + k];
                A.y[col] = v;
                if (A.out)
                {
                    const int32_t o = A.perm ? A.perm[col] : col;
                    A.out[o] = v;
                    if (UPDATE)
                    {
                        const double e = fabs(v);
                        if (e > A.max_err_abs || isnan(v)) flag |= DAECPP_B200_DIVERGED;
                        const double xa = fabs(A.x0[o]);
                        if ((xa > 1e-14 && __ddiv_rn(e, xa) > A.rtol) || e > A.atol)
                            flag |= e > A.floor ? (DAECPP_B200_NOT_CONVERGED | DAECPP_B200_ABOVE_FLOOR) : DAECPP_B200_NOT_CONVERGED;
                        A.xk[o] = __dadd_rn(A.xk[o], v);
                    }
                }
            }
        }
        stamp(A, unit, 3);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) *(volatile int32_t *)(A.done + unit) = A.epoch;
    }
    if (UPDATE && flag) atomicOr(&A.flags[2], flag);
}

__global__ void k_compose_entries(int64_t ne, const int32_t *__restrict__ ent_u, const int32_t *__restrict__ uJ,
                                  const int32_t *__restrict__ uM, const int32_t *__restrict__ srcJ,
                                  int32_t *__restrict__ sj, int32_t *__restrict__ sm)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ne) return;
    const int32_t u = ent_u[e];
    if (!uJ) { sj[e] = u; sm[e] = -1; return; } // matrix given directly on the union pattern
    const int32_t a = uJ[u];
    sj[e] = a < 0 ? -1 : (srcJ ? srcJ[a] : a);
    sm[e] = uM[u];
}

// entries of big fronts go straight into the (zeroed) global factor storage
__global__ void k_scatter_big(int64_t ne, const int32_t *__restrict__ sj, const int32_t *__restrict__ sm,
                              const uint32_t *__restrict__ dst, const double *__restrict__ jvals,
                              const double *__restrict__ mvals, double alpha, double *__restrict__ LU)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= ne) return;
    const double j = sj[e] >= 0 ? jvals[sj[e]] : 0.0;
    const double m = sm[e] >= 0 ? __dmul_rn(mvals[sm[e]], alpha) : 0.0;
    LU[dst[e]] = __dsub_rn(j, m);
}

UnitArgs base_args(daecpp_b200_ctx *c, const GroupLaunch &g)
{
    UnitArgs A{};
    A.blob = c->dS.blob.p; A.unit_blob = c->dS.unit_blob.p;
    A.begin = g.begin; A.count = g.count;
    A.ticket = reinterpret_cast<uint32_t *>(c->dS.sync_words.p);
    A.ticket_base = c->ticket_base;
    A.dep = c->dS.sync_words.p + 1;
    A.done = c->dS.sync_words.p + 1 + c->S.nunit;
    A.epoch = c->solve_epoch;
    A.flags = c->d_flags;
    A.LU = c->LU.p; A.CB = c->CB.p;
    A.trace = nullptr;
    return A;
}

int grid_for(const GroupLaunch &g, int resident) { return g.count < resident ? g.count : resident; }

} // namespace

// Shared-memory opt-in and resident-block counts for the persistent unit kernels.
int unit_setup(daecpp_b200_ctx *c)
{
    const Symbolic &S = c->S;
    if (S.nunit == 0) return DAECPP_B200_OK;
    int sms = 0;
    DAECPP_CUDA_TRY(c, cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    const int fs = S.factor_smem_bytes, ss = S.solve_smem_bytes;
    if (fs > 227 * 1024 || ss > 227 * 1024) { c->err = "unit exceeds shared memory"; return DAECPP_B200_ENOMEM; }
    {   // The opt-in is a property of the FUNCTION, shared by every context of the process on this device: it only ever
        // grows (another context with bigger units may be alive, or come back from the pool with its plan intact).
        static std::mutex mu;
        static int fs_max[64] = {0}, ss_max[64] = {0};
        std::lock_guard<std::mutex> lk(mu);
        const int d = c->device & 63;
        if (fs > fs_max[d])
        {
            DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_unit_factor, cudaFuncAttributeMaxDynamicSharedMemorySize, fs));
            fs_max[d] = fs;
        }
        if (ss > ss_max[d])
        {
            DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_unit_forward, cudaFuncAttributeMaxDynamicSharedMemorySize, ss));
            DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_unit_backward<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, ss));
            DAECPP_CUDA_TRY(c, cudaFuncSetAttribute(k_unit_backward<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, ss));
            ss_max[d] = ss;
        }
    }
    c->unit_threads_eff = c->unit_threads > 0 ? c->unit_threads : (c->S.budget_capped ? 128 : 256);
    int of = 0, os = 0, ob = 0;
    DAECPP_CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&of, k_unit_factor, c->unit_threads_eff, fs));
    DAECPP_CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&os, k_unit_forward, c->unit_threads_eff, ss));
    DAECPP_CUDA_TRY(c, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ob, k_unit_backward<true>, c->unit_threads_eff, ss));
    if (ob < os) os = ob;
    if (of < 1 || os < 1) { c->err = "unit kernels do not fit on an SM"; return DAECPP_B200_ENOMEM; }
    c->unit_grid_factor = sms * of;
    c->unit_grid_solve = sms * os;
    return DAECPP_B200_OK;
}

int compose_entries(daecpp_b200_ctx *c)
{
    const Symbolic &S = c->S;
    cudaStream_t st = c->stream;
    const int32_t *uJ = c->direct_csc ? nullptr : c->u_J.p;
    const int32_t *srcJ = (!c->direct_csc && c->jvals_from_triplets) ? c->J.src.p : nullptr;
    const int64_t ne = (int64_t)S.ent_u.size(), nb = (int64_t)S.big_u.size();
    if (ne)
        k_compose_entries<<<(int)((ne + 255) / 256), 256, 0, st>>>(ne, c->dS.ent_u.p, uJ, c->u_M.p, srcJ, c->dS.ent_sj.p,
                                                                  c->dS.ent_sm.p);
    if (nb)
        k_compose_entries<<<(int)((nb + 255) / 256), 256, 0, st>>>(nb, c->dS.big_u.p, uJ, c->u_M.p, srcJ, c->dS.big_sj.p,
                                                                  c->dS.big_sm.p);
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

static const double *jvals_of(daecpp_b200_ctx *c)
{
    if (c->direct_csc) return c->u_vals.p;
    return c->jvals_from_triplets ? c->J.A.p : c->J.vals.p;
}

int launch_scatter_big(daecpp_b200_ctx *c, double alpha, cudaStream_t st)
{
    const int64_t nb = (int64_t)c->S.big_u.size();
    if (!nb) return DAECPP_B200_OK;
    k_scatter_big<<<(int)((nb + 255) / 256), 256, 0, st>>>(nb, c->dS.big_sj.p, c->dS.big_sm.p, c->dS.big_dst.p,
                                                                 jvals_of(c), c->M.vals.p, alpha, c->LU.p);
    c->kernels += 1;
    return DAECPP_B200_OK;
}

int launch_unit_factor(daecpp_b200_ctx *c, const GroupLaunch &g, double alpha, cudaStream_t st)
{
    if (!st) st = c->stream;
    UnitArgs A = base_args(c, g);
    A.ent_sj = c->dS.ent_sj.p; A.ent_sm = c->dS.ent_sm.p; A.ent_dst = c->dS.ent_dst.p;
    A.jvals = jvals_of(c); A.mvals = c->M.vals.p;
    A.alpha = alpha; A.pivot_tol = c->pivot_tol;
    A.serial_dim = c->sopt.serial_dim;
    if (c->trace_units && c->trace.reserve((size_t)c->S.nunit * 12 + 4 + 256) == cudaSuccess)
    {
        A.trace = c->trace.p;
        if (g.begin == 0) { A.trace_levels = c->trace.p + (size_t)c->S.nunit * 12 + 4; A.trace_unit = c->trace_unit_index; }
    }
    const int grid = grid_for(g, c->unit_grid_factor);
    k_unit_factor<<<grid, c->unit_threads_eff, c->S.factor_smem_bytes, st>>>(A);
    c->ticket_base += (uint32_t)(g.count + grid);
    c->kernels += 1;
    return DAECPP_B200_OK;
}

int launch_unit_forward(daecpp_b200_ctx *c, const GroupLaunch &g, bool direct, cudaStream_t st)
{
    if (!st) st = c->stream;
    UnitArgs A = base_args(c, g);
    A.perm = direct ? c->dS.perm.p : nullptr;
    A.perm_rhs = direct ? c->dS.perm_rhs.p : nullptr;
    A.rhs = direct ? c->b : c->work;
    A.y = c->work; A.upd = c->upd.p;
    if (c->trace_units && c->trace.p) A.trace = c->trace.p + (size_t)c->S.nunit * 4;
    const int grid = grid_for(g, c->unit_grid_solve);
    k_unit_forward<<<grid, c->unit_threads_eff, c->S.solve_smem_bytes, st>>>(A);
    c->ticket_base += (uint32_t)(g.count + grid);
    c->kernels += 1;
    return DAECPP_B200_OK;
}

int launch_unit_backward(daecpp_b200_ctx *c, const GroupLaunch &g, bool direct, bool update, double atol, double rtol,
                         double max_err_abs, cudaStream_t st)
{
    if (!st) st = c->stream;
    UnitArgs A = base_args(c, g);
    A.perm = c->dS.perm.p;
    A.y = c->work;
    A.out = direct ? c->dx : nullptr;
    A.xk = c->xk; A.x0 = c->xh[0];
    A.atol = atol; A.rtol = rtol; A.max_err_abs = max_err_abs; A.floor = c->refine_floor;
    if (c->trace_units && c->trace.p) A.trace = c->trace.p + (size_t)c->S.nunit * 8;
    const int grid = grid_for(g, c->unit_grid_solve);
    if (direct && update) k_unit_backward<true><<<grid, c->unit_threads_eff, c->S.solve_smem_bytes, st>>>(A);
    else k_unit_backward<false><<<grid, c->unit_threads_eff, c->S.solve_smem_bytes, st>>>(A);
    c->ticket_base += (uint32_t)(g.count + grid);
    c->kernels += 1;
    return DAECPP_B200_OK;
}

} // namespace daecpp_b200
