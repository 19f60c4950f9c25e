// Triplet -> compressed-column assembly on the device.
//
// Replaces sparse_matrix::convert (dae-cpp/sparse-matrix.hpp:124-152, which drives
// Eigen::SparseMatrix::findOrInsertCoeff, Eigen/src/SparseCore/SparseMatrix.h:231-265) and the
// union-pattern subtraction `Jb -= M_*alpha` (dae-cpp/solver.hpp:523,
// Eigen/src/SparseCore/SparseCwiseBinaryOp.h:75-105).  Results are bit-identical to the
// reference: per slot the duplicates are added left to right in input order starting from
// +0.0, with round-to-nearest adds and no FMA contraction.
//
// Two paths:
//   cold  -- new (i,j) stream: stable LSD radix sort of (col,row) keys with the input index as
//            payload (cub::DeviceRadixSort, a toolkit primitive; once per pattern), segment
//            heads -> slots.  The sorted order is cached.
//   hot   -- the (i,j) stream equals the cached one (checked on the device): one gather-sum
//            kernel, values only.  12 B read per triplet + 8 B written per slot.
#include <cub/cub.cuh>

#include "ctx.hpp"

namespace daecpp_b200
{
namespace
{

constexpr int TPB = 256;
inline int blocks_for(int64_t count, int tpb = TPB) { return (int)((count + tpb - 1) / tpb); }

__global__ void k_check_and_key(const int32_t *__restrict__ ti, const int32_t *__restrict__ tj, int64_t nT, int32_t n,
                                uint64_t *__restrict__ keys, int32_t *__restrict__ pay, int32_t *__restrict__ flags)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nT) return;
    int32_t i = ti[k], j = tj[k];
    if ((uint32_t)i >= (uint32_t)n || (uint32_t)j >= (uint32_t)n) { atomicOr(&flags[0], 1); i = 0; j = 0; }
    keys[k] = ((uint64_t)(uint32_t)j << 32) | (uint32_t)i;
    pay[k] = (int32_t)k;
}

__global__ void k_compare_pattern(const int32_t *__restrict__ ti, const int32_t *__restrict__ tj,
                                  const int32_t *__restrict__ ci, const int32_t *__restrict__ cj, int64_t nT,
                                  int32_t *__restrict__ flags)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool bad = false;
    if (k < nT) bad = (ti[k] != ci[k]) | (tj[k] != cj[k]);
    if (__syncthreads_or(bad) && threadIdx.x == 0) atomicOr(&flags[1], 1);
}

__global__ void k_heads(const uint64_t *__restrict__ keys, int64_t count, int32_t *__restrict__ head)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= count) return;
    head[p] = (p == 0 || keys[p] != keys[p - 1]) ? 1 : 0;
}

// After the inclusive scan of heads: write per-segment start, low key word (inner index) and the
// compressed pointer array over the high key word.
__global__ void k_segments(const uint64_t *__restrict__ keys, const int32_t *__restrict__ scan, int64_t count, int32_t n,
                           int32_t *__restrict__ seg_ptr, int32_t *__restrict__ inner, int32_t *__restrict__ outer)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= count) return;
    const int32_t slot = scan[p] - 1;
    const uint64_t key = keys[p];
    const bool head = (p == 0) || (keys[p - 1] != key);
    if (head)
    {
        if (seg_ptr) seg_ptr[slot] = (int32_t)p;
        if (inner) inner[slot] = (int32_t)(key & 0xffffffffu);
        const int32_t col = (int32_t)(key >> 32);
        const int32_t prev = (p == 0) ? -1 : (int32_t)(keys[p - 1] >> 32);
        for (int32_t c = prev + 1; c <= col; ++c) outer[c] = slot;
    }
    if (p == count - 1)
    {
        const int32_t nseg = scan[p];
        const int32_t col = (int32_t)(key >> 32);
        for (int32_t c = col + 1; c <= n; ++c) outer[c] = nseg;
        if (seg_ptr) seg_ptr[nseg] = (int32_t)count;
    }
}

__global__ void k_fill_i32(int32_t *p, int64_t count, int32_t v)
{
    int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k < count) p[k] = v;
}

// vals[s] = 0.0 + A[src[p0]] + A[src[p0+1]] + ...   (input order; __dadd_rn forbids contraction)
__global__ void k_gather_sum(const double *__restrict__ A, const int32_t *__restrict__ src,
                             const int32_t *__restrict__ src_ptr, int64_t nnz, double *__restrict__ vals)
{
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= nnz) return;
    double v = 0.0;
    const int32_t e = src_ptr[s + 1];
    for (int32_t p = src_ptr[s]; p < e; ++p) v = __dadd_rn(v, A[src[p]]);
    vals[s] = v;
}

__global__ void k_gather_nodup(const double *__restrict__ A, const int32_t *__restrict__ src, int64_t nnz,
                               double *__restrict__ vals)
{
    int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s < nnz) vals[s] = __dadd_rn(0.0, A[src[s]]);
}

// keys of the slots of a compressed matrix: (major<<32 | minor) or swapped
__global__ void k_slot_keys(const int32_t *__restrict__ outer, const int32_t *__restrict__ inner, int32_t n,
                            uint64_t *__restrict__ keys, int32_t *__restrict__ pay, int32_t tag, int swap)
{
    int32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
    {
        const uint32_t r = (uint32_t)inner[p];
        keys[p] = swap ? (((uint64_t)r << 32) | (uint32_t)c) : (((uint64_t)(uint32_t)c << 32) | r);
        pay[p] = p | tag;
    }
}

__global__ void k_union_maps(const int32_t *__restrict__ pay, const int32_t *__restrict__ scan, int64_t count,
                             int32_t *__restrict__ uJ, int32_t *__restrict__ uM)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= count) return;
    const int32_t u = scan[p] - 1, v = pay[p];
    if (v < 0) uM[u] = v & 0x7fffffff; else uJ[u] = v;
}

__global__ void k_csr_cols(const uint64_t *__restrict__ keys, int64_t count, int32_t *__restrict__ cols)
{
    int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p < count) cols[p] = (int32_t)(keys[p] & 0xffffffffu);
}

int bits_for(int32_t n)
{
    int b = 1;
    while (b < 32 && (1ll << b) < (long long)n) ++b;
    return b;
}

// keys_a/pay_a [count] -> keys_b/pay_b sorted (stable); scan = inclusive count of segment heads.
// Returns the number of distinct keys through *nseg (synchronises the stream).
int sort_and_scan(daecpp_b200_ctx *c, int64_t count, int32_t n, int32_t *nseg)
{
    DAECPP_CUDA_TRY(c, c->keys_b.reserve(count));
    DAECPP_CUDA_TRY(c, c->pay_b.reserve(count));
    DAECPP_CUDA_TRY(c, c->scan.reserve(count + 1));
    const int end_bit = 32 + bits_for(n);
    size_t need = 0, need2 = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, need, c->keys_a.p, c->keys_b.p, c->pay_a.p, c->pay_b.p, (int)count, 0,
                                    end_bit, c->stream);
    cub::DeviceScan::InclusiveSum(nullptr, need2, c->scan.p, c->scan.p, (int)count, c->stream);
    if (need2 > need) need = need2;
    DAECPP_CUDA_TRY(c, c->cub_tmp.reserve(need + 256));
    size_t tmp_bytes = c->cub_tmp.cap;
    DAECPP_CUDA_TRY(c, cub::DeviceRadixSort::SortPairs(c->cub_tmp.p, tmp_bytes, c->keys_a.p, c->keys_b.p, c->pay_a.p,
                                                       c->pay_b.p, (int)count, 0, end_bit, c->stream));
    k_heads<<<blocks_for(count), TPB, 0, c->stream>>>(c->keys_b.p, count, c->scan.p);
    tmp_bytes = c->cub_tmp.cap;
    DAECPP_CUDA_TRY(c, cub::DeviceScan::InclusiveSum(c->cub_tmp.p, tmp_bytes, c->scan.p, c->scan.p, (int)count, c->stream));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(&c->h_flags[4], c->scan.p + (count - 1), sizeof(int32_t), cudaMemcpyDeviceToHost,
                                       c->stream));
    DAECPP_CUDA_TRY(c, cudaStreamSynchronize(c->stream));
    *nseg = c->h_flags[4];
    c->kernels += 4;
    return DAECPP_B200_OK;
}

} // namespace

int assemble_triplets(daecpp_b200_ctx *c, TripletSet &T, const double *A, const int32_t *ti, const int32_t *tj,
                      int64_t nT, bool lazy_vals)
{
    const int32_t n = c->n;
    cudaStream_t st = c->stream;
    if (nT < 0 || nT > 0x7fffff00ll) { c->err = "triplet count out of range"; return DAECPP_B200_EINVAL; }
    if (nT > 0 && (!A || !ti || !tj)) { c->err = "null triplet array"; return DAECPP_B200_EINVAL; }
    T.nT = nT;
    DAECPP_CUDA_TRY(c, T.outer.reserve((size_t)n + 1));
    if (nT == 0)
    {
        if (!(T.have_pattern && T.nT_cached == 0))
        {
            T.pattern_version++;
            T.have_pattern = true; T.nT_cached = 0; T.nnz = 0; T.no_duplicates = true;
            DAECPP_CUDA_TRY(c, cudaMemsetAsync(T.outer.p, 0, sizeof(int32_t) * ((size_t)n + 1), st));
        }
        return DAECPP_B200_OK;
    }
    DAECPP_CUDA_TRY(c, T.A.reserve(nT));
    DAECPP_CUDA_TRY(c, T.ti.reserve(nT));
    DAECPP_CUDA_TRY(c, T.tj.reserve(nT));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.ti.p, ti, sizeof(int32_t) * nT, cudaMemcpyHostToDevice, st));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.tj.p, tj, sizeof(int32_t) * nT, cudaMemcpyHostToDevice, st));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.A.p, A, sizeof(double) * nT, cudaMemcpyHostToDevice, st));
    c->flags_clean = false;
    DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags, 0, sizeof(int32_t) * 4, st));

    if (T.have_pattern && T.nT_cached == nT)
    { // optimistic hot path: verify the pattern and gather in the same submission
        k_compare_pattern<<<blocks_for(nT), TPB, 0, st>>>(T.ti.p, T.tj.p, T.ti_cached.p, T.tj_cached.p, nT, c->d_flags);
        c->kernels += 1;
        T.vals_valid = false;
        if (!lazy_vals)
        { // compressed-column values wanted now (mass matrix: the residual kernel reads them)
            int rc = regather_values(c, T);
            if (rc) return rc;
            T.vals_valid = true;
        }
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags, c->d_flags, sizeof(int32_t) * 4, cudaMemcpyDeviceToHost, st));
        DAECPP_CUDA_TRY(c, cudaStreamSynchronize(st));
        if (c->h_flags[1] == 0) return DAECPP_B200_OK;
        c->flags_clean = false;
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->d_flags, 0, sizeof(int32_t) * 4, st));
    }

    // ---- cold path: sort, segment, cache ------------------------------------------------
    c->n_cold += 1;
    DAECPP_CUDA_TRY(c, c->keys_a.reserve(nT));
    DAECPP_CUDA_TRY(c, c->pay_a.reserve(nT));
    k_check_and_key<<<blocks_for(nT), TPB, 0, st>>>(T.ti.p, T.tj.p, nT, n, c->keys_a.p, c->pay_a.p, c->d_flags);
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->h_flags, c->d_flags, sizeof(int32_t) * 4, cudaMemcpyDeviceToHost, st));
    int32_t nnz = 0;
    int rc = sort_and_scan(c, nT, n, &nnz);
    if (rc) return rc;
    if (c->h_flags[0])
    {
        T.have_pattern = false;
        c->err = "sparse matrix index out of boundaries";
        return DAECPP_B200_EINDEX;
    }
    T.nnz = nnz;
    T.no_duplicates = (nnz == nT);
    DAECPP_CUDA_TRY(c, T.src.reserve(nT));
    DAECPP_CUDA_TRY(c, T.src_ptr.reserve((size_t)nnz + 1));
    DAECPP_CUDA_TRY(c, T.inner.reserve(nnz));
    DAECPP_CUDA_TRY(c, T.vals.reserve(nnz));
    DAECPP_CUDA_TRY(c, T.ti_cached.reserve(nT));
    DAECPP_CUDA_TRY(c, T.tj_cached.reserve(nT));
    k_segments<<<blocks_for(nT), TPB, 0, st>>>(c->keys_b.p, c->scan.p, nT, n, T.src_ptr.p, T.inner.p, T.outer.p);
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.src.p, c->pay_b.p, sizeof(int32_t) * nT, cudaMemcpyDeviceToDevice, st));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.ti_cached.p, T.ti.p, sizeof(int32_t) * nT, cudaMemcpyDeviceToDevice, st));
    DAECPP_CUDA_TRY(c, cudaMemcpyAsync(T.tj_cached.p, T.tj.p, sizeof(int32_t) * nT, cudaMemcpyDeviceToDevice, st));
    k_gather_sum<<<blocks_for(nnz), TPB, 0, st>>>(T.A.p, T.src.p, T.src_ptr.p, nnz, T.vals.p);
    c->kernels += 3;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    T.have_pattern = true;
    T.vals_valid = true;
    T.nT_cached = nT;
    T.pattern_version++;
    return DAECPP_B200_OK;
}

int regather_values(daecpp_b200_ctx *c, TripletSet &T)
{
    if (!T.have_pattern) { c->err = "no cached pattern to re-gather"; return DAECPP_B200_EINVAL; }
    if (T.nnz == 0) return DAECPP_B200_OK;
    if (T.no_duplicates) k_gather_nodup<<<blocks_for(T.nnz), TPB, 0, c->stream>>>(T.A.p, T.src.p, T.nnz, T.vals.p);
    else k_gather_sum<<<blocks_for(T.nnz), TPB, 0, c->stream>>>(T.A.p, T.src.p, T.src_ptr.p, T.nnz, T.vals.p);
    c->kernels += 1;
    DAECPP_CUDA_TRY(c, cudaGetLastError());
    return DAECPP_B200_OK;
}

// Pattern of Jb = pattern(J) U pattern(M), slot maps u_J / u_M (-1 where absent).
int build_union(daecpp_b200_ctx *c)
{
    if (c->direct_csc) return DAECPP_B200_OK;
    TripletSet &J = c->J, &M = c->M;
    if (!J.have_pattern || !M.have_pattern) { c->err = "set_mass and set_jacobian must precede refactor"; return DAECPP_B200_EINVAL; }
    if (c->union_J_version == J.pattern_version && c->union_M_version == M.pattern_version) return DAECPP_B200_OK;
    const int32_t n = c->n;
    cudaStream_t st = c->stream;
    const int64_t total = J.nnz + M.nnz;
    DAECPP_CUDA_TRY(c, c->u_outer.reserve((size_t)n + 1));
    if (total == 0)
    {
        c->nnzU = 0;
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->u_outer.p, 0, sizeof(int32_t) * ((size_t)n + 1), st));
    }
    else
    {
        DAECPP_CUDA_TRY(c, c->keys_a.reserve(total));
        DAECPP_CUDA_TRY(c, c->pay_a.reserve(total));
        if (J.nnz) k_slot_keys<<<blocks_for(n), TPB, 0, st>>>(J.outer.p, J.inner.p, n, c->keys_a.p, c->pay_a.p, 0, 0);
        if (M.nnz) k_slot_keys<<<blocks_for(n), TPB, 0, st>>>(M.outer.p, M.inner.p, n, c->keys_a.p + J.nnz, c->pay_a.p + J.nnz,
                                                             (int32_t)0x80000000u, 0);
        int32_t nnzU = 0;
        int rc = sort_and_scan(c, total, n, &nnzU);
        if (rc) return rc;
        c->nnzU = nnzU;
        DAECPP_CUDA_TRY(c, c->u_inner.reserve(nnzU));
        DAECPP_CUDA_TRY(c, c->u_J.reserve(nnzU));
        DAECPP_CUDA_TRY(c, c->u_M.reserve(nnzU));
        DAECPP_CUDA_TRY(c, c->u_vals.reserve(nnzU));
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->u_J.p, 0xff, sizeof(int32_t) * nnzU, st));
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->u_M.p, 0xff, sizeof(int32_t) * nnzU, st));
        k_segments<<<blocks_for(total), TPB, 0, st>>>(c->keys_b.p, c->scan.p, total, n, nullptr, c->u_inner.p, c->u_outer.p);
        k_union_maps<<<blocks_for(total), TPB, 0, st>>>(c->pay_b.p, c->scan.p, total, c->u_J.p, c->u_M.p);
        c->kernels += 4;
        DAECPP_CUDA_TRY(c, cudaGetLastError());
    }
    c->union_J_version = J.pattern_version;
    c->union_M_version = M.pattern_version;
    c->have_symbolic = false;
    c->have_factor = false;
    return DAECPP_B200_OK;
}

// Row-major view of M: per row the slots in increasing column order, which is the order in
// which Eigen's column-major SpMV accumulates into each row (SparseDenseProduct.h:96-112).
int build_mass_csr(daecpp_b200_ctx *c)
{
    TripletSet &M = c->M;
    if (c->m_csr_version == M.pattern_version) return DAECPP_B200_OK;
    const int32_t n = c->n;
    cudaStream_t st = c->stream;
    DAECPP_CUDA_TRY(c, c->m_rowptr.reserve((size_t)n + 1));
    if (M.nnz == 0)
    {
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->m_rowptr.p, 0, sizeof(int32_t) * ((size_t)n + 1), st));
    }
    else
    {
        DAECPP_CUDA_TRY(c, c->keys_a.reserve(M.nnz));
        DAECPP_CUDA_TRY(c, c->pay_a.reserve(M.nnz));
        DAECPP_CUDA_TRY(c, c->m_rowslot.reserve(M.nnz));
        DAECPP_CUDA_TRY(c, c->m_rowcol.reserve(M.nnz));
        k_slot_keys<<<blocks_for(n), TPB, 0, st>>>(M.outer.p, M.inner.p, n, c->keys_a.p, c->pay_a.p, 0, 1);
        int32_t nseg = 0;
        int rc = sort_and_scan(c, M.nnz, n, &nseg);
        if (rc) return rc;
        // all keys are distinct, so scan[p] = p+1 and the "segments" are single slots; only the
        // pointer array over the high word (row) is needed
        k_segments<<<blocks_for(M.nnz), TPB, 0, st>>>(c->keys_b.p, c->scan.p, M.nnz, n, nullptr, nullptr, c->m_rowptr.p);
        k_csr_cols<<<blocks_for(M.nnz), TPB, 0, st>>>(c->keys_b.p, M.nnz, c->m_rowcol.p);
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->m_rowslot.p, c->pay_b.p, sizeof(int32_t) * M.nnz, cudaMemcpyDeviceToDevice, st));
        c->kernels += 3;
        DAECPP_CUDA_TRY(c, cudaGetLastError());
    }
    c->m_csr_version = M.pattern_version;
    return classify_mass(c);
}

// Row-major view of J's compressed-column slots (same construction as build_mass_csr), built on first use by
// the refinement step and cached per pattern.
int build_jac_csr(daecpp_b200_ctx *c)
{
    TripletSet &J = c->J;
    if (!J.have_pattern) { c->err = "no Jacobian pattern"; return DAECPP_B200_EINVAL; }
    if (c->j_csr_version == J.pattern_version) return DAECPP_B200_OK;
    const int32_t n = c->n;
    cudaStream_t st = c->stream;
    DAECPP_CUDA_TRY(c, c->j_rowptr.reserve((size_t)n + 1));
    if (J.nnz == 0)
        DAECPP_CUDA_TRY(c, cudaMemsetAsync(c->j_rowptr.p, 0, sizeof(int32_t) * ((size_t)n + 1), st));
    else
    {
        DAECPP_CUDA_TRY(c, c->keys_a.reserve(J.nnz));
        DAECPP_CUDA_TRY(c, c->pay_a.reserve(J.nnz));
        DAECPP_CUDA_TRY(c, c->j_rowslot.reserve(J.nnz));
        DAECPP_CUDA_TRY(c, c->j_rowcol.reserve(J.nnz));
        k_slot_keys<<<blocks_for(n), TPB, 0, st>>>(J.outer.p, J.inner.p, n, c->keys_a.p, c->pay_a.p, 0, 1);
        int32_t nseg = 0;
        int rc = sort_and_scan(c, J.nnz, n, &nseg);
        if (rc) return rc;
        k_segments<<<blocks_for(J.nnz), TPB, 0, st>>>(c->keys_b.p, c->scan.p, J.nnz, n, nullptr, nullptr, c->j_rowptr.p);
        k_csr_cols<<<blocks_for(J.nnz), TPB, 0, st>>>(c->keys_b.p, J.nnz, c->j_rowcol.p);
        DAECPP_CUDA_TRY(c, cudaMemcpyAsync(c->j_rowslot.p, c->pay_b.p, sizeof(int32_t) * J.nnz, cudaMemcpyDeviceToDevice, st));
        c->kernels += 3;
        DAECPP_CUDA_TRY(c, cudaGetLastError());
    }
    c->j_csr_version = J.pattern_version;
    return DAECPP_B200_OK;
}

} // namespace daecpp_b200
