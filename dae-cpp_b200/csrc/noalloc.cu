// daecpp_b200_variability_noalloc -- an entry point that allocates NOTHING (no cudaMalloc / cudaFree / allocating
// library routine), for callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses
// that forbid it.  It has no counterpart in the reference.
//
// What it computes: the maximum relative variability of dae-cpp/solver.hpp:703-717,
//     max_i |(x0_i - xk_i) / x0_i|   over the elements with |x0_i| > tol and |xk_i| > tol   (NaN counts as 0),
// i.e. exactly what k_accept (vector_ops.cu) measures behind daecpp_b200_accept_step -- without the history shift.
// k_accept accumulates into a scratch word the context allocated with cudaMalloc; this variant keeps its scratch
// (per-block partial maxima + a ticket counter) in a file-scope `static __device__` array of fixed size.
//
// The scratch array is shared by every call on a device and keeps the last call's contents: the ticket word is cleared
// by a memset node ahead of every launch (and again by the block that finishes last), the partial maxima of the blocks
// in use are overwritten before they are read.  Calls that overlap on the GPU (two streams, two host threads) RACE on it.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <mutex>

#include "../../include/daecpp_b200.h"

namespace
{
constexpr int NA_THREADS = 256;
constexpr int NA_PER_BLOCK = 4096;                                      // elements per block: 16 per thread
constexpr int NA_MAX_BLOCKS = 4096;                                     // => at most 16 777 216 elements
constexpr int64_t NA_MAX_N = (int64_t)NA_MAX_BLOCKS * NA_PER_BLOCK;
constexpr int NA_SCRATCH_WORDS = NA_MAX_BLOCKS + 32;                    // partial maxima, then the ticket word (own 256-byte line)
static_assert(NA_MAX_N == DAECPP_B200_NOALLOC_MAX_N, "the header states the limit");

// 33 024 bytes per GPU that loads the library, held until the process exits
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NA_SCRATCH_WORDS];

__global__ void __launch_bounds__(NA_THREADS) k_variability_noalloc(const double *__restrict__ xk, const double *__restrict__ x0, int64_t n,
                                                                    double tol, unsigned long long *__restrict__ scratch,
                                                                    double *__restrict__ out)
{
    __shared__ unsigned long long s_part[NA_THREADS / 32];
    __shared__ int s_last;
    const int t = threadIdx.x;
    const int64_t base = (int64_t)blockIdx.x * NA_PER_BLOCK;
    unsigned long long bits = 0ull; // non-negative doubles order like their bit patterns
#pragma unroll 4
    for (int k = 0; k < NA_PER_BLOCK / NA_THREADS; ++k)
    {
        const int64_t i = base + t + (int64_t)k * NA_THREADS;
        if (i < n)
        {
            const double a = xk[i], p = x0[i];
            double v = 0.0;
            if (fabs(p) > tol && fabs(a) > tol) v = fabs(__ddiv_rn(__dsub_rn(p, a), p));
            if (!(v == v)) v = 0.0; // the reference's `rel_change > variability` ignores NaN
            const unsigned long long b = (unsigned long long)__double_as_longlong(v);
            bits = b > bits ? b : bits;
        }
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
        bits = other > bits ? other : bits;
    }
    if ((t & 31) == 0) s_part[t >> 5] = bits;
    __syncthreads();
    if (t == 0)
    {
        for (int w = 1; w < NA_THREADS / 32; ++w) bits = s_part[w] > bits ? s_part[w] : bits;
        scratch[blockIdx.x] = bits;
        __threadfence();
        const unsigned long long ticket = atomicAdd(&scratch[NA_MAX_BLOCKS], 1ull);
        s_last = ticket == (unsigned long long)gridDim.x - 1ull;
    }
    __syncthreads();
    if (!s_last) return;
    // the block that finishes last folds the partial maxima of this launch (and only of this launch)
    __threadfence();
    bits = 0ull;
    for (int b = t; b < (int)gridDim.x; b += NA_THREADS)
    {
        const unsigned long long v = *(volatile unsigned long long *)&scratch[b];
        bits = v > bits ? v : bits;
    }
    for (int o = 16; o > 0; o >>= 1)
    {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, bits, o);
        bits = other > bits ? other : bits;
    }
    if ((t & 31) == 0) s_part[t >> 5] = bits;
    __syncthreads();
    if (t == 0)
    {
        for (int w = 1; w < NA_THREADS / 32; ++w) bits = s_part[w] > bits ? s_part[w] : bits;
        out[0] = __longlong_as_double((long long)bits);
        scratch[NA_MAX_BLOCKS] = 0ull; // ready for the next call even if its memset were skipped
    }
}

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mu;
unsigned long long *g_na_ptr[NA_MAX_DEVICES] = {nullptr}; // device address of the scratch array, per device

void na_error(char *buf, size_t cap, const char *msg)
{
    if (buf && cap) std::snprintf(buf, cap, "%s", msg);
}
} // namespace

extern "C" int daecpp_b200_variability_noalloc(const double *xk_dev, const double *x0_dev, int64_t n, double tolerance, double *result_dev,
                                               void *stream, char *errbuf, size_t errbuf_len)
{
    if (!xk_dev || !x0_dev || !result_dev || n < 0)
    {
        na_error(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: null pointer or negative size");
        return DAECPP_B200_EINVAL;
    }
    if (n > NA_MAX_N)
    {
        if (errbuf && errbuf_len)
            std::snprintf(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: n = %lld exceeds the fixed scratch (at most %lld elements)",
                          (long long)n, (long long)NA_MAX_N);
        return DAECPP_B200_EINVAL;
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES)
    {
        na_error(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: no current CUDA device");
        return DAECPP_B200_ECUDA;
    }
    unsigned long long *scratch = nullptr;
    {   // set-up, once per device, with that device current: in host code the array's NAME is a host-side stand-in, not
        // a device address -- the address comes from cudaGetSymbolAddress
        std::lock_guard<std::mutex> lk(g_na_mu);
        if (!g_na_ptr[dev])
        {
            void *p = nullptr;
            const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
            if (e != cudaSuccess || !p)
            {
                cudaGetLastError();
                na_error(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: cudaGetSymbolAddress failed");
                return DAECPP_B200_ECUDA;
            }
            g_na_ptr[dev] = static_cast<unsigned long long *>(p);
        }
        scratch = g_na_ptr[dev];
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int blocks = n == 0 ? 1 : (int)((n + NA_PER_BLOCK - 1) / NA_PER_BLOCK);
    // the ticket word must start at zero (a memset node: legal under stream capture, allocates nothing)
    if (cudaMemsetAsync(scratch + NA_MAX_BLOCKS, 0, sizeof(unsigned long long), st) != cudaSuccess)
    {
        cudaGetLastError();
        na_error(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: cudaMemsetAsync failed");
        return DAECPP_B200_ECUDA;
    }
    k_variability_noalloc<<<blocks, NA_THREADS, 0, st>>>(xk_dev, x0_dev, n, tolerance, scratch, result_dev);
    if (cudaGetLastError() != cudaSuccess)
    {
        na_error(errbuf, errbuf_len, "daecpp_b200_variability_noalloc: kernel launch failed");
        return DAECPP_B200_ECUDA;
    }
    return DAECPP_B200_OK;
}
