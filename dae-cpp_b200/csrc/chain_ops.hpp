// Per-thread arithmetic of the block-chain LU (chain.cu): the code every CUDA thread runs between
// two block-wide barriers, written against a thread index so that the same functions are walked
// on the CPU by tests/emu/chain_emulate.cpp in the GPU-less container.
//
// A level is a block-tridiagonal system  A_i x_{i-1} + D_i x_i + C_i x_{i+1} = r_i.  Rows are
// grouped into chunks of Q consecutive rows: the first Q-1 ("interior") are eliminated in order
// by one thread, the last one is the chunk's separator.  Eliminating the interior of a chunk
// whose left neighbour separator is row L and own separator is row R leaves
//     G  = entry (R, L)        Df = diagonal of R (updated from the left side)
//     SL = update of D_L       Hf = entry (L, R)
// i.e. one row of the next level's block-tridiagonal system over the separators:
//     A'_j = G_j,   D'_j = Df_j - SL_{j+1},   C'_j = Hf_{j+1}.
// Three levels are swept inside one CTA (T threads x Q0 rows, then Q2 threads x Q1 rows, then one
// thread x Q2 rows); what is left per CTA is one row of the grid-level reduced system.
// Factors kept per eliminated row k: L_k = A_{k+1} D_k^-1, LL_k = H_k D_k^-1 (multipliers of the
// next row and of row L), D_k^-1, C_k and the spike G_k (entry (k, L)).
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define CH_HD __host__ __device__ __forceinline__
#else
#define CH_HD inline
#endif

namespace daecpp_b200
{
namespace chain
{

template <int B> struct Cfg;
template <> struct Cfg<1> { static constexpr int T = 128, Q0 = 16, Q1 = 16, Q2 = 8; };
template <> struct Cfg<2> { static constexpr int T = 128, Q0 = 4, Q1 = 16, Q2 = 8; };
template <> struct Cfg<3> { static constexpr int T = 64, Q0 = 4, Q1 = 8, Q2 = 8; };

template <int B>
struct Dim
{
    using C = Cfg<B>;
    static constexpr int T = C::T, Q0 = C::Q0, Q1 = C::Q1, Q2 = C::Q2;
    static_assert(Q1 * Q2 == T, "level-1 chunks must tile the CTA");
    static constexpr int BB = B * B, NF = 5 * BB;
    static constexpr int ROWS = T * Q0;
    static constexpr int RP = ROWS + 1 + (ROWS + 1) / Q0 + 1; // staged rows incl. halo and bank padding
    static constexpr int F0 = NF * (Q0 - 1) * T, F1 = NF * (Q1 - 1) * Q2, F2 = NF * (Q2 - 1);
    static constexpr int TILE_F = F0 + F1 + F2 + BB; // + inverse of the top pivot
    static constexpr int Y0 = B * (Q0 - 1) * T, Y1 = B * (Q1 - 1) * Q2, Y2 = B * (Q2 - 1);
    static constexpr int TILE_Y = Y0 + Y1 + Y2;
};

// staged position of local row `row` (row = -1 is the halo); one pad double every Q0 rows keeps
// the T chunk-strided accesses of a warp on distinct shared-memory banks
template <int B> CH_HD int sidx(int row) { return (row + 1) + (row + 1) / Dim<B>::Q0; }

enum { FL = 0, FLL = 1, FDINV = 2, FC = 3, FG = 4 };

template <int B> struct Mat { double a[B * B]; };
template <int B> struct Vec { double a[B]; };

template <int B> CH_HD Mat<B> mzero() { Mat<B> r; for (int i = 0; i < B * B; ++i) r.a[i] = 0.0; return r; }
template <int B> CH_HD Mat<B> mident() { Mat<B> r = mzero<B>(); for (int i = 0; i < B; ++i) r.a[i * B + i] = 1.0; return r; }
template <int B> CH_HD Vec<B> vzero() { Vec<B> r; for (int i = 0; i < B; ++i) r.a[i] = 0.0; return r; }

template <int B> CH_HD Mat<B> mmul(const Mat<B> &X, const Mat<B> &Y)
{
    Mat<B> r;
#pragma unroll
    for (int i = 0; i < B; ++i)
#pragma unroll
        for (int j = 0; j < B; ++j)
        {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < B; ++k) s += X.a[i * B + k] * Y.a[k * B + j];
            r.a[i * B + j] = s;
        }
    return r;
}
template <int B> CH_HD void msub(Mat<B> &D, const Mat<B> &X) { for (int i = 0; i < B * B; ++i) D.a[i] -= X.a[i]; }
template <int B> CH_HD void madd(Mat<B> &D, const Mat<B> &X) { for (int i = 0; i < B * B; ++i) D.a[i] += X.a[i]; }
template <int B> CH_HD Mat<B> mneg(const Mat<B> &X) { Mat<B> r; for (int i = 0; i < B * B; ++i) r.a[i] = -X.a[i]; return r; }
template <int B> CH_HD Vec<B> mv(const Mat<B> &X, const Vec<B> &v)
{
    Vec<B> r;
#pragma unroll
    for (int i = 0; i < B; ++i)
    {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < B; ++k) s += X.a[i * B + k] * v.a[k];
        r.a[i] = s;
    }
    return r;
}
template <int B> CH_HD void vsub(Vec<B> &d, const Vec<B> &x) { for (int i = 0; i < B; ++i) d.a[i] -= x.a[i]; }
template <int B> CH_HD void vadd(Vec<B> &d, const Vec<B> &x) { for (int i = 0; i < B; ++i) d.a[i] += x.a[i]; }

// Reciprocal of a pivot.  Device: hardware seed + two Newton steps, branch-free -- the IEEE-rounded __drcp_rn carries
// a slow-path subroutine call that sits in the middle of every elimination step's dependency chain.  The result is
// within an ulp or two of 1/x; factors are never compared bit for bit (only dx / residuals are), and
// a zero or non-finite pivot is caught by bad_pivot, not here.
CH_HD double rcp(double x)
{
#if defined(__CUDA_ARCH__)
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    y = fma(y, fma(-x, y, 1.0), y);
    y = fma(y, fma(-x, y, 1.0), y);
    return y;
#else
    return 1.0 / x;
#endif
}
CH_HD bool bad_pivot(double p, double tol)
{
    const double a = fabs(p);
    return !(a > tol) || a > 1.7976931348623157e308;
}

// Inverse of the pivot block by Gauss-Jordan with partial (row) pivoting inside the block.  Rows of
// one block share their sparsity, so the row exchange stays inside the block row.
template <int B> CH_HD Mat<B> minv(const Mat<B> &D, double tol, int &fail)
{
    if (B == 1)
    {
        Mat<B> r;
        if (bad_pivot(D.a[0], tol)) fail = 1;
        r.a[0] = rcp(D.a[0]);
        return r;
    }
    double w[B][2 * B];
#pragma unroll
    for (int i = 0; i < B; ++i)
#pragma unroll
        for (int j = 0; j < B; ++j) { w[i][j] = D.a[i * B + j]; w[i][B + j] = (i == j) ? 1.0 : 0.0; }
#pragma unroll
    for (int k = 0; k < B; ++k)
    {
        int p = k;
        double best = fabs(w[k][k]);
#pragma unroll
        for (int i = k + 1; i < B; ++i)
            if (fabs(w[i][k]) > best) { best = fabs(w[i][k]); p = i; }
#pragma unroll
        for (int i = k + 1; i < B; ++i)
            if (i == p)
            {
#pragma unroll
                for (int j = 0; j < 2 * B; ++j) { const double tmp = w[k][j]; w[k][j] = w[i][j]; w[i][j] = tmp; }
            }
        if (bad_pivot(w[k][k], tol)) fail = 1;
        const double inv = rcp(w[k][k]);
#pragma unroll
        for (int j = 0; j < 2 * B; ++j) w[k][j] *= inv;
#pragma unroll
        for (int i = 0; i < B; ++i)
            if (i != k)
            {
                const double f = w[i][k];
#pragma unroll
                for (int j = 0; j < 2 * B; ++j) w[i][j] -= f * w[k][j];
            }
    }
    Mat<B> r;
#pragma unroll
    for (int i = 0; i < B; ++i)
#pragma unroll
        for (int j = 0; j < B; ++j) r.a[i * B + j] = w[i][B + j];
    return r;
}

// ---- factor storage of one chunk: element (k, field f, e) at p[(k*NF + f*BB + e)*STRIDE]; p already
// points at the chunk's lane, STRIDE is the number of chunks interleaved (T, Q2 or 1), so that with k
// unrolled every access is base + immediate
template <int B, int STRIDE> struct FacRef
{
    double *p;
    CH_HD void put(int k, int f, const Mat<B> &X) const
    {
#pragma unroll
        for (int e = 0; e < B * B; ++e) p[(size_t)(k * Dim<B>::NF + f * B * B + e) * STRIDE] = X.a[e];
    }
    CH_HD Mat<B> get(int k, int f) const
    {
        Mat<B> X;
#pragma unroll
        for (int e = 0; e < B * B; ++e) X.a[e] = p[(size_t)(k * Dim<B>::NF + f * B * B + e) * STRIDE];
        return X;
    }
};
template <int B, int STRIDE> struct YRef
{
    double *p;
    CH_HD void put(int k, const Vec<B> &v) const
    {
#pragma unroll
        for (int r = 0; r < B; ++r) p[(size_t)(k * B + r) * STRIDE] = v.a[r];
    }
    CH_HD Vec<B> get(int k) const
    {
        Vec<B> v;
#pragma unroll
        for (int r = 0; r < B; ++r) v.a[r] = p[(size_t)(k * B + r) * STRIDE];
        return v;
    }
};

// ---- row sources: rows are addressed relative to the chunk (k = 0 first interior row, k = -1 the
// row before the chunk, k = Q-1 the separator)
// level 0: the staged tile S[fe][RP] (fe = f*BB + e; f = 0 sub, 1 diag, 2 super)
template <int B> struct StagedRows
{
    const double *Sb; // S + sidx(first row of the chunk) - 1
    CH_HD Mat<B> get(int f, int k) const
    {
        Mat<B> X;
        const int p = (k + 1) + (k + 1) / Dim<B>::Q0; // chunks start on a multiple of Q0
#pragma unroll
        for (int e = 0; e < B * B; ++e) X.a[e] = Sb[(size_t)(f * B * B + e) * Dim<B>::RP + p];
        return X;
    }
    CH_HD Mat<B> A(int k) const { return get(0, k); }
    CH_HD Mat<B> D(int k) const { return get(1, k); }
    CH_HD Mat<B> C(int k) const { return get(2, k); }
    CH_HD Mat<B> Hinit() const { return get(2, -1); }
};
template <int B> CH_HD StagedRows<B> staged_rows(const double *S, int chunk) { return StagedRows<B>{S + chunk * (Dim<B>::Q0 + 1)}; }

// levels 1 and 2: the four per-chunk results of the level below, arrays [e][COUNT]
template <int B, int COUNT> struct ReducedRows
{
    const double *G, *Df, *SL, *Hf;
    int i0; // first row of the chunk
    CH_HD Mat<B> ld(const double *p, int i) const
    {
        Mat<B> X;
#pragma unroll
        for (int e = 0; e < B * B; ++e) X.a[e] = p[e * COUNT + i];
        return X;
    }
    CH_HD Mat<B> A(int k) const { return ld(G, i0 + k); }
    CH_HD Mat<B> D(int k) const
    {
        Mat<B> X = ld(Df, i0 + k);
        if (i0 + k + 1 < COUNT) msub<B>(X, ld(SL, i0 + k + 1));
        return X;
    }
    CH_HD Mat<B> C(int k) const { return (i0 + k + 1 < COUNT) ? ld(Hf, i0 + k + 1) : mzero<B>(); }
    CH_HD Mat<B> Hinit() const { return ld(Hf, i0); } // = C(-1); chunk 0 carries the level below
};
template <int B> CH_HD void st(double *p, int count, int i, const Mat<B> &X)
{
#pragma unroll
    for (int e = 0; e < B * B; ++e) p[e * count + i] = X.a[e];
}
template <int B> CH_HD void stv(double *p, int count, int i, const Vec<B> &v)
{
#pragma unroll
    for (int r = 0; r < B; ++r) p[r * count + i] = v.a[r];
}
template <int B> CH_HD Vec<B> ldv(const double *p, int count, int i)
{
    Vec<B> v;
#pragma unroll
    for (int r = 0; r < B; ++r) v.a[r] = p[r * count + i];
    return v;
}

// ---- elimination of one chunk's interior (QI rows).  The rows of step k+1 are fetched before the
// arithmetic of step k and the factor stores come last, so that the only thing a step waits for is
// the previous pivot: reciprocal, one multiply, one fused multiply-add.
template <int B> CH_HD double mabssum(const Mat<B> &X)
{
    double m = 0.0;
#pragma unroll
    for (int e = 0; e < B * B; ++e) m += fabs(X.a[e]);
    return m;
}
// tol: absolute pivot threshold; rel: a pivot below rel x (sum of the magnitudes of the row's own diagonal and
// super-diagonal blocks before elimination) has been cancelled to noise and counts as zero as well.  (Sums, not
// maxima: a double-precision max is a multi-instruction sequence on this machine, an add is one.)
template <int B, int QI, class Rows, class Fac>
CH_HD void chunk_factor(const Rows &src, const Fac &dst, double tol, double rel, int &fail, Mat<B> &G, Mat<B> &Dk, Mat<B> &SL, Mat<B> &H)
{
    G = src.A(0);
    Dk = src.D(0);
    H = src.Hinit();
    SL = mzero<B>();
    Mat<B> C = src.C(0), An = src.A(1), Dn = src.D(1);
    double scale = mabssum<B>(Dk);
#pragma unroll
    for (int k = 0; k < QI; ++k)
    {
        Mat<B> C1, A2, D2;
        if (k + 1 < QI) { C1 = src.C(k + 1); A2 = src.A(k + 2); D2 = src.D(k + 2); }
        const Mat<B> Dinv = minv<B>(Dk, fma(rel, scale + mabssum<B>(C), tol), fail);
        scale = mabssum<B>(Dn);
        const Mat<B> L = mmul<B>(An, Dinv);
        const Mat<B> LL = mmul<B>(H, Dinv);
        Dk = Dn;
        msub<B>(Dk, mmul<B>(L, C));
        dst.put(k, FL, L);
        dst.put(k, FLL, LL);
        dst.put(k, FDINV, Dinv);
        dst.put(k, FC, C);
        dst.put(k, FG, G);
        madd<B>(SL, mmul<B>(LL, G));
        G = mneg<B>(mmul<B>(L, G));
        H = mneg<B>(mmul<B>(LL, C));
        if (k + 1 < QI) { C = C1; An = A2; Dn = D2; }
    }
}

// forward sweep of one chunk: y_k, the separator's partial rhs and the update tL of row L's rhs.
// PRE: multipliers and right-hand sides are fetched up front (HBM operands: the sweep is a chain of
// dependent FMAs that must not wait on memory at every step); otherwise they are read step by step
// (shared-memory operands of the small levels, where registers are the scarcer resource).
template <int B, int QI, bool PRE, class Rhs, class Fac, class YDst>
CH_HD void chunk_forward(const Rhs &rsrc, const Fac &fac, const YDst &ydst, Vec<B> &ysep, Vec<B> &tL)
{
    tL = vzero<B>();
    if (PRE)
    {
        Mat<B> L[QI], LL[QI];
        Vec<B> r[QI + 1];
#pragma unroll
        for (int k = 0; k < QI; ++k) { L[k] = fac.get(k, FL); LL[k] = fac.get(k, FLL); }
#pragma unroll
        for (int k = 0; k <= QI; ++k) r[k] = rsrc.r(k);
        Vec<B> yk = r[0];
#pragma unroll
        for (int k = 0; k < QI; ++k)
        {
            ydst.put(k, yk);
            vadd<B>(tL, mv<B>(LL[k], yk));
            const Vec<B> ly = mv<B>(L[k], yk);
            yk = r[k + 1];
            vsub<B>(yk, ly);
        }
        ysep = yk;
    }
    else
    {
        Vec<B> yk = rsrc.r(0);
#pragma unroll
        for (int k = 0; k < QI; ++k)
        {
            const Mat<B> L = fac.get(k, FL), LL = fac.get(k, FLL);
            Vec<B> rn = rsrc.r(k + 1);
            ydst.put(k, yk);
            vadd<B>(tL, mv<B>(LL, yk));
            vsub<B>(rn, mv<B>(L, yk));
            yk = rn;
        }
        ysep = yk;
    }
}

// backward sweep of one chunk; xdst(k, x) receives the interior unknowns, last row first
template <int B, int QI, bool PRE, class Fac, class YSrc, class XDst>
CH_HD void chunk_backward(const Fac &fac, const YSrc &ysrc, const Vec<B> &xR, const Vec<B> &xL, XDst &xdst)
{
    Vec<B> xn = xR;
    if (PRE)
    {
        Mat<B> Di[QI], C[QI];
        Vec<B> y[QI];
#pragma unroll
        for (int k = 0; k < QI; ++k)
        {
            Di[k] = fac.get(k, FDINV);
            C[k] = fac.get(k, FC);
            y[k] = ysrc.get(k);
            vsub<B>(y[k], mv<B>(fac.get(k, FG), xL)); // the spike term does not depend on the sweep
        }
#pragma unroll
        for (int k = QI - 1; k >= 0; --k)
        {
            vsub<B>(y[k], mv<B>(C[k], xn));
            xn = mv<B>(Di[k], y[k]);
            xdst.put(k, xn);
        }
    }
    else
    {
#pragma unroll
        for (int k = QI - 1; k >= 0; --k)
        {
            const Mat<B> Di = fac.get(k, FDINV), C = fac.get(k, FC);
            Vec<B> y = ysrc.get(k);
            vsub<B>(y, mv<B>(fac.get(k, FG), xL));
            vsub<B>(y, mv<B>(C, xn));
            xn = mv<B>(Di, y);
            xdst.put(k, xn);
        }
    }
}

// ---- shared-memory images of one CTA -------------------------------------------------------
template <int B> struct FactorShared
{
    using D = Dim<B>;
    double S[3 * D::BB * D::RP];
    // The per-chunk results of levels 0 and 1 reuse the staged tile: level 0 keeps its results in
    // registers across one barrier before storing them (factor_phase0 / factor_store1).
    CH_HD double *G1() { return S; }
    CH_HD double *D1() { return S + D::BB * D::T; }
    CH_HD double *SL1() { return S + 2 * D::BB * D::T; }
    CH_HD double *H1() { return S + 3 * D::BB * D::T; }
    CH_HD double *G2() { return S + 4 * D::BB * D::T; }
    CH_HD double *D2() { return S + 4 * D::BB * D::T + D::BB * D::Q2; }
    CH_HD double *SL2() { return S + 4 * D::BB * D::T + 2 * D::BB * D::Q2; }
    CH_HD double *H2() { return S + 4 * D::BB * D::T + 3 * D::BB * D::Q2; }
    static_assert(4 * D::BB * (D::T + D::Q2) <= 3 * D::BB * D::RP, "level arrays must fit in the staged tile");
};
template <int B> struct SolveShared
{
    using D = Dim<B>;
    double R[B * D::RP]; // staged rhs, later the solution of the tile's rows
    double Y1[B * D::T], T1[B * D::T], X1[B * D::T];
    double Y2[B * D::Q2], T2[B * D::Q2], X2[B * D::Q2];
    double F12[D::F1 + D::F2 + D::BB]; // factors of the tile's levels 1 and 2 (+ top pivot inverse)
    double Y12[D::Y1 + D::Y2];         // forward results of levels 1 and 2
    double xL[B], xR[B];
};
// cooperative contiguous copy, loads batched ahead of the stores
CH_HD void copy_small(int t, int nt, double *dst, const double *src, int count)
{
    constexpr int U = 8;
    for (int i0 = t; i0 < count; i0 += nt * U)
    {
        double v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = i0 + u * nt; v[u] = i < count ? src[i] : 0.0; }
#pragma unroll
        for (int u = 0; u < U; ++u) { const int i = i0 + u * nt; if (i < count) dst[i] = v[u]; }
    }
}
template <int B> struct StagedRhs
{
    const double *Rb; // R + sidx(first row of the chunk) - 1
    CH_HD Vec<B> r(int k) const
    {
        Vec<B> v;
        const int p = (k + 1) + (k + 1) / Dim<B>::Q0;
#pragma unroll
        for (int q = 0; q < B; ++q) v.a[q] = Rb[(size_t)q * Dim<B>::RP + p];
        return v;
    }
};
template <int B, int COUNT> struct ReducedRhs
{
    const double *Y, *Tl;
    int i0;
    CH_HD Vec<B> r(int k) const
    {
        Vec<B> v = ldv<B>(Y, COUNT, i0 + k);
        if (i0 + k + 1 < COUNT) vsub<B>(v, ldv<B>(Tl, COUNT, i0 + k + 1));
        return v;
    }
};
template <int B> struct StagedX
{
    double *Rb;
    CH_HD void put(int k, const Vec<B> &x)
    {
        const int p = (k + 1) + (k + 1) / Dim<B>::Q0;
#pragma unroll
        for (int q = 0; q < B; ++q) Rb[(size_t)q * Dim<B>::RP + p] = x.a[q];
    }
};
template <int B> struct ArrayX
{
    double *X;
    int count, i0;
    CH_HD void put(int k, const Vec<B> &x) { stv<B>(X, count, i0 + k, x); }
};

// ---- staging (cooperative: thread t of nt) ---------------------------------------------------
CH_HD double mul_rn(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dmul_rn(a, b);
#else
    volatile double r = a * b;
    return r;
#endif
}
CH_HD double sub_rn(double a, double b)
{
#if defined(__CUDA_ARCH__)
    return __dsub_rn(a, b);
#else
    return a - b;
#endif
}
enum { TF_ANY = 1, TF_ONE = 2, TF_MASS = 4, TF_AFF = 8 };
// level 0: Jb = J - (M*alpha) (two roundings, dae-cpp/solver.hpp:523) gathered through the slot map
template <int B>
CH_HD void stage_matrix_first(int t, int nt, int tile, double *S, const int32_t *sj, const int32_t *sm, const int32_t *flags,
                              const double *jv, const double *mvals, double alpha)
{
    using D = Dim<B>;
    constexpr int PER = D::ROWS + 1, TOTAL = 3 * D::BB * PER, U = 12;
    const size_t base = (size_t)tile * TOTAL;
    const int32_t *fl = flags + tile * 3 * D::BB;
    // the map of a tile is one contiguous run: every load below is coalesced; all loads of a batch
    // are issued before the first shared-memory store so that they overlap
    for (int i0 = t; i0 < TOTAL; i0 += nt * U)
    {
        int32_t s[U], q[U];
        double j[U], m[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * nt;
            const bool ok = idx < TOTAL;
            s[u] = ok ? sj[base + idx] : -1;
            const int f = ok ? fl[idx / PER] : 0;
            q[u] = (f & TF_MASS) ? sm[base + idx] : -1;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            j[u] = s[u] >= 0 ? jv[s[u]] : (s[u] == -2 ? 1.0 : 0.0);
            m[u] = q[u] >= 0 ? mvals[q[u]] : 0.0;
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * nt;
            if (idx < TOTAL)
            {
                const int fe = idx / PER, rr = idx - fe * PER;
                S[(size_t)fe * D::RP + sidx<B>(rr - 1)] = sub_rn(j[u], q[u] >= 0 ? mul_rn(m[u], alpha) : 0.0);
            }
        }
    }
}
// level 0, affine tiles: when the value source of every field of the tile is an arithmetic
// progression in the row (the normal case: a functor that emits its triplets row by row, e.g.
// JacobianMatrixShape) the slot map is not read at all.  Per (tile, field element) a descriptor of
// DESC ints: {J index extrapolated to row 0 of the tile, J stride, M slot extrapolated, M stride,
// first / last row with a J entry, first / last row with an M entry, first row of the identity
// padding behind the end of the matrix}.  Rows outside the ranges are structural zeros (the matrix
// edge: no sub-diagonal in the first row, no super-diagonal in the last), so the boundary tiles are
// affine too.  All loads of a batch are independent.
constexpr int DESC = 10;
enum { DJ0 = 0, DJS = 1, DM0 = 2, DMS = 3, DJLO = 4, DJHI = 5, DMLO = 6, DMHI = 7, DONE = 8 };

template <int B>
CH_HD void stage_matrix_affine(int t, int nt, int tile, double *S, const int32_t *desc, const double *jv, const double *mvals,
                               double alpha)
{
    using D = Dim<B>;
    constexpr int PER = D::ROWS + 1, NFE = 3 * D::BB, U = (B == 1) ? 4 : (B == 2 ? 2 : 1);
    const int32_t *d = desc + (size_t)tile * NFE * DESC;
    for (int r0 = t; r0 < PER; r0 += nt * U)
    {
        double j[U][NFE], m[U][NFE];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int rr = r0 + u * nt;
#pragma unroll
            for (int fe = 0; fe < NFE; ++fe)
            {
                const int32_t *e = d + fe * DESC;
                j[u][fe] = (rr >= e[DJLO] && rr <= e[DJHI]) ? jv[e[DJ0] + e[DJS] * rr] : ((rr >= e[DONE] && rr < PER) ? 1.0 : 0.0);
                m[u][fe] = (rr >= e[DMLO] && rr <= e[DMHI]) ? mvals[e[DM0] + e[DMS] * rr] : 0.0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int rr = r0 + u * nt;
            if (rr >= PER) continue;
            const int p = sidx<B>(rr - 1);
#pragma unroll
            for (int fe = 0; fe < NFE; ++fe)
            {
                const int32_t *e = d + fe * DESC;
                S[(size_t)fe * D::RP + p] = sub_rn(j[u][fe], (rr >= e[DMLO] && rr <= e[DMHI]) ? mul_rn(m[u][fe], alpha) : 0.0);
            }
        }
    }
}
// levels >= 1: rows of the reduced system the level below produced, red = [4][rows][BB]
template <int B> CH_HD void stage_matrix_reduced(int t, int nt, int tile, int rows, double *S, const double *red)
{
    using D = Dim<B>;
    constexpr int U = (B == 1) ? 6 : 2;
    for (int r0 = t; r0 <= D::ROWS; r0 += nt * U)
    {
        double a[U][D::BB], d[U][D::BB], d2[U][D::BB], c[U][D::BB];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int i = tile * D::ROWS + r0 + u * nt - 1;
            const bool in = i >= 0 && i < rows, nx = in && i + 1 < rows;
#pragma unroll
            for (int e = 0; e < D::BB; ++e)
            {
                a[u][e] = in ? red[((size_t)0 * rows + i) * D::BB + e] : 0.0;
                d[u][e] = in ? red[((size_t)1 * rows + i) * D::BB + e] : ((i >= rows && e / B == e % B) ? 1.0 : 0.0);
                d2[u][e] = nx ? red[((size_t)2 * rows + i + 1) * D::BB + e] : 0.0;
                c[u][e] = nx ? red[((size_t)3 * rows + i + 1) * D::BB + e] : 0.0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int rr = r0 + u * nt;
            if (rr > D::ROWS) continue;
            const int p = sidx<B>(rr - 1);
#pragma unroll
            for (int e = 0; e < D::BB; ++e)
            {
                S[(size_t)(0 * D::BB + e) * D::RP + p] = a[u][e];
                S[(size_t)(1 * D::BB + e) * D::RP + p] = d[u][e] - d2[u][e];
                S[(size_t)(2 * D::BB + e) * D::RP + p] = c[u][e];
            }
        }
    }
}
template <int B>
CH_HD void stage_rhs_first(int t, int nt, int tile, int n, double *R, const double *b, const int32_t *perm)
{
    using D = Dim<B>;
    constexpr int TOTAL = B * D::ROWS, U = 8;
    for (int i0 = t; i0 < TOTAL; i0 += nt * U)
    {
        int64_t o[U];
        double v[U];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * nt;
            const int q = idx / D::ROWS, row = idx - q * D::ROWS;
            const int64_t k = ((int64_t)tile * D::ROWS + row) * B + q;
            o[u] = (idx < TOTAL && k < n) ? (perm ? (int64_t)perm[k] : k) : -1;
        }
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = o[u] >= 0 ? b[o[u]] : 0.0;
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int idx = i0 + u * nt;
            if (idx < TOTAL)
            {
                const int q = idx / D::ROWS, row = idx - q * D::ROWS;
                R[(size_t)q * D::RP + sidx<B>(row)] = v[u];
            }
        }
    }
}
template <int B> CH_HD void stage_rhs_reduced(int t, int nt, int tile, int rows, double *R, const double *rv)
{
    using D = Dim<B>;
    constexpr int U = (B == 1) ? 8 : 4;
    for (int r0 = t; r0 < D::ROWS; r0 += nt * U)
    {
        double y[U][B], tl[U][B];
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int i = tile * D::ROWS + r0 + u * nt;
            const bool in = r0 + u * nt < D::ROWS && i < rows, nx = in && i + 1 < rows;
#pragma unroll
            for (int q = 0; q < B; ++q)
            {
                y[u][q] = in ? rv[((size_t)0 * rows + i) * B + q] : 0.0;
                tl[u][q] = nx ? rv[((size_t)1 * rows + i + 1) * B + q] : 0.0;
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
        {
            const int row = r0 + u * nt;
            if (row >= D::ROWS) continue;
#pragma unroll
            for (int q = 0; q < B; ++q) R[(size_t)q * D::RP + sidx<B>(row)] = y[u][q] - tl[u][q];
        }
    }
}

// ---- the phases of one tile (thread index t; a block-wide barrier separates phases) ----------
// nvalid: rows of the tile that exist at this level (the rest is identity padding whose chunks
// are skipped consistently by factor and solve).
template <int B>
CH_HD void factor_phase0(int t, int nvalid, FactorShared<B> &sh, double *Ftile, double tol, double rel, int &fail, Mat<B> &G, Mat<B> &Dk,
                         Mat<B> &SL, Mat<B> &H)
{
    using D = Dim<B>;
    G = mzero<B>(); Dk = mident<B>(); SL = mzero<B>(); H = mzero<B>();
    if (t * D::Q0 < nvalid)
        chunk_factor<B, D::Q0 - 1>(staged_rows<B>(sh.S, t), FacRef<B, D::T>{Ftile + t}, tol, rel, fail, G, Dk, SL, H);
}
// after a barrier: every thread is done reading the staged tile
template <int B>
CH_HD void factor_store1(int t, FactorShared<B> &sh, const Mat<B> &G, const Mat<B> &Dk, const Mat<B> &SL, const Mat<B> &H)
{
    using D = Dim<B>;
    st<B>(sh.G1(), D::T, t, G); st<B>(sh.D1(), D::T, t, Dk); st<B>(sh.SL1(), D::T, t, SL); st<B>(sh.H1(), D::T, t, H);
}
template <int B> CH_HD void factor_phase1(int c, int nvalid, FactorShared<B> &sh, double *Ftile, double tol, double rel, int &fail)
{
    using D = Dim<B>;
    if (c >= D::Q2) return;
    Mat<B> G = mzero<B>(), Dk = mident<B>(), SL = mzero<B>(), H = mzero<B>();
    if (c * D::Q1 * D::Q0 < nvalid)
    {
        ReducedRows<B, D::T> src{sh.G1(), sh.D1(), sh.SL1(), sh.H1(), c * D::Q1};
        chunk_factor<B, D::Q1 - 1>(src, FacRef<B, D::Q2>{Ftile + D::F0 + c}, tol, rel, fail, G, Dk, SL, H);
    }
    st<B>(sh.G2(), D::Q2, c, G); st<B>(sh.D2(), D::Q2, c, Dk); st<B>(sh.SL2(), D::Q2, c, SL); st<B>(sh.H2(), D::Q2, c, H);
}
// thread 0 only.  red: this level's reduced system [4][ntiles][BB] (G, Df, SL, Hf).
template <int B>
CH_HD void factor_phase2(int tile, int ntiles, bool is_top, FactorShared<B> &sh, double *Ftile, double *red, double tol, double rel,
                         int &fail)
{
    using D = Dim<B>;
    Mat<B> G, Dk, SL, H;
    ReducedRows<B, D::Q2> src{sh.G2(), sh.D2(), sh.SL2(), sh.H2(), 0};
    chunk_factor<B, D::Q2 - 1>(src, FacRef<B, 1>{Ftile + D::F0 + D::F1}, tol, rel, fail, G, Dk, SL, H);
    // updates of the left tile's separator collected from the first chunk of every level
    for (int e = 0; e < B * B; ++e) SL.a[e] += sh.SL1()[e * D::T] + sh.SL2()[e * D::Q2];
    if (is_top)
    {
        const Mat<B> Dinv = minv<B>(Dk, tol, fail);
        for (int e = 0; e < B * B; ++e) Ftile[D::F0 + D::F1 + D::F2 + e] = Dinv.a[e];
    }
    for (int e = 0; e < B * B; ++e)
    {
        red[(size_t)(0 * ntiles + tile) * D::BB + e] = G.a[e];
        red[(size_t)(1 * ntiles + tile) * D::BB + e] = Dk.a[e];
        red[(size_t)(2 * ntiles + tile) * D::BB + e] = SL.a[e];
        red[(size_t)(3 * ntiles + tile) * D::BB + e] = H.a[e];
    }
}

// ---- forward solve --------------------------------------------------------------------------
// Before phase 0 the CTA copies the tile's level-1/2 factors into sh.F12 (load_small_factors);
// the forward results of those levels live in sh.Y12 (spilled to HBM between the forward and the
// backward kernel, see chain.cu).
template <int B> CH_HD void load_small_factors(int t, int nt, SolveShared<B> &sh, const double *Ftile)
{
    using D = Dim<B>;
    copy_small(t, nt, sh.F12, Ftile + D::F0, D::F1 + D::F2 + D::BB);
}
template <int B> CH_HD void forward_phase0(int t, int nvalid, SolveShared<B> &sh, double *Ftile, double *Ytile)
{
    using D = Dim<B>;
    Vec<B> ysep = vzero<B>(), tL = vzero<B>();
    if (t * D::Q0 < nvalid)
        chunk_forward<B, D::Q0 - 1, true>(StagedRhs<B>{sh.R + t * (D::Q0 + 1)}, FacRef<B, D::T>{Ftile + t}, YRef<B, D::T>{Ytile + t}, ysep,
                                    tL);
    stv<B>(sh.Y1, D::T, t, ysep); stv<B>(sh.T1, D::T, t, tL);
}
template <int B> CH_HD void forward_phase1(int c, int nvalid, SolveShared<B> &sh)
{
    using D = Dim<B>;
    if (c >= D::Q2) return;
    Vec<B> ysep = vzero<B>(), tL = vzero<B>();
    if (c * D::Q1 * D::Q0 < nvalid)
        chunk_forward<B, D::Q1 - 1, B == 1>(ReducedRhs<B, D::T>{sh.Y1, sh.T1, c * D::Q1}, FacRef<B, D::Q2>{sh.F12 + c},
                                    YRef<B, D::Q2>{sh.Y12 + c}, ysep, tL);
    stv<B>(sh.Y2, D::Q2, c, ysep); stv<B>(sh.T2, D::Q2, c, tL);
}
// thread 0 only.  rv: this level's reduced rhs [2][ntiles][B] (separator rhs, update of the left tile's).
template <int B> CH_HD void forward_phase2(int tile, int ntiles, bool is_top, SolveShared<B> &sh, double *rv)
{
    using D = Dim<B>;
    Vec<B> ysep, tL;
    chunk_forward<B, D::Q2 - 1, B == 1>(ReducedRhs<B, D::Q2>{sh.Y2, sh.T2, 0}, FacRef<B, 1>{sh.F12 + D::F1}, YRef<B, 1>{sh.Y12 + D::Y1},
                                ysep, tL);
    for (int q = 0; q < B; ++q) tL.a[q] += sh.T1[q * D::T] + sh.T2[q * D::Q2];
    for (int q = 0; q < B; ++q)
    {
        rv[(size_t)(0 * ntiles + tile) * B + q] = ysep.a[q];
        rv[(size_t)(1 * ntiles + tile) * B + q] = tL.a[q];
    }
    if (is_top)
    {
        Mat<B> Dinv;
        for (int e = 0; e < B * B; ++e) Dinv.a[e] = sh.F12[D::F1 + D::F2 + e];
        const Vec<B> x = mv<B>(Dinv, ysep);
        for (int q = 0; q < B; ++q) { sh.xR[q] = x.a[q]; sh.xL[q] = 0.0; }
    }
}

// ---- backward solve -------------------------------------------------------------------------
template <int B> CH_HD void backward_phase2(SolveShared<B> &sh)
{
    using D = Dim<B>;
    Vec<B> xR, xL;
    for (int q = 0; q < B; ++q) { xR.a[q] = sh.xR[q]; xL.a[q] = sh.xL[q]; }
    stv<B>(sh.X2, D::Q2, D::Q2 - 1, xR);
    ArrayX<B> xd{sh.X2, D::Q2, 0};
    chunk_backward<B, D::Q2 - 1, B == 1>(FacRef<B, 1>{sh.F12 + D::F1}, YRef<B, 1>{sh.Y12 + D::Y1}, xR, xL, xd);
}
template <int B> CH_HD void backward_phase1(int c, int nvalid, SolveShared<B> &sh)
{
    using D = Dim<B>;
    if (c >= D::Q2) return;
    const Vec<B> xR = ldv<B>(sh.X2, D::Q2, c);
    Vec<B> xL;
    if (c > 0) xL = ldv<B>(sh.X2, D::Q2, c - 1);
    else for (int q = 0; q < B; ++q) xL.a[q] = sh.xL[q];
    stv<B>(sh.X1, D::T, c * D::Q1 + D::Q1 - 1, xR);
    ArrayX<B> xd{sh.X1, D::T, c * D::Q1};
    if (c * D::Q1 * D::Q0 < nvalid)
        chunk_backward<B, D::Q1 - 1, B == 1>(FacRef<B, D::Q2>{sh.F12 + c}, YRef<B, D::Q2>{sh.Y12 + c}, xR, xL, xd);
    else
        for (int k = 0; k < D::Q1 - 1; ++k) xd.put(k, vzero<B>());
}
template <int B> CH_HD void backward_phase0(int t, int nvalid, SolveShared<B> &sh, double *Ftile, double *Ytile)
{
    using D = Dim<B>;
    const Vec<B> xR = ldv<B>(sh.X1, D::T, t);
    Vec<B> xL;
    if (t > 0) xL = ldv<B>(sh.X1, D::T, t - 1);
    else for (int q = 0; q < B; ++q) xL.a[q] = sh.xL[q];
    StagedX<B> xd{sh.R + t * (D::Q0 + 1)};
    xd.put(D::Q0 - 1, xR);
    if (t * D::Q0 < nvalid)
        chunk_backward<B, D::Q0 - 1, true>(FacRef<B, D::T>{Ftile + t}, YRef<B, D::T>{Ytile + t}, xR, xL, xd);
}

} // namespace chain
} // namespace daecpp_b200
