// BDF time-derivative operands shared by the residual kernels (vector_ops.cu) and the chain path's
// fused residual (chain.cu).  dxdt = alpha*xk - c1*x[0] + c2*x[1] - c3*x[2] + c4*x[3]
// (dae-cpp/solver.hpp:141-217), written with __dmul_rn/__dadd_rn/__dsub_rn: no FMA contraction.
#pragma once
#include <cstdint>

namespace daecpp_b200
{
struct Bdf
{
    double c[5];
    const double *xk, *x0, *x1, *x2, *x3;
    int order;
};

#ifdef __CUDACC__
__device__ __forceinline__ double bdf_one(const Bdf &s, double xk, double x0, double x1, double x2, double x3)
{
    if (s.order == 1) return __dmul_rn(__dsub_rn(xk, x0), s.c[0]);
    double r = __dmul_rn(s.c[0], xk);
    r = __dsub_rn(r, __dmul_rn(s.c[1], x0));
    r = __dadd_rn(r, __dmul_rn(s.c[2], x1));
    if (s.order >= 3) r = __dsub_rn(r, __dmul_rn(s.c[3], x2));
    if (s.order >= 4) r = __dadd_rn(r, __dmul_rn(s.c[4], x3));
    return r;
}
__device__ __forceinline__ double bdf_at(const Bdf &s, int64_t i)
{
    if (s.order == 1) return __dmul_rn(__dsub_rn(s.xk[i], s.x0[i]), s.c[0]);
    double r = __dmul_rn(s.c[0], s.xk[i]);
    r = __dsub_rn(r, __dmul_rn(s.c[1], s.x0[i]));
    r = __dadd_rn(r, __dmul_rn(s.c[2], s.x1[i]));
    if (s.order >= 3) r = __dsub_rn(r, __dmul_rn(s.c[3], s.x2[i]));
    if (s.order >= 4) r = __dadd_rn(r, __dmul_rn(s.c[4], s.x3[i]));
    return r;
}
#endif
} // namespace daecpp_b200
