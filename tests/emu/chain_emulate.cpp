// TEST INFRASTRUCTURE ONLY.  CPU walk over the block-chain plan (dae-cpp_b200/csrc/chain_host.cpp)
// executing, tile by tile and phase by phase, the SAME per-thread functions the CUDA kernels of
// chain.cu run (dae-cpp_b200/csrc/chain_ops.hpp), with a loop over the thread index in place of
// the CTA.  Validates detection, ordering, the slot map and the elimination hierarchy in the
// GPU-less build container; it is never linked into libdaecpp_b200.so.
#include <cstring>
#include <memory>
#include <vector>

#include "../../dae-cpp_b200/csrc/chain.hpp"
#include "../../dae-cpp_b200/csrc/chain_ops.hpp"

using namespace daecpp_b200::chain;

namespace
{

template <int B>
int run(const Plan &P, const double *vals, const double *b, double *x, int nrhs_repeat)
{
    using D = Dim<B>;
    const int nlev = (int)P.levels.size();
    std::vector<double> F((size_t)P.fac_doubles, 1e300), Y((size_t)P.y_doubles, 1e300), RED((size_t)P.red_doubles, 0.0),
        RV((size_t)P.rv_doubles, 0.0), X((size_t)P.x_doubles, 0.0);
    // compose: the analysed pattern is the matrix itself (sj = slot, no mass matrix)
    std::vector<int32_t> sm(P.ent_u.size(), -1);
    int fail = 0;
    auto fsh = std::make_unique<FactorShared<B>>();
    auto ssh = std::make_unique<SolveShared<B>>();

    for (int g = 0; g < nlev; ++g)
    {
        const Level &L = P.levels[g];
        const bool top = (g == nlev - 1);
        for (int tile = 0; tile < L.ntiles; ++tile)
        {
            std::memset(fsh.get(), 0, sizeof(FactorShared<B>));
            double *Ft = F.data() + L.fac0 + (size_t)tile * D::TILE_F;
            const int nvalid = std::min(D::ROWS, L.rows - tile * D::ROWS);
            for (int t = 0; t < D::T; ++t)
            {
                if (g == 0) stage_matrix_first<B>(t, D::T, tile, fsh->S, P.ent_u.data(), sm.data(), P.tile_flags.data(), vals, nullptr, 0.0);
                else stage_matrix_reduced<B>(t, D::T, tile, L.rows, fsh->S, RED.data() + P.levels[g - 1].red0);
            }
            std::vector<Mat<B>> rG(D::T), rD(D::T), rS(D::T), rH(D::T);
            for (int t = 0; t < D::T; ++t) factor_phase0<B>(t, nvalid, *fsh, Ft, 0.0, 1e-14, fail, rG[t], rD[t], rS[t], rH[t]);
            for (int t = 0; t < D::T; ++t) factor_store1<B>(t, *fsh, rG[t], rD[t], rS[t], rH[t]);
            for (int t = 0; t < D::T; ++t) factor_phase1<B>(t, nvalid, *fsh, Ft, 0.0, 1e-14, fail);
            factor_phase2<B>(tile, L.ntiles, top, *fsh, Ft, RED.data() + L.red0, 0.0, 1e-14, fail);
        }
    }
    if (fail) return 2;

    for (int rep = 0; rep < nrhs_repeat; ++rep)
    {
        for (int g = 0; g < nlev; ++g)
        {
            const Level &L = P.levels[g];
            const bool top = (g == nlev - 1);
            for (int tile = 0; tile < L.ntiles; ++tile)
            {
                double *Ft = F.data() + L.fac0 + (size_t)tile * D::TILE_F;
                double *Yt = Y.data() + L.y0 + (size_t)tile * D::TILE_Y;
                const int nvalid = std::min(D::ROWS, L.rows - tile * D::ROWS);
                for (int t = 0; t < D::T; ++t)
                {
                    if (g == 0) stage_rhs_first<B>(t, D::T, tile, P.n, ssh->R, b, P.identity_perm ? nullptr : P.perm.data());
                    else stage_rhs_reduced<B>(t, D::T, tile, L.rows, ssh->R, RV.data() + P.levels[g - 1].rv0);
                }
                for (int t = 0; t < D::T; ++t) load_small_factors<B>(t, D::T, *ssh, Ft);
                for (int t = 0; t < D::T; ++t) forward_phase0<B>(t, nvalid, *ssh, Ft, Yt);
                for (int t = 0; t < D::T; ++t) forward_phase1<B>(t, nvalid, *ssh);
                forward_phase2<B>(tile, L.ntiles, top, *ssh, RV.data() + L.rv0);
                for (int t = 0; t < D::T; ++t) copy_small(t, D::T, Yt + D::Y0, ssh->Y12, D::Y1 + D::Y2);
            }
        }
        for (int g = nlev - 1; g >= 0; --g)
        {
            const Level &L = P.levels[g];
            const bool top = (g == nlev - 1);
            for (int tile = 0; tile < L.ntiles; ++tile)
            {
                double *Ft = F.data() + L.fac0 + (size_t)tile * D::TILE_F;
                double *Yt = Y.data() + L.y0 + (size_t)tile * D::TILE_Y;
                const int nvalid = std::min(D::ROWS, L.rows - tile * D::ROWS);
                if (!top)
                {
                    const double *xup = X.data() + P.levels[g + 1].x0;
                    for (int q = 0; q < B; ++q)
                    {
                        ssh->xR[q] = xup[(size_t)tile * B + q];
                        ssh->xL[q] = tile > 0 ? xup[(size_t)(tile - 1) * B + q] : 0.0;
                    }
                }
                // (top: xR/xL were left in ssh by forward_phase2 of the single top tile)
                for (int t = 0; t < D::T; ++t) load_small_factors<B>(t, D::T, *ssh, Ft);
                for (int t = 0; t < D::T; ++t) copy_small(t, D::T, ssh->Y12, Yt + D::Y0, D::Y1 + D::Y2);
                backward_phase2<B>(*ssh);
                for (int t = 0; t < D::T; ++t) backward_phase1<B>(t, nvalid, *ssh);
                for (int t = 0; t < D::T; ++t) backward_phase0<B>(t, nvalid, *ssh, Ft, Yt);
                for (int row = 0; row < nvalid; ++row)
                    for (int q = 0; q < B; ++q)
                    {
                        const double v = ssh->R[(size_t)q * D::RP + sidx<B>(row)];
                        const int64_t k = ((int64_t)tile * D::ROWS + row) * B + q;
                        if (g == 0) { if (k < P.n) x[P.perm[(size_t)k]] = v; }
                        else X[(size_t)L.x0 + k] = v;
                    }
            }
        }
    }
    return 0;
}

} // namespace

extern "C" int chainemu_solve(int32_t n, const int32_t *outer, const int32_t *inner, const double *vals, const double *b,
                              double *x, int32_t max_b, double *stats /* >= 8 */)
{
    Plan P;
    if (!detect(n, outer, inner, max_b, P)) return 1;
    if (stats)
    {
        stats[0] = P.B; stats[1] = P.m; stats[2] = (double)P.levels.size(); stats[3] = P.identity_perm ? 1 : 0;
        stats[4] = (double)P.fac_doubles; stats[5] = P.levels[0].ntiles; stats[6] = (double)P.nnz_factor; stats[7] = 0;
    }
    int rc;
    if (P.B == 1) rc = run<1>(P, vals, b, x, 1);
    else if (P.B == 2) rc = run<2>(P, vals, b, x, 1);
    else rc = run<3>(P, vals, b, x, 1);
    return rc ? 10 + rc : 0;
}
