// Block-chain plan: once-per-pattern host analysis for matrices whose graph is a chain of small
// blocks (block-tridiagonal after a symmetric permutation): the banded Jacobians of 1-D problems
// -- flame-1D (block order 1), perovskite (block order 2 after interleaving P and Phi).
//
// Role replaced: Eigen::SparseLU::analyzePattern (Eigen/src/SparseLU/SparseLU.h:528-587) for
// this pattern class.  Where the general path (symbolic.cpp) ships per-front index tables to the
// device, a chain needs none: the elimination order is a fixed radix hierarchy
// (thread chunk -> CTA -> grid level), so the numeric kernels (chain.cu) move only matrix and
// factor values plus one slot map of the matrix entries.
#pragma once
#include <cstdint>
#include <vector>

namespace daecpp_b200
{
namespace chain
{

constexpr int MAX_B = 3;

struct Level
{
    int32_t rows = 0;    // block rows of this level's (block-)tridiagonal system
    int32_t ntiles = 0;  // CTAs = rows of the next level
    int64_t fac0 = 0;    // first double of this level's factor storage
    int64_t y0 = 0;      // first double of this level's forward-solve storage
    int64_t red0 = 0;    // first double of the reduced system this level produces [4][ntiles][B*B]
    int64_t rv0 = 0;     // first double of the reduced rhs this level produces [2][ntiles][B]
    int64_t x0 = 0;      // first double of this level's solution (levels >= 1) [rows][B]
};

struct Plan
{
    bool active = false;
    int32_t n = 0, B = 0;
    int32_t m = 0;              // block rows at level 0 = ceil(n / B)
    bool identity_perm = true;
    std::vector<int32_t> perm;  // perm[k] = original index at chain position k (size n)
    std::vector<Level> levels;  // levels.back() is the single-tile top level
    // slot map of level 0: for tile, field-element fe (f*B*B + r*B + c; f = 0 sub, 1 diag,
    // 2 super), staged row rr in [0, ROWS] (rr = 0 is the halo row before the tile):
    //   ent_u[(tile*3*B*B + fe)*(ROWS+1) + rr] = slot of the analysed pattern, -1 absent, -2 one
    std::vector<int32_t> ent_u;
    std::vector<int32_t> tile_flags; // [tile*3*B*B + fe]  bit0: any slot present, bit1: any -2
    int64_t fac_doubles = 0, y_doubles = 0, red_doubles = 0, rv_doubles = 0, x_doubles = 0;
    int64_t nnz_factor = 0;
};

// Try to recognise an n x n compressed-column pattern as a block chain of order <= max_b.
// Returns false (plan.active = false) when it is not one.
bool detect(int32_t n, const int32_t *outer, const int32_t *inner, int max_b, Plan &plan);

// geometry shared by host and device
int rows_per_tile(int B);
int tile_factor_doubles(int B);
int tile_y_doubles(int B);

} // namespace chain
} // namespace daecpp_b200
