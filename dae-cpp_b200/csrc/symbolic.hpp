// Once-per-pattern symbolic analysis for the multifrontal sparse LU (host side).
//
// Plays the role Eigen::SparseLU::analyzePattern plays in the reference
// (Eigen/src/SparseLU/SparseLU.h:528-587: COLAMD + column etree + postorder), but is built
// for a GPU numeric phase: a nested-dissection ordering that keeps the elimination tree
// shallow (O(log n) levels on chains and grids), a static (pivot-free) symbolic factorisation
// of pattern(A + A^T), supernodes = frontal matrices, a level schedule of the assembly tree
// and the scatter map from assembled-matrix slots into frontal storage.  It is computed once
// and reused by every numeric refactorisation until the sparsity pattern changes -- the
// reference redoes its analysis at every factorisation (dae-cpp/solver.hpp:537).
#pragma once
#include <cstdint>
#include <vector>

namespace daecpp_b200
{

struct SymbolicOptions
{
    int leaf_size = 2;      // stop dissecting below this many vertices
    int relax_zeros = 0;    // reserved: relaxed amalgamation budget
    int max_super = 256;    // widest supernode (pivot block) allowed
    int unit_smem_bytes = -1; // shared-memory budget of one subtree unit (-1: chosen from the problem
                              // size; 0: no units, every front goes through the per-level kernels)
    int serial_dim = 4;     // fronts up to this order are factored by one thread each
    int group8_dim = 8;     // local levels whose largest front is at most this order are swept eight lanes per front
    int warp_dim = 16;      // local levels of a unit whose largest front is at most this order are swept one warp per front
    int warp_dim_many = 16; // ... or at most this order when the level has at least warp_min_fronts fronts (enough to keep
    int warp_min_fronts = 4; //    the block's warps busy); larger fronts take the whole block, one after the other
    const double *weights = nullptr; // optional |a_ij| per slot of the analysed pattern: the transversal prefers big entries
    int transversal = 1;    // permute rows to a structurally zero-free diagonal first when the diagonal has holes
};

struct LaunchGroup
{
    int32_t kind = 0;           // 0: subtree units [begin,end) handed out by ticket; 1: big fronts big_sn[begin,end)
    int32_t begin = 0, end = 0;
    int32_t max_dim = 0;        // kind 1: largest front order in the group
    int32_t same_stage = 0;     // kind 0: every unit of the group is in the stage of the big-front group that follows it (independent of it)
};

struct Symbolic
{
    int32_t n = 0;
    std::vector<int32_t> perm;  // perm[k]  = original index eliminated k-th
    std::vector<int32_t> iperm; // iperm[v] = elimination position of original index v

    // supernodes / fronts, numbered in elimination order
    int32_t nsn = 0;
    std::vector<int32_t> sn_first;  // [nsn+1] first pivot column (permuted numbering)
    std::vector<int64_t> sn_rowptr; // [nsn+1] offsets into sn_rows (also offsets of update vectors)
    std::vector<int32_t> sn_rows;   // permuted indices below the pivot block, ascending
    std::vector<int32_t> sn_rel;    // same indexing: position of that row in the PARENT's index list
    std::vector<int32_t> sn_parent; // parent front or -1
    std::vector<int64_t> sn_lu;     // [nsn] offset of the front's factor storage:
                                    //   L panel  (s+r) x s, column major, ld = s+r  (F11 over F21)
                                    //   Ut panel  r x s,    column major, ld = r    (F12 transposed)
    std::vector<int64_t> sn_cb;     // [nsn] offset of the r x r contribution block (ld = r)
    std::vector<int32_t> child_ptr; // [nsn+1]
    std::vector<int32_t> child_idx; // children of each front, ascending

    // level schedule: fronts of level l are independent; children live in lower levels
    int32_t nlevels = 0;
    std::vector<int32_t> level_ptr; // [nlevels+1]
    std::vector<int32_t> level_sn;  // fronts ordered by (level, size class, id)

    // matrix slot -> factor storage.  a_dest[u] is the offset (in doubles) inside the LU storage
    // of slot u of the compressed-column pattern the analysis was given.
    std::vector<uint32_t> a_dest;

    // ---- subtree-unit schedule (schedule.hpp) ----------------------------------------------
    int32_t nunit = 0, nbig = 0, nstages = 0;
    std::vector<int32_t> blob;      // all unit blobs, concatenated
    std::vector<int32_t> unit_blob; // [nunit+1] word offsets into blob, ticket order
    std::vector<int32_t> ent_u;     // matrix slot (index into the pattern the analysis was given) ...
    std::vector<int32_t> ent_dst;   // ... and its position inside the owning unit's frontal arena
    std::vector<int32_t> big_u;     // matrix slots that belong to big fronts ...
    std::vector<uint32_t> big_dst;  // ... and their offset in the global factor storage
    std::vector<int32_t> big_sn;    // big fronts ordered by stage
    std::vector<LaunchGroup> groups;
    int64_t lu_big_begin = 0;       // factor storage of big fronts starts here (needs zeroing)
    int32_t factor_smem_bytes = 0, solve_smem_bytes = 0;
    int32_t n_serial_fronts = 0;

    int64_t lu_size = 0;  // doubles
    int64_t cb_size = 0;  // doubles
    int64_t upd_size = 0; // doubles (= sn_rows.size())
    int64_t nnz_factor = 0; // structural nonzeros of L+U (unit diagonal not counted twice)
    double flops = 0.0;
    int32_t max_front = 0, max_pivot_block = 0;
    int32_t n_missing_diag = 0; // structurally absent diagonal entries (become explicit zeros)
    // maximum transversal (only when the diagonal has holes): the analysed matrix is Q*A with row row_src[r] of A in
    // row r; the right-hand side is read through it.  Empty: Q = I.
    std::vector<int32_t> row_src;
    int32_t n_rows_moved = 0;

    bool budget_capped = false; // the automatic unit budget hit its ceiling (a large problem)
    int32_t s_of(int32_t f) const { return sn_first[f + 1] - sn_first[f]; }
    int32_t r_of(int32_t f) const { return (int32_t)(sn_rowptr[f + 1] - sn_rowptr[f]); }
};

// Analyse an n x n compressed-column pattern (outer[n+1], inner[nnz], rows ascending per
// column).  Returns false (and fills err) if the factor would not fit the 32-bit slot map.
bool analyse_pattern(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt,
                     Symbolic &S, const char **err);

} // namespace daecpp_b200
