// Subtree schedule of the multifrontal factorisation: layout of the per-unit integer "blob"
// shared by the host builder (symbolic.cpp), the CUDA kernels (subtree.cu) and the CPU walk
// used by the GPU-less tests (tests/emu/).
//
// A UNIT is a set of sibling subtrees of the assembly tree that is factored and solved by ONE
// thread block entirely out of shared memory: the frontal matrices of all its fronts live in a
// shared-memory arena, so contribution blocks never travel through HBM.  Only the contribution
// blocks of a unit's root fronts are written to global memory, for the unit that holds their
// parent.  Units are ordered children-first; a persistent kernel hands them out by ticket and
// a unit waits on a counter until the units below it have finished.
#pragma once
#include <cstdint>

namespace daecpp_b200
{
namespace sched
{
// ---- unit header (int32 words at the start of the unit's blob)
enum
{
    H_NF = 0,      // fronts in the unit
    H_NLEV,        // local levels
    H_WDOUBLES,    // frontal arena (doubles)
    H_LU0,         // first double of the unit's factor storage (uint32)
    H_LUCOUNT,     // doubles of factor storage (contiguous, schedule order)
    H_YDOUBLES,    // pivot unknowns of the unit (sum of s)
    H_UDOUBLES,    // update-vector arena (sum of r)
    H_NXR,         // external rows of the root fronts (sum of r over roots)
    H_ENT0,        // first matrix entry of the unit in ent_* arrays
    H_ENTCOUNT,    // number of matrix entries
    H_NDEP,        // child units that must finish first
    H_PARENT,      // parent unit (-1: none, -2: a big front outside the unit schedule)
    H_LEVELS,      // blob-relative offset of lev_ptr[nlev+1] followed by lev_group[nlev]
    H_FRONTS,      // blob-relative offset of the front records
    H_XR,          // blob-relative offset of the global (permuted) indices of the external rows
    H_NINTS,       // total int32 words of the blob
    H_SIZE
};
// ---- front record
enum
{
    F_COL0 = 0, // first pivot column, permuted numbering
    F_S,        // pivots
    F_R,        // rows below the pivot block
    F_WOFF,     // frontal matrix (s+r)^2, column major, inside the arena (doubles)
    F_LUOFF,    // factor storage offset RELATIVE to H_LU0 (L panel then transposed U panel)
    F_ROWS,     // blob-relative offset: rel[r] (position in the parent front), then rowsrc[r]
                //   (index into [y | xr] of that row's unknown, for the backward solve)
    F_CHILD,    // blob-relative offset of the child list
    F_NCHILD,   //   entries >= 0: local front index; < 0: external child, record at -1-v
    F_YOFF,     // offset of the pivots inside the unit's y arena
    F_UOFF,     // offset of the update vector inside the unit's u arena
    F_EXTCB,    // root fronts: global contribution-block offset (uint32), else 0xffffffff
    F_EXTUPD,   // root fronts: global update-vector offset, else -1
    F_SIZE
};
// ---- external child record (a root front of a child unit, or a big front), 4 words
enum
{
    X_CB = 0, // global contribution block offset (uint32), dense r x r
    X_R,      // its order
    X_REL,    // blob-relative offset of rel[r] into the parent front
    X_UPD,    // global update-vector offset
    X_SIZE
};
} // namespace sched
} // namespace daecpp_b200
