// Host side of the block-chain path (chain.hpp): recognise a block-tridiagonal pattern, fix the
// radix hierarchy and emit the level-0 slot map.  No reference code is followed: the reference
// hands every pattern to COLAMD + SuperLU-style column elimination
// (Eigen/src/SparseLU/SparseLU.h:528-814), which has no notion of a chain.
#include "chain.hpp"

#include <algorithm>
#include <cstdlib>

#include "chain_ops.hpp"

namespace daecpp_b200
{
namespace chain
{

int rows_per_tile(int B) { return B == 1 ? Dim<1>::ROWS : B == 2 ? Dim<2>::ROWS : Dim<3>::ROWS; }
int tile_factor_doubles(int B) { return B == 1 ? Dim<1>::TILE_F : B == 2 ? Dim<2>::TILE_F : Dim<3>::TILE_F; }
int tile_y_doubles(int B) { return B == 1 ? Dim<1>::TILE_Y : B == 2 ? Dim<2>::TILE_Y : Dim<3>::TILE_Y; }

namespace
{

struct Graph
{
    std::vector<int64_t> ptr;
    std::vector<int32_t> adj;
};

// adjacency of pattern(A + A^T) without the diagonal (duplicates allowed: only used for BFS)
Graph build_graph(int32_t n, const int32_t *outer, const int32_t *inner)
{
    Graph g;
    g.ptr.assign((size_t)n + 1, 0);
    for (int32_t j = 0; j < n; ++j)
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
        {
            const int32_t i = inner[p];
            if (i != j) { ++g.ptr[(size_t)i + 1]; ++g.ptr[(size_t)j + 1]; }
        }
    for (int32_t v = 0; v < n; ++v) g.ptr[(size_t)v + 1] += g.ptr[v];
    g.adj.resize((size_t)g.ptr[n]);
    std::vector<int64_t> fill(g.ptr.begin(), g.ptr.end() - 1);
    for (int32_t j = 0; j < n; ++j)
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
        {
            const int32_t i = inner[p];
            if (i != j) { g.adj[(size_t)fill[i]++] = j; g.adj[(size_t)fill[j]++] = i; }
        }
    return g;
}

// smallest block order b <= max_b such that every entry (i, j) has |pos(i)/b - pos(j)/b| <= 1
int block_order(int32_t n, const int32_t *outer, const int32_t *inner, const int32_t *pos, int max_b)
{
    int64_t w = 0;
    for (int32_t j = 0; j < n; ++j)
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
        {
            const int64_t d = std::llabs((int64_t)(pos ? pos[inner[p]] : inner[p]) - (int64_t)(pos ? pos[j] : j));
            if (d > w) w = d;
        }
    if (w > 2 * (int64_t)max_b - 1) return 0; // blocks b apart differ by at most 2b-1 positions
    for (int b = 1; b <= max_b; ++b)
    {
        bool ok = true;
        for (int32_t j = 0; j < n && ok; ++j)
        {
            const int32_t bj = (pos ? pos[j] : j) / b;
            for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
            {
                const int32_t bi = (pos ? pos[inner[p]] : inner[p]) / b;
                if (bi > bj + 1 || bj > bi + 1) { ok = false; break; }
            }
        }
        if (ok) return b;
    }
    return 0;
}

// breadth-first (Cuthill-McKee) numbering from a pseudo-peripheral vertex of every component,
// neighbours taken by increasing degree
void cuthill_mckee(int32_t n, const Graph &g, std::vector<int32_t> &order)
{
    order.clear();
    order.reserve((size_t)n);
    std::vector<int32_t> level((size_t)n, -1), queue, nb;
    std::vector<char> done((size_t)n, 0);
    auto deg = [&](int32_t v) { return g.ptr[(size_t)v + 1] - g.ptr[v]; };
    auto bfs = [&](int32_t s, std::vector<int32_t> &out) {
        out.clear();
        out.push_back(s);
        level[s] = 0;
        for (size_t h = 0; h < out.size(); ++h)
        {
            const int32_t v = out[h];
            nb.clear();
            for (int64_t p = g.ptr[v]; p < g.ptr[(size_t)v + 1]; ++p)
            {
                const int32_t w = g.adj[(size_t)p];
                if (level[w] < 0 && !done[w]) { level[w] = level[v] + 1; nb.push_back(w); }
            }
            std::sort(nb.begin(), nb.end(), [&](int32_t a, int32_t b) { return deg(a) != deg(b) ? deg(a) < deg(b) : a < b; });
            for (int32_t w : nb) out.push_back(w);
        }
    };
    for (int32_t s0 = 0; s0 < n; ++s0)
    {
        if (done[s0]) continue;
        int32_t s = s0;
        for (int pass = 0; pass < 4; ++pass)
        {
            bfs(s, queue);
            const int32_t far = level[queue.back()];
            int32_t best = queue.back();
            for (size_t k = queue.size(); k-- > 0 && level[queue[k]] == far;)
                if (deg(queue[k]) < deg(best) || (deg(queue[k]) == deg(best) && queue[k] < best)) best = queue[k];
            for (int32_t v : queue) level[v] = -1;
            if (pass == 3 || best == s) break;
            s = best;
        }
        bfs(s, queue);
        for (int32_t v : queue) { done[v] = 1; order.push_back(v); }
    }
}

template <int B>
void build_levels_and_map(int32_t n, const int32_t *outer, const int32_t *inner, Plan &P)
{
    using D = Dim<B>;
    P.m = (n + B - 1) / B;
    P.levels.clear();
    int64_t fac = 0, y = 0, red = 0, rv = 0, x = 0;
    int32_t rows = P.m;
    for (;;)
    {
        Level L;
        L.rows = rows;
        L.ntiles = (rows + D::ROWS - 1) / D::ROWS;
        if (L.ntiles < 1) L.ntiles = 1;
        L.fac0 = fac; fac += (int64_t)L.ntiles * D::TILE_F;
        L.y0 = y; y += (int64_t)L.ntiles * D::TILE_Y;
        L.red0 = red; red += (int64_t)4 * L.ntiles * D::BB;
        L.rv0 = rv; rv += (int64_t)2 * L.ntiles * B;
        L.x0 = x; x += (int64_t)rows * B;
        P.levels.push_back(L);
        if (L.ntiles == 1) break;
        rows = L.ntiles;
    }
    P.fac_doubles = fac; P.y_doubles = y; P.red_doubles = red; P.rv_doubles = rv; P.x_doubles = x;
    P.nnz_factor = (int64_t)5 * B * B * P.m;

    // slot map of level 0
    const Level &L0 = P.levels[0];
    const int64_t per_tile = (int64_t)3 * D::BB * (D::ROWS + 1);
    P.ent_u.assign((size_t)(per_tile * L0.ntiles), -1);
    P.tile_flags.assign((size_t)L0.ntiles * 3 * D::BB, 0);
    auto slot_of = [&](int32_t row, int32_t col) -> int32_t {
        const int32_t *lo = inner + outer[col], *hi = inner + outer[col + 1];
        const int32_t *it = std::lower_bound(lo, hi, row);
        return (it != hi && *it == row) ? (int32_t)(it - inner) : -1;
    };
    for (int32_t tile = 0; tile < L0.ntiles; ++tile)
        for (int f = 0; f < 3; ++f)
            for (int r = 0; r < B; ++r)
                for (int c = 0; c < B; ++c)
                {
                    const int fe = f * D::BB + r * B + c;
                    int32_t *dst = P.ent_u.data() + ((int64_t)tile * 3 * D::BB + fe) * (D::ROWS + 1);
                    int flags = 0;
                    for (int rr = 0; rr <= D::ROWS; ++rr)
                    {
                        const int64_t i = (int64_t)tile * D::ROWS + rr - 1; // block row
                        if (i < 0 || i >= P.m) { if (i >= P.m && f == 1 && r == c) { dst[rr] = -2; flags |= 2; } continue; }
                        const int64_t ib = i + f - 1;                      // block column
                        const int64_t row = i * B + r, col = ib * B + c;
                        if (row >= n)
                        {
                            if (f == 1 && r == c) { dst[rr] = -2; flags |= 2; }
                            continue;
                        }
                        if (ib < 0 || ib >= P.m || col >= n) continue;
                        const int32_t u = slot_of(P.perm[(size_t)row], P.perm[(size_t)col]);
                        dst[rr] = u;
                        if (u >= 0) flags |= 1;
                    }
                    P.tile_flags[(size_t)tile * 3 * D::BB + fe] = flags;
                }
}

} // namespace

bool detect(int32_t n, const int32_t *outer, const int32_t *inner, int max_b, Plan &P)
{
    P = Plan();
    if (n <= 0) return false;
    if (max_b > MAX_B) max_b = MAX_B;
    if (max_b < 1) return false;
    P.n = n;
    int b = block_order(n, outer, inner, nullptr, max_b);
    P.perm.resize((size_t)n);
    if (b)
    {
        for (int32_t k = 0; k < n; ++k) P.perm[k] = k;
        P.identity_perm = true;
    }
    else
    {
        Graph g = build_graph(n, outer, inner);
        std::vector<int32_t> order;
        cuthill_mckee(n, g, order);
        std::vector<int32_t> pos((size_t)n);
        for (int32_t k = 0; k < n; ++k) pos[order[k]] = k;
        b = block_order(n, outer, inner, pos.data(), max_b);
        if (!b) return false;
        P.perm = order;
        P.identity_perm = false;
    }
    if (b == 1)
    { // no pivoting inside a 1 x 1 block: a structurally missing diagonal goes to the general path
        for (int32_t k = 0; k < n; ++k)
        {
            const int32_t v = P.perm[k];
            const int32_t *lo = inner + outer[v], *hi = inner + outer[v + 1];
            if (!std::binary_search(lo, hi, v)) return false;
        }
    }
    P.B = b;
    if (b == 1) build_levels_and_map<1>(n, outer, inner, P);
    else if (b == 2) build_levels_and_map<2>(n, outer, inner, P);
    else build_levels_and_map<3>(n, outer, inner, P);
    P.active = true;
    return true;
}

} // namespace chain
} // namespace daecpp_b200
