// Host-side symbolic analysis (see symbolic.hpp).  No reference code is followed here: the
// reference delegates to Eigen's COLAMD + SuperLU-style interleaved symbolic/numeric
// factorisation (Eigen/src/SparseLU/SparseLU.h:528-814), which cannot be level-scheduled.
#include "symbolic.hpp"

#include "schedule.hpp"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <numeric>

namespace daecpp_b200
{
namespace
{

struct Graph
{
    int32_t n;
    std::vector<int64_t> ptr;
    std::vector<int32_t> adj;
};

// adjacency of pattern(A + A^T) without the diagonal, neighbours ascending and unique
Graph build_graph(int32_t n, const int32_t *outer, const int32_t *inner, int32_t &missing_diag)
{
    Graph g;
    g.n = n;
    std::vector<int64_t> cnt((size_t)n + 1, 0);
    std::vector<char> has_diag((size_t)n, 0);
    for (int32_t c = 0; c < n; ++c)
        for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
        {
            int32_t r = inner[p];
            if (r == c) { has_diag[c] = 1; continue; }
            cnt[r + 1]++; cnt[c + 1]++;
        }
    missing_diag = 0;
    for (int32_t c = 0; c < n; ++c) missing_diag += !has_diag[c];
    for (int32_t v = 0; v < n; ++v) cnt[v + 1] += cnt[v];
    std::vector<int32_t> raw((size_t)cnt[n]);
    std::vector<int64_t> fill(cnt.begin(), cnt.end() - 1);
    for (int32_t c = 0; c < n; ++c)
        for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
        {
            int32_t r = inner[p];
            if (r == c) continue;
            raw[fill[r]++] = c;
            raw[fill[c]++] = r;
        }
    g.ptr.assign((size_t)n + 1, 0);
    g.adj.reserve(raw.size());
    for (int32_t v = 0; v < n; ++v)
    {
        auto b = raw.begin() + cnt[v], e = raw.begin() + cnt[v + 1];
        std::sort(b, e);
        e = std::unique(b, e);
        g.adj.insert(g.adj.end(), b, e);
        g.ptr[v + 1] = (int64_t)g.adj.size();
    }
    return g;
}

// Nested dissection by level-structure bisection.  Vertices of the current piece live in
// work[begin,end); owner[v] == piece id marks membership.  Positions [lo, lo+size) of the
// elimination order belong to the piece; its separator takes the last ones.
struct Dissector
{
    const Graph &g;
    int leaf;
    std::vector<int32_t> work, owner, level, queue, perm;
    int32_t next_id = 1;

    Dissector(const Graph &g_, int leaf_) : g(g_), leaf(leaf_ < 1 ? 1 : leaf_)
    {
        work.resize(g.n); owner.assign(g.n, 0); level.assign(g.n, -1); queue.resize(g.n); perm.assign(g.n, -1);
        std::iota(work.begin(), work.end(), 0);
    }

    // BFS inside piece `id` from `root`; fills queue[0..count) in visit order and level[]; returns count
    int32_t bfs(int32_t root, int32_t id, int32_t &nlev)
    {
        int32_t head = 0, tail = 0;
        queue[tail++] = root; level[root] = 0;
        while (head < tail)
        {
            int32_t v = queue[head++];
            for (int64_t p = g.ptr[v]; p < g.ptr[v + 1]; ++p)
            {
                int32_t w = g.adj[p];
                if (owner[w] == id && level[w] < 0) { level[w] = level[v] + 1; queue[tail++] = w; }
            }
        }
        nlev = level[queue[tail - 1]] + 1;
        return tail;
    }
    void clear_levels(int32_t count) { for (int32_t k = 0; k < count; ++k) level[queue[k]] = -1; }

    struct Item { int32_t begin, end, lo, id; };

    void run()
    {
        std::vector<Item> stack;
        stack.push_back({0, g.n, 0, 0});
        std::vector<int32_t> lev_count;
        while (!stack.empty())
        {
            Item it = stack.back(); stack.pop_back();
            const int32_t size = it.end - it.begin;
            if (size <= 0) continue;
            if (size <= leaf) { order_leaf(it); continue; }
            // pseudo-peripheral root: repeat BFS from the farthest vertex while depth grows
            int32_t root = work[it.begin], nlev = 0, count = 0;
            count = bfs(root, it.id, nlev);
            if (count < size)
            { // disconnected piece: peel the reached component off, no separator needed
                split_component(it, count, stack);
                continue;
            }
            for (int iter = 0; iter < 4; ++iter)
            {
                // among the deepest level pick the vertex of smallest degree
                int32_t far = queue[count - 1];
                clear_levels(count);
                int32_t nlev2 = 0;
                count = bfs(far, it.id, nlev2);
                root = far;
                if (nlev2 <= nlev) { nlev = nlev2; break; }
                nlev = nlev2;
            }
            if (nlev < 3)
            { // (near-)clique: no useful separator, eliminate as a leaf
                clear_levels(count);
                order_leaf(it);
                continue;
            }
            lev_count.assign(nlev, 0);
            for (int32_t k = 0; k < count; ++k) lev_count[level[queue[k]]]++;
            // separator = the thinnest level whose prefix holds 40..60 % of the vertices,
            // falling back to the median level
            int32_t best = -1; int64_t acc = 0, best_skew = 0;
            for (int32_t l = 0; l < nlev; ++l)
            {
                int64_t before = acc; acc += lev_count[l];
                if (l == 0 || l == nlev - 1) continue;
                bool in_window = before * 10 >= (int64_t)size * 4 && before * 10 <= (int64_t)size * 6;
                if (!in_window) continue;
                // thinnest level wins; among equally thin ones the most balanced split (shallower tree)
                const int64_t skew = std::llabs(2 * before + lev_count[l] - (int64_t)size);
                if (best < 0 || lev_count[l] < lev_count[best] || (lev_count[l] == lev_count[best] && skew < best_skew))
                { best = l; best_skew = skew; }
            }
            if (best < 0)
            {
                acc = 0;
                for (int32_t l = 0; l < nlev; ++l) { acc += lev_count[l]; if (acc * 2 >= size) { best = l; break; } }
                if (best <= 0) best = 1;
                if (best >= nlev - 1) best = nlev - 2;
            }
            // queue is in BFS order = sorted by level: [A | S | B]
            int32_t nA = 0, nS = lev_count[best];
            for (int32_t l = 0; l < best; ++l) nA += lev_count[l];
            int32_t nB = size - nA - nS;
            int32_t idA = next_id++, idB = next_id++;
            // rewrite work[begin,end) as [A | B] and place S at the tail positions of the order
            for (int32_t k = 0; k < nA; ++k) { int32_t v = queue[k]; work[it.begin + k] = v; owner[v] = idA; }
            for (int32_t k = 0; k < nB; ++k) { int32_t v = queue[nA + nS + k]; work[it.begin + nA + k] = v; owner[v] = idB; }
            for (int32_t k = 0; k < nS; ++k) { int32_t v = queue[nA + k]; perm[it.lo + nA + nB + k] = v; owner[v] = -1; }
            clear_levels(count);
            stack.push_back({it.begin, it.begin + nA, it.lo, idA});
            stack.push_back({it.begin + nA, it.begin + nA + nB, it.lo + nA, idB});
        }
    }

    // The BFS from the first vertex did not reach the whole piece: label every connected
    // component (linear time) and queue each as its own piece -- no separator is needed
    // between them.  `count` vertices of the first component are in queue[] with levels set.
    void split_component(const Item &it, int32_t count, std::vector<Item> &stack)
    {
        clear_levels(count);
        std::vector<int32_t> verts(work.begin() + it.begin, work.begin() + it.end);
        int32_t out = it.begin;
        for (int32_t v : verts)
        {
            if (owner[v] != it.id) continue;
            int32_t nlev = 0;
            const int32_t id = next_id++;
            // bfs() tests owner == it.id, so relabel after the sweep
            const int32_t c = bfs(v, it.id, nlev);
            for (int32_t k = 0; k < c; ++k) { owner[queue[k]] = id; work[out + k] = queue[k]; }
            clear_levels(c);
            stack.push_back({out, out + c, it.lo + (out - it.begin), id});
            out += c;
        }
    }

    // leaves: eliminate in BFS order from a peripheral vertex (a chain leaf becomes a chain of
    // fronts whose fill only reaches the two separators)
    void order_leaf(const Item &it)
    {
        const int32_t size = it.end - it.begin;
        int32_t placed = 0;
        for (int32_t k = it.begin; k < it.end; ++k)
        {
            int32_t v = work[k];
            if (owner[v] != it.id) continue;
            int32_t nlev = 0;
            int32_t count = bfs(v, it.id, nlev);
            int32_t far = queue[count - 1];
            clear_levels(count);
            count = bfs(far, it.id, nlev);
            for (int32_t q = 0; q < count; ++q) { perm[it.lo + placed++] = queue[q]; owner[queue[q]] = -1; }
            clear_levels(count);
        }
        (void)size;
    }
};

} // namespace

static bool build_schedule(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt, Symbolic &S,
                           const std::vector<int32_t> &sn_of_col, const char **err);

static bool analyse_core(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt, Symbolic &S, const char **err);

// Maximum transversal: a row for every column such that the permuted matrix has a structurally zero-free
// diagonal (depth-first augmenting paths with a cheap first pass that keeps existing diagonal entries;
// Duff's MC21 idea).  The reference needs none: Eigen::SparseLU pivots numerically at factorisation time
// (Eigen/src/SparseLU/SparseLU_pivotL.h:62-131); a static-pivot factorisation must at least start from a
// pattern whose diagonal exists -- index-1 DAEs routinely write `0 = x_j - g(t)` in a row other than j.
// Returns false when the pattern is structurally singular.
static bool maximum_transversal(int32_t n, const int32_t *outer, const int32_t *inner, const double *w, std::vector<int32_t> &row_of_col)
{
    row_of_col.assign((size_t)n, -1);
    std::vector<int32_t> col_of_row((size_t)n, -1);
    // candidate rows of every column, heaviest first when magnitudes are known (a poor man's weighted matching:
    // the static elimination order has to live with the diagonal chosen here, so big entries are preferred --
    // the idea of MC64 / SuperLU_DIST's static pivoting, without the optimality)
    std::vector<int32_t> cand(inner, inner + outer[n]);
    if (w)
        for (int32_t j = 0; j < n; ++j)
            std::stable_sort(cand.begin() + outer[j], cand.begin() + outer[j + 1], [&](int32_t a, int32_t b) {
                const int32_t *lo = inner + outer[j];
                const double wa = w[outer[j] + (std::lower_bound(lo, inner + outer[j + 1], a) - lo)];
                const double wb = w[outer[j] + (std::lower_bound(lo, inner + outer[j + 1], b) - lo)];
                return wa > wb;
            });
    for (int32_t j = 0; j < n; ++j) // keep the diagonal wherever it exists
        if (std::binary_search(inner + outer[j], inner + outer[j + 1], j)) { row_of_col[j] = j; col_of_row[j] = j; }
    for (int32_t j = 0; j < n; ++j) // cheap assignment
    {
        if (row_of_col[j] >= 0) continue;
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p)
            if (col_of_row[cand[p]] < 0) { row_of_col[j] = cand[p]; col_of_row[cand[p]] = j; break; }
    }
    std::vector<int32_t> visited((size_t)n, -1), stack_col, stack_pos;
    for (int32_t j0 = 0; j0 < n; ++j0)
    {
        if (row_of_col[j0] >= 0) continue;
        // iterative depth-first search for an augmenting path from column j0
        stack_col.assign(1, j0); stack_pos.assign(1, outer[j0]);
        visited[j0] = j0;
        bool found = false;
        while (!stack_col.empty() && !found)
        {
            const int32_t j = stack_col.back();
            int32_t p = stack_pos.back();
            bool descended = false;
            for (; p < outer[j + 1]; ++p)
            {
                const int32_t i = cand[p];
                const int32_t jn = col_of_row[i];
                if (jn < 0)
                {   // free row: flip the path
                    int32_t row = i;
                    for (size_t d = stack_col.size(); d-- > 0;)
                    {
                        const int32_t cj = stack_col[d];
                        const int32_t prev = row_of_col[cj];
                        row_of_col[cj] = row; col_of_row[row] = cj;
                        row = prev;
                    }
                    found = true;
                    break;
                }
                if (visited[jn] != j0)
                {
                    visited[jn] = j0;
                    stack_pos.back() = p + 1;
                    stack_col.push_back(jn); stack_pos.push_back(outer[jn]);
                    descended = true;
                    break;
                }
            }
            if (found) break;
            if (!descended) { stack_col.pop_back(); stack_pos.pop_back(); }
        }
        if (!found) return false;
    }
    return true;
}

bool analyse_pattern(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt,
                     Symbolic &S, const char **err)
{
    int32_t missing = 0;
    for (int32_t j = 0; j < n; ++j)
        if (!std::binary_search(inner + outer[j], inner + outer[j + 1], j)) ++missing;
    std::vector<int32_t> row_of_col;
    if (missing == 0 || !opt.transversal || !maximum_transversal(n, outer, inner, opt.weights, row_of_col))
        return analyse_core(n, outer, inner, opt, S, err); // (a structurally singular pattern fails numerically, like Eigen's would)
    // Row i of the matrix is moved to position rowmap[i]: A' = Q A has a zero-free diagonal.  The analysis runs on
    // A'; matrix slots are mapped back, and the right-hand side is read through row_src (Q A x = Q b).
    std::vector<int32_t> rowmap((size_t)n);
    for (int32_t j = 0; j < n; ++j) rowmap[row_of_col[j]] = j;
    const int64_t nnz = outer[n];
    std::vector<int32_t> inner2((size_t)nnz), slot2((size_t)nnz);
    std::vector<std::pair<int32_t, int32_t>> colbuf;
    for (int32_t j = 0; j < n; ++j)
    {
        colbuf.clear();
        for (int32_t p = outer[j]; p < outer[j + 1]; ++p) colbuf.emplace_back(rowmap[inner[p]], p);
        std::sort(colbuf.begin(), colbuf.end());
        for (size_t k = 0; k < colbuf.size(); ++k) { inner2[(size_t)outer[j] + k] = colbuf[k].first; slot2[(size_t)outer[j] + k] = colbuf[k].second; }
    }
    if (!analyse_core(n, outer, inner2.data(), opt, S, err)) return false;
    for (int32_t &u : S.ent_u) u = slot2[(size_t)u];
    for (int32_t &u : S.big_u) u = slot2[(size_t)u];
    S.row_src.assign((size_t)n, 0);
    for (int32_t i = 0; i < n; ++i) S.row_src[(size_t)rowmap[i]] = i;
    S.n_missing_diag = missing;
    S.n_rows_moved = 0;
    for (int32_t i = 0; i < n; ++i) S.n_rows_moved += rowmap[i] != i;
    return true;
}

static bool analyse_core(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt,
                         Symbolic &S, const char **err)
{
    S = Symbolic();
    S.n = n;
    Graph g = build_graph(n, outer, inner, S.n_missing_diag);

    // ---- fill- and depth-reducing ordering ------------------------------------------------
    {
        Dissector d(g, opt.leaf_size);
        d.run();
        S.perm.swap(d.perm);
    }
    S.iperm.assign(n, -1);
    for (int32_t k = 0; k < n; ++k) S.iperm[S.perm[k]] = k;

    // ---- supernodal symbolic factorisation of pattern(A+A^T) under that ordering ------------
    // Columns are visited in elimination order.  A column either extends the open supernode
    // (it is the parent of the supernode's last column, has no other child, and brings no new
    // row) or opens a new one whose row set is the union of its own adjacency and its
    // children's row sets.
    std::vector<int32_t> stamp((size_t)n, -1);
    std::vector<int32_t> first_child((size_t)n, -1), next_sibling; // closed supernodes hanging off a column
    std::vector<int32_t> sn_of_col((size_t)n, -1);
    std::vector<std::vector<int32_t>> rows; // per supernode (moved into the flat array afterwards)
    std::vector<int32_t> sn_first;
    std::vector<int32_t> tmp;
    rows.reserve((size_t)n / 2 + 16);
    int32_t open = -1; // index of the open supernode

    auto close_open = [&]() {
        if (open < 0) return;
        const std::vector<int32_t> &r = rows[open];
        if (!r.empty())
        {
            int32_t pc = r[0];
            if ((int32_t)next_sibling.size() <= open) next_sibling.resize(open + 1, -1);
            next_sibling[open] = first_child[pc];
            first_child[pc] = open;
        }
        open = -1;
    };

    for (int32_t j = 0; j < n; ++j)
    {
        const int32_t v = S.perm[j];
        bool merged = false;
        if (open >= 0 && first_child[j] < 0 && !rows[open].empty() && rows[open][0] == j &&
            (j - sn_first[open]) < opt.max_super)
        {
            // candidate: check adj_{>j}(j) is inside rows(open) \ {j}
            std::vector<int32_t> &r = rows[open];
            for (int32_t w : r) stamp[w] = j;
            bool subset = true;
            for (int64_t p = g.ptr[v]; p < g.ptr[v + 1]; ++p)
            {
                int32_t w = S.iperm[g.adj[p]];
                if (w > j && stamp[w] != j) { subset = false; break; }
            }
            if (subset)
            {
                r.erase(r.begin());
                sn_of_col[j] = open;
                merged = true;
            }
        }
        if (merged) continue;
        close_open();
        // new supernode starting at j
        tmp.clear();
        for (int64_t p = g.ptr[v]; p < g.ptr[v + 1]; ++p)
        {
            int32_t w = S.iperm[g.adj[p]];
            if (w > j && stamp[w] != -2 - j) { stamp[w] = -2 - j; tmp.push_back(w); }
        }
        for (int32_t c = first_child[j]; c >= 0; c = next_sibling[c])
            for (int32_t w : rows[c])
                if (w > j && stamp[w] != -2 - j) { stamp[w] = -2 - j; tmp.push_back(w); }
        std::sort(tmp.begin(), tmp.end());
        open = (int32_t)rows.size();
        rows.emplace_back(tmp);
        sn_first.push_back(j);
        sn_of_col[j] = open;
        if ((int32_t)next_sibling.size() <= open) next_sibling.resize(open + 1, -1);
    }
    close_open();

    const int32_t nsn = (int32_t)rows.size();
    S.nsn = nsn;
    S.sn_first = sn_first;
    S.sn_first.push_back(n);
    S.sn_rowptr.assign((size_t)nsn + 1, 0);
    for (int32_t f = 0; f < nsn; ++f) S.sn_rowptr[f + 1] = S.sn_rowptr[f] + (int64_t)rows[f].size();
    S.sn_rows.resize((size_t)S.sn_rowptr[nsn]);
    S.sn_parent.assign(nsn, -1);
    for (int32_t f = 0; f < nsn; ++f)
    {
        std::copy(rows[f].begin(), rows[f].end(), S.sn_rows.begin() + S.sn_rowptr[f]);
        if (!rows[f].empty()) S.sn_parent[f] = sn_of_col[rows[f][0]];
        std::vector<int32_t>().swap(rows[f]);
    }
    std::vector<std::vector<int32_t>>().swap(rows);

    // children lists (ascending) and level schedule
    S.child_ptr.assign((size_t)nsn + 1, 0);
    for (int32_t f = 0; f < nsn; ++f) if (S.sn_parent[f] >= 0) S.child_ptr[S.sn_parent[f] + 1]++;
    for (int32_t f = 0; f < nsn; ++f) S.child_ptr[f + 1] += S.child_ptr[f];
    S.child_idx.resize((size_t)S.child_ptr[nsn]);
    {
        std::vector<int32_t> fillp(S.child_ptr.begin(), S.child_ptr.end() - 1);
        for (int32_t f = 0; f < nsn; ++f) if (S.sn_parent[f] >= 0) S.child_idx[fillp[S.sn_parent[f]]++] = f;
    }
    std::vector<int32_t> lvl((size_t)nsn, 0);
    int32_t nlev = nsn ? 1 : 0;
    for (int32_t f = 0; f < nsn; ++f)
    { // parents have larger ids than children, so one ascending sweep suffices
        int32_t p = S.sn_parent[f];
        if (p >= 0 && lvl[p] < lvl[f] + 1) { lvl[p] = lvl[f] + 1; }
        if (lvl[f] + 1 > nlev) nlev = lvl[f] + 1;
    }
    S.nlevels = nlev; // height of the assembly tree (statistics only)

    double flops = 0; int64_t nnzf = 0;
    for (int32_t f = 0; f < nsn; ++f)
    {
        const int64_t s = S.s_of(f), r = S.r_of(f);
        flops += 2.0 / 3.0 * (double)s * s * s + 2.0 * (double)s * s * r + 2.0 * (double)s * r * r;
        nnzf += s * s + 2 * s * r;
        if (s + r > S.max_front) S.max_front = (int32_t)(s + r);
        if (s > S.max_pivot_block) S.max_pivot_block = (int32_t)s;
    }
    S.flops = flops; S.nnz_factor = nnzf; S.upd_size = (int64_t)S.sn_rows.size();
    if (S.upd_size >= 0x7ffffff0ll)
    {
        if (err) *err = "front row lists exceed the 32-bit schedule";
        return false;
    }

    // relative indices: position of every row of a front inside its parent's index list
    S.sn_rel.assign(S.sn_rows.size(), -1);
    {
        std::vector<int32_t> pos((size_t)n, -1);
        for (int32_t p = 0; p < nsn; ++p)
        {
            if (S.child_ptr[p] == S.child_ptr[p + 1]) continue;
            const int32_t first = S.sn_first[p], last = S.sn_first[p + 1] - 1, s = last - first + 1;
            for (int64_t q = S.sn_rowptr[p]; q < S.sn_rowptr[p + 1]; ++q) pos[S.sn_rows[q]] = s + (int32_t)(q - S.sn_rowptr[p]);
            for (int32_t ci = S.child_ptr[p]; ci < S.child_ptr[p + 1]; ++ci)
            {
                const int32_t c = S.child_idx[ci];
                for (int64_t q = S.sn_rowptr[c]; q < S.sn_rowptr[c + 1]; ++q)
                {
                    const int32_t w = S.sn_rows[q];
                    S.sn_rel[q] = (w <= last) ? (w - first) : pos[w];
                }
            }
        }
    }

    if (!build_schedule(n, outer, inner, opt, S, sn_of_col, err)) return false;
    return true;
}

// --------------------------------------------------------------------------------------------
// Partition the assembly tree into shared-memory units and big fronts, order them children
// first, lay the factor storage out in that order and emit the per-unit blobs (schedule.hpp).
static bool build_schedule(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt, Symbolic &S,
                           const std::vector<int32_t> &sn_of_col, const char **err)
{
    using namespace sched;
    const int32_t nsn = S.nsn;
    int64_t budget = opt.unit_smem_bytes;
    auto dim = [&](int32_t f) { return (int64_t)S.s_of(f) + S.r_of(f); };
    auto wsz = [&](int32_t f) { int64_t d = dim(f); return (d * d) | 1; };           // odd: no bank aliasing
    auto lusz = [&](int32_t f) { int64_t s = S.s_of(f), r = S.r_of(f); return (s + r) * s + r * s; };
    auto own_bytes = [&](int32_t f) {
        const int64_t s = S.s_of(f), r = S.r_of(f);
        const int64_t ints = F_SIZE + (S.child_ptr[f + 1] - S.child_ptr[f]) + 2 * r + 2;
        const int64_t fac = wsz(f), sol = s + 2 * r + lusz(f);
        return 4 * ints + 8 * (fac > sol ? fac : sol);
    };

    if (budget < 0)
    { // auto: enough units to occupy every SM a few times over, each as large as that allows
        int64_t total = 0;
        for (int32_t f = 0; f < nsn; ++f) total += own_bytes(f);
        budget = total / (4 * 148);
        if (budget < 6 * 1024) budget = 6 * 1024;
        // Large problems: 40 KB units swept by 128 threads (ctx: unit_threads = 0) measured 3 % faster on C4 than 64 KB units
        // swept by 256 -- a unit is a chain of local levels with a few lanes busy, so more, leaner resident units win
        S.budget_capped = budget > 40 * 1024;
        if (S.budget_capped) budget = 40 * 1024;
    }
    else
        S.budget_capped = false;

    // ---- 1. cuts ---------------------------------------------------------------------------
    std::vector<char> cut((size_t)nsn, 0), big((size_t)nsn, 0);
    std::vector<int64_t> acc((size_t)nsn, 0);
    const int64_t header_bytes = 4 * (H_SIZE + 4);
    std::vector<std::pair<int64_t, int32_t>> kids;
    for (int32_t f = 0; f < nsn; ++f)
    {
        int64_t own = own_bytes(f);
        const int32_t cb = S.child_ptr[f], ce = S.child_ptr[f + 1];
        int64_t all_cut = own;
        for (int32_t ci = cb; ci < ce; ++ci) all_cut += 4 * (X_SIZE + (int64_t)S.r_of(S.child_idx[ci]));
        if (budget <= 0 || all_cut + header_bytes > budget)
        {
            big[f] = 1; cut[f] = 1;
            for (int32_t ci = cb; ci < ce; ++ci) cut[S.child_idx[ci]] = 1;
            acc[f] = 0;
            continue;
        }
        int64_t a = own;
        kids.clear();
        for (int32_t ci = cb; ci < ce; ++ci)
        {
            const int32_t c = S.child_idx[ci];
            if (cut[c]) a += 4 * (X_SIZE + (int64_t)S.r_of(c));
            else { a += acc[c]; kids.push_back({acc[c], c}); }
        }
        if (a + header_bytes > budget)
        { // cut ALL children: units then come in layers of the tree and the chain of dependent
          // units from a leaf to the root has length height / (levels per unit), not height
            for (const auto &kc : kids)
            {
                const int32_t c = kc.second;
                a -= kc.first;
                a += 4 * (X_SIZE + (int64_t)S.r_of(c));
                cut[c] = 1;
            }
        }
        acc[f] = a;
        if (S.sn_parent[f] < 0) cut[f] = 1;
    }

    // ---- 2. units = bins of cut, non-big siblings ---------------------------------------------
    std::vector<int32_t> unit_of((size_t)nsn, -1), root_of((size_t)nsn, -1);
    int32_t nunit = 0;
    {
        // cut children of the same parent are contiguous in child_idx; parentless roots are
        // visited in id order
        int64_t fill = 0; int32_t cur_parent = -2;
        auto place = [&](int32_t c, int32_t parent) {
            if (big[c]) return;
            if (parent != cur_parent || fill + acc[c] + header_bytes > budget) { nunit++; fill = 0; cur_parent = parent; }
            fill += acc[c];
            unit_of[c] = nunit - 1;
        };
        for (int32_t f = 0; f < nsn; ++f) if (S.sn_parent[f] < 0) place(f, -1);
        for (int32_t p = 0; p < nsn; ++p)
            for (int32_t ci = S.child_ptr[p]; ci < S.child_ptr[p + 1]; ++ci)
                if (cut[S.child_idx[ci]]) place(S.child_idx[ci], p);
    }
    for (int32_t f = nsn - 1; f >= 0; --f)
    {
        if (cut[f]) { root_of[f] = f; continue; }
        unit_of[f] = unit_of[S.sn_parent[f]];
        root_of[f] = root_of[S.sn_parent[f]];
    }

    // ---- 3. stages (units and big fronts are the nodes), ticket order, launch groups -----------
    // node id: unit u -> u ; big front f -> nunit + f
    auto node_of = [&](int32_t f) { return big[f] ? nunit + f : unit_of[f]; };
    std::vector<int32_t> stage((size_t)nunit + nsn, 0), ndep((size_t)nunit, 0), unit_parent((size_t)nunit, -1);
    {
        std::vector<int32_t> last_child_unit((size_t)nunit + nsn, -1);
        for (int32_t f = 0; f < nsn; ++f)
        {
            const int32_t me = node_of(f);
            for (int32_t ci = S.child_ptr[f]; ci < S.child_ptr[f + 1]; ++ci)
            {
                const int32_t c = S.child_idx[ci];
                if (!cut[c]) continue;
                const int32_t cn = node_of(c);
                if (stage[me] < stage[cn] + 1) stage[me] = stage[cn] + 1;
                if (cn < nunit)
                {
                    unit_parent[cn] = me < nunit ? me : -2;
                    if (me < nunit && last_child_unit[me] != cn) { ndep[me]++; last_child_unit[me] = cn; }
                }
            }
        }
    }
    int32_t nstage = 0;
    for (int32_t u = 0; u < nunit; ++u) if (stage[u] + 1 > nstage) nstage = stage[u] + 1;
    S.nbig = 0;
    for (int32_t f = 0; f < nsn; ++f) if (big[f]) { S.nbig++; if (stage[nunit + f] + 1 > nstage) nstage = stage[nunit + f] + 1; }
    std::vector<int32_t> order((size_t)nunit), rank((size_t)nunit);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int32_t a, int32_t b) { return stage[a] < stage[b]; });
    for (int32_t t = 0; t < nunit; ++t) rank[order[t]] = t;
    S.big_sn.clear();
    {
        std::vector<int32_t> bigs;
        for (int32_t f = 0; f < nsn; ++f) if (big[f]) bigs.push_back(f);
        std::stable_sort(bigs.begin(), bigs.end(), [&](int32_t a, int32_t b) { return stage[nunit + a] < stage[nunit + b]; });
        S.big_sn = bigs;
    }
    S.groups.clear();
    {
        int32_t tu = 0, tb = 0, open_begin = 0;
        for (int32_t st = 0; st < nstage; ++st)
        {
            while (tu < nunit && stage[order[tu]] == st) ++tu;
            const int32_t b0 = tb;
            int32_t md = 0;
            while (tb < S.nbig && stage[nunit + S.big_sn[tb]] == st) { md = std::max<int32_t>(md, (int32_t)dim(S.big_sn[tb])); ++tb; }
            if (tb > b0)
            {
                if (tu > open_begin) S.groups.push_back({0, open_begin, tu, 0, stage[order[open_begin]] == st ? 1 : 0});
                open_begin = tu;
                S.groups.push_back({1, b0, tb, md, 0});
            }
        }
        if (nunit > open_begin) S.groups.push_back({0, open_begin, nunit, 0, 0});
    }
    S.nunit = nunit;
    S.nstages = nstage;

    // ---- 4. fronts of every unit in (local level, id) order -------------------------------------
    std::vector<int32_t> llev((size_t)nsn, 0);
    for (int32_t f = 0; f < nsn; ++f)
    {
        if (big[f]) continue;
        const int32_t p = S.sn_parent[f];
        if (!cut[f] && llev[p] < llev[f] + 1) llev[p] = llev[f] + 1;
    }
    std::vector<int32_t> ucount((size_t)nunit + 1, 0);
    for (int32_t f = 0; f < nsn; ++f) if (!big[f]) ucount[rank[unit_of[f]] + 1]++;
    for (int32_t t = 0; t < nunit; ++t) ucount[t + 1] += ucount[t];
    std::vector<int32_t> ufront((size_t)ucount[nunit]);
    {
        std::vector<int32_t> fillp(ucount.begin(), ucount.end() - 1);
        for (int32_t f = 0; f < nsn; ++f) if (!big[f]) ufront[fillp[rank[unit_of[f]]]++] = f;
        for (int32_t t = 0; t < nunit; ++t)
            std::stable_sort(ufront.begin() + ucount[t], ufront.begin() + ucount[t + 1],
                             [&](int32_t a, int32_t b) { return llev[a] < llev[b]; });
    }

    // ---- 5. factor storage in schedule order: units first, big fronts last ----------------------
    S.sn_lu.assign(nsn, 0); S.sn_cb.assign(nsn, -1);
    int64_t lu = 0, cbsz = 0;
    for (int32_t q = 0; q < ucount[nunit]; ++q) { const int32_t f = ufront[q]; S.sn_lu[f] = lu; lu += lusz(f); }
    S.lu_big_begin = lu;
    for (int32_t f : S.big_sn) { S.sn_lu[f] = lu; lu += lusz(f); }
    for (int32_t f = 0; f < nsn; ++f)
        if (cut[f] && S.r_of(f) > 0) { S.sn_cb[f] = cbsz; cbsz += (int64_t)S.r_of(f) * S.r_of(f); }
    S.lu_size = lu; S.cb_size = cbsz;
    if (lu >= (int64_t)0xFFFFFFF0u || cbsz >= (int64_t)0xFFFFFFF0u)
    {
        if (err) *err = "factor storage exceeds the 32-bit slot map";
        return false;
    }

    // ---- 6. blobs ------------------------------------------------------------------------------
    S.blob.clear(); S.unit_blob.assign((size_t)nunit + 1, 0);
    std::vector<int32_t> local_idx((size_t)nsn, -1), woff((size_t)nsn, 0), yoff((size_t)nsn, 0), xroff((size_t)nsn, 0);
    int64_t max_fac = 0, max_sol = 0;
    S.n_serial_fronts = 0;
    for (int32_t t = 0; t < nunit; ++t)
    {
        const int32_t u = order[t];
        const int32_t fb = ucount[t], fe = ucount[t + 1], nf = fe - fb;
        const size_t base = S.blob.size();
        if (base + 64 >= (size_t)0x7fffff00)
        {
            if (err) *err = "unit schedule exceeds 32-bit offsets";
            return false;
        }
        S.unit_blob[t] = (int32_t)base;
        S.blob.resize(base + H_SIZE, 0);
        int32_t nl = 0;
        for (int32_t q = fb; q < fe; ++q) nl = std::max(nl, llev[ufront[q]] + 1);
        // level table
        const int32_t lev_off = (int32_t)(S.blob.size() - base);
        S.blob.resize(S.blob.size() + 2 * (size_t)nl + 1, 0);
        {
            int32_t *lp = S.blob.data() + base + lev_off;
            int32_t *lg = lp + nl + 1;
            std::vector<int32_t> maxd((size_t)nl, 0);
            for (int32_t q = fb; q < fe; ++q)
            {
                const int32_t f = ufront[q];
                lp[llev[f] + 1]++;
                maxd[llev[f]] = std::max<int32_t>(maxd[llev[f]], (int32_t)dim(f));
            }
            for (int32_t l = 0; l < nl; ++l) lp[l + 1] += lp[l];
            for (int32_t l = 0; l < nl; ++l)
            {
                const int32_t d = maxd[l];
                const int32_t cnt = lp[l + 1] - lp[l];
                const bool warp = d <= opt.warp_dim || (d <= opt.warp_dim_many && cnt >= opt.warp_min_fronts);
                lg[l] = d <= opt.serial_dim ? 1 : (d <= opt.group8_dim ? 8 : (warp ? 32 : 0));
                if (lg[l] == 1) S.n_serial_fronts += lp[l + 1] - lp[l];
            }
        }
        // arenas
        int64_t w = 0, y = 0, ud = 0, xr = 0, lucount = 0;
        for (int32_t q = fb; q < fe; ++q)
        {
            const int32_t f = ufront[q];
            local_idx[f] = q - fb;
            woff[f] = (int32_t)w; w += wsz(f);
            yoff[f] = (int32_t)y; y += S.s_of(f);
            ud += S.r_of(f);
            lucount += lusz(f);
            if (cut[f]) { xroff[f] = (int32_t)xr; xr += S.r_of(f); }
        }
        // front records
        const int32_t fr_off = (int32_t)(S.blob.size() - base);
        S.blob.resize(S.blob.size() + (size_t)nf * F_SIZE, 0);
        int64_t uacc = 0;
        for (int32_t q = fb; q < fe; ++q)
        {
            const int32_t f = ufront[q];
            const int32_t s = S.s_of(f), r = S.r_of(f);
            const int32_t nch = S.child_ptr[f + 1] - S.child_ptr[f];
            const int64_t rp = S.sn_rowptr[f];
            // rows: rel[r], rowsrc[r]
            const int32_t rows_off = (int32_t)(S.blob.size() - base);
            S.blob.resize(S.blob.size() + 2 * (size_t)r);
            {
                int32_t *rel = S.blob.data() + base + rows_off, *src = rel + r;
                const int32_t root = root_of[f];
                const int32_t *rr = S.sn_rows.data() + S.sn_rowptr[root];
                const int32_t nrr = S.r_of(root);
                for (int32_t a = 0; a < r; ++a)
                {
                    const int32_t wcol = S.sn_rows[rp + a];
                    rel[a] = S.sn_rel[rp + a];
                    const int32_t g = sn_of_col[wcol];
                    if (!big[g] && unit_of[g] == u && local_idx[g] >= 0 && g > f && root_of[g] == root)
                        src[a] = yoff[g] + (wcol - S.sn_first[g]);
                    else
                        src[a] = (int32_t)y + xroff[root] + (int32_t)(std::lower_bound(rr, rr + nrr, wcol) - rr);
                }
            }
            // children
            const int32_t child_off = (int32_t)(S.blob.size() - base);
            S.blob.resize(S.blob.size() + (size_t)nch);
            for (int32_t k = 0; k < nch; ++k)
            {
                const int32_t c = S.child_idx[S.child_ptr[f] + k];
                if (!cut[c]) { S.blob[base + child_off + k] = local_idx[c]; continue; }
                const int32_t rc = S.r_of(c);
                const int32_t xo = (int32_t)(S.blob.size() - base);
                S.blob.resize(S.blob.size() + X_SIZE + (size_t)rc);
                int32_t *x = S.blob.data() + base + xo;
                x[X_CB] = (int32_t)(uint32_t)(rc > 0 ? S.sn_cb[c] : 0);
                x[X_R] = rc;
                x[X_REL] = xo + X_SIZE;
                x[X_UPD] = (int32_t)S.sn_rowptr[c];
                for (int32_t a = 0; a < rc; ++a) x[X_SIZE + a] = S.sn_rel[S.sn_rowptr[c] + a];
                S.blob[base + child_off + k] = -xo;
            }
            int32_t *F = S.blob.data() + base + fr_off + (size_t)(q - fb) * F_SIZE;
            F[F_COL0] = S.sn_first[f]; F[F_S] = s; F[F_R] = r; F[F_WOFF] = woff[f];
            F[F_LUOFF] = (int32_t)(S.sn_lu[f] - S.sn_lu[ufront[fb]]);
            F[F_ROWS] = rows_off; F[F_CHILD] = child_off; F[F_NCHILD] = nch;
            F[F_YOFF] = yoff[f]; F[F_UOFF] = (int32_t)uacc;
            F[F_EXTCB] = (cut[f] && r > 0) ? (int32_t)(uint32_t)S.sn_cb[f] : -1;
            F[F_EXTUPD] = (cut[f] && r > 0) ? (int32_t)rp : -1;
            uacc += r;
        }
        // external rows of the roots (global permuted indices)
        const int32_t xr_off = (int32_t)(S.blob.size() - base);
        S.blob.resize(S.blob.size() + (size_t)xr);
        for (int32_t q = fb; q < fe; ++q)
        {
            const int32_t f = ufront[q];
            if (!cut[f]) continue;
            for (int32_t a = 0; a < S.r_of(f); ++a) S.blob[base + xr_off + xroff[f] + a] = S.sn_rows[S.sn_rowptr[f] + a];
        }
        while ((S.blob.size() - base) % 4) S.blob.push_back(0); // 16-byte granules for vector loads
        int32_t *H = S.blob.data() + base;
        H[H_NF] = nf; H[H_NLEV] = nl; H[H_WDOUBLES] = (int32_t)w;
        H[H_LU0] = (int32_t)(uint32_t)S.sn_lu[ufront[fb]]; H[H_LUCOUNT] = (int32_t)lucount;
        H[H_YDOUBLES] = (int32_t)y; H[H_UDOUBLES] = (int32_t)ud; H[H_NXR] = (int32_t)xr;
        H[H_NDEP] = ndep[u]; H[H_PARENT] = unit_parent[u] >= 0 ? rank[unit_parent[u]] : unit_parent[u];
        H[H_LEVELS] = lev_off; H[H_FRONTS] = fr_off; H[H_XR] = xr_off;
        H[H_NINTS] = (int32_t)(S.blob.size() - base);
        const int64_t ints_b = 4 * (int64_t)H[H_NINTS];
        max_fac = std::max(max_fac, ints_b + 8 * w);
        max_sol = std::max(max_sol, ints_b + 8 * (y + xr + ud + lucount));
    }
    S.unit_blob[nunit] = (int32_t)S.blob.size();
    S.factor_smem_bytes = (int32_t)max_fac + 64;
    S.solve_smem_bytes = (int32_t)max_sol + 64;

    // ---- 7. matrix slots -> frontal positions -----------------------------------------------------
    const int64_t nnz = outer[n];
    std::vector<int32_t> ecount((size_t)nunit + 1, 0);
    std::vector<int32_t> slot_front((size_t)nnz);
    std::vector<int32_t> slot_pos((size_t)nnz); // unit fronts: index inside the frontal matrix; big: unused
    S.big_u.clear(); S.big_dst.clear();
    for (int32_t c = 0; c < n; ++c)
    {
        const int32_t pj = S.iperm[c];
        for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
        {
            const int32_t pi = S.iperm[inner[p]];
            const int32_t k = pi < pj ? pi : pj;
            const int32_t f = sn_of_col[k];
            const int32_t first = S.sn_first[f], last = S.sn_first[f + 1] - 1;
            const int64_t s = last - first + 1, r = S.r_of(f), d = s + r;
            const int32_t *rb = S.sn_rows.data() + S.sn_rowptr[f];
            int64_t ii, jj;
            if (pj <= last)
            {
                jj = pj - first;
                ii = (pi <= last) ? (pi - first) : s + (std::lower_bound(rb, rb + r, pi) - rb);
            }
            else
            {
                ii = pi - first;
                jj = s + (std::lower_bound(rb, rb + r, pj) - rb);
            }
            slot_front[p] = f;
            if (big[f])
            {
                const int64_t off = (jj < s) ? S.sn_lu[f] + jj * d + ii : S.sn_lu[f] + d * s + ii * r + (jj - s);
                S.big_u.push_back(p);
                S.big_dst.push_back((uint32_t)off);
            }
            else
            {
                slot_pos[p] = (int32_t)(woff[f] + jj * d + ii);
                ecount[rank[unit_of[f]] + 1]++;
            }
        }
    }
    for (int32_t t = 0; t < nunit; ++t) ecount[t + 1] += ecount[t];
    S.ent_u.assign((size_t)ecount[nunit], 0); S.ent_dst.assign((size_t)ecount[nunit], 0);
    {
        std::vector<int32_t> fillp(ecount.begin(), ecount.end() - 1);
        for (int64_t p = 0; p < nnz; ++p)
        {
            const int32_t f = slot_front[p];
            if (big[f]) continue;
            const int32_t q = fillp[rank[unit_of[f]]]++;
            S.ent_u[q] = (int32_t)p; S.ent_dst[q] = slot_pos[p];
        }
    }
    for (int32_t t = 0; t < nunit; ++t)
    {
        int32_t *H = S.blob.data() + S.unit_blob[t];
        H[H_ENT0] = ecount[t]; H[H_ENTCOUNT] = ecount[t + 1] - ecount[t];
    }
    S.a_dest.clear();
    return true;
}

} // namespace daecpp_b200
