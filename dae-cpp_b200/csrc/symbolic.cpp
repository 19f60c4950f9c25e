// Host-side symbolic analysis (see symbolic.hpp).  No reference code is followed here: the
// reference delegates to Eigen's COLAMD + SuperLU-style interleaved symbolic/numeric
// factorisation (Eigen/src/SparseLU/SparseLU.h:528-814), which cannot be level-scheduled.
#include "symbolic.hpp"

#include <algorithm>
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
            int32_t best = -1; int64_t acc = 0;
            for (int32_t l = 0; l < nlev; ++l)
            {
                int64_t before = acc; acc += lev_count[l];
                if (l == 0 || l == nlev - 1) continue;
                bool in_window = before * 10 >= (int64_t)size * 4 && before * 10 <= (int64_t)size * 6;
                if (in_window && (best < 0 || lev_count[l] < lev_count[best])) best = l;
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

bool analyse_pattern(int32_t n, const int32_t *outer, const int32_t *inner, const SymbolicOptions &opt,
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
    S.nlevels = nlev;
    S.level_ptr.assign((size_t)nlev + 1, 0);
    for (int32_t f = 0; f < nsn; ++f) S.level_ptr[lvl[f] + 1]++;
    for (int32_t l = 0; l < nlev; ++l) S.level_ptr[l + 1] += S.level_ptr[l];
    S.level_sn.resize(nsn);
    {
        std::vector<int32_t> fillp(S.level_ptr.begin(), S.level_ptr.end() - 1);
        for (int32_t f = 0; f < nsn; ++f) S.level_sn[fillp[lvl[f]]++] = f;
    }

    // storage offsets, relative indices, statistics
    S.sn_lu.assign(nsn, 0); S.sn_cb.assign(nsn, 0);
    int64_t lu = 0, cb = 0; double flops = 0; int64_t nnzf = 0;
    for (int32_t f = 0; f < nsn; ++f)
    {
        const int64_t s = S.s_of(f), r = S.r_of(f);
        S.sn_lu[f] = lu; lu += (s + r) * s + r * s;
        S.sn_cb[f] = cb; cb += r * r;
        flops += 2.0 / 3.0 * (double)s * s * s + 2.0 * (double)s * s * r + 2.0 * (double)s * r * r;
        nnzf += s * s + 2 * s * r;
        if (s + r > S.max_front) S.max_front = (int32_t)(s + r);
        if (s > S.max_pivot_block) S.max_pivot_block = (int32_t)s;
    }
    S.lu_size = lu; S.cb_size = cb; S.upd_size = (int64_t)S.sn_rows.size(); S.flops = flops; S.nnz_factor = nnzf;
    if (lu >= (int64_t)0xFFFFFFF0u)
    {
        if (err) *err = "factor storage exceeds the 32-bit slot map";
        return false;
    }
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

    // matrix slot -> factor storage
    const int64_t nnz = outer[n];
    S.a_dest.resize((size_t)nnz);
    for (int32_t c = 0; c < n; ++c)
    {
        const int32_t pj = S.iperm[c];
        for (int32_t p = outer[c]; p < outer[c + 1]; ++p)
        {
            const int32_t pi = S.iperm[inner[p]];
            const int32_t k = pi < pj ? pi : pj;
            const int32_t f = sn_of_col[k];
            const int32_t first = S.sn_first[f], last = S.sn_first[f + 1] - 1;
            const int64_t s = last - first + 1, r = S.r_of(f);
            const int32_t *rb = S.sn_rows.data() + S.sn_rowptr[f];
            int64_t off;
            if (pj <= last)
            { // column inside the pivot block: L panel
                int64_t rowpos;
                if (pi <= last) rowpos = pi - first;
                else rowpos = s + (std::lower_bound(rb, rb + r, pi) - rb);
                off = S.sn_lu[f] + (int64_t)(pj - first) * (s + r) + rowpos;
            }
            else
            { // row inside the pivot block, column beyond it: Ut panel (transposed)
                int64_t colpos = std::lower_bound(rb, rb + r, pj) - rb;
                off = S.sn_lu[f] + (s + r) * s + (int64_t)(pi - first) * r + colpos;
            }
            S.a_dest[p] = (uint32_t)off;
        }
    }
    return true;
}

} // namespace daecpp_b200
