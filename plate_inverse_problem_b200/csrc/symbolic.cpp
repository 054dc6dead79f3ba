// Host-side symbolic analysis: nested-dissection ordering (BFS level-structure separators,
// constrained minimum degree inside the leaves), elimination tree, postorder, column
// structures, relaxed supernodes, and the static multifrontal schedule for the GPU.
// See symbolic.h for what it replaces in the reference.
#include "symbolic.h"

#include <algorithm>
#include <cstring>
#include <numeric>
#include <stdexcept>
#include <string>

namespace jpb {
namespace {

// ---------------------------------------------------------------- graph of A + A^T
void build_graph(int N, const int64_t* ap, const int64_t* ai, std::vector<int64_t>& xadj,
                 std::vector<int>& adj) {
    std::vector<int64_t> cnt(N + 1, 0);
    for (int j = 0; j < N; ++j) {
        if (ap[j + 1] < ap[j]) throw std::runtime_error("add_mat: indptr not monotone");
        for (int64_t p = ap[j]; p < ap[j + 1]; ++p) {
            int64_t i = ai[p];
            if (i < 0 || i >= N) throw std::runtime_error("add_mat: row index out of range");
            if (i != j) { cnt[i + 1]++; cnt[j + 1]++; }
        }
    }
    for (int i = 0; i < N; ++i) cnt[i + 1] += cnt[i];
    std::vector<int> raw(cnt[N]);
    std::vector<int64_t> pos(cnt.begin(), cnt.end() - 1);
    for (int j = 0; j < N; ++j)
        for (int64_t p = ap[j]; p < ap[j + 1]; ++p) {
            int i = (int)ai[p];
            if (i != j) { raw[pos[i]++] = j; raw[pos[j]++] = i; }
        }
    xadj.assign(N + 1, 0);
    adj.clear();
    adj.reserve(raw.size() / 2 + 16);
    for (int v = 0; v < N; ++v) {
        auto b = raw.begin() + cnt[v], e = raw.begin() + cnt[v + 1];
        std::sort(b, e);
        auto u = std::unique(b, e);
        adj.insert(adj.end(), b, u);
        xadj[v + 1] = (int64_t)adj.size();
    }
}

// ---------------------------------------------------------------- nested dissection
struct ND {
    int N;
    const std::vector<int64_t>& xadj;
    const std::vector<int>& adj;
    const SymbolicOptions& opt;
    std::vector<int> part;    // region id of a vertex; -1 = separator / already placed
    std::vector<int> lvl;     // BFS level scratch
    std::vector<int> stamp;   // visit stamps
    std::vector<int> loc;     // global -> local scratch for the leaf ordering
    int cur_stamp = 0;
    int next_region = 1;
    std::vector<int> order;   // elimination sequence (old vertex ids)

    ND(int n, const std::vector<int64_t>& xa, const std::vector<int>& a,
       const SymbolicOptions& o)
        : N(n), xadj(xa), adj(a), opt(o), part(n, 0), lvl(n, 0), stamp(n, 0), loc(n, -1) {
        order.reserve(n);
    }

    int degree_in(int v, int region) const {
        int d = 0;
        for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) d += part[adj[p]] == region;
        return d;
    }

    // BFS inside `region` from root; fills queue (visit order) and lptr (level starts).
    void bfs(int root, int region, std::vector<int>& queue, std::vector<int>& lptr) {
        ++cur_stamp;
        queue.clear();
        lptr.clear();
        queue.push_back(root);
        stamp[root] = cur_stamp;
        lvl[root] = 0;
        lptr.push_back(0);
        size_t head = 0;
        while (head < queue.size()) {
            size_t end = queue.size();
            for (; head < end; ++head) {
                int v = queue[head];
                for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) {
                    int u = adj[p];
                    if (part[u] == region && stamp[u] != cur_stamp) {
                        stamp[u] = cur_stamp;
                        lvl[u] = lvl[v] + 1;
                        queue.push_back(u);
                    }
                }
            }
            if (queue.size() > end) lptr.push_back((int)end);
        }
        lptr.push_back((int)queue.size());
    }

    // constrained minimum degree on a small region: region vertices are eliminated, their
    // outside neighbours (all belong to enclosing separators, numbered later) only count
    // towards degrees.
    void leaf_order(const std::vector<int>& verts, int region) {
        const int n = (int)verts.size();
        if (n == 0) return;
        if (n <= 2) {
            for (int v : verts) { order.push_back(v); part[v] = -1; }
            return;
        }
        std::vector<int> ext;
        for (int i = 0; i < n; ++i) loc[verts[i]] = i;
        int nloc = n;
        for (int v : verts)
            for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) {
                int u = adj[p];
                if (loc[u] < 0) { loc[u] = nloc++; ext.push_back(u); }
            }
        const int W = (nloc + 63) / 64;
        std::vector<uint64_t> bits((size_t)n * W, 0);
        auto row = [&](int i) { return bits.data() + (size_t)i * W; };
        for (int i = 0; i < n; ++i) {
            int v = verts[i];
            uint64_t* r = row(i);
            for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) {
                int l = loc[adj[p]];
                r[l >> 6] |= 1ull << (l & 63);
            }
        }
        auto popcnt = [&](const uint64_t* r) {
            int c = 0;
            for (int w = 0; w < W; ++w) c += __builtin_popcountll(r[w]);
            return c;
        };
        std::vector<int> deg(n);
        std::vector<char> done(n, 0);
        for (int i = 0; i < n; ++i) deg[i] = popcnt(row(i));
        for (int step = 0; step < n; ++step) {
            int best = -1;
            for (int i = 0; i < n; ++i)
                if (!done[i] && (best < 0 || deg[i] < deg[best])) best = i;
            done[best] = 1;
            order.push_back(verts[best]);
            const uint64_t* rb = row(best);
            for (int w = 0; w < W; ++w) {
                uint64_t m = rb[w];
                while (m) {
                    int l = (w << 6) + __builtin_ctzll(m);
                    m &= m - 1;
                    if (l >= n || done[l]) continue;
                    uint64_t* ru = row(l);
                    for (int x = 0; x < W; ++x) ru[x] |= rb[x];
                    ru[l >> 6] &= ~(1ull << (l & 63));
                    ru[best >> 6] &= ~(1ull << (best & 63));
                    deg[l] = popcnt(ru);
                }
            }
        }
        for (int v : verts) { loc[v] = -1; part[v] = -1; }
        for (int u : ext) loc[u] = -1;
    }

    void dissect(std::vector<int>& verts, int region) {
        const int n = (int)verts.size();
        if (n <= opt.leaf_size) { leaf_order(verts, region); return; }

        std::vector<int> queue, lptr;
        bfs(verts[0], region, queue, lptr);
        if ((int)queue.size() < n) {
            // disconnected: order every component on its own (no separator needed)
            std::vector<std::vector<int>> comps;
            std::vector<int> rids;
            {
                int r = next_region++;
                for (int v : queue) part[v] = r;
                comps.push_back(queue);
                rids.push_back(r);
            }
            for (int v : verts) {
                if (part[v] != region) continue;
                bfs(v, region, queue, lptr);
                int r = next_region++;
                for (int u : queue) part[u] = r;
                comps.push_back(queue);
                rids.push_back(r);
            }
            for (size_t c = 0; c < comps.size(); ++c) dissect(comps[c], rids[c]);
            return;
        }

        // pseudo-peripheral root
        int root = verts[0];
        int depth = (int)lptr.size() - 1;
        for (int it = 0; it < 6; ++it) {
            int lastb = lptr[lptr.size() - 2], laste = lptr[lptr.size() - 1];
            int cand = queue[lastb], cd = degree_in(cand, region);
            for (int q = lastb + 1; q < laste; ++q) {
                int d = degree_in(queue[q], region);
                if (d < cd) { cd = d; cand = queue[q]; }
            }
            std::vector<int> q2, l2;
            bfs(cand, region, q2, l2);
            int d2 = (int)l2.size() - 1;
            if (d2 > depth) { depth = d2; root = cand; queue.swap(q2); lptr.swap(l2); }
            else break;
        }
        bfs(root, region, queue, lptr);   // lvl/stamp must describe the chosen root
        const int nl = (int)lptr.size() - 1;   // number of levels
        if (nl < 3) { leaf_order(verts, region); return; }

        // separator = smallest level inside the balance window (else the median level)
        int best = -1;
        int64_t best_sz = 0;
        for (int l = 1; l + 1 < nl; ++l) {
            int left = lptr[l], right = n - lptr[l + 1];
            if (std::min(left, right) < opt.balance * n) continue;
            int sz = lptr[l + 1] - lptr[l];
            if (best < 0 || sz < best_sz) { best = l; best_sz = sz; }
        }
        if (best < 0) {
            best = 1;
            for (int l = 1; l + 1 < nl; ++l)
                if (lptr[l + 1] >= n / 2) { best = l; break; }
            best = std::max(1, std::min(best, nl - 2));
        }
        std::vector<int> A, B, S;
        A.assign(queue.begin(), queue.begin() + lptr[best]);
        B.assign(queue.begin() + lptr[best + 1], queue.end());
        for (int q = lptr[best]; q < lptr[best + 1]; ++q) {
            int v = queue[q];
            bool touches = false;
            for (int64_t p = xadj[v]; p < xadj[v + 1] && !touches; ++p) {
                int u = adj[p];
                touches = part[u] == region && stamp[u] == cur_stamp && lvl[u] == best + 1;
            }
            if (touches) S.push_back(v); else A.push_back(v);
        }
        // note: lvl/stamp describe the last bfs() call, which produced `queue`
        int ra = next_region++, rb = next_region++;
        for (int v : A) part[v] = ra;
        for (int v : B) part[v] = rb;
        for (int v : S) part[v] = -1;
        verts.clear();
        verts.shrink_to_fit();
        dissect(A, ra);
        dissect(B, rb);
        for (int v : S) order.push_back(v);
    }
};

}  // namespace

void nested_dissection_order(int N, const std::vector<int64_t>& xadj,
                             const std::vector<int>& adj, const SymbolicOptions& opt,
                             std::vector<int>& perm) {
    ND nd(N, xadj, adj, opt);
    std::vector<int> all(N);
    std::iota(all.begin(), all.end(), 0);
    if (N > 0) nd.dissect(all, 0);
    if ((int)nd.order.size() != N) throw std::runtime_error("nested dissection lost vertices");
    perm = nd.order;
}

void analyze(int N, const int64_t* ap, const int64_t* ai, const SymbolicOptions& opt,
             Symbolic& S) {
    if (N <= 0) throw std::runtime_error("add_mat: N must be positive");
    S = Symbolic();
    S.N = N;
    S.nnz = ap[N];
    std::vector<int64_t> xadj;
    std::vector<int> adj;
    build_graph(N, ap, ai, xadj, adj);

    // ---- ordering
    std::vector<int> perm;
    nested_dissection_order(N, xadj, adj, opt, perm);
    std::vector<int> pinv(N, -1);
    for (int i = 0; i < N; ++i) pinv[perm[i]] = i;
    for (int i = 0; i < N; ++i)
        if (pinv[i] < 0) throw std::runtime_error("ordering is not a permutation");

    // ---- elimination tree (Liu) in the permuted numbering
    auto etree = [&](const std::vector<int>& pm, const std::vector<int>& pi,
                     std::vector<int>& parent) {
        parent.assign(N, -1);
        std::vector<int> anc(N, -1);
        for (int j = 0; j < N; ++j) {
            int v = pm[j];
            for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) {
                int i = pi[adj[p]];
                while (i != -1 && i < j) {
                    int nx = anc[i];
                    anc[i] = j;
                    if (nx == -1) parent[i] = j;
                    i = nx;
                }
            }
        }
    };
    std::vector<int> parent;
    etree(perm, pinv, parent);

    // ---- postorder, then relabel
    {
        std::vector<int> head(N, -1), next(N, -1);
        for (int j = N - 1; j >= 0; --j)
            if (parent[j] >= 0) { next[j] = head[parent[j]]; head[parent[j]] = j; }
        std::vector<int> post;
        post.reserve(N);
        std::vector<int> stack;
        for (int r = 0; r < N; ++r) {
            if (parent[r] != -1) continue;
            stack.push_back(r);
            while (!stack.empty()) {
                int v = stack.back();
                int c = head[v];
                if (c == -1) { post.push_back(v); stack.pop_back(); }
                else { head[v] = next[c]; stack.push_back(c); }
            }
        }
        std::vector<int> perm2(N);
        for (int i = 0; i < N; ++i) perm2[i] = perm[post[i]];
        perm.swap(perm2);
        for (int i = 0; i < N; ++i) pinv[perm[i]] = i;
        etree(perm, pinv, parent);
    }
    S.perm = perm;
    S.pinv = pinv;

    // ---- column structures of L (strictly below the diagonal), sorted
    std::vector<std::vector<int>> cs(N);
    {
        std::vector<int> head(N, -1), next(N, -1), mark(N, -1);
        for (int j = N - 1; j >= 0; --j)
            if (parent[j] >= 0) { next[j] = head[parent[j]]; head[parent[j]] = j; }
        for (int j = 0; j < N; ++j) {
            std::vector<int>& s = cs[j];
            mark[j] = j;
            int v = perm[j];
            for (int64_t p = xadj[v]; p < xadj[v + 1]; ++p) {
                int i = pinv[adj[p]];
                if (i > j && mark[i] != j) { mark[i] = j; s.push_back(i); }
            }
            for (int c = head[j]; c != -1; c = next[c])
                for (int i : cs[c])
                    if (mark[i] != j) { mark[i] = j; s.push_back(i); }
            std::sort(s.begin(), s.end());
            if (!s.empty() && parent[j] != s[0])
                throw std::runtime_error("symbolic: etree / structure mismatch");
        }
    }

    // ---- fundamental supernodes + relaxed amalgamation (stack of [a,b] column ranges)
    struct SN { int a, b; double z; };
    std::vector<SN> st;
    auto rows_of = [&](const SN& s) { return (int64_t)cs[s.b].size(); };
    int j = 0;
    while (j < N) {
        int a = j;
        while (j + 1 < N && parent[j] == j + 1 && cs[j].size() == cs[j + 1].size() + 1) ++j;
        SN cur{a, j, 0.0};
        ++j;
        while (!st.empty()) {
            const SN& c = st.back();
            int pc = parent[c.b];
            if (pc < cur.a || pc > cur.b) break;
            int64_t kc = c.b - c.a + 1, mc = rows_of(c);
            int64_t kp = cur.b - cur.a + 1, mp = rows_of(cur);
            int64_t sp = kp + mp;
            int64_t k = kc + kp, s = kc + sp;
            double newz = (double)kc * (double)(sp - mc);
            double z = c.z + cur.z + newz;
            double total = 0.5 * (double)k * (double)(k + 1) + (double)k * (double)(s - k);
            double frac = z / total;
            bool ok = k <= opt.relax_small ||
                      (k <= opt.relax_mid && frac < opt.relax_mid_frac) ||
                      (k <= opt.relax_big && frac < opt.relax_big_frac) ||
                      frac < opt.relax_any_frac;
            if (!ok) break;
            cur.a = c.a;
            cur.z = z;
            st.pop_back();
        }
        st.push_back(cur);
    }

    // ---- fronts
    const int nf = (int)st.size();
    S.nfronts = nf;
    S.f_c0.resize(nf); S.f_k.resize(nf); S.f_s.resize(nf); S.f_off.resize(nf); S.f_ld.resize(nf);
    S.f_voff.resize(nf); S.f_rowptr.assign(nf + 1, 0); S.f_parent.assign(nf, -1);
    S.f_level.assign(nf, 0);
    std::vector<int> sn_of(N);
    for (int f = 0; f < nf; ++f) {
        const SN& s = st[f];
        S.f_c0[f] = s.a;
        S.f_k[f] = s.b - s.a + 1;
        S.f_s[f] = S.f_k[f] + (int)cs[s.b].size();
        S.f_rowptr[f + 1] = S.f_rowptr[f] + S.f_s[f];
        for (int c = s.a; c <= s.b; ++c) sn_of[c] = f;
    }
    S.rowidx.resize(S.f_rowptr[nf]);
    int64_t off = 0, voff = 0;
    for (int f = 0; f < nf; ++f) {
        const SN& s = st[f];
        int* r = S.rowidx.data() + S.f_rowptr[f];
        for (int c = s.a; c <= s.b; ++c) *r++ = c;
        for (int i : cs[s.b]) *r++ = i;
        S.f_off[f] = off;
        S.f_voff[f] = (int)voff;
        const int64_t ss = S.f_s[f], kk = S.f_k[f];
        S.f_ld[f] = (int)((ss + 1) & ~(int64_t)1);
        off += ss * S.f_ld[f];
        voff += ss;
        S.nnz_lu += kk * (2 * ss - kk);
        for (int64_t q = 0; q < kk; ++q) {
            double m = (double)(ss - q - 1);
            S.flops += 6.0 * m + 8.0 * m * m;
        }
        S.max_s = std::max(S.max_s, (int)ss);
        S.max_k = std::max(S.max_k, (int)kk);
        if (!cs[s.b].empty()) S.f_parent[f] = sn_of[cs[s.b][0]];
    }
    if (voff > INT32_MAX) throw std::runtime_error("symbolic: vector workspace too large");
    S.front_slots = off;
    S.vec_slots = voff;
    std::vector<std::vector<int>>().swap(cs);

    // children lists, levels
    S.f_childptr.assign(nf + 1, 0);
    for (int f = 0; f < nf; ++f)
        if (S.f_parent[f] >= 0) {
            if (S.f_parent[f] <= f) throw std::runtime_error("symbolic: parent before child");
            S.f_childptr[S.f_parent[f] + 1]++;
        }
    for (int f = 0; f < nf; ++f) S.f_childptr[f + 1] += S.f_childptr[f];
    S.children.resize(S.f_childptr[nf]);
    {
        std::vector<int> pos(S.f_childptr.begin(), S.f_childptr.end() - 1);
        for (int f = 0; f < nf; ++f)
            if (S.f_parent[f] >= 0) S.children[pos[S.f_parent[f]]++] = f;
    }
    for (int f = 0; f < nf; ++f)
        if (S.f_parent[f] >= 0)
            S.f_level[S.f_parent[f]] = std::max(S.f_level[S.f_parent[f]], S.f_level[f] + 1);
    S.nlevels = 0;
    for (int f = 0; f < nf; ++f) S.nlevels = std::max(S.nlevels, S.f_level[f] + 1);
    S.levelptr.assign(S.nlevels + 1, 0);
    for (int f = 0; f < nf; ++f) S.levelptr[S.f_level[f] + 1]++;
    for (int l = 0; l < S.nlevels; ++l) S.levelptr[l + 1] += S.levelptr[l];
    S.levelfronts.resize(nf);
    {
        std::vector<int> pos(S.levelptr.begin(), S.levelptr.end() - 1);
        for (int f = 0; f < nf; ++f) S.levelfronts[pos[S.f_level[f]]++] = f;
        for (int l = 0; l < S.nlevels; ++l)
            std::stable_sort(S.levelfronts.begin() + S.levelptr[l],
                             S.levelfronts.begin() + S.levelptr[l + 1],
                             [&](int x, int y) { return S.f_s[x] > S.f_s[y]; });
    }

    // ---- child -> parent maps
    S.f_reloff.assign(nf, -1);
    {
        int64_t nrel = 0;
        for (int f = 0; f < nf; ++f)
            if (S.f_parent[f] >= 0) { S.f_reloff[f] = nrel; nrel += S.f_s[f] - S.f_k[f]; }
        S.relmap.assign(nrel, -1);
        for (int f = 0; f < nf; ++f) {
            int p = S.f_parent[f];
            if (p < 0) continue;
            const int* rc = S.rowidx.data() + S.f_rowptr[f];
            const int* rp = S.rowidx.data() + S.f_rowptr[p];
            int sc = S.f_s[f], kc = S.f_k[f], sp = S.f_s[p];
            int q = 0;
            for (int i = kc; i < sc; ++i) {
                while (q < sp && rp[q] < rc[i]) ++q;
                if (q >= sp || rp[q] != rc[i])
                    throw std::runtime_error("symbolic: child row missing from parent front");
                S.relmap[S.f_reloff[f] + (i - kc)] = q;
            }
        }
    }
    S.rowidx_orig.resize(S.rowidx.size());
    for (size_t i = 0; i < S.rowidx.size(); ++i) S.rowidx_orig[i] = perm[S.rowidx[i]];

    // ---- A entry -> (front, local slot), grouped by front
    if (S.nnz > INT32_MAX) throw std::runtime_error("symbolic: nnz exceeds int32");
    std::vector<int> owner(S.nnz), lslot(S.nnz);
    S.f_entptr.assign(nf + 1, 0);
    for (int c = 0; c < N; ++c) {
        for (int64_t e = ap[c]; e < ap[c + 1]; ++e) {
            int i = pinv[ai[e]], jj = pinv[c];
            int f = sn_of[std::min(i, jj)];
            const int* r = S.rowidx.data() + S.f_rowptr[f];
            int k = S.f_k[f], s = S.f_s[f], c0 = S.f_c0[f];
            auto local = [&](int x) -> int {
                if (x < c0 + k) return x - c0;
                const int* it = std::lower_bound(r + k, r + s, x);
                if (it == r + s || *it != x) throw std::runtime_error("symbolic: entry outside front");
                return (int)(it - r);
            };
            int64_t ls = (int64_t)local(i) * S.f_ld[f] + local(jj);
            if (ls > INT32_MAX) throw std::runtime_error("symbolic: front too large");
            lslot[e] = (int)ls;
            owner[e] = f;
            S.f_entptr[f + 1]++;
        }
    }
    for (int f = 0; f < nf; ++f) S.f_entptr[f + 1] += S.f_entptr[f];
    S.ent_src.resize(S.nnz);
    S.ent_slot.resize(S.nnz);
    {
        std::vector<int64_t> pos(S.f_entptr.begin(), S.f_entptr.end() - 1);
        for (int64_t e = 0; e < S.nnz; ++e) {
            int64_t q = pos[owner[e]]++;
            S.ent_src[q] = (int)e;
            S.ent_slot[q] = lslot[e];
        }
    }
}

}  // namespace jpb
