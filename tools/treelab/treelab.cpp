// treelab.cpp — CPU laboratory for the BVH the GPU builder makes (tools only; not part of the product, never linked into it).
//
// The B200 box is a scarce resource, tree QUALITY is not a GPU question: nodes visited and triangles tested per ray depend only on the
// tree and the traversal order, so they can be measured on the host. This tool restates, in plain C++:
//   builders   lbvh     the Morton / Karras radix tree vermilion_b200/csrc/bvh_build.cu builds today
//              ploc:R   parallel locally-ordered clustering (Meister & Bittner 2018) with search radius R over the Morton order
//              sah      full-sweep top-down SAH (quality yardstick; not a GPU candidate)
//   collapses  greedy   today's 2->8 collapse: open the child with the largest area, leaves of <= 3 triangles
//              dp:Cp:Pm the SAH-optimal collapse of Ylitie, Karras & Laine 2017 (cost tables C(n,i), leaf cost Cp per triangle, <= Pm per leaf)
//   traversal  the octant-ordered BVH8 walk of traverse.cuh (8-bit quantised child boxes, hit bits taken at the parent's visit, popped
//              groups are not re-tested) — counting nodes and triangles only
// and prints nodes / triangles per ray, B_ray and the SAH cost, for C3's uniform rays or for C2/C4-like path rays
// (camera rays of the bench camera + up to four cosine bounces, the room's walls as planes).
//
//   g++ -O2 -fopenmp -std=c++17 tools/treelab/treelab.cpp vermilion_b200/host/scenes.cpp -o /tmp/treelab
//   /tmp/treelab soup:1000000 uniform:200000 lbvh greedy
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <numeric>
#include <random>
#include <string>
#include <vector>
#include <omp.h>

#include "../../vermilion_b200/host/scenes.h"

struct V3 { float x, y, z; };
static inline V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
static inline V3 normalize(V3 a) { float l = std::sqrt(dot(a, a)); return l > 0 ? a * (1.f / l) : a; }

struct Box {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = FLT_MAX; hi[a] = -FLT_MAX; } }
    void grow(const Box& b) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], b.lo[a]); hi[a] = std::max(hi[a], b.hi[a]); } }
    float area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; }
};
static inline Box merged(const Box& a, const Box& b) { Box r = a; r.grow(b); return r; }

struct Tri { V3 v0, e1, e2; };
struct Scene {
    std::vector<Tri> tris;
    std::vector<Box> boxes;
    Box bounds;
};

// binary tree: nodes [0, n-1) internal in any order, leaves are triangle indices encoded as ~tri
struct BNode { Box box; int left, right; uint32_t count; float area; };
struct BTree { std::vector<BNode> nodes; int root; };
static inline bool is_leaf(int c) { return c < 0; }
static inline int leaf_tri(int c) { return ~c; }

// ---------------------------------------------------------------- Morton order
static uint64_t spread21(uint64_t x) {
    x &= 0x1FFFFFull;
    x = (x | x << 32) & 0x1F00000000FFFFull; x = (x | x << 16) & 0x1F0000FF0000FFull; x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull; x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
static int g_emc = 0;
static void morton_order(const Scene& sc, std::vector<uint64_t>& keys, std::vector<uint32_t>& order) {
    const size_t n = sc.tris.size();
    keys.resize(n); order.resize(n);
    std::vector<uint64_t> k(n);
#pragma omp parallel for
    for (size_t i = 0; i < n; ++i) {
        uint64_t q[3];
        for (int a = 0; a < 3; ++a) {
            float c = 0.5f * (sc.boxes[i].lo[a] + sc.boxes[i].hi[a]), ext = sc.bounds.hi[a] - sc.bounds.lo[a];
            float f = ext > 0 ? (c - sc.bounds.lo[a]) / ext : 0.f;
            f = std::min(std::max(f, 0.f), 1.f);
            q[a] = (uint64_t)std::min(f * 2097152.f, 2097151.f);
        }
        k[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
        if (g_emc > 0) { // extended Morton code (Vinkler, Bittner, Havran 2017): a box-size bit after every g_emc-th xyz triple
            float dx = sc.boxes[i].hi[0] - sc.boxes[i].lo[0], dy = sc.boxes[i].hi[1] - sc.boxes[i].lo[1], dz = sc.boxes[i].hi[2] - sc.boxes[i].lo[2];
            float ex = sc.bounds.hi[0] - sc.bounds.lo[0], ey = sc.bounds.hi[1] - sc.bounds.lo[1], ez = sc.bounds.hi[2] - sc.bounds.lo[2];
            float diag = std::sqrt(dx * dx + dy * dy + dz * dz) / std::sqrt(ex * ex + ey * ey + ez * ez);
            uint64_t sq = (uint64_t)std::min(diag * 1048576.f, 1048575.f); // 20 bits, large boxes first... (MSB = large)
            uint64_t out = 0; int ob = 63, sb = 19;
            for (int t = 20; t >= 0 && ob >= 0; --t) { // xyz triples from the top
                for (int a = 0; a < 3 && ob >= 0; ++a) { out |= ((q[a] >> t) & 1ull) << ob; --ob; }
                if (((20 - t) % g_emc) == g_emc - 1 && sb >= 0 && ob >= 0) { out |= ((sq >> sb) & 1ull) << ob; --ob; --sb; }
            }
            k[i] = out;
        }
    }
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return k[a] < k[b]; });
    for (size_t i = 0; i < n; ++i) keys[i] = k[order[i]];
}

// ---------------------------------------------------------------- builders
static int lbvh_delta(const std::vector<uint64_t>& keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __builtin_clz((uint32_t)i ^ (uint32_t)j);
    return __builtin_clzll(a ^ b);
}
static BTree build_lbvh(const Scene& sc) {
    std::vector<uint64_t> keys; std::vector<uint32_t> order;
    morton_order(sc, keys, order);
    const int n = (int)sc.tris.size();
    BTree t; t.nodes.resize(std::max(n - 1, 1)); t.root = 0;
    if (n == 1) { t.root = ~0; return t; }
    std::vector<int> parent(2 * n - 1, -1);
#pragma omp parallel for
    for (int i = 0; i < n - 1; ++i) { // Karras 2012, as k_hierarchy
        int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
        int dmin = lbvh_delta(keys, n, i, i - d);
        int lmax = 2;
        while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
        int l = 0;
        for (int tt = lmax >> 1; tt >= 1; tt >>= 1) if (lbvh_delta(keys, n, i, i + (l + tt) * d) > dmin) l += tt;
        int j = i + l * d, dnode = lbvh_delta(keys, n, i, j), s = 0, tt = l;
        do { tt = (tt + 1) >> 1; if (lbvh_delta(keys, n, i, i + (s + tt) * d) > dnode) s += tt; } while (tt > 1);
        int gamma = i + s * d + std::min(d, 0), lo = std::min(i, j), hi = std::max(i, j);
        t.nodes[i].left = (lo == gamma) ? ~(int)order[gamma] : gamma;
        t.nodes[i].right = (hi == gamma + 1) ? ~(int)order[gamma + 1] : gamma + 1;
    }
    // boxes bottom-up (recursive post-order with an explicit stack)
    std::vector<int> st{0}, post;
    while (!st.empty()) { int x = st.back(); st.pop_back(); post.push_back(x); for (int c : {t.nodes[x].left, t.nodes[x].right}) if (!is_leaf(c)) st.push_back(c); }
    for (size_t k = post.size(); k-- > 0;) {
        BNode& nd = t.nodes[post[k]];
        nd.box.reset(); nd.count = 0;
        for (int c : {nd.left, nd.right}) {
            if (is_leaf(c)) { nd.box.grow(sc.boxes[leaf_tri(c)]); nd.count += 1; }
            else { nd.box.grow(t.nodes[c].box); nd.count += t.nodes[c].count; }
        }
        nd.area = nd.box.area();
    }
    return t;
}

static BTree build_ploc(const Scene& sc, int R) {
    std::vector<uint64_t> keys; std::vector<uint32_t> order;
    morton_order(sc, keys, order);
    const int n = (int)sc.tris.size();
    BTree t; t.nodes.reserve(std::max(n - 1, 1));
    std::vector<int> cl(n), next; std::vector<Box> cb(n), nb;
    for (int i = 0; i < n; ++i) { cl[i] = ~(int)order[i]; cb[i] = sc.boxes[order[i]]; }
    std::vector<int> nn;
    int iters = 0;
    while (cl.size() > 1) {
        const int m = (int)cl.size();
        nn.assign(m, -1);
#pragma omp parallel for schedule(static, 4096)
        for (int i = 0; i < m; ++i) {
            float best = FLT_MAX; int bj = -1;
            for (int j = std::max(0, i - R); j <= std::min(m - 1, i + R); ++j) {
                if (j == i) continue;
                float a = merged(cb[i], cb[j]).area();
                if (a < best) { best = a; bj = j; }
            }
            nn[i] = bj;
        }
        next.clear(); nb.clear();
        for (int i = 0; i < m; ++i) {
            const int j = nn[i];
            if (nn[j] == i) {
                if (i < j) {
                    BNode nd; nd.left = cl[i]; nd.right = cl[j]; nd.box = merged(cb[i], cb[j]); nd.area = nd.box.area();
                    nd.count = (is_leaf(cl[i]) ? 1u : t.nodes[cl[i]].count) + (is_leaf(cl[j]) ? 1u : t.nodes[cl[j]].count);
                    t.nodes.push_back(nd);
                    next.push_back((int)t.nodes.size() - 1); nb.push_back(nd.box);
                }
            } else { next.push_back(cl[i]); nb.push_back(cb[i]); }
        }
        cl.swap(next); cb.swap(nb);
        ++iters;
    }
    t.root = cl[0];
    fprintf(stderr, "ploc: %d iterations\n", iters);
    return t;
}

static BTree build_sah(const Scene& sc) {
    const int n = (int)sc.tris.size();
    BTree t; t.nodes.reserve(std::max(n - 1, 1));
    std::vector<uint32_t> idx(n); std::iota(idx.begin(), idx.end(), 0u);
    std::vector<float> right_area(n + 1);
    // binned SAH (32 bins) top-down, explicit recursion
    std::function<int(int, int)> rec = [&](int lo, int hi) -> int {
        if (hi - lo == 1) return ~(int)idx[lo];
        Box cb, bb; cb.reset(); bb.reset();
        for (int i = lo; i < hi; ++i) {
            const Box& b = sc.boxes[idx[i]]; bb.grow(b);
            Box c; for (int a = 0; a < 3; ++a) c.lo[a] = c.hi[a] = 0.5f * (b.lo[a] + b.hi[a]);
            cb.grow(c);
        }
        int best_axis = -1, best_bin = -1; float best_cost = FLT_MAX;
        const int NB = 32;
        for (int axis = 0; axis < 3; ++axis) {
            float ext = cb.hi[axis] - cb.lo[axis];
            if (!(ext > 0)) continue;
            Box bins[NB]; int cnt[NB];
            for (int b = 0; b < NB; ++b) { bins[b].reset(); cnt[b] = 0; }
            for (int i = lo; i < hi; ++i) {
                const Box& b = sc.boxes[idx[i]];
                int k = std::min(NB - 1, (int)((0.5f * (b.lo[axis] + b.hi[axis]) - cb.lo[axis]) / ext * NB));
                bins[k].grow(b); ++cnt[k];
            }
            float ra[NB]; int rc[NB]; Box acc; acc.reset(); int c = 0;
            for (int b = NB - 1; b > 0; --b) { acc.grow(bins[b]); c += cnt[b]; ra[b] = c ? acc.area() : 0.f; rc[b] = c; }
            acc.reset(); c = 0;
            for (int b = 0; b + 1 < NB; ++b) {
                acc.grow(bins[b]); c += cnt[b];
                if (c == 0 || rc[b + 1] == 0) continue;
                float cost = acc.area() * c + ra[b + 1] * rc[b + 1];
                if (cost < best_cost) { best_cost = cost; best_axis = axis; best_bin = b; }
            }
        }
        int mid;
        if (best_axis < 0) mid = (lo + hi) / 2;
        else {
            float ext = cb.hi[best_axis] - cb.lo[best_axis];
            auto it = std::partition(idx.begin() + lo, idx.begin() + hi, [&](uint32_t i) {
                const Box& b = sc.boxes[i];
                int k = std::min(NB - 1, (int)((0.5f * (b.lo[best_axis] + b.hi[best_axis]) - cb.lo[best_axis]) / ext * NB));
                return k <= best_bin;
            });
            mid = (int)(it - idx.begin());
            if (mid == lo || mid == hi) mid = (lo + hi) / 2;
        }
        const int me = (int)t.nodes.size();
        t.nodes.emplace_back();
        int l = rec(lo, mid), r = rec(mid, hi);
        BNode& nd = t.nodes[me];
        nd.left = l; nd.right = r; nd.box = bb; nd.area = bb.area(); nd.count = (uint32_t)(hi - lo);
        return me;
    };
    t.root = rec(0, n);
    return t;
}

// ---------------------------------------------------------------- wide tree
struct WNode {
    int child[8];          // >= 0: wide node index; -1 empty; <= -2: leaf, triangles at tri_first[s] .. +ntri[s]
    int tri_first[8], ntri[8];
    Box qbox[8];           // conservative dequantised boxes (what the traversal tests)
};
struct WTree { std::vector<WNode> nodes; std::vector<uint32_t> leaf_tris; int depth = 0; double sah = 0; };

struct CollapseOpt { bool dp = false; float c_prim = 0.3f; int p_max = 3; };

static void subtree_tris(const BTree& t, int c, std::vector<uint32_t>& out) {
    if (is_leaf(c)) { out.push_back((uint32_t)leaf_tri(c)); return; }
    subtree_tris(t, t.nodes[c].left, out); subtree_tris(t, t.nodes[c].right, out);
}
static inline const Box& cbox(const Scene& sc, const BTree& t, int c) { return is_leaf(c) ? sc.boxes[leaf_tri(c)] : t.nodes[c].box; }
static inline uint32_t ccount(const BTree& t, int c) { return is_leaf(c) ? 1u : t.nodes[c].count; }

// DP tables (Ylitie et al. 2017, eq. 1-3): cost[n][i-1], i = 1..7; decision: for i = 1: 0 leaf / 1 internal; split k for distribute
struct DpTab { float c[7]; uint8_t k[8]; uint8_t leaf; }; // k[j-1] for distribute(n, j), j = 2..8 -> stored at index j-1 (1..7)

static WTree collapse(const Scene& sc, const BTree& t, const CollapseOpt& opt) {
    WTree w;
    const int nint = (int)t.nodes.size();
    std::vector<DpTab> tab;
    const float root_area = is_leaf(t.root) ? 1.f : t.nodes[t.root].area;
    auto leaf_cost = [&](int c, int i) -> float { (void)i; return cbox(sc, t, c).area() / root_area * opt.c_prim; };
    if (opt.dp) {
        tab.resize(nint);
        // post-order
        std::vector<int> st, post;
        if (!is_leaf(t.root)) st.push_back(t.root);
        while (!st.empty()) { int x = st.back(); st.pop_back(); post.push_back(x); for (int c : {t.nodes[x].left, t.nodes[x].right}) if (!is_leaf(c)) st.push_back(c); }
        auto C = [&](int c, int i) -> float { return is_leaf(c) ? leaf_cost(c, i) : tab[c].c[i - 1]; };
        for (size_t q = post.size(); q-- > 0;) {
            const int n = post[q];
            const BNode& nd = t.nodes[n];
            DpTab& T = tab[n];
            float dist[9]; // distribute(n, j), j = 2..8
            for (int j = 2; j <= 8; ++j) {
                float best = FLT_MAX; int bk = 1;
                for (int k = 1; k < j; ++k) {
                    if (k > 7 || j - k > 7) continue;
                    float v = C(nd.left, k) + C(nd.right, j - k);
                    if (v < best) { best = v; bk = k; }
                }
                dist[j] = best; T.k[j - 1] = (uint8_t)bk;
            }
            const float a = nd.area / root_area;
            const float c_leaf = (int)nd.count <= opt.p_max ? a * nd.count * opt.c_prim : FLT_MAX;
            const float c_int = dist[8] + a; // C_node = 1
            T.leaf = c_leaf <= c_int;
            T.c[0] = std::min(c_leaf, c_int);
            for (int i = 2; i <= 7; ++i) T.c[i - 1] = std::min(dist[i], T.c[i - 2]);
        }
    }
    // children of the wide node rooted at binary node b
    auto gather = [&](int b, int* ch) -> int {
        int nch = 0;
        if (!opt.dp) {
            ch[0] = t.nodes[b].left; ch[1] = t.nodes[b].right; nch = 2;
            while (nch < 8) {
                int best = -1; float ba = -1.f;
                for (int k = 0; k < nch; ++k) if (!is_leaf(ch[k]) && (int)t.nodes[ch[k]].count > 3 && t.nodes[ch[k]].area > ba) { ba = t.nodes[ch[k]].area; best = k; }
                if (best < 0) break;
                const BNode& nd = t.nodes[ch[best]];
                ch[best] = nd.left; ch[nch++] = nd.right;
            }
            return nch;
        }
        // DP: follow the decisions of distribute(b, 8)
        std::function<void(int, int)> emit = [&](int c, int i) { // represent subtree c with at most i roots
            if (is_leaf(c)) { ch[nch++] = c; return; }
            const DpTab& T = tab[c];
            // C(c,i) = min(distribute(c,i), C(c,i-1)) for i >= 2: find the i' <= i that realises the minimum
            int ii = i;
            while (ii >= 2 && T.c[ii - 1] == T.c[ii - 2]) --ii;
            if (ii == 1) { ch[nch++] = c; return; }
            const int k = T.k[ii - 1];
            emit(t.nodes[c].left, k); emit(t.nodes[c].right, ii - k);
        };
        const int k = tab[b].k[7];
        emit(t.nodes[b].left, k); emit(t.nodes[b].right, 8 - k);
        return nch;
    };
    auto child_is_leaf = [&](int c) -> bool {
        if (is_leaf(c)) return true;
        if (!opt.dp) return (int)t.nodes[c].count <= 3;
        return tab[c].leaf != 0;
    };
    struct Item { int b, widx, depth; };
    std::vector<Item> queue;
    if (is_leaf(t.root)) { // single triangle
        WNode nd{}; for (int s = 0; s < 8; ++s) nd.child[s] = -1;
        nd.child[0] = -2; nd.tri_first[0] = 0; nd.ntri[0] = 1; nd.qbox[0] = sc.boxes[leaf_tri(t.root)];
        w.leaf_tris.push_back((uint32_t)leaf_tri(t.root)); w.nodes.push_back(nd); w.depth = 1;
        return w;
    }
    w.nodes.emplace_back();
    queue.push_back({t.root, 0, 1});
    for (size_t qi = 0; qi < queue.size(); ++qi) {
        const Item it = queue[qi];
        w.depth = std::max(w.depth, it.depth);
        int ch[8];
        const int nch = gather(it.b, ch);
        const Box& nb = t.nodes[it.b].box;
        const float ncx = 0.5f * (nb.lo[0] + nb.hi[0]), ncy = 0.5f * (nb.lo[1] + nb.hi[1]), ncz = 0.5f * (nb.lo[2] + nb.hi[2]);
        float cost[8][8];
        for (int k = 0; k < nch; ++k) {
            const Box& b = cbox(sc, t, ch[k]);
            float dx = 0.5f * (b.lo[0] + b.hi[0]) - ncx, dy = 0.5f * (b.lo[1] + b.hi[1]) - ncy, dz = 0.5f * (b.lo[2] + b.hi[2]) - ncz;
            for (int s = 0; s < 8; ++s) cost[k][s] = ((s & 4) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 1) ? dz : -dz);
        }
        int slot_child[8]; for (int s = 0; s < 8; ++s) slot_child[s] = INT32_MIN;
        uint32_t cd = 0, sd = 0;
        for (int i2 = 0; i2 < nch; ++i2) {
            float bc = -FLT_MAX; int bk = -1, bs = -1;
            for (int k = 0; k < nch; ++k) { if (cd >> k & 1) continue; for (int s = 0; s < 8; ++s) { if (sd >> s & 1) continue; if (cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; } } }
            if (bk < 0) { for (int k = 0; k < nch && bk < 0; ++k) if (!(cd >> k & 1)) bk = k; for (int s = 0; s < 8 && bs < 0; ++s) if (!(sd >> s & 1)) bs = s; }
            slot_child[bs] = ch[bk]; cd |= 1u << bk; sd |= 1u << bs;
        }
        // quantisation frame (as k_collapse)
        float scale[3];
        for (int a = 0; a < 3; ++a) {
            float s = (nb.hi[a] - nb.lo[a]) / 255.f; s = std::nextafter(s, FLT_MAX);
            int e; std::frexp(s, &e); float p2 = std::ldexp(1.f, e); if (std::ldexp(1.f, e - 1) >= s) p2 = std::ldexp(1.f, e - 1);
            scale[a] = std::max(p2, FLT_MIN);
        }
        WNode nd{};
        for (int s = 0; s < 8; ++s) {
            const int c = slot_child[s];
            nd.child[s] = -1; nd.ntri[s] = 0; nd.tri_first[s] = 0;
            if (c == INT32_MIN) continue;
            const Box& b = cbox(sc, t, c);
            for (int a = 0; a < 3; ++a) {
                float fl = std::floor((b.lo[a] - nb.lo[a]) / scale[a]), fh = std::ceil((b.hi[a] - nb.lo[a]) / scale[a]);
                fl = std::min(std::max(fl, 0.f), 255.f); fh = std::min(std::max(fh, 0.f), 255.f);
                nd.qbox[s].lo[a] = nb.lo[a] + (fl - 1.f / 256) * scale[a]; nd.qbox[s].hi[a] = nb.lo[a] + (fh + 1.f / 256) * scale[a];
            }
            if (child_is_leaf(c)) {
                std::vector<uint32_t> tt; subtree_tris(t, c, tt);
                nd.child[s] = -2; nd.tri_first[s] = (int)w.leaf_tris.size(); nd.ntri[s] = (int)tt.size();
                w.leaf_tris.insert(w.leaf_tris.end(), tt.begin(), tt.end());
                w.sah += b.area() / root_area * tt.size() * 0.3;
            } else {
                nd.child[s] = (int)w.nodes.size();
                w.nodes.emplace_back();
                queue.push_back({c, nd.child[s], it.depth + 1});
                w.sah += b.area() / root_area;
            }
        }
        w.nodes[it.widx] = nd;
    }
    return w;
}

// ---------------------------------------------------------------- traversal (counting)
struct Ray { V3 o, d; };
static inline bool tri_hit(const Tri& T, V3 o, V3 d, float* t) {
    V3 p = cross(d, T.e2); float det = dot(T.e1, p);
    if (std::fabs(det) <= 1e-8f) return false;
    float inv = 1 / det; V3 tv = o - T.v0; float u = dot(tv, p) * inv;
    if (u < 0 || u > 1) return false;
    V3 q = cross(tv, T.e1); float v = dot(d, q) * inv;
    if (v < 0 || u + v > 1) return false;
    float dist = dot(T.e2, q) * inv;
    if (dist > 0) { *t = dist; return true; }
    return false;
}
static inline float safe_inv(float x) { return 1.f / (std::fabs(x) > 1e-20f ? x : std::copysign(1e-20f, x)); }

struct Hit { float t; int tri; };
// g_order: 0 = octant order (traverse.cuh today); 1 = the NEAREST hit child first, the others in octant order; 2 = all hit children of a
// node in near-to-far order (pending lists kept sorted per node; popped groups still not re-tested); 3 = octant order but the two
// nearest... (unused)
static int g_order = 0;
static Hit trace(const Scene& sc, const WTree& w, const Ray& r, uint64_t* n_nodes, uint64_t* n_tris) {
    Hit h{999999999.f, -1};
    const float idx = safe_inv(r.d.x), idy = safe_inv(r.d.y), idz = safe_inv(r.d.z);
    const uint32_t oct_inv = (r.d.x >= 0 ? 4u : 0u) | (r.d.y >= 0 ? 2u : 0u) | (r.d.z >= 0 ? 1u : 0u);
    struct Group { int node; uint32_t bits; uint8_t order[8]; int nord; float tmin; float tn[8]; }; // bits: pending internal hits, bit (slot ^ oct_inv); order: sorted slots (mode 2)
    Group stack[256]; int sp = 0;
    Group cur{-1, 0, {0}, 0, 0.f, {0}};
    int next_node = 0; // root
    while (true) {
        if (next_node >= 0) {
            const WNode& nd = w.nodes[next_node];
            ++*n_nodes;
            uint32_t bits = 0;
            float tns[8]; int hs[8]; int nh = 0;
            for (int s = 0; s < 8; ++s) {
                if (nd.child[s] == -1) continue;
                const Box& b = nd.qbox[s];
                float x0 = (b.lo[0] - r.o.x) * idx, x1 = (b.hi[0] - r.o.x) * idx, y0 = (b.lo[1] - r.o.y) * idy, y1 = (b.hi[1] - r.o.y) * idy;
                float z0 = (b.lo[2] - r.o.z) * idz, z1 = (b.hi[2] - r.o.z) * idz;
                float tn = std::max(std::max(std::min(x0, x1), std::min(y0, y1)), std::max(std::min(z0, z1), 0.f));
                float tf = std::min(std::min(std::max(x0, x1), std::max(y0, y1)), std::min(std::max(z0, z1), h.t));
                if (!(tn <= tf)) continue;
                if (nd.child[s] >= 0) { bits |= 1u << (s ^ oct_inv); tns[nh] = tn; hs[nh++] = s; }
                else {
                    for (int k = 0; k < nd.ntri[s]; ++k) {
                        const int ti = (int)w.leaf_tris[nd.tri_first[s] + k];
                        ++*n_tris;
                        float t;
                        if (tri_hit(sc.tris[ti], r.o, r.d, &t) && t < h.t) { h.t = t; h.tri = ti; }
                    }
                }
            }
            cur = {next_node, bits, {0}, 0, FLT_MAX, {0}};
            for (int k = 0; k < nh; ++k) { cur.tmin = std::min(cur.tmin, tns[k]); cur.tn[hs[k]] = tns[k]; }
            if (g_order == 5 && nh) { // bound of the group that stays behind: min entry distance over the hits except the octant-first one
                int first = 0; for (int k = 1; k < nh; ++k) if ((hs[k] ^ (int)oct_inv) > (hs[first] ^ (int)oct_inv)) first = k;
                cur.tmin = FLT_MAX; for (int k = 0; k < nh; ++k) if (k != first) cur.tmin = std::min(cur.tmin, tns[k]);
                cur.nord = -1; // marks "first child not taken yet"
            }
            if (g_order == 2) { // sorted far -> near in order[], popped from the back
                int id[8]; for (int k = 0; k < nh; ++k) id[k] = k;
                std::sort(id, id + nh, [&](int a, int b) { return tns[a] > tns[b] || (tns[a] == tns[b] && (hs[a] ^ (int)oct_inv) < (hs[b] ^ (int)oct_inv)); });
                for (int k = 0; k < nh; ++k) cur.order[k] = (uint8_t)hs[id[k]];
                cur.nord = nh;
            } else if (g_order == 1 && nh > 1) {
                int bk = 0; for (int k = 1; k < nh; ++k) if (tns[k] < tns[bk]) bk = k;
                cur.order[0] = (uint8_t)hs[bk]; cur.nord = 1; // the nearest goes first
            }
        }
        int slot;
        if (g_order == 2) {
            if (cur.nord == 0) { if (sp == 0) break; cur = stack[--sp]; }
            slot = cur.order[--cur.nord];
            if (cur.nord) stack[sp++] = cur;
            next_node = w.nodes[cur.node].child[slot];
            cur.nord = 0;
            continue;
        }
        if (cur.bits == 0) {
            if (sp == 0) break;
            cur = stack[--sp];
            if (g_order == 3 && cur.tmin > h.t) { cur.bits = 0; next_node = -1; continue; } // the whole group lies behind the current hit
        }
        if (g_order == 5) {
            if (cur.nord == -1) cur.nord = 0;                                     // the first child of a fresh group is always taken
            else if (cur.tmin > h.t) { cur.bits = 0; next_node = -1; continue; }  // everything left in the group lies behind the hit
        }
        if (g_order == 4) { // octant order, every child re-checked against the current hit when it is taken
            const int bit = 31 - __builtin_clz(cur.bits); cur.bits &= ~(1u << bit); slot = bit ^ (int)oct_inv;
            if (cur.bits) stack[sp++] = cur;
            if (cur.tn[slot] > h.t) { cur.bits = 0; next_node = -1; continue; }
            next_node = w.nodes[cur.node].child[slot]; cur.bits = 0;
            continue;
        }
        if (cur.nord == 1) { slot = cur.order[0]; cur.nord = 0; cur.bits &= ~(1u << (slot ^ (int)oct_inv)); }
        else { const int bit = 31 - __builtin_clz(cur.bits); cur.bits &= ~(1u << bit); slot = bit ^ (int)oct_inv; }
        const int child = w.nodes[cur.node].child[slot];
        if (cur.bits) stack[sp++] = cur;
        next_node = child;
        cur.bits = 0;
    }
    return h;
}

// yardstick: the same tree walked in exact near-to-far order of the child boxes with the entry distance re-checked at pop time
// (what a per-child sorted stack would give; TL_IDEAL=1)
static Hit trace_ideal(const Scene& sc, const WTree& w, const Ray& r, uint64_t* n_nodes, uint64_t* n_tris) {
    Hit h{999999999.f, -1};
    const float idx = safe_inv(r.d.x), idy = safe_inv(r.d.y), idz = safe_inv(r.d.z);
    struct E { int node; float tn; };
    E stack[512]; int sp = 0;
    stack[sp++] = {0, 0.f};
    while (sp) {
        E e = stack[--sp];
        if (e.tn > h.t) continue;
        const WNode& nd = w.nodes[e.node];
        ++*n_nodes;
        E loc[8]; int nl = 0;
        for (int s = 0; s < 8; ++s) {
            if (nd.child[s] == -1) continue;
            const Box& b = nd.qbox[s];
            float x0 = (b.lo[0] - r.o.x) * idx, x1 = (b.hi[0] - r.o.x) * idx, y0 = (b.lo[1] - r.o.y) * idy, y1 = (b.hi[1] - r.o.y) * idy;
            float z0 = (b.lo[2] - r.o.z) * idz, z1 = (b.hi[2] - r.o.z) * idz;
            float tn = std::max(std::max(std::min(x0, x1), std::min(y0, y1)), std::max(std::min(z0, z1), 0.f));
            float tf = std::min(std::min(std::max(x0, x1), std::max(y0, y1)), std::min(std::max(z0, z1), h.t));
            if (!(tn <= tf)) continue;
            if (nd.child[s] >= 0) loc[nl++] = {nd.child[s], tn};
            else for (int k = 0; k < nd.ntri[s]; ++k) {
                const int ti = (int)w.leaf_tris[nd.tri_first[s] + k];
                ++*n_tris;
                float t;
                if (tri_hit(sc.tris[ti], r.o, r.d, &t) && t < h.t) { h.t = t; h.tri = ti; }
            }
        }
        std::sort(loc, loc + nl, [](const E& a, const E& b) { return a.tn > b.tn; }); // farthest first -> nearest on top
        for (int k = 0; k < nl; ++k) stack[sp++] = loc[k];
    }
    return h;
}
static bool g_ideal = false;
static inline Hit trace_any(const Scene& sc, const WTree& w, const Ray& r, uint64_t* n, uint64_t* t) { return g_ideal ? trace_ideal(sc, w, r, n, t) : trace(sc, w, r, n, t); }

// ---------------------------------------------------------------- scenes & rays
static Scene make_scene(const std::string& spec) {
    std::vector<float> pos;
    auto arg = [&](size_t k, long def) { size_t p = 0; std::string s = spec; for (size_t i = 0; i < k; ++i) { p = s.find(':'); if (p == std::string::npos) return def; s = s.substr(p + 1); } return atol(s.c_str()); };
    if (spec.rfind("soup", 0) == 0) { uint64_t n = (uint64_t)arg(1, 1000000); pos.resize(9 * n); vmn_gen_soup(pos.data(), nullptr, nullptr, n, 1234); }
    else if (spec.rfind("spheres", 0) == 0) { int lvl = (int)arg(1, 6); uint64_t n = vmn_gen_spheres_count(4, 3, lvl); pos.resize(9 * n); vmn_gen_spheres(pos.data(), nullptr, nullptr, 4, 3, lvl, 150.f); }
    else if (spec.rfind("city", 0) == 0) { int g = (int)arg(1, 16), tess = (int)arg(2, 23); uint64_t n = vmn_gen_city_count(g, g, tess, 2024); pos.resize(9 * n); vmn_gen_city(pos.data(), nullptr, nullptr, g, g, tess, 2024); }
    else { fprintf(stderr, "unknown scene %s\n", spec.c_str()); exit(1); }
    Scene sc; const size_t n = pos.size() / 9;
    sc.tris.resize(n); sc.boxes.resize(n); sc.bounds.reset();
    for (size_t i = 0; i < n; ++i) {
        const float* p = &pos[9 * i];
        V3 a{p[0], p[1], p[2]}, b{p[3], p[4], p[5]}, c{p[6], p[7], p[8]};
        sc.tris[i] = {a, b - a, c - a};
        Box bx; for (int k = 0; k < 3; ++k) { bx.lo[k] = std::min(std::min(p[k], p[3 + k]), p[6 + k]); bx.hi[k] = std::max(std::max(p[k], p[3 + k]), p[6 + k]); }
        sc.boxes[i] = bx; sc.bounds.grow(bx);
    }
    return sc;
}

// glm-style camera of pathtracer.cpp:216-221 / 251-280 (as csrc/api.cu, wavefront.cu), jitter-free pixel centres
static void rotate(float m[16], float angle, float ax, float ay, float az) {
    const float c = std::cos(angle), s = std::sin(angle);
    const float x = ax, y = ay, z = az, tx = x * (1 - c), ty = y * (1 - c), tz = z * (1 - c);
    float R[3][3] = {{c + tx * x, tx * y + s * z, tx * z - s * y}, {ty * x - s * z, c + ty * y, ty * z + s * x}, {tz * x + s * y, tz * y - s * x, c + tz * z}};
    float out[16];
    for (int col = 0; col < 3; ++col) for (int r = 0; r < 4; ++r) out[col * 4 + r] = m[r] * R[col][0] + m[4 + r] * R[col][1] + m[8 + r] * R[col][2];
    for (int r = 0; r < 4; ++r) out[12 + r] = m[12 + r];
    memcpy(m, out, sizeof(out));
}

struct PathCfg { int w, h, spp; V3 pos; float rot_deg[3]; float room_x0, room_x1, room_y1, room_z0, room_z1; };

// rays of a FIXED-mode frame: camera ray + up to 4 cosine bounces; walls of the reference room as planes. The tree under test traces them.
static void path_rays(const Scene& sc, const WTree& w, const PathCfg& cfg, uint64_t* rays, uint64_t* nodes, uint64_t* tris) {
    float m[16]; for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.f : 0.f;
    const float PI = 3.1415926535f;
    rotate(m, -cfg.rot_deg[1] * PI / 180, 0, 1, 0); rotate(m, -cfg.rot_deg[0] * PI / 180, 1, 0, 0); rotate(m, cfg.rot_deg[2] * PI / 180, 0, 0, 1);
    const float sx = 3.6f, sy = 3.6f * cfg.h / cfg.w, dist = 6.f;
    uint64_t R = 0, N = 0, T = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : R, N, T)
    for (int p = 0; p < cfg.w * cfg.h; ++p) {
        std::mt19937 rng(p * 9781 + 7);
        std::uniform_real_distribution<float> U(0.f, 1.f);
        for (int s = 0; s < cfg.spp; ++s) {
            float hx = ((p % cfg.w) + U(rng)) / cfg.w - 0.5f, hy = ((p / cfg.w) + U(rng)) / cfg.h - 0.5f;
            float g0 = hx * sx, g1 = -hy * sy, g2 = -dist;
            V3 d = normalize({m[0] * g0 + m[4] * g1 + m[8] * g2, m[1] * g0 + m[5] * g1 + m[9] * g2, m[2] * g0 + m[6] * g1 + m[10] * g2});
            V3 o = cfg.pos;
            for (int seg = 0; seg < 5; ++seg) {
                ++R;
                Hit h = trace_any(sc, w, {o, d}, &N, &T);
                // walls
                float tw = FLT_MAX; V3 nw{0, 0, 0};
                auto plane = [&](float num, float den, V3 n) { if (den != 0) { float t = num / den; if (t > 1e-3f && t < tw) { tw = t; nw = n; } } };
                plane(0.f - o.y, d.y, {0, 1, 0}); plane(cfg.room_y1 - o.y, d.y, {0, -1, 0});
                plane(cfg.room_x0 - o.x, d.x, {1, 0, 0}); plane(cfg.room_x1 - o.x, d.x, {-1, 0, 0});
                plane(cfg.room_z0 - o.z, d.z, {0, 0, 1}); plane(cfg.room_z1 - o.z, d.z, {0, 0, -1});
                float t; V3 n;
                if (h.tri >= 0 && h.t < tw) { t = h.t; n = normalize(cross(sc.tris[h.tri].e1, sc.tris[h.tri].e2)); }
                else if (tw < FLT_MAX) { t = tw; n = nw; }
                else break;
                if (dot(n, d) > 0) n = n * -1.f;
                V3 P = o + d * t;
                const float r1 = 2 * PI * U(rng), r2 = U(rng), r2s = std::sqrt(r2);
                V3 u = normalize(cross(std::fabs(n.x) > .1f ? V3{0, 1, 0} : V3{1, 0, 0}, n)), v = cross(n, u);
                d = normalize(u * (std::cos(r1) * r2s) + v * (std::sin(r1) * r2s) + n * std::sqrt(1 - r2));
                o = P - (d * -1.f) * 0.f + n * 0.001f;
            }
        }
    }
    *rays = R; *nodes = N; *tris = T;
}

int main(int argc, char** argv) {
    if (argc < 5) { fprintf(stderr, "usage: treelab <scene> <rays> <builder> <collapse>\n  scene: soup:N | spheres[:level] | city:G[:tess]\n  rays: uniform:N | path:W:H:spp[:c2|c4]\n  builder: lbvh | ploc:R | sah\n  collapse: greedy | dp:Cprim:Pmax\n"); return 1; }
    const std::string sspec = argv[1], rspec = argv[2], bspec = argv[3], cspec = argv[4];
    g_ideal = getenv("TL_IDEAL") && getenv("TL_IDEAL")[0] == '1';
    g_order = getenv("TL_ORDER") ? atoi(getenv("TL_ORDER")) : 0;
    g_emc = getenv("TL_EMC") ? atoi(getenv("TL_EMC")) : 0;
    double t0 = omp_get_wtime();
    Scene sc = make_scene(sspec);
    double t1 = omp_get_wtime();
    BTree bt;
    if (bspec == "lbvh") bt = build_lbvh(sc);
    else if (bspec.rfind("ploc", 0) == 0) bt = build_ploc(sc, bspec.size() > 5 ? atoi(bspec.c_str() + 5) : 16);
    else if (bspec == "sah") bt = build_sah(sc);
    else { fprintf(stderr, "unknown builder\n"); return 1; }
    double t2 = omp_get_wtime();
    double sah2 = 0; for (auto& n : bt.nodes) sah2 += n.area; sah2 /= bt.nodes[is_leaf(bt.root) ? 0 : bt.root].area;
    CollapseOpt co;
    if (cspec.rfind("dp", 0) == 0) { co.dp = true; sscanf(cspec.c_str(), "dp:%f:%d", &co.c_prim, &co.p_max); }
    WTree w = collapse(sc, bt, co);
    double t3 = omp_get_wtime();
    uint64_t rays = 0, nodes = 0, tris = 0, hits = 0;
    if (rspec.rfind("uniform", 0) == 0) {
        const uint64_t n = (uint64_t)atol(rspec.c_str() + 8);
        std::vector<float> r8(8 * n);
        vmn_gen_rays_uniform(r8.data(), n, 4321, 0.f, 1000.f);
        uint64_t N = 0, T = 0, H = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : N, T, H)
        for (uint64_t i = 0; i < n; ++i) {
            Ray r{{r8[8 * i], r8[8 * i + 1], r8[8 * i + 2]}, {r8[8 * i + 4], r8[8 * i + 5], r8[8 * i + 6]}};
            Hit h = trace_any(sc, w, r, &N, &T);
            H += h.tri >= 0;
        }
        rays = n; nodes = N; tris = T; hits = H;
    } else {
        PathCfg cfg{}; char which[16] = "c2";
        sscanf(rspec.c_str(), "path:%d:%d:%d:%15s", &cfg.w, &cfg.h, &cfg.spp, which);
        if (!strcmp(which, "c4")) { cfg.pos = {0.f, 900.f, 1950.f}; cfg.rot_deg[0] = 25.f; }
        else cfg.pos = {0.f, 500.f, 1900.f};
        cfg.room_x0 = -2000; cfg.room_x1 = 2000; cfg.room_y1 = 1000; cfg.room_z0 = -2000; cfg.room_z1 = 2000;
        path_rays(sc, w, cfg, &rays, &nodes, &tris);
    }
    double t4 = omp_get_wtime();
    const double npr = (double)nodes / rays, tpr = (double)tris / rays;
    printf("{\"scene\": \"%s\", \"tris\": %zu, \"rays\": \"%s\", \"builder\": \"%s\", \"collapse\": \"%s\", \"nodes_per_ray\": %.3f, \"tris_per_ray\": %.3f, "
           "\"bytes_per_ray\": %.0f, \"wide_nodes\": %zu, \"wide_depth\": %d, \"sah_binary\": %.2f, \"sah_wide\": %.3f, \"hit_rate\": %.4f, "
           "\"gen_s\": %.1f, \"build_s\": %.1f, \"collapse_s\": %.1f, \"trace_s\": %.1f}\n",
           sspec.c_str(), sc.tris.size(), rspec.c_str(), bspec.c_str(), cspec.c_str(), npr, tpr, 48 + 80 * npr + 48 * tpr, w.nodes.size(), w.depth, sah2,
           w.sah, (double)hits / std::max<uint64_t>(rays, 1), t1 - t0, t2 - t1, t3 - t2, t4 - t3);
    return 0;
}
