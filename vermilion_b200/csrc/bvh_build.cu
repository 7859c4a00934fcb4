// bvh_build.cu — GPU BVH construction (replaces BVH::build, reference core/accelerators/bvh.cpp:179-279).
//
// The reference builds a binary midpoint-split tree serially on the host. This builder produces a
// DIFFERENT tree with the same closest hits (tie order excepted), entirely on the GPU:
//   1. k_prim_bounds   per-triangle box + scene box (atomic min/max on order-preserving float bits)
//   2. k_morton        63-bit Morton code of the box centre, 21 bits per axis
//   3. radix sort      (key = Morton code, value = triangle id)
//   4. k_leaves        leaf boxes + Morton-ordered 48-byte triangle records
//   5. k_hierarchy     Karras 2012 binary radix tree, one thread per internal node
//   6. k_refit         bottom-up boxes and subtree triangle counts, second-arrival atomics; the same climb fills the cost table of
//                      the SAH-optimal wide collapse (Ylitie, Karras & Laine 2017, "Efficient incoherent ray traversal on GPUs through
//                      compressed wide BVHs", section 3.2): C(n, i) = cheapest way to cover the subtree of n with at most i wide-node
//                      slots, i = 1..7, and the split that realises it
//   7. k_collapse      level-synchronous collapse of the binary tree into 8-wide nodes following those decisions (which subtrees become
//                      the <= 8 children, which of them are leaves of <= 3 triangles), octant-ordered child slots, 8-bit quantised
//                      child boxes -> 80-byte WideNode, leaf triangles re-laid out contiguously per node.
// Measured on the host restatement of this builder and of the traversal (tools/treelab): against the greedy "open the largest child"
// collapse of round 1 the triangles tested per ray drop from 18.5 to 8.0 on the C3 soup and from 1.55 to 1.0 on C2 at the same number
// of node visits; a PLOC or full-sweep SAH binary tree under the same collapse is WORSE than the Morton tree for the octant-ordered
// traversal (city: 13.8 nodes per ray with LBVH, 16.4 PLOC, 15.4 SAH), so the radix tree stays. VMN_COLLAPSE=greedy selects the old rule (A/B).
#include "engine.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <algorithm>
#include <cfloat>
#include <vector>

namespace vmn {

// order-preserving float <-> uint mapping for atomicMin/atomicMax
__device__ __forceinline__ uint32_t f2ord(float f) { uint32_t u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ __forceinline__ float ord2f(uint32_t u) {
    uint32_t v = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(v);
#else
    float f; memcpy(&f, &v, 4); return f;
#endif
}

__global__ void k_init_bounds(uint32_t* sb) {
    if (threadIdx.x < 3) sb[threadIdx.x] = 0xFFFFFFFFu;      // min
    else if (threadIdx.x < 6) sb[threadIdx.x] = 0u;          // max
}

// pos: 9 floats per triangle (v0,v1,v2)
__global__ void k_prim_bounds(const float* __restrict__ pos, uint32_t n, float4* __restrict__ pmin, float4* __restrict__ pmax,
                              uint32_t* __restrict__ sb) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n) {
        const float* p = pos + 9ull * i;
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float x0 = p[a], x1 = p[3 + a], x2 = p[6 + a];
            lo[a] = fminf(fminf(x0, x1), x2);
            hi[a] = fmaxf(fmaxf(x0, x1), x2);
        }
        pmin[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
        pmax[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    // warp reduce, one atomic per warp per component
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float l = lo[a], h = hi[a];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            l = fminf(l, __shfl_xor_sync(0xffffffffu, l, o));
            h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, o));
        }
        if ((threadIdx.x & 31) == 0 && l <= h) { atomicMin(&sb[a], f2ord(l)); atomicMax(&sb[3 + a], f2ord(h)); }
    }
}

__device__ __forceinline__ uint64_t spread21(uint64_t x) { // 21 bits -> every third bit
    x &= 0x1FFFFFull;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void k_morton(const float4* __restrict__ pmin, const float4* __restrict__ pmax, uint32_t n, const uint32_t* __restrict__ sb,
                         uint64_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float sl[3] = {ord2f(sb[0]), ord2f(sb[1]), ord2f(sb[2])}, sh[3] = {ord2f(sb[3]), ord2f(sb[4]), ord2f(sb[5])};
    float4 a = pmin[i], b = pmax[i];
    float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
    uint64_t q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float ext = sh[k] - sl[k];
        float f = ext > 0.f ? (c[k] - sl[k]) / ext : 0.f;
        f = fminf(fmaxf(f, 0.f), 1.f);
        q[k] = (uint64_t)fminf(f * 2097152.f, 2097151.f);
    }
    keys[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    vals[i] = i;
}

// leaves of the binary tree + Morton-ordered triangle records
__global__ void k_leaves(const float* __restrict__ pos, const uint32_t* __restrict__ sorted_ids, uint32_t n,
                         const float4* __restrict__ pmin, const float4* __restrict__ pmax, float4* __restrict__ bmin,
                         float4* __restrict__ bmax, TriRec* __restrict__ tris_sorted, uint32_t* __restrict__ count) {
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    uint32_t id = sorted_ids[k];
    uint32_t node = (n - 1) + k;
    bmin[node] = pmin[id];
    bmax[node] = pmax[id];
    count[node] = 1;
    const float* p = pos + 9ull * id;
    TriRec t;
    t.r0 = make_float4(p[0], p[1], p[2], __uint_as_float(id));
    t.r1 = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], 0.f);
    t.r2 = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], 0.f);
    tris_sorted[k] = t;
}

// records indexed by ORIGINAL id: intersection record (for getNormal's barycentrics) and shading record
__global__ void k_orig_records(const float* __restrict__ pos, const float* __restrict__ nrm, const float* __restrict__ uv, uint32_t n,
                               TriRec* __restrict__ tris_orig, ShadeRec* __restrict__ shade) {
    uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= n) return;
    const float* p = pos + 9ull * id;
    TriRec t;
    t.r0 = make_float4(p[0], p[1], p[2], __uint_as_float(id));
    t.r1 = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], 0.f);
    t.r2 = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], 0.f);
    tris_orig[id] = t;
    float nn[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0}, tt[6] = {0, 0, 0, 0, 0, 0};
    if (nrm) for (int k = 0; k < 9; ++k) nn[k] = nrm[9ull * id + k];
    if (uv) for (int k = 0; k < 6; ++k) tt[k] = uv[6ull * id + k];
    ShadeRec s;
    s.s0 = make_float4(nn[0], nn[1], nn[2], tt[0]);
    s.s1 = make_float4(nn[3], nn[4], nn[5], tt[1]);
    s.s2 = make_float4(nn[6], nn[7], nn[8], tt[2]);
    s.s3 = make_float4(tt[3], tt[4], tt[5], 0.f);
    shade[id] = s;
}

// Karras 2012. delta(i,j) = common prefix length of keys i and j, ties broken by index.
__device__ __forceinline__ int lbvh_delta(const uint64_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(a ^ b));
}

__global__ void k_hierarchy(const uint64_t* __restrict__ keys, int n, int2* __restrict__ child, int* __restrict__ parent,
                            int* __restrict__ first) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (lbvh_delta(keys, n, i, i + 1) - lbvh_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = lbvh_delta(keys, n, i, i - d);
    int lmax = 2;
    while (lbvh_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (lbvh_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = lbvh_delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (lbvh_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    first[i] = lo;
    int left = (lo == gamma) ? (n - 1) + gamma : gamma;
    int right = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    child[i] = make_int2(left, right);
    parent[left] = i;
    parent[right] = i;
    if (i == 0) parent[0] = -1;
}

// Cost table row of one internal binary node (32 B): c[i-1] = C(n, i), i = 1..7; k: bits [3(j-2), 3(j-2)+3) = how many of the j slots of
// distribute(n, j) go to the left child, j = 2..8; bit 31 = "n as a single slot is a LEAF" (else an internal wide node).
struct DpRow { float c[7]; uint32_t k; };
#define VMN_LEAF_MAX 3u // triangles per leaf slot (unary 3-bit count in meta)

__device__ __forceinline__ float box_area(float4 a, float4 b) {
    float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
    return dx * dy + dy * dz + dz * dx;
}

// C(child, 1..7): a triangle costs its box area x c_prim whatever the budget
__device__ __forceinline__ void dp_load(const DpRow* __restrict__ tab, int c, int nint, float area, float c_prim, float out[7]) {
    if (c >= nint) {
#pragma unroll
        for (int i = 0; i < 7; ++i) out[i] = area * c_prim;
    } else {
        const float4 a = __ldcg(reinterpret_cast<const float4*>(tab + c)), b = __ldcg(reinterpret_cast<const float4*>(tab + c) + 1);
        out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w; out[4] = b.x; out[5] = b.y; out[6] = b.z;
    }
}

__global__ void k_refit(int n, const int2* __restrict__ child, const int* __restrict__ parent, float4* bmin, float4* bmax,
                        uint32_t* count, uint32_t* __restrict__ flags, DpRow* tab, float c_prim) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int nint = n - 1;
    int node = parent[(n - 1) + k];
    while (node >= 0) {
        if (atomicAdd(&flags[node], 1u) == 0u) return; // first arrival stops; the second sees both children
        __threadfence();
        int2 c = child[node];
        float4 a0 = __ldcg(&bmin[c.x]), a1 = __ldcg(&bmax[c.x]), b0 = __ldcg(&bmin[c.y]), b1 = __ldcg(&bmax[c.y]);
        const float4 lo = make_float4(fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z), 0.f);
        const float4 hi = make_float4(fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z), 0.f);
        const uint32_t cnt = __ldcg(&count[c.x]) + __ldcg(&count[c.y]);
        __stcg(&bmin[node], lo);
        __stcg(&bmax[node], hi);
        __stcg(&count[node], cnt);
        // cost table (Ylitie et al. 2017, eq. 1-3; C_node = 1, C_prim = c_prim, areas left unnormalised: the decisions are scale-free)
        float L[7], R[7];
        dp_load(tab, c.x, nint, box_area(a0, a1), c_prim, L);
        dp_load(tab, c.y, nint, box_area(b0, b1), c_prim, R);
        float dist[9];
        uint32_t kp = 0;
#pragma unroll
        for (int j = 2; j <= 8; ++j) {
            float best = 3.0e38f; int bk = 1;
#pragma unroll
            for (int q = 1; q < j; ++q) {
                if (q > 7 || j - q > 7) continue;
                const float v = L[q - 1] + R[j - q - 1];
                if (v < best) { best = v; bk = q; }
            }
            dist[j] = best;
            kp |= (uint32_t)bk << (3 * (j - 2));
        }
        const float area = box_area(lo, hi);
        const float c_leaf = cnt <= VMN_LEAF_MAX ? area * (float)cnt * c_prim : 3.0e38f;
        const float c_int = dist[8] + area;
        if (c_leaf <= c_int) kp |= 0x80000000u;
        float cc[7];
        cc[0] = fminf(c_leaf, c_int);
#pragma unroll
        for (int i = 2; i <= 7; ++i) cc[i - 1] = fminf(dist[i], cc[i - 2]);
        __stcg(reinterpret_cast<float4*>(tab + node), make_float4(cc[0], cc[1], cc[2], cc[3]));
        __stcg(reinterpret_cast<float4*>(tab + node) + 1, make_float4(cc[4], cc[5], cc[6], __uint_as_float(kp)));
        __threadfence();
        node = parent[node];
    }
}


// ---- collapse to the compressed 8-wide BVH ----------------------------------------------------
struct CollapseCtl { uint32_t n_nodes, n_tris, q_out, pad; };

// One thread turns one binary subtree root into one WideNode and enqueues its internal children.
// dp != 0: the <= 8 children are the slots distribute(root, 8) of the cost table hands out, and a child is a leaf where the table says
// so; dp == 0: the round-1 rule (open the child with the largest area until there are 8; a subtree of <= 3 triangles is a leaf).
__device__ __forceinline__ bool child_internal(int c, int nint, const uint32_t* __restrict__ count, const DpRow* __restrict__ tab, int dp) {
    if (c >= nint) return false;
    return dp ? (tab[c].k & 0x80000000u) == 0u : count[c] > VMN_LEAF_MAX;
}

__global__ void k_collapse(const uint2* __restrict__ q_in, uint32_t n_in, uint2* __restrict__ q_out, CollapseCtl* ctl, int n,
                           const int2* __restrict__ child, const int* __restrict__ first, const uint32_t* __restrict__ count,
                           const float4* __restrict__ bmin, const float4* __restrict__ bmax, const TriRec* __restrict__ tris_sorted,
                           WideNode* __restrict__ nodes, LeafTri* __restrict__ tris, const DpRow* __restrict__ tab, int dp, uint32_t* err) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_in) return;
    const int bnode = (int)q_in[w].x;
    const uint32_t widx = q_in[w].y;
    const int nint = n - 1; // binary nodes >= nint are leaves

    int ch[8];
    int nch = 0;
    if (bnode >= nint) { ch[0] = bnode; nch = 1; } // single-triangle scene
    else if (dp) {
        // follow the table: (subtree, budget) pairs on a small stack; a pair with budget i is covered by the i' <= i slots that realise
        // C(c, i) — i' is where the running minimum last improved — and i' == 1 means "this subtree is one child"
        int st_node[8], st_budget[8], sp = 0;
        { const int2 c = child[bnode]; const uint32_t k = (tab[bnode].k >> 18) & 7u; // distribute(root, 8)
          st_node[0] = c.y; st_budget[0] = 8 - (int)k; st_node[1] = c.x; st_budget[1] = (int)k; sp = 2; }
        while (sp) {
            --sp;
            const int c = st_node[sp];
            int i = st_budget[sp];
            if (c >= nint || i <= 1) { ch[nch++] = c; continue; }
            const DpRow T = tab[c];
            while (i >= 2 && T.c[i - 1] == T.c[i - 2]) --i;
            if (i == 1) { ch[nch++] = c; continue; }
            const int k = (int)((T.k >> (3 * (i - 2))) & 7u);
            const int2 cc = child[c];
            st_node[sp] = cc.y; st_budget[sp] = i - k; ++sp;
            st_node[sp] = cc.x; st_budget[sp] = k; ++sp;
        }
    } else {
        int2 c = child[bnode]; ch[0] = c.x; ch[1] = c.y; nch = 2;
        while (nch < 8) {
            int best = -1; float bestA = -1.f;
            for (int k = 0; k < nch; ++k) {
                int c = ch[k];
                if (c < nint && count[c] > VMN_LEAF_MAX) {
                    float a = box_area(bmin[c], bmax[c]);
                    if (a > bestA) { bestA = a; best = k; }
                }
            }
            if (best < 0) break;
            int2 c = child[ch[best]];
            ch[best] = c.x; ch[nch++] = c.y;
        }
    }

    const float4 nlo = bmin[bnode], nhi = bmax[bnode];
    const float ncx = 0.5f * (nlo.x + nhi.x), ncy = 0.5f * (nlo.y + nhi.y), ncz = 0.5f * (nlo.z + nhi.z);

    // greedy octant-ordered slot assignment: slot bit2/1/0 set <=> child sits towards +x/+y/+z
    float cost[8][8];
    for (int k = 0; k < nch; ++k) {
        float4 a = bmin[ch[k]], b = bmax[ch[k]];
        float dx = 0.5f * (a.x + b.x) - ncx, dy = 0.5f * (a.y + b.y) - ncy, dz = 0.5f * (a.z + b.z) - ncz;
        for (int s = 0; s < 8; ++s)
            cost[k][s] = ((s & 4) ? dx : -dx) + ((s & 2) ? dy : -dy) + ((s & 1) ? dz : -dz);
    }
    int slot_child[8];
    for (int s = 0; s < 8; ++s) slot_child[s] = -1;
    uint32_t child_done = 0, slot_done = 0;
    for (int it = 0; it < nch; ++it) {
        float bc = -FLT_MAX; int bk = -1, bs = -1;
        for (int k = 0; k < nch; ++k) {
            if (child_done >> k & 1u) continue;
            for (int s = 0; s < 8; ++s) {
                if (slot_done >> s & 1u) continue;
                if (cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; }
            }
        }
        if (bk < 0) { // all costs NaN/-inf: fill first free
            for (int k = 0; k < nch && bk < 0; ++k) if (!(child_done >> k & 1u)) bk = k;
            for (int s = 0; s < 8 && bs < 0; ++s) if (!(slot_done >> s & 1u)) bs = s;
        }
        slot_child[bs] = ch[bk]; child_done |= 1u << bk; slot_done |= 1u << bs;
    }

    // quantisation frame
    float ext[3] = {nhi.x - nlo.x, nhi.y - nlo.y, nhi.z - nlo.z};
    uint32_t eb[3]; float scale[3];
    for (int a = 0; a < 3; ++a) {
        float s = __fdiv_ru(ext[a], 255.f);
        uint32_t e = (__float_as_uint(s) + 0x7FFFFFu) >> 23; // next power of two >= s
        e = min(max(e, 1u), 254u);
        eb[a] = e; scale[a] = __uint_as_float(e << 23);
    }
    const float pl[3] = {nlo.x, nlo.y, nlo.z};

    uint32_t imask = 0, n_int = 0, n_leaf_tris = 0;
    for (int s = 0; s < 8; ++s) {
        int c = slot_child[s];
        if (c < 0) continue;
        if (child_internal(c, nint, count, tab, dp)) { imask |= 1u << s; ++n_int; }
        else n_leaf_tris += count[c];
    }
    uint32_t cbase = n_int ? atomicAdd(&ctl->n_nodes, n_int) : 0u;
    uint32_t tbase = n_leaf_tris ? atomicAdd(&ctl->n_tris, n_leaf_tris) : 0u;
    uint32_t qbase = n_int ? atomicAdd(&ctl->q_out, n_int) : 0u;

    uint8_t meta[8], qlo[3][8], qhi[3][8];
    uint32_t toff = 0, irank = 0;
    for (int s = 0; s < 8; ++s) {
        int c = slot_child[s];
        if (c < 0) {
            meta[s] = 0;
            for (int a = 0; a < 3; ++a) { qlo[a][s] = 255; qhi[a][s] = 0; }
            continue;
        }
        float4 a4 = bmin[c], b4 = bmax[c];
        float lo[3] = {a4.x, a4.y, a4.z}, hi[3] = {b4.x, b4.y, b4.z};
        for (int a = 0; a < 3; ++a) {
            float fl = floorf(__fsub_rd(lo[a], pl[a]) / scale[a]);
            float fh = ceilf(__fsub_ru(hi[a], pl[a]) / scale[a]);
            qlo[a][s] = (uint8_t)fminf(fmaxf(fl, 0.f), 255.f);
            qhi[a][s] = (uint8_t)fminf(fmaxf(fh, 0.f), 255.f);
        }
        if (imask >> s & 1u) {
            meta[s] = (uint8_t)((1u << 5) | (24u + (uint32_t)s));
            VMN_CHECK(err, qbase + irank < (uint32_t)n && cbase + irank < (uint32_t)n);
            q_out[qbase + irank] = make_uint2((uint32_t)c, cbase + irank);
            ++irank;
        } else {
            uint32_t cnt = count[c];
            uint32_t f = (c < nint) ? (uint32_t)first[c] : (uint32_t)(c - nint);
            meta[s] = (uint8_t)((((1u << cnt) - 1u) << 5) | toff);
            for (uint32_t j = 0; j < cnt; ++j) {
                VMN_CHECK(err, f + j < (uint32_t)n && tbase + toff + j < (uint32_t)n);
                const TriRec t = tris_sorted[f + j]; LeafTri& o = tris[tbase + toff + j]; o.r0 = t.r0; o.r1 = t.r1; o.r2 = t.r2;
            }
            toff += cnt;
        }
    }
    auto pack4 = [](const uint8_t* b) { return (uint32_t)b[0] | ((uint32_t)b[1] << 8) | ((uint32_t)b[2] << 16) | ((uint32_t)b[3] << 24); };
    WideNode nd{};
    nd.q0 = make_uint4(__float_as_uint(pl[0]), __float_as_uint(pl[1]), __float_as_uint(pl[2]), eb[0] | (eb[1] << 8) | (eb[2] << 16) | (imask << 24));
    nd.q1 = make_uint4(cbase, tbase, pack4(meta), pack4(meta + 4));
    nd.q2 = make_uint4(pack4(qlo[0]), pack4(qlo[0] + 4), pack4(qlo[1]), pack4(qlo[1] + 4));
    nd.q3 = make_uint4(pack4(qlo[2]), pack4(qlo[2] + 4), pack4(qhi[0]), pack4(qhi[0] + 4));
    nd.q4 = make_uint4(pack4(qhi[1]), pack4(qhi[1] + 4), pack4(qhi[2]), pack4(qhi[2] + 4));
    VMN_CHECK(err, widx < (uint32_t)n && toff <= 24u && nch <= 8);
    nodes[widx] = nd;
}

static inline uint32_t cdiv(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

// The build in two parts. build_geometry needs the positions only (bounds, Morton order, hierarchy, cost table, collapse): it is what
// vmn_scene_set_triangles runs while the normals and uvs are still crossing PCIe on the copy stream. build_records turns positions +
// normals + uvs into the by-original-id records k_shade gathers; vmn_scene_commit runs it once everything has arrived.
vmn_status build_geometry(vmn_ctx* c) {
    const uint64_t n64 = c->n_tris;
    if (n64 >= 0x7FFFFFF0ull) return fail(c, VMN_ERR_INVALID, "too many triangles (%llu)", (unsigned long long)n64);
    const uint32_t n = (uint32_t)n64;
    cudaStream_t st = c->stream;
    c->scene = SceneDev{};
    c->scene.err = c->errflag.as<uint32_t>();
    c->scene.n_tris = n;
    c->n_wide = 0;
    c->trav_bytes = 0;
    c->stats.n_triangles = n; c->stats.n_nodes = 0; c->stats.n_wide_nodes = 0;
    c->stats.node_bytes = 80; c->stats.tri_bytes = 48; // payload bytes
    if (n == 0) return VMN_OK;

    VMN_CUDA(c, cudaEventRecord(c->ev[0], st));
    const uint32_t nnodes = 2 * n - 1;
    // build scratch lives in the context and only grows: a re-commit of a same-sized scene (per-frame rebuild, bench e2e) does not
    // touch cudaMalloc/cudaFree (cudaFree synchronises the device; 15 of them per build cost more than the build itself)
    DevBuf &pmin = c->build_scratch[0], &pmax = c->build_scratch[1], &keys = c->build_scratch[2], &keys2 = c->build_scratch[3],
           &vals = c->build_scratch[4], &vals2 = c->build_scratch[5], &parent = c->build_scratch[6], &first = c->build_scratch[7],
           &count = c->build_scratch[8], &flags = c->build_scratch[9], &sb = c->build_scratch[10], &tmp = c->build_scratch[11],
           &qa = c->build_scratch[12], &qb = c->build_scratch[13], &ctl = c->build_scratch[14], &dptab = c->build_scratch[15];
    // collapse rule and its leaf-cost constant (node cost = 1): env knobs for A/B, read at every commit
    const char* e_rule = getenv("VMN_COLLAPSE");
    const int dp = !(e_rule && !strcmp(e_rule, "greedy"));
    const char* e_cp = getenv("VMN_DP_CPRIM");
    const float c_prim = e_cp ? (float)atof(e_cp) : 0.4f;
    vmn_status s = VMN_OK;
    auto cleanup = [&]() {};
#define TRY(x) do { s = (x); if (s != VMN_OK) { cleanup(); return s; } } while (0)
#define dev_alloc dev_reserve /* grow-only inside the build */
#define TRYCUDA(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { cleanup(); return fail(c, VMN_ERR_CUDA, "%s: %s", #x, cudaGetErrorString(e__)); } } while (0)
    TRY(dev_alloc(c, pmin, 16ull * n)); TRY(dev_alloc(c, pmax, 16ull * n));
    TRY(dev_alloc(c, keys, 8ull * n)); TRY(dev_alloc(c, keys2, 8ull * n));
    TRY(dev_alloc(c, vals, 4ull * n)); TRY(dev_alloc(c, vals2, 4ull * n));
    TRY(dev_alloc(c, parent, 4ull * nnodes)); TRY(dev_alloc(c, first, 4ull * n));
    TRY(dev_alloc(c, count, 4ull * nnodes)); TRY(dev_alloc(c, flags, 4ull * n));
    TRY(dev_alloc(c, sb, 32)); TRY(dev_alloc(c, ctl, sizeof(CollapseCtl)));
    TRY(dev_alloc(c, dptab, sizeof(DpRow) * (uint64_t)n));
    TRY(dev_alloc(c, c->b_min, 16ull * nnodes)); TRY(dev_alloc(c, c->b_max, 16ull * nnodes));
    TRY(dev_alloc(c, c->b_child, 8ull * n));
    TRY(dev_alloc(c, c->tris_sorted, sizeof(TriRec) * (uint64_t)n));
    // the two arrays the traversal reads — leaf triangles, then the wide nodes — share one allocation, so that a single L2 access-policy
    // window (api.cu: apply_l2_policy) can cover both: [tris: n records][nodes: capacity n, n_wide used]
    TRY(dev_alloc(c, c->tris, (sizeof(LeafTri) + sizeof(WideNode)) * (uint64_t)n));
    WideNode* const wide_nodes = reinterpret_cast<WideNode*>(c->tris.as<char>() + sizeof(LeafTri) * (uint64_t)n);
    TRY(dev_alloc(c, c->tris_orig, sizeof(TriRec) * (uint64_t)n));
    TRY(dev_alloc(c, c->shade, sizeof(ShadeRec) * (uint64_t)n));

    const uint32_t B = 256;
    const float* pos = c->pos.as<float>();
    k_init_bounds<<<1, 32, 0, st>>>(sb.as<uint32_t>());
    k_prim_bounds<<<cdiv(n, B), B, 0, st>>>(pos, n, pmin.as<float4>(), pmax.as<float4>(), sb.as<uint32_t>());
    k_morton<<<cdiv(n, B), B, 0, st>>>(pmin.as<float4>(), pmax.as<float4>(), n, sb.as<uint32_t>(), keys.as<uint64_t>(), vals.as<uint32_t>());
    size_t tmp_bytes = 0;
    TRYCUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys.as<uint64_t>(), keys2.as<uint64_t>(), vals.as<uint32_t>(),
                                            vals2.as<uint32_t>(), (int)n, 0, 63, st));
    TRY(dev_alloc(c, tmp, tmp_bytes));
    TRYCUDA(cub::DeviceRadixSort::SortPairs(tmp.p, tmp_bytes, keys.as<uint64_t>(), keys2.as<uint64_t>(), vals.as<uint32_t>(),
                                            vals2.as<uint32_t>(), (int)n, 0, 63, st));
    k_leaves<<<cdiv(n, B), B, 0, st>>>(pos, vals2.as<uint32_t>(), n, pmin.as<float4>(), pmax.as<float4>(), c->b_min.as<float4>(),
                                       c->b_max.as<float4>(), c->tris_sorted.as<TriRec>(), count.as<uint32_t>());
    TRYCUDA(cudaMemsetAsync(flags.p, 0, 4ull * n, st));
    if (n > 1) {
        k_hierarchy<<<cdiv(n - 1, B), B, 0, st>>>(keys2.as<uint64_t>(), (int)n, c->b_child.as<int2>(), parent.as<int>(), first.as<int>());
        k_refit<<<cdiv(n, B), B, 0, st>>>((int)n, c->b_child.as<int2>(), parent.as<int>(), c->b_min.as<float4>(), c->b_max.as<float4>(),
                                          count.as<uint32_t>(), flags.as<uint32_t>(), dptab.as<DpRow>(), c_prim);
    }
    TRYCUDA(cudaGetLastError());

    // level-synchronous collapse; queue capacity = number of internal binary nodes (+1)
    TRY(dev_alloc(c, qa, 8ull * n)); TRY(dev_alloc(c, qb, 8ull * n));
    CollapseCtl h{1u, 0u, 0u, 0u};
    uint2 root = make_uint2(0u, 0u); // binary node 0 is the root (for n == 1 it is the only leaf)
    TRYCUDA(cudaMemcpyAsync(qa.p, &root, sizeof(root), cudaMemcpyHostToDevice, st));
    TRYCUDA(cudaMemcpyAsync(ctl.p, &h, sizeof(h), cudaMemcpyHostToDevice, st));
    uint32_t n_in = 1, levels = 0;
    DevBuf* qi = &qa; DevBuf* qo = &qb;
    while (n_in) {
        ++levels;
        k_collapse<<<cdiv(n_in, 128), 128, 0, st>>>(qi->as<uint2>(), n_in, qo->as<uint2>(), ctl.as<CollapseCtl>(), (int)n, c->b_child.as<int2>(),
                                                    first.as<int>(), count.as<uint32_t>(), c->b_min.as<float4>(), c->b_max.as<float4>(),
                                                    c->tris_sorted.as<TriRec>(), wide_nodes, c->tris.as<LeafTri>(), dptab.as<DpRow>(), dp, c->errflag.as<uint32_t>());
        TRYCUDA(cudaMemcpyAsync(&h, ctl.p, sizeof(h), cudaMemcpyDeviceToHost, st));
        TRYCUDA(cudaStreamSynchronize(st));
        n_in = h.q_out;
        h.q_out = 0;
        TRYCUDA(cudaMemcpyAsync(ctl.p, &h, sizeof(h), cudaMemcpyHostToDevice, st));
        std::swap(qi, qo);
    }
    if (h.n_tris != n) { cleanup(); return fail(c, VMN_ERR_STATE, "collapse placed %u of %u triangles", h.n_tris, n); }
    // the BVH8 traversal keeps at most one stack entry per level: a tree deeper than the stack is refused here rather than traversed wrongly
    // (VMN_DEBUG_NO_DEPTH_CHECK=1 lets a too-deep tree through: the test hook for the device-side overflow report)
    static const bool skip_depth_check = [] { const char* e = getenv("VMN_DEBUG_NO_DEPTH_CHECK"); return e && e[0] == '1'; }();
    if (levels > kTravStackCap && !skip_depth_check) { cleanup(); return fail(c, VMN_ERR_STATE, "BVH8 depth %u exceeds the traversal stack (%u entries)", levels, kTravStackCap); }
    c->stats.wide_depth = levels;
    c->n_wide = h.n_nodes;
    VMN_CUDA(c, cudaEventRecord(c->ev[1], st));
    VMN_CUDA(c, cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
    cleanup();
#undef TRY
#undef dev_alloc
#undef TRYCUDA

    c->scene.nodes = wide_nodes;
    c->scene.n_wide = (uint32_t)c->n_wide;
    c->trav_bytes = sizeof(LeafTri) * (uint64_t)n + sizeof(WideNode) * (uint64_t)c->n_wide;
    c->scene.tris = c->tris.as<LeafTri>();
    c->scene.tris_sorted = c->tris_sorted.as<TriRec>();
    c->scene.tris_orig = c->tris_orig.as<TriRec>();
    c->scene.shade = c->shade.as<ShadeRec>();
    c->scene.bvh2.bmin = c->b_min.as<float4>();
    c->scene.bvh2.bmax = c->b_max.as<float4>();
    c->scene.bvh2.child = c->b_child.as<int2>();
    c->scene.bvh2.n_prims = n;
    c->stats.n_nodes = nnodes;
    c->stats.n_wide_nodes = c->n_wide;
    c->stats.build_ms = ms;
    return VMN_OK;
}

vmn_status build_records(vmn_ctx* c) {
    const uint32_t n = (uint32_t)c->n_tris;
    if (n == 0) return VMN_OK;
    cudaStream_t st = c->stream;
    VMN_CUDA(c, cudaEventRecord(c->ev[0], st));
    k_orig_records<<<cdiv(n, 256), 256, 0, st>>>(c->pos.as<float>(), c->have_nrm ? c->nrm.as<float>() : nullptr, c->have_uv ? c->uv.as<float>() : nullptr, n,
                                                 c->tris_orig.as<TriRec>(), c->shade.as<ShadeRec>());
    VMN_CUDA(c, cudaGetLastError());
    VMN_CUDA(c, cudaEventRecord(c->ev[1], st));
    VMN_CUDA(c, cudaStreamSynchronize(st));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]);
    c->stats.build_ms += ms;
    return VMN_OK;
}

vmn_status build_bvh(vmn_ctx* c) {
    const vmn_status s = build_geometry(c);
    return s != VMN_OK ? s : build_records(c);
}

} // namespace vmn
