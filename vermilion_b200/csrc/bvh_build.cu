// bvh_build.cu — GPU BVH construction (replaces BVH::build, reference core/accelerators/bvh.cpp:179-279).
//
// The reference builds a binary midpoint-split tree serially on the host. This builder produces a
// DIFFERENT tree with the same closest hits (tie order excepted), entirely on the GPU:
//   1. k_prim_bounds   per-triangle box + scene box (atomic min/max on order-preserving float bits)
//   2. k_morton        63-bit Morton code of the box centre, 21 bits per axis
//   3. radix sort      (key = Morton code, value = triangle id)
//   4. k_leaves        leaf boxes + Morton-ordered 48-byte triangle records
//   5. k_hierarchy     Karras 2012 binary radix tree, one thread per internal node
//   6. k_refit         bottom-up boxes and subtree triangle counts, second-arrival atomics
//   6b. top-level SAH  (experiment, env VMN_TOP_SAH=1, off by default — see top_sah_enabled) the top of the tree (the <= 1024 internal
//                      nodes with the largest boxes) is replaced by a surface-area-heuristic tree built on the host over the subtrees
//                      hanging below it (k_area_hist, k_top_collect, SahTop, k_top_scatter)
//   7. k_collapse      level-synchronous collapse of the binary tree into 8-wide nodes
//                      (greedy largest-surface-area expansion), octant-ordered child slots,
//                      8-bit quantised child boxes -> 80-byte WideNode, leaf triangles re-laid out
//                      contiguously per node.
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

__global__ void k_refit(int n, const int2* __restrict__ child, const int* __restrict__ parent, float4* bmin, float4* bmax,
                        uint32_t* count, uint32_t* __restrict__ flags) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    int node = parent[(n - 1) + k];
    while (node >= 0) {
        if (atomicAdd(&flags[node], 1u) == 0u) return; // first arrival stops; the second sees both children
        __threadfence();
        int2 c = child[node];
        float4 a0 = __ldcg(&bmin[c.x]), a1 = __ldcg(&bmax[c.x]), b0 = __ldcg(&bmin[c.y]), b1 = __ldcg(&bmax[c.y]);
        __stcg(&bmin[node], make_float4(fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z), 0.f));
        __stcg(&bmax[node], make_float4(fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z), 0.f));
        __stcg(&count[node], __ldcg(&count[c.x]) + __ldcg(&count[c.y]));
        __threadfence();
        node = parent[node];
    }
}


// ---- top-level SAH rebuild -----------------------------------------------------------------------
// Half-octave bin of a box area relative to the root's: 0 = as large as the root, 127 = tiny. Areas shrink monotonically from parent to
// child (nested boxes), so "bin <= b" selects a connected top part T of the tree; the nodes hanging directly below T are the cut.
__device__ __forceinline__ float box_area(float4 a, float4 b) {
    float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
    return dx * dy + dy * dz + dz * dx;
}
__device__ __forceinline__ int area_bin(float a, float root_area) {
    if (!(a > 0.f) || !(root_area > 0.f)) return 127;
    if (a >= root_area) return 0;
    const float v = -2.f * log2f(a / root_area);
    return v >= 127.f ? 127 : (int)v;
}

__global__ void k_area_hist(int nint, const float4* __restrict__ bmin, const float4* __restrict__ bmax, uint32_t* __restrict__ hist) {
    __shared__ uint32_t sh[128];
    if (threadIdx.x < 128) sh[threadIdx.x] = 0;
    __syncthreads();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const float ra = box_area(bmin[0], bmax[0]);
    if (i < nint) atomicAdd(&sh[area_bin(box_area(bmin[i], bmax[i]), ra)], 1u);
    __syncthreads();
    if (threadIdx.x < 128 && sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
}

struct CutRec { uint32_t id, count; float lo[3], hi[3]; };        // a subtree below the top part
struct TopRec { uint32_t slot, left, right, count; float lo[3], hi[3]; }; // a rebuilt top node, written into the slot of an old one

__global__ void k_top_collect(int n, int bin_max, const int* __restrict__ parent, const float4* __restrict__ bmin,
                              const float4* __restrict__ bmax, const uint32_t* __restrict__ count, uint32_t* __restrict__ ctr,
                              uint32_t* __restrict__ top_ids, CutRec* __restrict__ cut, uint32_t cap) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int nint = n - 1;
    if (i >= 2 * n - 1) return;
    const float ra = box_area(bmin[0], bmax[0]);
    const float4 lo = bmin[i], hi = bmax[i];
    const bool in_top = i < nint && area_bin(box_area(lo, hi), ra) <= bin_max;
    if (in_top) {
        const uint32_t k = atomicAdd(&ctr[0], 1u);
        if (k < cap) top_ids[k] = (uint32_t)i;
        return;
    }
    const int p = parent[i];
    if (p < 0 || area_bin(box_area(bmin[p], bmax[p]), ra) > bin_max) return;
    const uint32_t k = atomicAdd(&ctr[1], 1u);
    if (k > cap) return;
    CutRec r;
    r.id = (uint32_t)i; r.count = count[i];
    r.lo[0] = lo.x; r.lo[1] = lo.y; r.lo[2] = lo.z; r.hi[0] = hi.x; r.hi[1] = hi.y; r.hi[2] = hi.z;
    cut[k] = r;
}

__global__ void k_top_scatter(uint32_t m, const TopRec* __restrict__ rec, int2* __restrict__ child, float4* __restrict__ bmin,
                              float4* __restrict__ bmax, uint32_t* __restrict__ count) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= m) return;
    const TopRec r = rec[k];
    child[r.slot] = make_int2((int)r.left, (int)r.right);
    bmin[r.slot] = make_float4(r.lo[0], r.lo[1], r.lo[2], 0.f);
    bmax[r.slot] = make_float4(r.hi[0], r.hi[1], r.hi[2], 0.f);
    count[r.slot] = r.count;
}

// ---- collapse to the compressed 8-wide BVH ----------------------------------------------------
#define VMN_LEAF_MAX 3u // triangles per leaf slot (unary 3-bit count in meta)

struct CollapseCtl { uint32_t n_nodes, n_tris, q_out, pad; };

// One thread turns one binary subtree root into one WideNode and enqueues its internal children.
__global__ void k_collapse(const uint2* __restrict__ q_in, uint32_t n_in, uint2* __restrict__ q_out, CollapseCtl* ctl, int n,
                           const int2* __restrict__ child, const int* __restrict__ first, const uint32_t* __restrict__ count,
                           const float4* __restrict__ bmin, const float4* __restrict__ bmax, const TriRec* __restrict__ tris_sorted,
                           WideNode* __restrict__ nodes, LeafTri* __restrict__ tris) {
    uint32_t w = blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n_in) return;
    const int bnode = (int)q_in[w].x;
    const uint32_t widx = q_in[w].y;
    const int nint = n - 1; // binary nodes >= nint are leaves

    int ch[8];
    int nch;
    if (bnode >= nint) { ch[0] = bnode; nch = 1; } // single-triangle scene
    else { int2 c = child[bnode]; ch[0] = c.x; ch[1] = c.y; nch = 2; }
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
        if (c < nint && count[c] > VMN_LEAF_MAX) { imask |= 1u << s; ++n_int; }
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
            q_out[qbase + irank] = make_uint2((uint32_t)c, cbase + irank);
            ++irank;
        } else {
            uint32_t cnt = count[c];
            uint32_t f = (c < nint) ? (uint32_t)first[c] : (uint32_t)(c - nint);
            meta[s] = (uint8_t)((((1u << cnt) - 1u) << 5) | toff);
            for (uint32_t j = 0; j < cnt; ++j) { const TriRec t = tris_sorted[f + j]; LeafTri& o = tris[tbase + toff + j]; o.r0 = t.r0; o.r1 = t.r1; o.r2 = t.r2; }
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
    nodes[widx] = nd;
}

// Host side of the top-level rebuild: a binary SAH tree (full sweep on all three axes, cost = area x triangle count) over the cut
// subtrees. Internal nodes take the slots of the old top nodes in preorder, so the root stays binary node 0. Below depth 20 the split is
// the count median along the widest centroid axis, which bounds the depth of the rebuilt part (the traversal stacks are finite).
namespace {
struct HBox {
    float lo[3], hi[3];
    void reset() { for (int a = 0; a < 3; ++a) { lo[a] = FLT_MAX; hi[a] = -FLT_MAX; } }
    void grow(const float* l, const float* h) { for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], l[a]); hi[a] = std::max(hi[a], h[a]); } }
    float area() const { float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2]; return dx * dy + dy * dz + dz * dx; }
};

struct SahTop {
    const std::vector<CutRec>& cut;
    const std::vector<uint32_t>& slots;
    std::vector<TopRec>& out;
    std::vector<uint32_t> idx;
    std::vector<float> right_area;
    std::vector<uint64_t> right_count;
    size_t next_slot = 0;

    SahTop(const std::vector<CutRec>& c, const std::vector<uint32_t>& s, std::vector<TopRec>& o) : cut(c), slots(s), out(o) {
        idx.resize(cut.size());
        for (size_t i = 0; i < idx.size(); ++i) idx[i] = (uint32_t)i;
        right_area.resize(cut.size() + 1); right_count.resize(cut.size() + 1);
    }
    void sort_axis(size_t lo, size_t hi, int axis) {
        std::sort(idx.begin() + lo, idx.begin() + hi, [&](uint32_t a, uint32_t b) {
            const float ca = cut[a].lo[axis] + cut[a].hi[axis], cb = cut[b].lo[axis] + cut[b].hi[axis];
            return ca < cb || (ca == cb && cut[a].id < cut[b].id);
        });
    }
    // returns the binary node id of the subtree over idx[lo, hi)
    uint32_t build(size_t lo, size_t hi, int depth, HBox* box, uint64_t* count) {
        if (hi - lo == 1) {
            const CutRec& r = cut[idx[lo]];
            box->reset(); box->grow(r.lo, r.hi); *count = r.count;
            return r.id;
        }
        const size_t k = next_slot++;
        out.emplace_back();
        int best_axis = 0; size_t best_pos = lo + (hi - lo) / 2;
        if (depth < 20) {
            float best_cost = FLT_MAX;
            for (int axis = 0; axis < 3; ++axis) {
                sort_axis(lo, hi, axis);
                HBox b; b.reset(); uint64_t c = 0;
                for (size_t i = hi; i-- > lo;) { b.grow(cut[idx[i]].lo, cut[idx[i]].hi); c += cut[idx[i]].count; right_area[i] = b.area(); right_count[i] = c; }
                b.reset(); c = 0;
                for (size_t i = lo; i + 1 < hi; ++i) {
                    b.grow(cut[idx[i]].lo, cut[idx[i]].hi); c += cut[idx[i]].count;
                    const float cost = b.area() * (float)c + right_area[i + 1] * (float)right_count[i + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_pos = i + 1; }
                }
            }
            if (best_axis != 2) sort_axis(lo, hi, best_axis);
        } else {
            HBox cb; cb.reset(); uint64_t total = 0;
            for (size_t i = lo; i < hi; ++i) {
                const CutRec& r = cut[idx[i]];
                const float c3[3] = {r.lo[0] + r.hi[0], r.lo[1] + r.hi[1], r.lo[2] + r.hi[2]};
                cb.grow(c3, c3); total += r.count;
            }
            for (int a = 1; a < 3; ++a) if (cb.hi[a] - cb.lo[a] > cb.hi[best_axis] - cb.lo[best_axis]) best_axis = a;
            sort_axis(lo, hi, best_axis);
            uint64_t c = 0; best_pos = lo + 1;
            for (size_t i = lo; i + 1 < hi; ++i) { c += cut[idx[i]].count; best_pos = i + 1; if (2 * c >= total) break; }
        }
        HBox bl, br; uint64_t cl = 0, cr = 0;
        const uint32_t left = build(lo, best_pos, depth + 1, &bl, &cl);
        const uint32_t right = build(best_pos, hi, depth + 1, &br, &cr);
        *box = bl; box->grow(br.lo, br.hi); *count = cl + cr;
        TopRec& r = out[k];
        r.slot = slots[k]; r.left = left; r.right = right;
        // a rebuilt node never becomes a leaf slot of the wide tree: its triangles are not contiguous in the sorted order
        r.count = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(*count, 4), 0xFFFFFFFFull);
        for (int a = 0; a < 3; ++a) { r.lo[a] = box->lo[a]; r.hi[a] = box->hi[a]; }
        return r.slot;
    }
};
constexpr uint32_t kTopCap = 1024; // internal nodes in the rebuilt top part
} // namespace

// Off by default (env VMN_TOP_SAH=1 turns it on, read at every commit). Measured on B200 (profiles/r01u_stats_*.jsonl): C2 spheres 4.37 ->
// 4.15 nodes per ray, k_extend -3 %; but the C4 city goes from 15.2 nodes + 5.3 triangles per ray to 22.1 + 8.9 (k_extend +49 %): the
// traversal visits children in a fixed octant order, which matches the 2x2x2 arrangement that three Morton median splits produce and
// is a poor front-to-back proxy for the slab-like partitions the SAH picks.
static bool top_sah_enabled() {
    const char* e = getenv("VMN_TOP_SAH");
    return e && e[0] == '1';
}

// replaces the top of the binary tree in place; leaves it untouched when the tree is too small or degenerate
static vmn_status rebuild_top(vmn_ctx* c, uint32_t n, DevBuf& scratch, const int* parent, uint32_t* count) {
    cudaStream_t st = c->stream;
    if (n < 64 || !top_sah_enabled()) return VMN_OK;
    const size_t off_hist = 0, off_ctr = 512, off_top = 1024, off_cut = off_top + 4 * kTopCap, off_rec = off_cut + sizeof(CutRec) * (kTopCap + 2);
    vmn_status s = dev_reserve(c, scratch, off_rec + sizeof(TopRec) * kTopCap);
    if (s != VMN_OK) return s;
    char* base = scratch.as<char>();
    VMN_CUDA(c, cudaMemsetAsync(base, 0, 1024, st));
    const float4* bmin = c->b_min.as<float4>(); const float4* bmax = c->b_max.as<float4>();
    k_area_hist<<<(n - 1 + 255) / 256, 256, 0, st>>>((int)n - 1, bmin, bmax, reinterpret_cast<uint32_t*>(base + off_hist));
    uint32_t hist[128];
    VMN_CUDA(c, cudaMemcpyAsync(hist, base + off_hist, sizeof(hist), cudaMemcpyDeviceToHost, st));
    VMN_CUDA(c, cudaStreamSynchronize(st));
    int bin_max = -1; uint64_t cum = 0;
    for (int b = 0; b < 128; ++b) { cum += hist[b]; if (cum > kTopCap) break; bin_max = b; }
    if (bin_max < 0) return VMN_OK; // more than kTopCap boxes as large as the root's: nothing sensible to rebuild
    k_top_collect<<<(2 * n - 1 + 255) / 256, 256, 0, st>>>((int)n, bin_max, parent, bmin, bmax, count, reinterpret_cast<uint32_t*>(base + off_ctr),
                                                           reinterpret_cast<uint32_t*>(base + off_top), reinterpret_cast<CutRec*>(base + off_cut), kTopCap);
    uint32_t ctr[2] = {0, 0};
    VMN_CUDA(c, cudaMemcpyAsync(ctr, base + off_ctr, sizeof(ctr), cudaMemcpyDeviceToHost, st));
    VMN_CUDA(c, cudaStreamSynchronize(st));
    if (ctr[0] < 2 || ctr[0] > kTopCap || ctr[1] != ctr[0] + 1) return VMN_OK;
    std::vector<uint32_t> slots(ctr[0]);
    std::vector<CutRec> cut(ctr[1]);
    VMN_CUDA(c, cudaMemcpyAsync(slots.data(), base + off_top, 4 * slots.size(), cudaMemcpyDeviceToHost, st));
    VMN_CUDA(c, cudaMemcpyAsync(cut.data(), base + off_cut, sizeof(CutRec) * cut.size(), cudaMemcpyDeviceToHost, st));
    VMN_CUDA(c, cudaStreamSynchronize(st));
    std::sort(slots.begin(), slots.end()); // the atomics hand out positions in any order: sort for a reproducible tree (slot 0 = root first)
    std::sort(cut.begin(), cut.end(), [](const CutRec& a, const CutRec& b) { return a.id < b.id; });
    if (slots[0] != 0u) return VMN_OK;
    std::vector<TopRec> recs;
    recs.reserve(slots.size());
    SahTop sah(cut, slots, recs);
    HBox box; uint64_t total = 0;
    sah.build(0, cut.size(), 0, &box, &total);
    if (recs.size() != slots.size()) return fail(c, VMN_ERR_STATE, "top-level rebuild produced %zu of %zu nodes", recs.size(), slots.size());
    VMN_CUDA(c, cudaMemcpyAsync(base + off_rec, recs.data(), sizeof(TopRec) * recs.size(), cudaMemcpyHostToDevice, st));
    k_top_scatter<<<((uint32_t)recs.size() + 127) / 128, 128, 0, st>>>((uint32_t)recs.size(), reinterpret_cast<const TopRec*>(base + off_rec),
                                                                     c->b_child.as<int2>(), c->b_min.as<float4>(), c->b_max.as<float4>(), count);
    VMN_CUDA(c, cudaStreamSynchronize(st)); // recs is a host temporary
    return VMN_OK;
}

static inline uint32_t cdiv(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

vmn_status build_bvh(vmn_ctx* c) {
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
           &qa = c->build_scratch[12], &qb = c->build_scratch[13], &ctl = c->build_scratch[14], &top = c->build_scratch[15];
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
    k_orig_records<<<cdiv(n, B), B, 0, st>>>(pos, c->have_nrm ? c->nrm.as<float>() : nullptr, c->have_uv ? c->uv.as<float>() : nullptr, n, c->tris_orig.as<TriRec>(), c->shade.as<ShadeRec>());
    TRYCUDA(cudaMemsetAsync(flags.p, 0, 4ull * n, st));
    if (n > 1) {
        k_hierarchy<<<cdiv(n - 1, B), B, 0, st>>>(keys2.as<uint64_t>(), (int)n, c->b_child.as<int2>(), parent.as<int>(), first.as<int>());
        k_refit<<<cdiv(n, B), B, 0, st>>>((int)n, c->b_child.as<int2>(), parent.as<int>(), c->b_min.as<float4>(), c->b_max.as<float4>(),
                                          count.as<uint32_t>(), flags.as<uint32_t>());
    }
    TRYCUDA(cudaGetLastError());
    if (n > 1) TRY(rebuild_top(c, n, top, parent.as<int>(), count.as<uint32_t>()));

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
                                                    c->tris_sorted.as<TriRec>(), wide_nodes, c->tris.as<LeafTri>());
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

} // namespace vmn
