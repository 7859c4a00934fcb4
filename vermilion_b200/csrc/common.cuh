// common.cuh — shared device/host definitions of the B200 engine.
//
// Arithmetic contract: this library is compiled with -fmad=false, so every `a*b+c` written
// with * and + is two IEEE roundings, exactly as the reference's x86-64 build (no FMA) does.
// Where a fused multiply-add is wanted (box slabs, where only conservativeness matters, not the
// bit pattern) it is written explicitly as __fmaf_rn. Division and sqrt are IEEE (-prec-div,
// -prec-sqrt defaults), denormals are kept (-ftz=false default).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/vermilion_b200.h"

#define VMN_SM_COUNT_DEFAULT 148

namespace vmn {

// ---- tiny vec3 with glm 0.9.9.4 evaluation order (func_geometric.inl) --------------------------
struct V3 {
    float x, y, z;
    __host__ __device__ V3() {}
    __host__ __device__ V3(float a, float b, float c) : x(a), y(b), z(c) {}
};
__host__ __device__ inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
__host__ __device__ inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
__host__ __device__ inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
__host__ __device__ inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
__host__ __device__ inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
// dot = (x*x' + y*y') + z*z'   (func_geometric.inl:48-55)
__host__ __device__ inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// cross (func_geometric.inl:68-79)
__host__ __device__ inline V3 cross(V3 x, V3 y) {
    return V3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y);
}
// normalize = v * inversesqrt(dot(v,v)), inversesqrt = 1/sqrt (func_geometric.inl:82-90, func_exponential.inl:134-139)
__device__ inline V3 normalize(V3 v) { return v * (1.0f / sqrtf(dot(v, v))); }
__device__ inline float length(V3 v) { return sqrtf(dot(v, v)); }

// ---- device-side scene layout ---------------------------------------------------------------
// Triangle record for intersection, 48 B = 3 x float4 (three LDG.128):
//   r0 = (v0.x, v0.y, v0.z, as_float(original id))
//   r1 = (e1.x, e1.y, e1.z, 0)   e1 = v1 - v0   (the same fp32 subtraction triangle.cpp:16 performs per ray)
//   r2 = (e2.x, e2.y, e2.z, 0)   e2 = v2 - v0   (triangle.cpp:17)
struct TriRec { float4 r0, r1, r2; };
// Shading record, 64 B, indexed by ORIGINAL triangle id: vertex normals and uvs (triangle.cpp:56-87)
//   s0 = (n0.xyz, uv0.x) s1 = (n1.xyz, uv0.y) s2 = (n2.xyz, uv1.x) s3 = (uv1.y, uv2.x, uv2.y, 0)
struct ShadeRec { float4 s0, s1, s2, s3; };

// Compressed 8-wide BVH node, 80 B = 5 x 16 B (five LDG.128), 16-byte aligned.
//   q0 = (p.x, p.y, p.z, [ex | ey<<8 | ez<<16 | imask<<24])     quantisation frame + internal-child mask
//   q1 = (child_base, tri_base, meta[0..3], meta[4..7])
//   q2 = (qlo.x[0..3], qlo.x[4..7], qlo.y[0..3], qlo.y[4..7])
//   q3 = (qlo.z[0..3], qlo.z[4..7], qhi.x[0..3], qhi.x[4..7])
//   q4 = (qhi.y[0..3], qhi.y[4..7], qhi.z[0..3], qhi.z[4..7])
// child box = p + q * 2^(e-127) per axis. meta[i]: 0 = empty; internal = (1<<5) | (24+i);
// leaf = (unary triangle count 1|3|7) << 5 | offset of its first triangle from tri_base (0..23).
// VMN_NODE_PAD / VMN_TRI_PAD (in 16-byte units) pad the records of the two traversal arrays to a sector-friendly stride
// (experiment knobs; the algorithmic byte counts of the roofline always use the 80 / 48 payload bytes).
#ifndef VMN_NODE_PAD
#define VMN_NODE_PAD 0
#endif
#ifndef VMN_TRI_PAD
#define VMN_TRI_PAD 1 // 64-byte leaf triangles = exactly two 32-byte sectors: +3.7 % on C3, +1 % on C2
#endif
struct WideNode {
    uint4 q0, q1, q2, q3, q4;
#if VMN_NODE_PAD
    uint4 pad[VMN_NODE_PAD];
#endif
};
// triangle record of the BVH8 leaf-order array (same payload as TriRec)
struct LeafTri {
    float4 r0, r1, r2;
#if VMN_TRI_PAD
    float4 pad[VMN_TRI_PAD];
#endif
};
constexpr uint32_t kNodeStride16 = sizeof(WideNode) / 16, kLeafTriStride16 = sizeof(LeafTri) / 16;
static_assert(sizeof(WideNode) == 80 + 16 * VMN_NODE_PAD, "WideNode payload must be 80 bytes");
static_assert(sizeof(TriRec) == 48, "TriRec must be 48 bytes");

// Binary LBVH (build intermediate and the BVH2 baseline traversal). Node i < n-1 internal,
// leaf of sorted primitive k is node (n-1)+k. Boxes as two float4 (min.xyz,_) (max.xyz,_).
struct Bvh2 {
    const float4* bmin;
    const float4* bmax;
    const int2* child; // internal nodes only
    uint32_t n_prims;
};

// Traversal stacks (BVH8: one entry per wide level with pending siblings; BVH2: one per binary level) hold this many entries;
// build_bvh() measures the depth of both trees and refuses to commit one that could overflow them.
#ifndef VMN_STACK_CAP
#define VMN_STACK_CAP 128 // (the `tinystack` test variant of csrc/Makefile builds with 3 to reach the overflow paths)
#endif
constexpr uint32_t kTravStackCap = VMN_STACK_CAP;
#define VMN_DEVERR_STACK 1u  // device error word: a traversal push was dropped (result of that ray is unreliable)
#define VMN_DEVERR_BOUNDS 2u // an index left its array (only raised by the VMN_CHECKED build, see below)

// compute-sanitizer is closed on this GPU pool, so the library carries its own memcheck: the `checked` variant of csrc/Makefile
// (-DVMN_CHECKED=1) tests every data-dependent index (node, triangle, ray, queue slot, stack entry) against the size of its array before
// the access and raises VMN_DEVERR_BOUNDS in the device error word, which the next synchronising entry point turns into VMN_ERR_STATE.
// tools/sanitize_small.py runs every kernel of the library under that build (tests/test_gpu_robustness.py).
#ifndef VMN_CHECKED
#define VMN_CHECKED 0
#endif
#if VMN_CHECKED
#define VMN_CHECK(errp, cond) do { if (!(cond)) atomicOr((errp), VMN_DEVERR_BOUNDS); } while (0)
#else
#define VMN_CHECK(errp, cond) ((void)0)
#endif

constexpr uint32_t kMaxSpheres = 16; // vmn_scene_set_spheres refuses more

struct SceneDev {
    uint32_t* err;           // device error word owned by the context (VMN_DEVERR_*), checked at every synchronising entry point
    const WideNode* nodes;   // BVH8
    const LeafTri* tris;     // in BVH8 leaf order
    const TriRec* tris_sorted; // in Morton order (BVH2 leaves)
    const ShadeRec* shade;   // by original id
    const TriRec* tris_orig; // by original id (for barycentrics / getNormal)
    Bvh2 bvh2;
    uint32_t n_tris;
    uint32_t n_wide;         // wide nodes in use
    uint32_t n_spheres;
    const float* tex;        // texture slot 0 or nullptr
    uint32_t tex_w, tex_h, tex_c;
    const float* tex1;       // texture slot 1 (BruteForceTracer's albedo) or nullptr
    uint32_t tex1_w, tex1_h, tex1_c;
};

// Device record of one analytic sphere, 64 B, laid out for the loop of raycast_finish: the first two 16-byte words are all a solve needs
// (centre, flags with VMN_SPHD_EMITTER, and (double)(float)(r * r) — the reference's `(double)(rad * rad)` computed once on the host with
// the same float multiply); emission and the point normals are measured from are only read for an accepted sphere.
#define VMN_SPHD_EMITTER 0x100u // emission != (0,0,0): what meshEngine.cpp's accepted emitter blocks test implicitly by assigning Le
struct __align__(16) SphereDev { float cx, cy, cz; uint32_t flags; double r2; float r, pad0; float ex, ey, ez, pad1; float nx, ny, nz, pad2; };

// which pixels this context renders: a contiguous range or interleaved square tiles
struct PixelMap {
    uint32_t width, height;
    uint64_t begin;       // contiguous: first pixel
    uint32_t tile, rank, world, tiles_x;
    uint32_t tile_log2p1; // log2(tile) + 1 when the tile size is a power of two (the usual 64: shifts instead of three divides), else 0
};
inline uint32_t tile_log2p1(uint32_t tile) { // host side, when a PixelMap is made
    if (tile == 0 || (tile & (tile - 1))) return 0;
    uint32_t l = 0;
    while ((1u << l) < tile) ++l;
    return l + 1;
}

__device__ __forceinline__ bool map_pixel(const PixelMap& m, uint64_t lp, uint64_t* pixel) {
    if (m.tile == 0) { *pixel = m.begin + lp; return *pixel < (uint64_t)m.width * m.height; }
    const uint32_t tt = m.tile * m.tile;
    uint64_t j; uint32_t w, wx, wy;
    if (m.tile_log2p1) { // power-of-two tile: shifts and masks
        const uint32_t sh = m.tile_log2p1 - 1u;
        j = lp >> (2u * sh); w = (uint32_t)lp & (tt - 1u); wx = w & (m.tile - 1u); wy = w >> sh;
    } else {
        if ((lp >> 32) == 0) { const uint32_t j32 = (uint32_t)lp / tt; j = j32; w = (uint32_t)lp - j32 * tt; } // the usual case: 32-bit divides
        else { j = lp / tt; w = (uint32_t)(lp % tt); }
        wy = w / m.tile; wx = w - wy * m.tile;
    }
    const uint64_t t = j * m.world + m.rank;
    uint32_t ty, tx;
    if ((t >> 32) == 0) { ty = (uint32_t)t / m.tiles_x; tx = (uint32_t)t - ty * m.tiles_x; }
    else { ty = (uint32_t)(t / m.tiles_x); tx = (uint32_t)(t % m.tiles_x); }
    const uint32_t x = tx * m.tile + wx, y = ty * m.tile + wy;
    if (x >= m.width || y >= m.height) return false;
    *pixel = (uint64_t)y * m.width + x;
    return true;
}

// Tile mode only: map_pixel plus the local tile index and the pixel's rank among the LIVE pixels of its tile (row-major over the part of
// the tile inside the image) — what k_raygen needs to place a path in the queue without an atomic (live pixels before it = prefix[tile] + rank).
__device__ __forceinline__ bool map_pixel_tile(const PixelMap& m, uint32_t lp, uint64_t* pixel, uint32_t* tile_j, uint32_t* rank_in_tile) {
    const uint32_t tt = m.tile * m.tile;
    uint32_t j, w, wx, wy;
    if (m.tile_log2p1) { const uint32_t sh = m.tile_log2p1 - 1u; j = lp >> (2u * sh); w = lp & (tt - 1u); wx = w & (m.tile - 1u); wy = w >> sh; }
    else { j = lp / tt; w = lp - j * tt; wy = w / m.tile; wx = w - wy * m.tile; }
    const uint64_t t = (uint64_t)j * m.world + m.rank;
    uint32_t ty, tx;
    if ((t >> 32) == 0) { ty = (uint32_t)t / m.tiles_x; tx = (uint32_t)t - ty * m.tiles_x; }
    else { ty = (uint32_t)(t / m.tiles_x); tx = (uint32_t)(t % m.tiles_x); }
    const uint32_t x = tx * m.tile + wx, y = ty * m.tile + wy;
    const uint32_t left = m.width - tx * m.tile, lw = left < m.tile ? left : m.tile; // live width of this tile
    *tile_j = j; *rank_in_tile = wy * lw + wx;
    if (x >= m.width || y >= m.height) return false;
    *pixel = (uint64_t)y * m.width + x;
    return true;
}

// ---- Philox4x32-10, key (seed_lo, seed_hi), counter (pixel_lo, pixel_hi, sample, block) ------------
__host__ __device__ inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
#ifdef __CUDA_ARCH__
        uint32_t h0 = __umulhi(M0, c[0]), l0 = M0 * c[0];
        uint32_t h1 = __umulhi(M1, c[2]), l1 = M1 * c[2];
#else
        uint64_t p0 = (uint64_t)M0 * c[0], p1 = (uint64_t)M1 * c[2];
        uint32_t h0 = (uint32_t)(p0 >> 32), l0 = (uint32_t)p0, h1 = (uint32_t)(p1 >> 32), l1 = (uint32_t)p1;
#endif
        uint32_t n0 = h1 ^ c[1] ^ k0, n1 = l1, n2 = h0 ^ c[3] ^ k1, n3 = l0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += W0; k1 += W1;
    }
}

// Sequential view of the stream: draw index -> (block = draw>>2, lane = draw&3), one block cached.
struct Rng {
    uint64_t seed, pixel;
    uint32_t sample, draw, cached_block;
    uint32_t cache[4];
    __device__ Rng(uint64_t s, uint64_t p, uint32_t smp, uint32_t first_draw)
        : seed(s), pixel(p), sample(smp), draw(first_draw), cached_block(0xFFFFFFFFu) {}
    __device__ uint32_t next_u32() {
        uint32_t block = draw >> 2, lane = draw & 3u;
        if (block != cached_block) {
            cache[0] = (uint32_t)pixel; cache[1] = (uint32_t)(pixel >> 32); cache[2] = sample; cache[3] = block;
            philox4x32_10(cache, (uint32_t)seed, (uint32_t)(seed >> 32));
            cached_block = block;
        }
        ++draw;
        return lane == 0 ? cache[0] : lane == 1 ? cache[1] : lane == 2 ? cache[2] : cache[3];
    }
    __device__ void skip(uint32_t n) { draw += n; } // draws the reference makes and never uses: advance the counter, generate nothing
    __device__ double next01() { return (double)next_u32() * (1.0 / 4294967296.0); }
    __device__ float jitter() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f) * 0.5f; }
};

// The four numbers of one bounce-loop iteration: Philox block `iteration` of stream (seed, pixel, sample), consumed in order.
struct BounceRng {
    uint32_t r[4];
    uint32_t k;
    __device__ void init(uint64_t seed, uint64_t pixel, uint32_t sample, uint32_t iteration) {
        r[0] = (uint32_t)pixel; r[1] = (uint32_t)(pixel >> 32); r[2] = sample; r[3] = iteration;
        philox4x32_10(r, (uint32_t)seed, (uint32_t)(seed >> 32));
        k = 0;
    }
    __device__ double next01() {
        const uint32_t v = k == 0 ? r[0] : k == 1 ? r[1] : k == 2 ? r[2] : r[3];
        ++k;
        return (double)v * (1.0 / 4294967296.0);
    }
};

} // namespace vmn
