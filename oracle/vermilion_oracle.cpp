/*
 * vermilion_oracle.cpp — CPU restatement of Vermilion's ray-tracing hot path.
 *
 * TEST INFRASTRUCTURE ONLY (see vermilion_oracle.h). Parity status: PINNED
 * against oracle/_ref/libvermref.so (the reference's own sources compiled
 * unmodified) by tests/test_oracle_vs_ref.py, and against tests/golden/.
 *
 * Build: g++ -std=c++17 -O2 -ffp-contract=off -fopenmp (IEEE semantics, no FMA
 * contraction, no reassociation) — the "ieee" variant SURVEY.md §8c names as
 * THE oracle. All file:line citations are into /root/reference/.
 *
 * No glm here: the few glm operations on the path are restated with the exact
 * scalar evaluation order of glm 0.9.9.4 (extern/glm/glm/detail/func_geometric.inl:
 * dot = (x*x' + y*y') + z*z' :48-55, cross :68-79, normalize = v * (1/sqrt(dot)) :82-90,
 * func_exponential.inl:134-139; min/max func_common.inl:16-29).
 */
#include "vermilion_oracle.h"

#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <random>
#include <vector>
#include <algorithm>
#include <omp.h>

namespace {

// ----------------------------------------------------------------------------
// glm restatement
// ----------------------------------------------------------------------------
struct V3 {
    float x, y, z;
    V3() : x(0), y(0), z(0) {}
    V3(float a, float b, float c) : x(a), y(b), z(c) {}
    float operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
};
struct V4 { float x, y, z, w; };

inline V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, V3 b) { return V3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
inline V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
// func_geometric.inl:48-55
inline float dot(V3 a, V3 b) { V3 t = a * b; return t.x + t.y + t.z; }
// func_geometric.inl:68-79
inline V3 cross(V3 x, V3 y) { return V3(x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y); }
// func_geometric.inl:8-14
inline float length(V3 v) { return std::sqrt(dot(v, v)); }
// func_geometric.inl:82-90 + func_exponential.inl:134-139
inline V3 normalize(V3 v) { return v * (1.0f / std::sqrt(dot(v, v))); }
// func_common.inl:16-29 — glm::min(x,y) = (y<x)?y:x ; glm::max(x,y) = (x<y)?y:x
inline float gmin(float x, float y) { return (y < x) ? y : x; }
inline float gmax(float x, float y) { return (x < y) ? y : x; }

// ----------------------------------------------------------------------------
// Philox4x32-10 (Salmon et al. 2011) — the counter-based RNG the GPU path uses.
// ----------------------------------------------------------------------------
inline void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)M0 * c[0];
        uint64_t p1 = (uint64_t)M1 * c[2];
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c[0] = n0; c[1] = n1; c[2] = n2; c[3] = n3;
        k0 += W0; k1 += W1;
    }
}

struct RngPhilox {
    uint64_t seed, pixel;
    uint32_t sample, draw;
    uint32_t cache[4];
    uint32_t cached_block;
    RngPhilox(uint64_t s, uint64_t p) : seed(s), pixel(p), sample(0), draw(0), cached_block(0xFFFFFFFFu) {}
    void begin_sample(uint32_t s) { sample = s; draw = 0; cached_block = 0xFFFFFFFFu; }
    // iteration i (1-based) of the bounce loop draws from Philox block i: at most four numbers are used per iteration (roulette, mirror
    // test, two angles; the reference's three unused noise draws only advance the counter), so the GPU generates one block per bounce
    void begin_bounce(uint32_t i) { draw = 4u * i; }
    uint32_t next_u32() {
        uint32_t block = draw >> 2, lane = draw & 3u;
        if (block != cached_block) {
            cache[0] = (uint32_t)pixel; cache[1] = (uint32_t)(pixel >> 32); cache[2] = sample; cache[3] = block;
            philox4x32_10(cache, (uint32_t)seed, (uint32_t)(seed >> 32));
            cached_block = block;
        }
        ++draw;
        return cache[lane];
    }
    double next01() { return (double)next_u32() * (1.0 / 4294967296.0); }
    float jitter() { return (float)(next_u32() >> 8) * (1.0f / 16777216.0f) * 0.5f; }
};

// std::mt19937_64 exactly as the reference uses it: Radiance draws through
// uniform_real_distribution<double>(0,1) (pathtracer.cpp:23), the pixel loop
// through uniform_real_distribution<float>(0,0.5) (pathtracer.cpp:230).
struct RngMt64 {
    std::mt19937_64 eng;
    explicit RngMt64(uint64_t s) : eng(s) {}
    void begin_sample(uint32_t) {}
    void begin_bounce(uint32_t) {}
    double next01() { std::uniform_real_distribution<double> d(0.0, 1.0); return d(eng); }
    float jitter() { std::uniform_real_distribution<float> d(0, 0.5); return d(eng); }
};

// ----------------------------------------------------------------------------
// Accelerators
// ----------------------------------------------------------------------------
struct Ray { // Ray.h:4-11
    V3 o, d, inv_d;
    Ray(V3 o_, V3 d_) : o(o_), d(d_), inv_d(V3(1, 1, 1) / d_) {}
};

struct BBox { // bbox.h:6-17
    V3 min, max, extent;
    BBox() {}
    BBox(V3 mn, V3 mx) : min(mn), max(mx) { extent = max - min; }
    explicit BBox(V3 p) : min(p), max(p) { extent = max - min; }
    // bbox.cpp:11-27: std::min(a,b) = (b<a)?b:a ; std::max(a,b) = (a<b)?b:a
    static V3 min3(V3 a, V3 b) { return V3(std::min(a.x, b.x), std::min(a.y, b.y), std::min(a.z, b.z)); }
    static V3 max3(V3 a, V3 b) { return V3(std::max(a.x, b.x), std::max(a.y, b.y), std::max(a.z, b.z)); }
    void expand(V3 p) { min = min3(min, p); max = max3(max, p); extent = max - min; }          // bbox.cpp:29-33
    void expand(const BBox& b) { min = min3(min, b.min); max = max3(max, b.max); extent = max - min; } // bbox.cpp:35-39
    uint32_t maxDimension() const { // bbox.cpp:41-46 (quirk: z is compared with y only)
        uint32_t r = 0;
        if (extent.y > extent.x) r = 1;
        if (extent.z > extent.y) r = 2;
        return r;
    }
    bool intersect(const Ray& ray, float* tnear, float* tfar) const { // bbox.cpp:70-83
        V3 t0s = (min - ray.o) * ray.inv_d;
        V3 t1s = (max - ray.o) * ray.inv_d;
        V3 ts(gmin(t0s.x, t1s.x), gmin(t0s.y, t1s.y), gmin(t0s.z, t1s.z));
        V3 tl(gmax(t0s.x, t1s.x), gmax(t0s.y, t1s.y), gmax(t0s.z, t1s.z));
        *tnear = std::max(std::max(ts.x, ts.y), ts.z);
        *tfar = std::min(std::min(tl.x, tl.y), tl.z);
        return *tnear <= *tfar;
    }
};

struct Tri { // triangle.h:5-9
    V3 v0, v1, v2, n0, n1, n2;
    float uv0[2], uv1[2], uv2[2];

    bool intersect(const Ray& ray, float* t) const { // triangle.cpp:4-54
        V3 rot = ray.d, pos = ray.o;
        V3 e1 = v1 - v0;
        V3 e2 = v2 - v0;
        V3 pvec = cross(rot, e2);
        float det = dot(e1, pvec);
        if (det < 1e-8 && det > -1e-8) return false; // compared as double, :25
        float inv_det = 1 / det;
        V3 tvec = pos - v0;
        float u = dot(tvec, pvec) * inv_det;
        if (u < 0 || u > 1) return false;
        V3 qvec = cross(tvec, e1);
        float v = dot(rot, qvec) * inv_det;
        if (v < 0 || u + v > 1) return false;
        float dist = dot(e2, qvec) * inv_det;
        if (dist > 0.0f) { *t = dist; return true; }
        return false;
    }
    // triangle.cpp:56-87 (the cross-product weights at :62-65 are dead stores)
    V3 getNormal(V3 hit, float* uv, float* bary) const {
        V3 f0 = v1 - v0;
        V3 f1 = v2 - v0;
        V3 f2 = hit - v0;
        float d00 = dot(f0, f0);
        float d01 = dot(f0, f1);
        float d11 = dot(f1, f1);
        float d20 = dot(f2, f0);
        float d21 = dot(f2, f1);
        float denom = d00 * d11 - d01 * d01;
        float w1 = (d11 * d20 - d01 * d21) / denom;
        float w2 = (d00 * d21 - d01 * d20) / denom;
        float w0 = 1 - w1 - w2;
        V3 n = n0 * w0 + n1 * w1 + n2 * w2;
        if (uv) {
            uv[0] = uv0[0] * w0 + uv1[0] * w1 + uv2[0] * w2;
            uv[1] = uv0[1] * w0 + uv1[1] * w1 + uv2[1] * w2;
        }
        if (bary) { bary[0] = w1; bary[1] = w2; }
        return -n;
    }
    BBox getBBox() const { // triangle.cpp:107-114
        return BBox(BBox::min3(BBox::min3(v0, v1), v2), BBox::max3(BBox::max3(v0, v1), v2));
    }
    V3 getCentroid() const { return (v0 + v1 + v2) * 0.333f; } // triangle.cpp:116-119
};

struct FlatNode { BBox bbox; uint32_t start, nPrims, rightOffset; }; // bvh.h:11-14

struct Bvh {
    std::vector<uint32_t> prims; // permuted triangle ids (the reference permutes Object* : bvh.cpp:249-255)
    std::vector<FlatNode> nodes;
    uint32_t nNodes = 0, nLeafs = 0, leafSize = 4;

    void build(const std::vector<Tri>& tris) { // bvh.cpp:179-279
        struct Entry { uint32_t parent, start, end; };
        std::vector<Entry> todo(128);
        uint32_t stackptr = 0;
        const uint32_t Untouched = 0xffffffff, TouchedTwice = 0xfffffffd;
        prims.resize(tris.size());
        for (size_t i = 0; i < tris.size(); ++i) prims[i] = (uint32_t)i;
        if (tris.empty()) return;
        todo[stackptr].start = 0; todo[stackptr].end = (uint32_t)tris.size(); todo[stackptr].parent = 0xfffffffc;
        stackptr++;
        FlatNode node;
        nodes.reserve(tris.size() * 2);
        while (stackptr > 0) {
            Entry bnode = todo[--stackptr];
            uint32_t start = bnode.start, end = bnode.end, nPrims = end - start;
            nNodes++;
            node.start = start; node.nPrims = nPrims; node.rightOffset = Untouched;
            BBox bb(tris[prims[start]].getBBox());
            BBox bc(tris[prims[start]].getCentroid());
            for (uint32_t p = start + 1; p < end; ++p) {
                bb.expand(tris[prims[p]].getBBox());
                bc.expand(tris[prims[p]].getCentroid());
            }
            node.bbox = bb;
            if (nPrims <= leafSize) { node.rightOffset = 0; nLeafs++; }
            nodes.push_back(node);
            if (bnode.parent != 0xfffffffc) {
                nodes[bnode.parent].rightOffset--;
                if (nodes[bnode.parent].rightOffset == TouchedTwice)
                    nodes[bnode.parent].rightOffset = nNodes - 1 - bnode.parent;
            }
            if (node.rightOffset == 0) continue;
            uint32_t split_dim = bc.maxDimension();
            float split_coord = .5f * (bc.min[split_dim] + bc.max[split_dim]);
            uint32_t mid = start;
            for (uint32_t i = start; i < end; ++i) {
                if (tris[prims[i]].getCentroid()[split_dim] < split_coord) { std::swap(prims[i], prims[mid]); ++mid; }
            }
            if (mid == start || mid == end) mid = start + (end - start) / 2;
            if (stackptr + 2 > todo.size()) todo.resize(todo.size() * 2); // the reference overflows here (UB); we grow
            todo[stackptr].start = mid; todo[stackptr].end = end; todo[stackptr].parent = nNodes - 1; stackptr++;
            todo[stackptr].start = start; todo[stackptr].end = mid; todo[stackptr].parent = nNodes - 1; stackptr++;
        }
    }

    // bvh.cpp:47-145. Returns triangle id or 0xFFFFFFFF; *t_out as the reference leaves it.
    template <bool COUNT>
    uint32_t closest(const std::vector<Tri>& tris, const Ray& ray, float* t_out, uint64_t* cnodes, uint64_t* ctris) const {
        float best_t = 999999999.f;
        uint32_t best = 0xFFFFFFFFu;
        if (nodes.empty()) { *t_out = best_t; return best; }
        float bbhits[4] = {};
        int32_t closer, other;
        struct Trav { uint32_t i; float mint; };
        Trav todo[64];
        int32_t stackptr = 0;
        todo[stackptr].i = 0; todo[stackptr].mint = -9999999.f;
        while (stackptr >= 0) {
            int ni = (int)todo[stackptr].i;
            float near = todo[stackptr].mint;
            stackptr--;
            const FlatNode& node = nodes[ni];
            if (near > best_t) continue;
            if (COUNT) ++*cnodes;
            if (node.rightOffset == 0) {
                for (uint32_t o = 0; o < node.nPrims; ++o) {
                    float ct;
                    uint32_t id = prims[node.start + o];
                    if (COUNT) ++*ctris;
                    if (tris[id].intersect(ray, &ct)) {
                        if (ct < best_t) { best_t = ct; best = id; }
                    }
                }
            } else {
                bool hitc0 = nodes[ni + 1].bbox.intersect(ray, bbhits, bbhits + 1);
                bool hitc1 = nodes[ni + node.rightOffset].bbox.intersect(ray, bbhits + 2, bbhits + 3);
                if (hitc0 && hitc1) {
                    closer = ni + 1; other = ni + (int32_t)node.rightOffset;
                    if (bbhits[2] < bbhits[0]) {
                        std::swap(bbhits[0], bbhits[2]); std::swap(bbhits[1], bbhits[3]); std::swap(closer, other);
                    }
                    if (stackptr + 2 >= 64) { fprintf(stderr, "oracle: traversal stack overflow (reference UB)\n"); abort(); }
                    ++stackptr; todo[stackptr].i = (uint32_t)other; todo[stackptr].mint = bbhits[2];
                    ++stackptr; todo[stackptr].i = (uint32_t)closer; todo[stackptr].mint = bbhits[0];
                } else if (hitc0) {
                    ++stackptr; todo[stackptr].i = (uint32_t)(ni + 1); todo[stackptr].mint = bbhits[0];
                } else if (hitc1) {
                    ++stackptr; todo[stackptr].i = (uint32_t)(ni + (int32_t)node.rightOffset); todo[stackptr].mint = bbhits[2];
                }
            }
        }
        *t_out = best_t;
        return best;
    }
};

struct Texture { std::vector<float> data; uint32_t w = 0, h = 0, c = 0; };

} // namespace

struct orc_scene {
    std::vector<Tri> tris;
    Bvh bvh;
    std::vector<Texture> textures;
    bool spheres = true;
};

namespace {

// meshEngine.cpp:21-46
inline void texture_sample(const Texture& tx, const float uv[2], float out[4]) {
    float sx = uv[0] - std::floor(uv[0]);
    float sy = uv[1] - std::floor(uv[1]);
    uint32_t mx = (uint32_t)std::round(sx * (tx.w - 1));
    uint32_t my = (uint32_t)std::round(sy * (tx.h - 1));
    const float* p = &tx.data[(size_t)(my * tx.w + mx) * tx.c];
    out[0] = out[1] = out[2] = out[3] = 0.f;
    switch (tx.c) {
    case 1: out[0] = out[1] = out[2] = out[3] = p[0]; break;
    case 2: out[0] = p[0]; out[1] = p[1]; break;
    case 3: out[0] = p[0]; out[1] = p[1]; out[2] = p[2]; break;
    case 4: out[0] = p[0]; out[1] = p[1]; out[2] = p[2]; out[3] = p[3]; break;
    default: break;
    }
}

// meshEngine.cpp:182-194
inline float sphereIntersect(V3 pos, V3 rot, V3 p, const float rad) {
    V3 op = p - pos;
    double t;
    double eps = 1e-4;
    double b = dot(op, rot);
    double det = b * b - dot(op, op) + rad * rad;
    if (det < 0) return 0;
    else det = std::sqrt(det);
    return (t = b - det) > eps ? t : ((t = b + det) > eps ? t : 0);
}

struct RayCastOut {
    bool hit = false, material = false;
    V3 P, N, Le;
    float dist = 0.f;
    float uv[2] = {0.f, 0.f};
    uint32_t tri = 0xFFFFFFFFu;
};

// meshEngine.cpp:239-509
inline RayCastOut ray_cast(const orc_scene& sc, V3 rayStart, V3 rayDirection) {
    RayCastOut r;
    float nearestHit = INFINITY, testHit = 0.f;
    int32_t hitMeshIndex = -1;
    Ray ray(rayStart, rayDirection);
    float bt;
    uint32_t id = sc.bvh.closest<false>(sc.tris, ray, &bt, nullptr, nullptr); // :361
    if (id != 0xFFFFFFFFu) {
        nearestHit = bt;
        V3 hp = ray.o + ray.d * bt; // bvh.cpp:140
        r.N = normalize(sc.tris[id].getNormal(hp, r.uv, nullptr)); // :369
        hitMeshIndex = 0;
        r.tri = id;
    }
    if (sc.spheres) {
        // emitter A :377-387
        testHit = sphereIntersect(rayStart, rayDirection, V3(15, 140, 25), 3.5f);
        if (testHit > 0.f && testHit < nearestHit) {
            nearestHit = testHit;
            r.Le = V3(0, .5, 1.0) * 15.f;
            r.N = -normalize(rayStart + (rayDirection * nearestHit) - V3(-55, 350, -150));
        }
        // emitter B :410-420
        testHit = sphereIntersect(rayStart, rayDirection, V3(0, 3300, 1300), 250.f);
        if (testHit > 0.f && testHit < nearestHit) {
            nearestHit = testHit;
            r.Le = V3(1.0, 1.0, 1.0) * 15.2f;
            r.N = normalize(rayStart + (rayDirection * nearestHit) - V3(500, 800, 1300));
        }
        // walls :424-500 — centres are computed in double and narrowed by the vec3 ctor
        auto sizeOfSpheres = 1e7 * 5;
        struct Wall { V3 c; float sign; };
        const Wall walls[6] = {
            {V3(0, (float)-sizeOfSpheres, 0), 1.f},             // :444
            {V3(0, (float)(sizeOfSpheres + 1000), 0), 1.f},     // :452
            {V3((float)(-sizeOfSpheres + 2000), 0, 0), -1.f},   // :463
            {V3((float)(sizeOfSpheres - 2000), 0, 0), -1.f},    // :472
            {V3(0, 0, (float)(-sizeOfSpheres + 2000)), -1.f},   // :483
            {V3(0, 0, (float)(sizeOfSpheres - 2000)), 1.f},     // :492
        };
        for (int i = 0; i < 6; ++i) {
            testHit = sphereIntersect(rayStart, rayDirection, walls[i].c, (float)sizeOfSpheres);
            if (testHit > 0.f && testHit < nearestHit) {
                nearestHit = testHit;
                V3 n = normalize(rayStart + (rayDirection * nearestHit) - walls[i].c);
                r.N = walls[i].sign < 0 ? -n : n;
            }
        }
    }
    r.material = hitMeshIndex >= 0; // :502-503
    r.P = rayStart + (rayDirection * nearestHit);
    r.dist = nearestHit;
    r.hit = nearestHit < INFINITY;
    return r;
}

// pathtracer.cpp:21-198 (mode COMPAT) and the documented FIXED estimator:
// identical control flow and draw order, but r2 = xi (not 10*xi) and the path
// ends after max_segments RayCast calls instead of by roulette.
template <class Rng>
inline V4 radiance(const orc_scene& sc, V3 rStart, V3 rDir, Rng& rng, int mode, uint32_t max_segments, uint32_t* nrays) {
    V4 accumColour{0, 0, 0, -100};
    V4 accumRadiance{1, 1, 1, 1};
    short depth = 0;
    uint32_t segs = 0;
    const float r2scale = (mode == ORC_MODE_FIXED) ? 1.f : 10.f;
    while (1) {
        RayCastOut h = ray_cast(sc, rStart, rDir);
        ++segs;
        rng.begin_bounce(segs);
        if (nrays) ++*nrays;
        if (!h.hit) return accumColour; // :37-41
        accumColour.x += accumRadiance.x * h.Le.x; // :43
        accumColour.y += accumRadiance.y * h.Le.y;
        accumColour.z += accumRadiance.z * h.Le.z;
        accumColour.w += accumRadiance.w * 0.f;
        if (depth == 0) accumColour.w = h.dist; // :44-47
        if (length(h.Le) > 1.f) return accumColour; // :52
        if (mode == ORC_MODE_FIXED) {
            ++depth;
            if (segs >= max_segments) return accumColour;
        } else {
            if (++depth > 5 && (rng.next01() > 0.95f || depth > 1000)) return accumColour; // :56-59
        }
        float sampleColour[4] = {1.f, 1.f, 1.f, 1.f}; // :62-80
        if (h.material && !sc.textures.empty()) texture_sample(sc.textures[0], h.uv, sampleColour);
        V3 hitNormal = h.N, hitLocation = h.P;
        if (h.material && rng.next01() >= 0.96) { // :98-109
            rng.next01(); rng.next01(); rng.next01(); // gx, gy, gz (unused noise)
            V3 newStart = hitLocation - rDir * 0.001f;
            rDir = normalize(rDir - hitNormal * 2.f * dot(hitNormal, rDir));
            rStart = newStart;
        } else if (h.material) { // :151-165
            accumRadiance.x *= sampleColour[0]; accumRadiance.y *= sampleColour[1];
            accumRadiance.z *= sampleColour[2]; accumRadiance.w *= sampleColour[3];
            float r1 = 2 * M_PI * rng.next01();
            float r2 = r2scale * rng.next01();
            float r2s = sqrt(r2);
            V3 w = dot(hitNormal, rDir) < 0.f ? hitNormal : hitNormal * -1.f;
            V3 u = normalize(cross(fabs(w.x) > .1 ? V3(0, 1, 0) : V3(1, 0, 0), w));
            V3 v = cross(w, u);
            V3 d = normalize(u * float(cos(r1)) * r2s + v * float(sin(r1)) * r2s + w * float(sqrt(1 - r2)));
            rStart = hitLocation - rDir * 0.001f;
            rDir = d;
        } else { // :166-190
            double r1 = 2 * M_PI * rng.next01();
            double r2 = r2scale * rng.next01();
            float r2s = sqrt(r2);
            rng.next01(); rng.next01(); rng.next01(); // gx, gy, gz (unused noise)
            V3 w = dot(hitNormal, rDir) < 0.f ? hitNormal : hitNormal * -1.f;
            V3 u = normalize(cross(fabs(w.x) > .1 ? V3(0, 1, 0) : V3(1, 0, 0), w));
            V3 v = cross(w, u);
            V3 d = normalize(u * float(cos(r1)) * r2s + v * float(sin(r1)) * r2s + w * float(sqrt(1 - r2)));
            rStart = hitLocation - rDir * 0.001f;
            rDir = d;
        }
    }
}

// glm::rotate (extern/glm/glm/ext/matrix_transform.inl:18-46), column-major m[col][row]
struct M4 { float m[4][4]; };
inline M4 rotate(const M4& m, float angle, V3 v) {
    const float a = angle;
    const float c = std::cos(a);
    const float s = std::sin(a);
    V3 axis = normalize(v);
    V3 temp = axis * (1.f - c); // (T(1)-c) * axis
    float R[3][3];
    R[0][0] = c + temp.x * axis.x;
    R[0][1] = temp.x * axis.y + s * axis.z;
    R[0][2] = temp.x * axis.z - s * axis.y;
    R[1][0] = temp.y * axis.x - s * axis.z;
    R[1][1] = c + temp.y * axis.y;
    R[1][2] = temp.y * axis.z + s * axis.x;
    R[2][0] = temp.z * axis.x + s * axis.y;
    R[2][1] = temp.z * axis.y - s * axis.x;
    R[2][2] = c + temp.z * axis.z;
    M4 out;
    for (int col = 0; col < 3; ++col)
        for (int r = 0; r < 4; ++r)
            out.m[col][r] = m.m[0][r] * R[col][0] + m.m[1][r] * R[col][1] + m.m[2][r] * R[col][2];
    for (int r = 0; r < 4; ++r) out.m[3][r] = m.m[3][r];
    return out;
}

// pathtracer.cpp:216-221
inline M4 camera_matrix(const orc_camera& cam) {
    M4 I;
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) I.m[c][r] = (c == r) ? 1.f : 0.f;
    M4 t = rotate(I, cam.rotation[1], V3(0, 1, 0));
    t = rotate(t, cam.rotation[0], V3(1, 0, 0));
    t = rotate(t, cam.rotation[2], V3(0, 0, 1));
    return t;
}

// pathtracer.cpp:251-280
inline void camera_ray(const orc_camera& cam, const M4& M, uint64_t p, uint32_t sampleX, uint32_t sampleY, float jx, float jy,
                       V3* o, V3* d) {
    float homogenousX = ((float(p % cam.width) + (sampleX * 0.5f - 0.5f) + jx - 0.25) / cam.width) - 0.5;
    float homogenousY = ((float(p / cam.width) + (sampleY * 0.5f - 0.5f) + jy - 0.25) / cam.height) - 0.5;
    float cameraBackXCm = homogenousX * cam.sensor_x;
    float cameraBackYCm = homogenousY * cam.sensor_y;
    float g[4] = {cameraBackXCm, -cameraBackYCm, -cam.dist_to_film, 1.0f};
    // mat4 * vec4, extern/glm/glm/detail/type_mat4x4.inl:561-572: (m0*x + m1*y) + (m2*z + m3*w)
    float rd[4];
    for (int r = 0; r < 4; ++r) rd[r] = (M.m[0][r] * g[0] + M.m[1][r] * g[1]) + (M.m[2][r] * g[2] + M.m[3][r] * g[3]);
    V3 dir(rd[0] / rd[3], rd[1] / rd[3], rd[2] / rd[3]);
    *d = normalize(dir);
    *o = V3(cam.position[0], cam.position[1], cam.position[2]);
}

template <class Rng>
uint64_t render_pixel(const orc_scene& sc, const orc_camera& cam, const M4& M, int mode, uint64_t p, uint32_t s0, uint32_t s1,
                      uint32_t max_segments, Rng& rng, float* px5, float* raw4) {
    uint64_t nrays = 0;
    V4 accum{0, 0, 0, 0};
    uint32_t nTotalSamples = 0;
    if (mode == ORC_MODE_COMPAT) { // pathtracer.cpp:242-312
        for (uint16_t sampleX = 0; sampleX < 2; ++sampleX) {
            for (uint16_t sampleY = 0; sampleY < 2; ++sampleY) {
                for (uint32_t sampleZ = 0; sampleZ < (cam.spp / 4); ++sampleZ) {
                    ++nTotalSamples;
                    rng.begin_sample(nTotalSamples - 1);
                    float jx = rng.jitter();
                    float jy = rng.jitter();
                    V3 o, d;
                    camera_ray(cam, M, p, sampleX, sampleY, jx, jy, &o, &d);
                    uint32_t nr = 0;
                    V4 s = radiance(sc, o, d, rng, mode, max_segments, &nr);
                    nrays += nr;
                    accum.x += s.x; accum.y += s.y; accum.z += s.z; accum.w += s.w;
                    if (nTotalSamples > sqrt(cam.spp)) {
                        float fn = float(nTotalSamples), fn1 = float(nTotalSamples + 1);
                        V3 a(accum.x / fn, accum.y / fn, accum.z / fn);
                        V3 b((accum.x + s.x) / fn1, (accum.y + s.y) / fn1, (accum.z + s.z) / fn1);
                        if (fabs(length(a - b)) < 0.00001f) break;
                    }
                }
            }
        }
    } else {
        // FIXED: every sample index in [s0,s1) of the spp-long stratified sequence, same
        // loop nesting as the reference (sampleX outer, sampleY, sampleZ inner), no early-out.
        uint32_t per = cam.spp / 4;
        for (uint32_t s = s0; s < s1; ++s) {
            uint32_t sampleX = per ? (s / (2 * per)) & 1u : 0, sampleY = per ? (s / per) & 1u : 0;
            ++nTotalSamples;
            rng.begin_sample(s);
            float jx = rng.jitter();
            float jy = rng.jitter();
            V3 o, d;
            camera_ray(cam, M, p, sampleX, sampleY, jx, jy, &o, &d);
            uint32_t nr = 0;
            V4 sv = radiance(sc, o, d, rng, mode, max_segments, &nr);
            nrays += nr;
            accum.x += sv.x; accum.y += sv.y; accum.z += sv.z; accum.w += sv.w;
        }
    }
    if (raw4) { raw4[0] = accum.x; raw4[1] = accum.y; raw4[2] = accum.z; raw4[3] = (float)nTotalSamples; }
    // pathtracer.cpp:318-323
    px5[0] = std::max(std::min(accum.x / nTotalSamples, 1.f), 0.f);
    px5[1] = std::max(std::min(accum.y / nTotalSamples, 1.f), 0.f);
    px5[2] = std::max(std::min(accum.z / nTotalSamples, 1.f), 0.f);
    px5[3] = 1.f;
    px5[4] = nTotalSamples;
    return nrays;
}

// ----------------------------------------------------------------------------
// BruteForceTracer::Render (integrators.cpp:9-186): direct light from a point light,
// optional normal-map perturbation, one mirror-direction probe, its own early-out.
// ----------------------------------------------------------------------------
// std::mt19937 exactly as integrators.cpp:30,35 uses it: ONE engine shared by all pixels
// (the reference races on it under OpenMP; with one thread the pixels draw in order),
// uniform_real_distribution<float>(0, 0.5).
struct RngMt32Shared {
    std::mt19937& eng;
    explicit RngMt32Shared(std::mt19937& e) : eng(e) {}
    void begin_sample(uint32_t) {}
    float jitter() { std::uniform_real_distribution<float> d(0, 0.5); return d(eng); }
};

// integrators.cpp:170: `abs` there resolves to ::abs(int) (no float overload is visible in that TU), so the float
// sum is truncated to int first. x86 cvttss2si returns INT_MIN for NaN / out-of-range, and abs(INT_MIN) stays INT_MIN.
inline float abs_via_int(float x) {
    int32_t i = (std::fabs(x) < 2147483648.f) ? (int32_t)x : INT32_MIN;
    uint32_t a = i < 0 ? 0u - (uint32_t)i : (uint32_t)i;
    return (float)(int32_t)a;
}

inline float std_min(float a, float b) { return (b < a) ? b : a; } // std::min(a, b)
inline float std_max(float a, float b) { return (a < b) ? b : a; } // std::max(a, b)

template <class Rng>
uint64_t render_pixel_bf(const orc_scene& sc, const orc_camera& cam, const M4& M, uint64_t p, uint32_t flags, Rng& rng, float* px5,
                         uint32_t* samples_taken, float* raw6 = nullptr) {
    const V3 lightLocation(500, 1100, 2000); // :16
    uint64_t nrays = 0;
    V4 accum{0, 0, 0, 0}, last{0, 0, 0, 0}, cur{0, 0, 0, 0};
    uint32_t nTotalSamples = 0;
    float hitDistance = 0.f; // uninitialised in the reference when spp == 0; every RayCast call writes it
    for (uint16_t sample = 0; sample < cam.spp; ++sample) { // :59
        ++nTotalSamples;
        rng.begin_sample(sample);
        float jx = rng.jitter();
        float homogenousX = ((float(p % cam.width) + jx - 0.25) / cam.width) * 2 - 1; // :65
        float jy = rng.jitter();
        float homogenousY = ((float(p / cam.width) + jy - 0.25) / cam.height) * 2 - 1; // :66
        float cameraBackXCm = homogenousX * cam.sensor_x * 0.5; // :73
        float cameraBackYCm = homogenousY * cam.sensor_y * 0.5;
        float g[4] = {cameraBackXCm, -cameraBackYCm, -cam.dist_to_film, 1.0f};
        float rd[4];
        for (int r = 0; r < 4; ++r) rd[r] = (M.m[0][r] * g[0] + M.m[1][r] * g[1]) + (M.m[2][r] * g[2] + M.m[3][r] * g[3]);
        V3 dir = normalize(V3(rd[0] / rd[3], rd[1] / rd[3], rd[2] / rd[3]));
        RayCastOut h = ray_cast(sc, V3(cam.position[0], cam.position[1], cam.position[2]), dir); // :81
        ++nrays;
        hitDistance = h.dist;
        if (h.hit) {
            V3 lightDirection = normalize(lightLocation - h.P); // :83-84
            V3 hitNormal = h.N;
            float vNDL = dot(lightDirection, normalize(hitNormal)); // :88
            if (sc.textures.size() > 0) { // :98-106
                float n4[4];
                texture_sample(sc.textures[0], h.uv, n4);
                hitNormal = hitNormal + V3(n4[0], n4[1], n4[2]);
                vNDL = dot(lightDirection, normalize(hitNormal));
            }
            if (flags & ORC_BF_SHADOW) { // the commented-out :109-110, RayCastCollision semantics (meshEngine.cpp:196-207)
                Ray sr(h.P, lightDirection);
                float st;
                uint32_t id = sc.bvh.closest<false>(sc.tris, sr, &st, nullptr, nullptr);
                ++nrays;
                if (id != 0xFFFFFFFFu && st > 1e-3) vNDL *= 0.00f;
            }
            V3 secondBounceDirection = -lightDirection - (hitNormal * 2.f) * dot(hitNormal, -lightDirection); // :119
            RayCastOut h2 = ray_cast(sc, h.P, secondBounceDirection); // :121 (its outputs only feed `resp`, which is never used)
            ++nrays;
            if (!h2.hit) { vNDL *= 0.9f; vNDL += 0.1f; } // :133-137
            if (sc.textures.size() > 1) { // :141-147
                float c4[4];
                texture_sample(sc.textures[1], h.uv, c4);
                cur = V4{c4[0] * vNDL, c4[1] * vNDL, c4[2] * vNDL, 1.0f};
            } else {
                cur = V4{0.890196078f * vNDL, 0.258823529f * vNDL, 0.203921569f * vNDL, 1.0f}; // :150-155
            }
            accum.x += cur.x; accum.y += cur.y; accum.z += cur.z; accum.w += cur.w; // :158
        }
        const float fn = (float)nTotalSamples;
        if (nTotalSamples > 2) { // :167-172
            last.x -= accum.x / fn; last.y -= accum.y / fn; last.z -= accum.z / fn; last.w -= accum.w / fn;
            float mag = abs_via_int(last.x + last.y + last.z + last.w);
            if (mag < 0.001f) break;
        }
        last = V4{accum.x / fn, accum.y / fn, accum.z / fn, accum.w / fn}; // :173
    }
    px5[0] = std_max(std_min(accum.x / nTotalSamples, 1.f), 0.f); // :177-181
    px5[1] = std_max(std_min(accum.y / nTotalSamples, 1.f), 0.f);
    px5[2] = std_max(std_min(accum.z / nTotalSamples, 1.f), 0.f);
    px5[3] = accum.w / nTotalSamples;
    px5[4] = hitDistance;
    if (samples_taken) *samples_taken = nTotalSamples;
    if (raw6) { raw6[0] = accum.x; raw6[1] = accum.y; raw6[2] = accum.z; raw6[3] = (float)nTotalSamples; raw6[4] = accum.w; raw6[5] = hitDistance; }
    return nrays;
}

inline int clamp_threads(int t) { int m = omp_get_max_threads(); return (t <= 0 || t > m) ? m : t; }

} // namespace

// ----------------------------------------------------------------------------
// C API
// ----------------------------------------------------------------------------
extern "C" {

orc_scene* orc_scene_create(const float* pos, const float* nrm, const float* uv, uint64_t ntri, uint32_t leaf_size) {
    orc_scene* s = new orc_scene();
    s->tris.resize(ntri);
    for (uint64_t i = 0; i < ntri; ++i) {
        Tri& t = s->tris[i];
        const float* p = pos + 9 * i;
        t.v0 = V3(p[0], p[1], p[2]); t.v1 = V3(p[3], p[4], p[5]); t.v2 = V3(p[6], p[7], p[8]);
        if (nrm) { const float* n = nrm + 9 * i; t.n0 = V3(n[0], n[1], n[2]); t.n1 = V3(n[3], n[4], n[5]); t.n2 = V3(n[6], n[7], n[8]); }
        if (uv) { const float* q = uv + 6 * i; t.uv0[0] = q[0]; t.uv0[1] = q[1]; t.uv1[0] = q[2]; t.uv1[1] = q[3]; t.uv2[0] = q[4]; t.uv2[1] = q[5]; }
        else { t.uv0[0] = t.uv0[1] = t.uv1[0] = t.uv1[1] = t.uv2[0] = t.uv2[1] = 0.f; }
    }
    s->bvh.leafSize = leaf_size ? leaf_size : 4;
    s->bvh.build(s->tris);
    return s;
}
void orc_scene_destroy(orc_scene* s) { delete s; }
void orc_scene_set_texture(orc_scene* s, const float* data, uint32_t w, uint32_t h, uint32_t c) {
    Texture t; t.w = w; t.h = h; t.c = c; t.data.assign(data, data + (size_t)w * h * c);
    s->textures.clear(); s->textures.push_back(std::move(t));
}
void orc_scene_add_texture(orc_scene* s, const float* data, uint32_t w, uint32_t h, uint32_t c) { // MeshEngine::bindTexture appends (meshEngine.cpp:91)
    Texture t; t.w = w; t.h = h; t.c = c; t.data.assign(data, data + (size_t)w * h * c);
    s->textures.push_back(std::move(t));
}
void orc_scene_set_spheres(orc_scene* s, int enabled) { s->spheres = enabled != 0; }
void orc_scene_stats(const orc_scene* s, uint32_t* n_nodes, uint32_t* n_leafs) {
    if (n_nodes) *n_nodes = s->bvh.nNodes;
    if (n_leafs) *n_leafs = s->bvh.nLeafs;
}

void orc_intersect(const orc_scene* s, const orc_ray* rays, uint64_t n, orc_hit* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Ray r(V3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), V3(rays[i].d[0], rays[i].d[1], rays[i].d[2]));
        float t;
        uint32_t id = s->bvh.closest<false>(s->tris, r, &t, nullptr, nullptr);
        out[i].t = t; out[i].tri = id; out[i].u = 0.f; out[i].v = 0.f;
        if (id != 0xFFFFFFFFu) {
            float bary[2];
            V3 hp = r.o + r.d * t;
            s->tris[id].getNormal(hp, nullptr, bary);
            out[i].u = bary[0]; out[i].v = bary[1];
        }
    }
}

void orc_intersect_stats(const orc_scene* s, const orc_ray* rays, uint64_t n, uint64_t* nodes, uint64_t* tris) {
    uint64_t cn = 0, ct = 0;
#pragma omp parallel for schedule(dynamic, 4096) reduction(+ : cn, ct)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Ray r(V3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), V3(rays[i].d[0], rays[i].d[1], rays[i].d[2]));
        float t; uint64_t a = 0, b = 0;
        s->bvh.closest<true>(s->tris, r, &t, &a, &b);
        cn += a; ct += b;
    }
    *nodes = cn; *tris = ct;
}

void orc_occluded(const orc_scene* s, const orc_ray* rays, uint64_t n, uint8_t* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) { // meshEngine.cpp:196-207
        Ray r(V3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), V3(rays[i].d[0], rays[i].d[1], rays[i].d[2]));
        float t;
        uint32_t id = s->bvh.closest<false>(s->tris, r, &t, nullptr, nullptr);
        out[i] = (id != 0xFFFFFFFFu && t > 1e-3) ? 1 : 0;
    }
}

void orc_raycast_many(const orc_scene* s, const orc_ray* rays, uint64_t n, orc_raycast* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        RayCastOut r = ray_cast(*s, V3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), V3(rays[i].d[0], rays[i].d[1], rays[i].d[2]));
        orc_raycast& o = out[i];
        o.P[0] = r.P.x; o.P[1] = r.P.y; o.P[2] = r.P.z; o.dist = r.dist;
        o.N[0] = r.N.x; o.N[1] = r.N.y; o.N[2] = r.N.z;
        o.flags = (r.hit ? ORC_RC_HIT : 0u) | (r.material ? ORC_RC_MATERIAL : 0u);
        o.uv[0] = r.uv[0]; o.uv[1] = r.uv[1];
        o.Le[0] = r.Le.x; o.Le[1] = r.Le.y; o.Le[2] = r.Le.z;
        o.tri = r.tri; o.pad[0] = o.pad[1] = 0;
    }
}

void orc_radiance_many(const orc_scene* s, const orc_ray* rays, uint64_t n, uint64_t seed, int rng_kind, int mode, float* out,
                       uint32_t* rays_cast, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        V3 o(rays[i].o[0], rays[i].o[1], rays[i].o[2]), d(rays[i].d[0], rays[i].d[1], rays[i].d[2]);
        uint32_t nr = 0;
        V4 v;
        if (rng_kind == ORC_RNG_MT64) { RngMt64 rng(seed + (uint64_t)i); v = radiance(*s, o, d, rng, mode, 5, &nr); }
        else { RngPhilox rng(seed, (uint64_t)i); rng.begin_sample(0); rng.draw = 2; v = radiance(*s, o, d, rng, mode, 5, &nr); }
        out[4 * i + 0] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z; out[4 * i + 3] = v.w;
        if (rays_cast) rays_cast[i] = nr;
    }
}

void orc_camera_matrix(const orc_camera* cam, float* m16) {
    M4 M = camera_matrix(*cam);
    for (int c = 0; c < 4; ++c) for (int r = 0; r < 4; ++r) m16[c * 4 + r] = M.m[c][r];
}

void orc_camera_ray(const orc_camera* cam, uint64_t p, uint32_t sx, uint32_t sy, float jx, float jy, orc_ray* out) {
    M4 M = camera_matrix(*cam);
    V3 o, d;
    camera_ray(*cam, M, p, sx, sy, jx, jy, &o, &d);
    out->o[0] = o.x; out->o[1] = o.y; out->o[2] = o.z; out->pad0 = 0;
    out->d[0] = d.x; out->d[1] = d.y; out->d[2] = d.z; out->pad1 = 0;
}

uint64_t orc_render(const orc_scene* s, const orc_camera* cam, int mode, int rng_kind, uint64_t seed, uint64_t p0, uint64_t p1,
                    uint32_t s0, uint32_t s1, uint32_t max_segments, float* image, float* raw_sum, int threads) {
    threads = clamp_threads(threads);
    M4 M = camera_matrix(*cam);
    uint64_t total = 0;
    if (max_segments == 0) max_segments = 5;
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) reduction(+ : total)
    for (int64_t p = (int64_t)p0; p < (int64_t)p1; ++p) {
        float* px = image + 5 * (size_t)p;
        float* raw = raw_sum ? raw_sum + 4 * (size_t)p : nullptr;
        if (rng_kind == ORC_RNG_MT64) { RngMt64 rng(seed + (uint64_t)p); total += render_pixel(*s, *cam, M, mode, (uint64_t)p, s0, s1, max_segments, rng, px, raw); }
        else { RngPhilox rng(seed, (uint64_t)p); total += render_pixel(*s, *cam, M, mode, (uint64_t)p, s0, s1, max_segments, rng, px, raw); }
    }
    return total;
}

uint64_t orc_render_bruteforce(const orc_scene* s, const orc_camera* cam, int rng_kind, uint64_t seed, uint32_t flags, uint64_t p0,
                                uint64_t p1, float* image, uint32_t* samples_taken, float* raw, int threads) {
    M4 M = camera_matrix(*cam);
    uint64_t total = 0;
    if (rng_kind == ORC_RNG_MT32_SHARED) { // sequential by construction: one engine, pixels in order
        std::mt19937 eng((uint32_t)seed);
        RngMt32Shared rng(eng);
        for (uint64_t p = p0; p < p1; ++p)
            total += render_pixel_bf(*s, *cam, M, p, flags, rng, image + 5 * (size_t)p, samples_taken ? samples_taken + p : nullptr, raw ? raw + 6 * (size_t)p : nullptr);
        return total;
    }
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads) reduction(+ : total)
    for (int64_t p = (int64_t)p0; p < (int64_t)p1; ++p) {
        RngPhilox rng(seed, (uint64_t)p);
        total += render_pixel_bf(*s, *cam, M, (uint64_t)p, flags, rng, image + 5 * (size_t)p, samples_taken ? samples_taken + p : nullptr, raw ? raw + 6 * (size_t)p : nullptr);
    }
    return total;
}

void orc_philox(uint64_t seed, uint64_t pixel, uint32_t sample, uint32_t block, uint32_t out[4]) {
    out[0] = (uint32_t)pixel; out[1] = (uint32_t)(pixel >> 32); out[2] = sample; out[3] = block;
    philox4x32_10(out, (uint32_t)seed, (uint32_t)(seed >> 32));
}

int orc_max_threads(void) { return omp_get_max_threads(); }

} // extern "C"
