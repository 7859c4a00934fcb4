// wavefront.cu — the path-tracing wavefront: raygen -> (extend -> shade)* -> accumulate -> resolve.
//
// Replaces PathTracer::Render's pixel loop and Radiance() (reference core/integrators/pathtracer.cpp:
// 21-198, 200-328). Stages are separate kernels connected by device-resident ray queues:
//   k_raygen      one thread per (pixel, sample) of the round: camera ray (pathtracer.cpp:251-280), meta word 0
//   k_extend      persistent warps, BVH8 closest hit per queued ray -> (t, triangle id)      [the hot kernel]
//   k_shade       the rest of MeshEngine::RayCast (normals, 8 spheres) + one Radiance iteration; survivors are
//                 appended to the next queue with one atomicAdd per warp (ballot/popc compaction)
//   k_accumulate  per pixel, adds the round's finished samples to the film in sample order (deterministic,
//                 no float atomics); in COMPAT mode also runs the reference's adaptive early-out state machine
//   k_resolve     film sums -> RGBAZ (pathtracer.cpp:318-323, camera.cpp:106-113)
// Path state (radiance, throughput) is SoA indexed by slot; queue entries carry the slot and the path's meta word (depth,
// throughput-written bit, segments). Queue counters live in device memory, so a FIXED-mode round needs no host synchronisation.
#include <cstdlib>
#include "engine.cuh"
#include "traverse.cuh"

namespace vmn {

constexpr int kWfBlock = 128;

struct CamDev {
    float m[16];
    float px, py, pz;
    float dist, sx, sy;
    uint32_t width, height, spp;
};

// pathtracer.cpp:251-280 (types follow the C++ promotions: the 0.25 / 0.5 literals are double)
__device__ __forceinline__ void camera_ray(const CamDev& c, uint64_t p, uint32_t sampleX, uint32_t sampleY, float jx, float jy, V3* o, V3* d) {
    uint64_t py = p / c.width, px = p - py * c.width;
    if ((p >> 32) == 0) { const uint32_t y32 = (uint32_t)p / c.width; py = y32; px = (uint32_t)p - y32 * c.width; } // one 32-bit divide
    float hx = (float)((((double)(float(px) + (sampleX * 0.5f - 0.5f) + jx) - 0.25) / (double)c.width) - 0.5);
    float hy = (float)((((double)(float(py) + (sampleY * 0.5f - 0.5f) + jy) - 0.25) / (double)c.height) - 0.5);
    float bx = hx * c.sx, by = hy * c.sy;
    float g0 = bx, g1 = -by, g2 = -c.dist, g3 = 1.0f;
    float rd[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) rd[r] = (c.m[r] * g0 + c.m[4 + r] * g1) + (c.m[8 + r] * g2 + c.m[12 + r] * g3);
    V3 dir(rd[0] / rd[3], rd[1] / rd[3], rd[2] / rd[3]);
    *d = normalize(dir);
    *o = V3(c.px, c.py, c.pz);
}

struct PathPool {
    uint64_t capacity = 0;
    DevBuf ray[2];    // 32 B per entry: (o.xyz, slot) (d.xyz, meta = depth 0..14 | throughput-written bit 15 | segments 16..31)
    DevBuf hit;       // float2 per queue entry: (t, original triangle id)
    DevBuf radiance;  // float4 per slot (rgb, first-hit distance / -100)
    DevBuf thr;       // float4 per slot
    DevBuf meta;      // vmn_radiance: final meta word per slot (segments cast); BRUTEFORCE: uint2 per pixel (nTotalSamples, done)
    DevBuf ctl;       // uint32 counters: [0..1] queue counts, [8..] claim counters (unsigned long long)
    DevBuf cstate;    // COMPAT per-local-pixel stratum state
    DevBuf bf_last;   // BRUTEFORCE per-local-pixel lastSampleColour (float4)
    DevBuf tile_prefix; // tile mode, FIXED: live pixels in the local tiles before tile j (uint32 per local tile), see k_raygen
    uint32_t tile_live = 0;            // live pixels in all local tiles
    uint32_t tile_key[5] = {0, 0, 0, 0, 0}; // (width, height, tile, rank, world) the table was made for
    uint32_t* h_count = nullptr; // pinned
    std::vector<cudaEvent_t> events;
    size_t ev_used = 0;
};

void pool_destroy(vmn_ctx* c) {
    PathPool* p = c->pool;
    if (!p) return;
    for (DevBuf* b : {&p->ray[0], &p->ray[1], &p->hit, &p->radiance, &p->thr, &p->meta, &p->ctl, &p->cstate, &p->bf_last, &p->tile_prefix}) dev_free(*b);
    if (p->h_count) cudaFreeHost(p->h_count);
    for (auto e : p->events) cudaEventDestroy(e);
    delete p;
    c->pool = nullptr;
}

static vmn_status pool_ensure(vmn_ctx* c, uint64_t paths) {
    if (!c->pool) c->pool = new PathPool();
    PathPool* p = c->pool;
    vmn_status s;
    if (paths > p->capacity) {
        // A pool at or above half of max_paths is allocated at max_paths at once: two frames whose (pixels x samples per batch) products
        // differ by a percent must not free and reallocate the whole pool inside a render (r02z: C2 then C4 at 8 GPUs, 132.7 M then
        // 133.7 M paths of a 134.2 M cap = 15 GB freed and allocated again, +97 ms on the timed C4 frame).
        const uint64_t cap = c->max_paths < 0xFFFFFFF0ull ? c->max_paths : 0xFFFFFFF0ull;
        if (paths <= cap && paths * 2 >= cap) paths = cap;
        if ((s = dev_alloc(c, p->ray[0], 32 * paths)) != VMN_OK) return s;
        if ((s = dev_alloc(c, p->ray[1], 32 * paths)) != VMN_OK) return s;
        if ((s = dev_alloc(c, p->hit, 8 * paths)) != VMN_OK) return s;
        if ((s = dev_alloc(c, p->radiance, 16 * paths)) != VMN_OK) return s;
        if ((s = dev_alloc(c, p->thr, 16 * paths)) != VMN_OK) return s;
        if ((s = dev_alloc(c, p->meta, 8 * paths)) != VMN_OK) return s;
        p->capacity = paths;
    }
    if (!p->ctl.p && (s = dev_alloc(c, p->ctl, 4096)) != VMN_OK) return s;
    if (!p->h_count) VMN_CUDA(c, cudaMallocHost(&p->h_count, 64));
    return VMN_OK;
}

// ---- raygen ----------------------------------------------------------------------------------------
struct RoundParams {
    uint64_t seed;
    uint64_t n_local;   // local pixels in this chunk
    uint64_t lp0;       // first local pixel of the chunk
    uint32_t s0, batch; // first sample index, samples per pixel this round
    uint32_t per;       // spp / 4 (stratum length)
    int32_t mode;
    uint32_t max_segments;
};

constexpr uint32_t kThrWritten = 0x8000u; // path meta bit 15: thr[slot] holds the throughput (otherwise it is exactly (1,1,1,1))

#ifndef VMN_RAYGEN_DIRECT
#define VMN_RAYGEN_DIRECT 1
#endif
__global__ void __launch_bounds__(kWfBlock) k_raygen(CamDev cam, PixelMap pm, RoundParams rp, const uint32_t* __restrict__ cstate,
                                                     float4* __restrict__ rays, uint32_t* __restrict__ qcount, float4* __restrict__ radiance,
                                                     uint32_t* __restrict__ meta, const uint32_t* __restrict__ tile_prefix, uint32_t tile_live) {
    const uint64_t idx = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t total = rp.n_local * rp.batch;
    bool live = idx < total;
    uint64_t pixel = 0;
    uint32_t sample = 0, sxy = 0;
    uint32_t qdirect = 0; // tile mode with a prefix table: this path's queue position
    if (live) {
        // (pixel, sample) of a slot with a 32-bit divide whenever the round fits 32 bits (queue slots are 32-bit, so it always does)
        uint64_t lp; uint32_t k;
        if ((total >> 32) == 0) { k = (uint32_t)idx / (uint32_t)rp.n_local; lp = (uint32_t)idx - k * (uint32_t)rp.n_local; }
        else { lp = idx % rp.n_local; k = (uint32_t)(idx / rp.n_local); }
        sample = rp.s0 + k;
#if VMN_RAYGEN_DIRECT
        if (tile_prefix) {
            uint32_t tj, rk;
            live = map_pixel_tile(pm, (uint32_t)lp, &pixel, &tj, &rk);
            qdirect = k * tile_live + __ldg(&tile_prefix[tj]) + rk;
        } else
#endif
        live = map_pixel(pm, rp.lp0 + lp, &pixel);
        if (live) {
            if (rp.mode == VMN_MODE_COMPAT) {
                const uint32_t st = cstate[lp] >> 30; // stratum 0..3, 4 (encoded as done flag below) never reaches here
                live = (cstate[lp] & 0x20000000u) == 0u;
                sxy = st;
            } else {
                sxy = rp.per ? ((sample / rp.per) & 3u) : 0u; // sampleX = bit 1, sampleY = bit 0 (loop nesting of pathtracer.cpp:242-247)
            }
        }
        // the path's meta word (depth, segments, kThrWritten) travels in the w of the ray's second float4 and starts at 0; radiance[idx] is
        // first written by the path's first k_shade iteration; throughput = (1,1,1,1) until a textured bounce writes it
    }
    V3 o(0, 0, 0), d(0, 0, 0);
    if (live) {
        Rng rng(rp.seed, pixel, sample, 0);
        const float jx = rng.jitter(), jy = rng.jitter();
        camera_ray(cam, pixel, (sxy >> 1) & 1u, sxy & 1u, jx, jy, &o, &d);
    }
#if VMN_RAYGEN_DIRECT
    // A FIXED round over a contiguous pixel range has no dead slots: path idx goes to queue slot idx, no atomics, rays in pixel order.
    // Tiles: the live pixels before a path are known in closed form from a per-tile prefix table (only the image's last tile row and
    // column are partial), so its queue position is too.
    if (rp.mode == VMN_MODE_FIXED && (pm.tile == 0u || tile_prefix)) {
        const uint64_t q = pm.tile == 0u ? idx : (uint64_t)qdirect;
        if (idx == 0) *qcount = pm.tile == 0u ? (uint32_t)total : tile_live * rp.batch;
        if (live) {
            st_stream(&rays[2ull * q], make_float4(o.x, o.y, o.z, __uint_as_float((uint32_t)idx)));
            st_stream(&rays[2ull * q + 1], make_float4(d.x, d.y, d.z, 0.f));
        }
        return;
    }
#endif
    // warp-aggregated append
    const uint32_t ballot = __ballot_sync(0xffffffffu, live);
    if (ballot == 0u) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (uint32_t)(__ffs(ballot) - 1)) base = atomicAdd(qcount, (uint32_t)__popc(ballot));
    base = __shfl_sync(0xffffffffu, base, __ffs(ballot) - 1);
    if (live) {
        const uint32_t q = base + __popc(ballot & ((1u << lane) - 1u));
        st_stream(&rays[2ull * q], make_float4(o.x, o.y, o.z, __uint_as_float((uint32_t)idx)));
        st_stream(&rays[2ull * q + 1], make_float4(d.x, d.y, d.z, 0.f));
    }
}

// ---- extend ----------------------------------------------------------------------------------------
#ifndef VMN_EXTEND_MINB
#define VMN_EXTEND_MINB 8
#endif
template <bool COUNT>
__global__ void __launch_bounds__(kWfBlock, VMN_EXTEND_MINB) k_extend(SceneDev sc, const float4* __restrict__ rays, const uint32_t* __restrict__ qcount,
                                                     float2* __restrict__ hits, unsigned long long* next, TravCounters* gcnt,
                                                     unsigned long long* total_rays, uint32_t refill_min, uint32_t* __restrict__ qcount_other) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kWfBlock];
    __shared__ __align__(16) uint32_t wscratch_all[VMN_WSCRATCH * (kWfBlock / 32)];
    uint2* sstack = sstack_all + threadIdx.x;
    const uint64_t n = *qcount;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        atomicAdd(total_rays, (unsigned long long)n);
        *qcount_other = 0u; // the other queue is dead while this one is traced; the stage after this kernel appends to it from zero
    }
    TravCounters cnt{0ull, 0ull};
    persistent_trace<COUNT>(sc, n, next, pick_chunk(n, gridDim.x * (kWfBlock / 32)), refill_min, sstack, kWfBlock, wscratch_all + VMN_WSCRATCH * (threadIdx.x >> 5), &cnt,
                            [&](uint64_t i, V3* o, V3* d) {
                                const float4 a = ld_stream(&rays[2 * i]), b = ld_stream(&rays[2 * i + 1]);
                                *o = V3(a.x, a.y, a.z); *d = V3(b.x, b.y, b.z);
                            },
                            [&](uint64_t i, const Trav& s) {
                                st_stream(&hits[i], make_float2(s.best_t, __uint_as_float(leaf_original_id(sc.tris, s.best))));
                            }, rays);
    if (COUNT) { atomicAdd(&gcnt->nodes, cnt.nodes); atomicAdd(&gcnt->tris, cnt.tris); }
}

// ---- shade -----------------------------------------------------------------------------------------
// VermiTexture::Sample (meshEngine.cpp:21-46)
__device__ __forceinline__ void texture_sample(const float* __restrict__ tex, uint32_t tw, uint32_t th, uint32_t tc, float u, float v, float out[4]) {
    float sx = u - floorf(u), sy = v - floorf(v);
    uint32_t mx = (uint32_t)roundf(sx * (float)(tw - 1u));
    uint32_t my = (uint32_t)roundf(sy * (float)(th - 1u));
    const float* p = tex + (size_t)(my * tw + mx) * tc;
    out[0] = out[1] = out[2] = out[3] = 0.f;
    switch (tc) {
    case 1: out[0] = out[1] = out[2] = out[3] = p[0]; break;
    case 2: out[0] = p[0]; out[1] = p[1]; break;
    case 3: out[0] = p[0]; out[1] = p[1]; out[2] = p[2]; break;
    case 4: out[0] = p[0]; out[1] = p[1]; out[2] = p[2]; out[3] = p[3]; break;
    default: break;
    }
}
__device__ __forceinline__ void texture_sample(const SceneDev& sc, float u, float v, float out[4]) {
    texture_sample(sc.tex, sc.tex_w, sc.tex_h, sc.tex_c, u, v, out);
}

// cosine-ish hemisphere direction of pathtracer.cpp:155-162 / :175-186 around normal N for incoming rDir
__device__ __forceinline__ V3 diffuse_dir(V3 N, V3 rDir, float c1, float s1, float r2s, float sq) {
    V3 w = dot(N, rDir) < 0.f ? N : N * -1.f;
    V3 u = normalize(cross((double)fabsf(w.x) > .1 ? V3(0, 1, 0) : V3(1, 0, 0), w));
    V3 v = cross(w, u);
    return normalize(u * c1 * r2s + v * s1 * r2s + w * sq);
}

// The part of one Radiance() iteration (pathtracer.cpp:54-196) that follows the emission bookkeeping: roulette / segment cap, the Philox
// block of this bounce, texture lookup, mirror or diffuse continuation. Shared by the wavefront stage (k_shade) and the one-launch kernels
// for small frames (k_mega_compat), so that both produce the same bits. `segs` is the 1-based index of this iteration (already
// incremented); pixel_of(&pixel, &sample) names the Philox stream and is only evaluated for a path that is still alive.
template <class PixelOf>
__device__ __forceinline__ bool radiance_bounce(const SceneDev& sc, int mode, uint32_t max_segments, uint64_t seed, const RayCastOut& h, V3 rDir,
                                                uint32_t& depth, uint32_t segs, float4& tp, bool& thr_dirty, V3* nStart_out, V3* nDir_out, PixelOf pixel_of) {
    bool alive = true;
    V3 nStart(0, 0, 0), nDir(0, 0, 0);
    uint64_t pixel; uint32_t sample;
    pixel_of(&pixel, &sample);
    // iteration `segs` of the path draws from Philox block `segs` (block 0 holds the two pixel jitters): one block per bounce,
    // generated once by every lane that still needs numbers; at most four are used (roulette, mirror test, two angles)
    BounceRng rng;
    if (mode == VMN_MODE_FIXED) {
        ++depth;
        if (segs >= max_segments) alive = false;
        else rng.init(seed, pixel, sample, segs);
    } else {
        ++depth;
        rng.init(seed, pixel, sample, segs);
        if (depth > 5 && (rng.next01() > 0.95f || depth > 1000)) alive = false; // :56-59
    }
    if (alive) {
        const float r2scale = (mode == VMN_MODE_FIXED) ? 1.f : 10.f;
        float sc4[4] = {1.f, 1.f, 1.f, 1.f};
        if (h.material && sc.tex) texture_sample(sc, h.uv0, h.uv1, sc4); // :62-80
        if (h.material && rng.next01() >= 0.96) { // mirror :98-109 (its three unused noise draws need no numbers)
            nStart = h.P - rDir * 0.001f;
            nDir = normalize(rDir - h.N * 2.f * dot(h.N, rDir));
        } else {
            // diffuse off a triangle (:151-165) or off a sphere (:166-190): the two branches differ in where the draws are narrowed
            // to float, and share one sincos + hemisphere frame (a warp almost always holds both kinds of hit)
            const double xi1 = rng.next01(), xi2 = rng.next01();
            double r1d;
            float r2s, sq;
            if (h.material) {
                if (sc.tex) { tp.x *= sc4[0]; tp.y *= sc4[1]; tp.z *= sc4[2]; tp.w *= sc4[3]; thr_dirty = true; } // without a texture: * 1
                const float r1 = (float)(2 * 3.14159265358979323846 * xi1);
                const float r2 = (float)((double)r2scale * xi2);
                r2s = sqrtf(r2); sq = sqrtf(1 - r2);
                r1d = (double)r1; // cos(r1), sin(r1) of pathtracer.cpp:160 take the float r1 promoted to double
            } else {
                r1d = 2 * 3.14159265358979323846 * xi1;
                const double r2 = (double)r2scale * xi2;
                r2s = (float)sqrt(r2); sq = (float)sqrt(1 - r2); // (:172-174 draw three unused numbers)
            }
            double s1, c1;
            sincos(r1d, &s1, &c1);
            nDir = diffuse_dir(h.N, rDir, (float)c1, (float)s1, r2s, sq);
            nStart = h.P - rDir * 0.001f;
        }
    }
    *nStart_out = nStart; *nDir_out = nDir;
    return alive;
}

// One iteration of Radiance()'s loop (pathtracer.cpp:36-196) for every queued ray.
#ifndef VMN_SHADE_MINB
#define VMN_SHADE_MINB 9 // 56 registers: nine 128-thread CTAs per SM (8 -> 9: shade -1.3 %; 10 spills)
#endif
__global__ void __launch_bounds__(kWfBlock, VMN_SHADE_MINB) k_shade(SceneDev sc, const SphereDev* __restrict__ sph, PixelMap pm, RoundParams rp,
                                                    const float4* __restrict__ rays, const float2* __restrict__ hits,
                                                    const uint32_t* __restrict__ qcount_in, float4* __restrict__ rays_out,
                                                    uint32_t* __restrict__ qcount_out, float4* __restrict__ radiance, float4* __restrict__ thr,
                                                    uint32_t* __restrict__ meta, int radiance_query) {
    const uint32_t n = *qcount_in;
    const uint32_t lane = threadIdx.x & 31u;
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; base < n; base += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = base + lane;
        bool alive = false;
        V3 nStart(0, 0, 0), nDir(0, 0, 0);
        uint32_t slot = 0, mt_out = 0;
        if (i < n) {
            const float4 ra = ld_stream(&rays[2 * i]), rb = ld_stream(&rays[2 * i + 1]);
            const float2 hr = ld_stream(&hits[i]);
            slot = __float_as_uint(ra.w);
            VMN_CHECK(sc.err, (uint64_t)slot < rp.n_local * (rp.batch ? rp.batch : 1u));
            V3 rStart(ra.x, ra.y, ra.z), rDir(rb.x, rb.y, rb.z);
            const uint32_t mt = __float_as_uint(rb.w); // path meta rides in the ray record: no dependent gather, no scatter per bounce
            // the last segment of a FIXED path (every ray of that launch) ends here: no normal, uv or position wanted
            const bool last_seg = rp.mode == VMN_MODE_FIXED && (mt >> 16) + 1u >= rp.max_segments;
            const RayCastOut h = raycast_finish(sc, sph, rStart, rDir, hr.x, __float_as_uint(hr.y), !last_seg);
            // Path state traffic: the throughput stays exactly (1,1,1,1) until a textured diffuse bounce scales it, and the radiance only
            // changes on an emitter hit or at the first hit — so thr[] is not touched while kThrWritten is clear, and radiance[] is only
            // read-modified-written when the sum can change (x + 1*0 = x; with a written throughput it may be non-finite, so always then).
            uint32_t depth = mt & 0x7FFFu, segs = mt >> 16;
            const bool thr_written = (mt & kThrWritten) != 0u;
            float4 tp = thr_written ? thr[slot] : make_float4(1.f, 1.f, 1.f, 1.f);
            bool thr_dirty = false;
            ++segs;
            alive = true;
            if (!h.hit) alive = false; // :37-41
            if (alive) {
                if (thr_written || depth == 0 || h.Le.x != 0.f || h.Le.y != 0.f || h.Le.z != 0.f) {
                    // the first iteration of a path starts from accumColour = (0,0,0,-100) (pathtracer.cpp:26): nothing to read
                    float4 acc = depth == 0 ? make_float4(0.f, 0.f, 0.f, -100.f) : radiance[slot];
                    acc.x += tp.x * h.Le.x; acc.y += tp.y * h.Le.y; acc.z += tp.z * h.Le.z; acc.w += tp.w * 0.f; // :43
                    if (depth == 0) acc.w = h.dist; // :44-47
                    radiance[slot] = acc;
                }
                if (length(h.Le) > 1.f) alive = false; // :52
            } else if (depth == 0) {
                radiance[slot] = make_float4(0.f, 0.f, 0.f, -100.f); // a primary ray that leaves the scene (:37-41 returns the initial value)
            }
            if (alive)
                alive = radiance_bounce(sc, rp.mode, rp.max_segments, rp.seed, h, rDir, depth, segs, tp, thr_dirty, &nStart, &nDir, [&](uint64_t* pixel, uint32_t* sample) {
                    if (radiance_query) { *pixel = slot; *sample = 0; return; }
                    // slot = k * n_local + lp with a 32-bit slot: one 32-bit divide (a 64-bit one was 4 % of this kernel's instructions)
                    uint32_t k = 0, lp = slot;
                    if ((rp.n_local >> 32) == 0) { k = slot / (uint32_t)rp.n_local; lp = slot - k * (uint32_t)rp.n_local; }
                    map_pixel(pm, rp.lp0 + lp, pixel);
                    *sample = rp.s0 + k;
                });
            if (thr_dirty) thr[slot] = tp;
            mt_out = (depth & 0x7FFFu) | ((thr_written || thr_dirty) ? kThrWritten : 0u) | (segs << 16);
            if (!alive && radiance_query) meta[slot] = mt_out; // vmn_radiance reports the segments of each finished path
        }
        const uint32_t ballot = __ballot_sync(0xffffffffu, alive);
        if (ballot) {
            const int leader = __ffs(ballot) - 1;
            uint32_t qb = 0;
            if ((int)lane == leader) qb = atomicAdd(qcount_out, (uint32_t)__popc(ballot));
            qb = __shfl_sync(0xffffffffu, qb, leader);
            if (alive) {
                const uint32_t q = qb + __popc(ballot & ((1u << lane) - 1u));
                VMN_CHECK(sc.err, q < n); // survivors never outnumber the rays that came in
                st_stream(&rays_out[2ull * q], make_float4(nStart.x, nStart.y, nStart.z, __uint_as_float(slot)));
                st_stream(&rays_out[2ull * q + 1], make_float4(nDir.x, nDir.y, nDir.z, __uint_as_float(mt_out)));
            }
        }
    }
}

// ---- accumulate / resolve --------------------------------------------------------------------------
// FIXED: film[pixel] += the round's `batch` samples, in sample order.
// COMPAT (batch = 1): additionally the early-out state machine of pathtracer.cpp:242-312.
__global__ void __launch_bounds__(256) k_accumulate(PixelMap pm, RoundParams rp, uint32_t spp, const float4* __restrict__ radiance,
                                                    float4* __restrict__ film, uint32_t* __restrict__ cstate,
                                                    uint32_t* __restrict__ live_count) {
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= rp.n_local) return;
    uint64_t pixel;
    if (!map_pixel(pm, rp.lp0 + lp, &pixel)) return;
    float4 f = film[pixel];
    if (rp.mode == VMN_MODE_FIXED) {
        for (uint32_t k = 0; k < rp.batch; ++k) {
            const float4 s = radiance[(uint64_t)k * rp.n_local + lp];
            f.x += s.x; f.y += s.y; f.z += s.z; f.w += 1.f;
        }
        film[pixel] = f;
        return;
    }
    uint32_t cs = cstate[lp];
    if (cs & 0x20000000u) return; // done earlier
    const float4 s = radiance[lp];
    f.x += s.x; f.y += s.y; f.z += s.z; f.w += 1.f;
    film[pixel] = f;
    uint32_t st = cs >> 30, sz = cs & 0x1FFFFFFFu;
    const uint32_t n = (uint32_t)f.w;
    bool brk = false;
    if ((double)n > sqrt((double)spp)) { // :290
        const float fn = (float)n, fn1 = (float)(n + 1);
        V3 a(f.x / fn, f.y / fn, f.z / fn), b((f.x + s.x) / fn1, (f.y + s.y) / fn1, (f.z + s.z) / fn1);
        if (fabsf(length(a - b)) < 0.00001f) brk = true;
    }
    ++sz;
    bool done = false;
    if (brk || sz >= rp.per) { sz = 0; if (st == 3) done = true; else ++st; }
    cstate[lp] = (st << 30) | (done ? 0x20000000u : 0u) | sz;
    if (!done) atomicAdd(live_count, 1u);
}

// std::max(std::min(x, 1.f), 0.f) as pathtracer.cpp:319-321 and integrators.cpp:177-179 evaluate it: std::min(a,b) = (b < a) ? b : a, std::max(a,b) = (a < b) ? b : a,
// so a NaN mean stays NaN (fminf/fmaxf would drop it)
__device__ __forceinline__ float clamp01_std(float x) {
    const float m = (1.f < x) ? 1.f : x;
    return (m < 0.f) ? 0.f : m;
}

// pathtracer.cpp:318-323 + camera.cpp:106-113
__global__ void k_resolve(const float4* __restrict__ film, float* __restrict__ out, uint64_t npix) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 f = film[p];
    float* o = out + 5 * p;
    if (f.w == 0.f) { o[0] = o[1] = o[2] = o[3] = o[4] = 0.f; return; } // pixel not rendered by anyone
    o[0] = clamp01_std(f.x / f.w);
    o[1] = clamp01_std(f.y / f.w);
    o[2] = clamp01_std(f.z / f.w);
    o[3] = 1.f;
    o[4] = f.w;
}

// integrators.cpp:176-183 + camera.cpp:106-113: BRUTEFORCE film (sum rgb, n) + (sum alpha, last primary distance) -> RGBAZ
__global__ void k_resolve_bf(const float4* __restrict__ film, const float2* __restrict__ aux, float* __restrict__ out, uint64_t npix) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 f = film[p];
    const float2 a = aux[p];
    float* o = out + 5 * p;
    if (f.w == 0.f) { o[0] = o[1] = o[2] = o[3] = o[4] = 0.f; return; } // pixel not rendered by anyone
    o[0] = clamp01_std(f.x / f.w);
    o[1] = clamp01_std(f.y / f.w);
    o[2] = clamp01_std(f.z / f.w);
    o[3] = a.x / f.w;
    o[4] = a.y;
}

// Camera::saveFrame's conversion loop (camera.cpp:138-175, RGBAZ): floor(x * 255) per channel of the resolved pixel; depth = the
// resolved pixel's fifth float (samples for PathTracer films, distance for BRUTEFORCE films)
__global__ void k_pack_rgba8(const float4* __restrict__ film, const float2* __restrict__ aux, uchar4* __restrict__ rgba, float* __restrict__ depth,
                             uint64_t npix) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npix) return;
    const float4 f = film[p];
    float r = 0.f, g = 0.f, b = 0.f, a = 0.f, z = f.w;
    if (f.w != 0.f) {
        if (aux) {
            const float2 x = aux[p];
            r = clamp01_std(f.x / f.w); g = clamp01_std(f.y / f.w); b = clamp01_std(f.z / f.w); a = x.x / f.w; z = x.y;
        } else {
            r = clamp01_std(f.x / f.w); g = clamp01_std(f.y / f.w); b = clamp01_std(f.z / f.w); a = 1.f;
        }
    }
    rgba[p] = make_uchar4((unsigned char)floorf(r * 255), (unsigned char)floorf(g * 255), (unsigned char)floorf(b * 255), (unsigned char)floorf(a * 255));
    if (depth) depth[p] = z;
}

__global__ void k_camera_rays(CamDev cam, const uint64_t* __restrict__ pixel, const uint32_t* __restrict__ sx, const uint32_t* __restrict__ sy,
                              const float* __restrict__ jx, const float* __restrict__ jy, uint64_t n, float4* __restrict__ out) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    V3 o, d;
    camera_ray(cam, pixel[i], sx[i], sy[i], jx[i], jy[i], &o, &d);
    out[2 * i] = make_float4(o.x, o.y, o.z, 0.f);
    out[2 * i + 1] = make_float4(d.x, d.y, d.z, 0.f);
}

// vmn_radiance: turn caller rays into queue entries with slot = index
__global__ void k_radiance_init(const float4* __restrict__ in, uint64_t n, float4* __restrict__ rays, float4* __restrict__ radiance,
                                uint32_t* __restrict__ meta) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 a = in[2 * i], b = in[2 * i + 1];
    a.w = __uint_as_float((uint32_t)i); b.w = 0.f;
    rays[2 * i] = a; rays[2 * i + 1] = b;
    meta[i] = 0u;
}

__global__ void k_radiance_out(const float4* __restrict__ radiance, const uint32_t* __restrict__ meta, uint64_t n, float4* __restrict__ out,
                               uint32_t* __restrict__ rays_cast) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = radiance[i];
    rays_cast[i] = meta[i] >> 16;
}

// ---- host orchestration ----------------------------------------------------------------------------
static inline uint32_t cdiv64(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

static CamDev cam_dev(const vmn_ctx* c) {
    CamDev cd;
    memcpy(cd.m, c->cam_m, sizeof(cd.m));
    cd.px = c->cam.position[0]; cd.py = c->cam.position[1]; cd.pz = c->cam.position[2];
    cd.dist = c->cam.dist_to_film; cd.sx = c->cam.sensor_x; cd.sy = c->cam.sensor_y;
    cd.width = c->cam.width; cd.height = c->cam.height; cd.spp = c->cam.spp;
    return cd;
}

// per-stage device timing, summed after the stream drains. Consecutive stages share ONE event (the end of a stage is the beginning of the
// next): an event record is a stream operation of its own, and a frame is a chain of short launches — at two events per kernel plus a
// 4-byte memset per bounce every launch carried about 21 us of stream overhead (profiles/r02j_maxpaths.txt). flush() closes the open stage
// before work that belongs to no stage (memsets, copies) is enqueued.
struct StageClock {
    vmn_ctx* c; PathPool* p;
    struct Span { size_t e0, e1; int stage; };
    std::vector<Span> spans;
    bool open = false;
    cudaEvent_t get() {
        if (p->ev_used == p->events.size()) { cudaEvent_t e; cudaEventCreate(&e); p->events.push_back(e); }
        return p->events[p->ev_used++];
    }
    void begin(int stage) {
        const size_t e = p->ev_used;
        cudaEventRecord(get(), c->stream);
        if (open) spans.back().e1 = e;
        spans.push_back({e, e, stage});
        open = false;
    }
    void end() { open = true; } // closed by the next begin() or by flush()
    void flush() {
        if (!open) return;
        spans.back().e1 = p->ev_used;
        cudaEventRecord(get(), c->stream);
        open = false;
    }
    void collect(double ms[4]) {
        flush();
        for (auto& s : spans) {
            float t = 0.f;
            if (s.e1 != s.e0 && cudaEventElapsedTime(&t, p->events[s.e0], p->events[s.e1]) == cudaSuccess) ms[s.stage] += t;
        }
    }
};
enum { ST_RAYGEN = 0, ST_EXTEND = 1, ST_SHADE = 2, ST_RESOLVE = 3 };

static int extend_grid(vmn_ctx* c, bool count) {
    int per_sm = 0;
    // experiment knob: VMN_CARVEOUT=<percent of the unified L1/shared array given to shared memory> for k_extend (default: driver's choice)
    static const int carve = [] { const char* e = getenv("VMN_CARVEOUT"); return e ? atoi(e) : -1; }();
    if (carve >= 0) {
        cudaFuncSetAttribute(k_extend<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        cudaFuncSetAttribute(k_extend<true>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
    }
    cudaError_t e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend<true>, kWfBlock, 0)
                          : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_extend<false>, kWfBlock, 0);
    if (e != cudaSuccess || per_sm < 1) per_sm = 4;
    return c->sm_count * per_sm;
}

// ctl layout: [0,32) queue counts (qc[0], qc[1], qc[4] = live pixels); [64, 64+8*kMaxClaims) claim counters — both zeroed every
// round; [kOffFrame, +32) frame-level counters (nodes, tris, rays), zeroed once per call.
constexpr uint32_t kMaxClaims = 480;
constexpr size_t kOffFrame = 4000;

// k_extend over queue `cur`; claim counter index `it`
static vmn_status launch_extend(vmn_ctx* c, PathPool* p, int cur, uint32_t it, StageClock* clk, int grid_extend) {
    uint32_t* qc = p->ctl.as<uint32_t>();
    unsigned long long* claims = reinterpret_cast<unsigned long long*>(p->ctl.as<char>() + 64);
    TravCounters* gcnt = reinterpret_cast<TravCounters*>(p->ctl.as<char>() + kOffFrame);
    unsigned long long* total = reinterpret_cast<unsigned long long*>(p->ctl.as<char>() + kOffFrame + 16);
    const bool count = (c->flags & VMN_CFG_COUNTERS) != 0;
    if (clk) clk->begin(ST_EXTEND);
    if (count)
        k_extend<true><<<grid_extend, kWfBlock, 0, c->stream>>>(c->scene, p->ray[cur].as<float4>(), qc + cur, p->hit.as<float2>(), claims + it, gcnt, total, refill_min(c), qc + (cur ^ 1));
    else
        k_extend<false><<<grid_extend, kWfBlock, 0, c->stream>>>(c->scene, p->ray[cur].as<float4>(), qc + cur, p->hit.as<float2>(), claims + it, gcnt, total, refill_min(c), qc + (cur ^ 1));
    if (clk) clk->end();
    ++c->stats.kernel_launches;
    return VMN_OK;
}

// one extend + shade iteration over queue `cur`; claim counter index `it`
static vmn_status bounce(vmn_ctx* c, PathPool* p, const PixelMap& pm, const RoundParams& rp, int cur, uint32_t it, int radiance_query,
                         StageClock* clk, int grid_extend) {
    uint32_t* qc = p->ctl.as<uint32_t>();
    vmn_status s = launch_extend(c, p, cur, it, clk, grid_extend); // (k_extend also zeroes the count of the queue the shade stage appends to)
    if (s != VMN_OK) return s;
    if (clk) clk->begin(ST_SHADE);
    k_shade<<<c->sm_count * VMN_SHADE_MINB, kWfBlock, 0, c->stream>>>(c->scene, c->spheres.as<SphereDev>(), pm, rp, p->ray[cur].as<float4>(), p->hit.as<float2>(),
                                                         qc + cur, p->ray[cur ^ 1].as<float4>(), qc + (cur ^ 1), p->radiance.as<float4>(),
                                                         p->thr.as<float4>(), p->meta.as<uint32_t>(), radiance_query);
    if (clk) clk->end();
    VMN_CUDA(c, cudaGetLastError());
    ++c->stats.kernel_launches;
    return VMN_OK;
}


vmn_status launch_radiance(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, uint64_t seed, int mode, uint32_t max_segments, float* d_out,
                           uint32_t* d_rays_cast) {
    c->stats.kernel_launches = 0; c->stats.rays_traced = 0;
    if (n == 0) return VMN_OK;
    if (n >= 0xFFFFFFFFull) return fail(c, VMN_ERR_INVALID, "too many rays");
    vmn_status s = pool_ensure(c, n);
    if (s != VMN_OK) return s;
    PathPool* p = c->pool;
    uint32_t* qc = p->ctl.as<uint32_t>();
    VMN_CUDA(c, cudaMemsetAsync(p->ctl.p, 0, 4096, c->stream));
    k_radiance_init<<<cdiv64(n, 256), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(d_rays), n, p->ray[0].as<float4>(), p->radiance.as<float4>(),
                                                           p->meta.as<uint32_t>());
    const uint32_t n32 = (uint32_t)n;
    VMN_CUDA(c, cudaMemcpyAsync(qc, &n32, 4, cudaMemcpyHostToDevice, c->stream));
    PixelMap pm{}; pm.width = 0xFFFFFFFFu; pm.height = 1;
    RoundParams rp{}; rp.seed = seed; rp.n_local = n; rp.mode = mode; rp.max_segments = max_segments; rp.batch = 1;
    const int grid = extend_grid(c, (c->flags & VMN_CFG_COUNTERS) != 0);
    int cur = 0;
    uint64_t live = n, rays = 0;
    for (uint32_t it = 0; live; ++it) {
        if (it && it % kMaxClaims == 0) VMN_CUDA(c, cudaMemsetAsync(p->ctl.as<char>() + 64, 0, 8 * kMaxClaims, c->stream));
        rays += live;
        if ((s = bounce(c, p, pm, rp, cur, it % kMaxClaims, 1, nullptr, grid)) != VMN_OK) return s;
        cur ^= 1;
        VMN_CUDA(c, cudaMemcpyAsync(p->h_count, qc + cur, 4, cudaMemcpyDeviceToHost, c->stream));
        VMN_CUDA(c, cudaStreamSynchronize(c->stream));
        live = p->h_count[0];
    }
    k_radiance_out<<<cdiv64(n, 256), 256, 0, c->stream>>>(p->radiance.as<float4>(), p->meta.as<uint32_t>(), n, reinterpret_cast<float4*>(d_out), d_rays_cast);
    VMN_CUDA(c, cudaGetLastError());
    c->stats.rays_traced = rays; c->stats.paths_traced = n; c->stats.kernel_launches += 2;
    return VMN_OK;
}

vmn_status launch_camera_rays(vmn_ctx* c, const uint64_t* d_pixel, const uint32_t* d_sx, const uint32_t* d_sy, const float* d_jx,
                              const float* d_jy, uint64_t n, vmn_ray* d_out) {
    if (n == 0) return VMN_OK;
    k_camera_rays<<<cdiv64(n, 256), 256, 0, c->stream>>>(cam_dev(c), d_pixel, d_sx, d_sy, d_jx, d_jy, n, reinterpret_cast<float4*>(d_out));
    VMN_CUDA(c, cudaGetLastError());
    return VMN_OK;
}

vmn_status film_resolve(vmn_ctx* c, const float* d_sum, const float* d_aux, float* d_rgbaz, uint64_t npix) {
    if (npix == 0) return VMN_OK;
    if (d_aux)
        k_resolve_bf<<<cdiv64(npix, 256), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(d_sum), reinterpret_cast<const float2*>(d_aux), d_rgbaz, npix);
    else
        k_resolve<<<cdiv64(npix, 256), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(d_sum), d_rgbaz, npix);
    VMN_CUDA(c, cudaGetLastError());
    return VMN_OK;
}

vmn_status film_pack_rgba8(vmn_ctx* c, const float* d_sum, const float* d_aux, uint8_t* d_rgba, float* d_depth, uint64_t npix) {
    if (npix == 0) return VMN_OK;
    k_pack_rgba8<<<cdiv64(npix, 256), 256, 0, c->stream>>>(reinterpret_cast<const float4*>(d_sum), reinterpret_cast<const float2*>(d_aux),
                                                           reinterpret_cast<uchar4*>(d_rgba), d_depth, npix);
    VMN_CUDA(c, cudaGetLastError());
    return VMN_OK;
}

// this context's share of the frame: interleaved tiles or a contiguous pixel range
vmn_status make_pixel_map(vmn_ctx* c, const vmn_render_params* prm, PixelMap* out, uint64_t* n_local_total) {
    const uint64_t npix = c->film_pixels;
    PixelMap pm{};
    pm.width = c->cam.width; pm.height = c->cam.height;
    if (prm->tile_size) {
        if (!prm->tile_world || prm->tile_rank >= prm->tile_world) return fail(c, VMN_ERR_INVALID, "bad tile rank/world");
        pm.tile = prm->tile_size; pm.rank = prm->tile_rank; pm.world = prm->tile_world;
        pm.tiles_x = (pm.width + pm.tile - 1) / pm.tile;
        pm.tile_log2p1 = tile_log2p1(pm.tile);
        const uint64_t tiles = (uint64_t)pm.tiles_x * ((pm.height + pm.tile - 1) / pm.tile);
        const uint64_t mine = tiles > pm.rank ? (tiles - pm.rank + pm.world - 1) / pm.world : 0;
        *n_local_total = mine * pm.tile * pm.tile;
    } else {
        const uint64_t pe = prm->pixel_end ? prm->pixel_end : npix;
        if (prm->pixel_begin > pe || pe > npix) return fail(c, VMN_ERR_INVALID, "bad pixel range");
        pm.begin = prm->pixel_begin;
        *n_local_total = pe - prm->pixel_begin;
    }
    *out = pm;
    return VMN_OK;
}

static bool mega_wanted(const vmn_ctx* c, uint64_t n_local);
static vmn_status render_mega(vmn_ctx* c, const vmn_render_params* prm, const PixelMap& pm, uint64_t n_local);

vmn_status render_frame(vmn_ctx* c, const vmn_render_params* prm) {
    const uint64_t npix = c->film_pixels;
    const uint32_t spp = c->cam.spp;
    PixelMap pm{};
    uint64_t n_local_total;
    vmn_status s = make_pixel_map(c, prm, &pm, &n_local_total);
    if (s != VMN_OK) return s;
    const int mode = prm->mode;
    uint32_t s_begin = 0, s_end = spp;
    if (mode == VMN_MODE_FIXED) {
        s_begin = prm->sample_begin; s_end = prm->sample_end ? prm->sample_end : spp;
        if (s_begin > s_end) return fail(c, VMN_ERR_INVALID, "bad sample range");
    }
    c->stats.kernel_launches = 0; c->stats.rays_traced = 0; c->stats.paths_traced = 0;
    c->stats.raygen_ms = c->stats.extend_ms = c->stats.shade_ms = c->stats.resolve_ms = c->stats.total_ms = 0;
    if ((prm->flags & VMN_RENDER_ACCUMULATE) && mode != VMN_MODE_FIXED) return fail(c, VMN_ERR_INVALID, "VMN_RENDER_ACCUMULATE needs VMN_MODE_FIXED");
    if (!(prm->flags & VMN_RENDER_ACCUMULATE)) VMN_CUDA(c, cudaMemsetAsync(c->film.p, 0, 16 * npix, c->stream));
    if (n_local_total == 0 || s_begin == s_end) return VMN_OK;
    if (mode == VMN_MODE_COMPAT && mega_wanted(c, n_local_total)) {
        if (spp / 4 == 0) return VMN_OK; // spp < 4: the reference takes no samples at all
        return render_mega(c, prm, pm, n_local_total);
    }

    // queue slots and counts are 32-bit: never more than 2^32 - 16 paths in flight, whatever max_paths says
    const uint64_t cap = c->max_paths < 0xFFFFFFF0ull ? c->max_paths : 0xFFFFFFF0ull;
    const uint64_t chunk = n_local_total < cap ? n_local_total : cap;
    uint32_t batch = 1;
    if (mode == VMN_MODE_FIXED) {
        batch = (uint32_t)(cap / chunk);
        if (batch > s_end - s_begin) batch = s_end - s_begin;
        if (batch < 1) batch = 1;
    }
    if ((s = pool_ensure(c, chunk * batch)) != VMN_OK) return s;
    PathPool* p = c->pool;
    if (mode == VMN_MODE_COMPAT && (s = dev_reserve(c, p->cstate, 4 * chunk)) != VMN_OK) return s;
    p->ev_used = 0;
    StageClock clk{c, p, {}};
    uint32_t* qc = p->ctl.as<uint32_t>();
    const CamDev cd = cam_dev(c);
    const uint32_t max_seg = prm->max_segments ? prm->max_segments : 5;
    const bool count = (c->flags & VMN_CFG_COUNTERS) != 0;
    const int grid = extend_grid(c, count);
    clk.flush();
    VMN_CUDA(c, cudaMemsetAsync(p->ctl.as<char>() + kOffFrame, 0, 32, c->stream));

    // FIXED rounds over tiles: the per-tile prefix of live pixels that lets k_raygen place every path without an atomic (single chunk only;
    // made once per (image, tile, rank, world) and kept in the pool)
    const uint32_t* tile_prefix = nullptr;
#if VMN_RAYGEN_DIRECT
    if (mode == VMN_MODE_FIXED && pm.tile && chunk == n_local_total && n_local_total < 0xFFFFFFFFull) {
        const uint32_t key[5] = {pm.width, pm.height, pm.tile, pm.rank, pm.world};
        const uint32_t tt = pm.tile * pm.tile, ntiles = (uint32_t)(n_local_total / tt);
        if (memcmp(key, p->tile_key, sizeof(key)) != 0 || !p->tile_prefix.p) {
            std::vector<uint32_t> pre(ntiles);
            uint32_t run = 0;
            for (uint32_t j = 0; j < ntiles; ++j) {
                const uint64_t t = (uint64_t)j * pm.world + pm.rank;
                const uint32_t ty = (uint32_t)(t / pm.tiles_x), tx = (uint32_t)(t % pm.tiles_x);
                const uint32_t lw = std::min(pm.tile, pm.width - tx * pm.tile), lh = std::min(pm.tile, pm.height - ty * pm.tile);
                pre[j] = run;
                run += lw * lh;
            }
            if ((s = dev_reserve(c, p->tile_prefix, 4ull * ntiles)) != VMN_OK) return s;
            VMN_CUDA(c, cudaMemcpyAsync(p->tile_prefix.p, pre.data(), 4ull * ntiles, cudaMemcpyHostToDevice, c->stream)); // pageable: staged before it returns
            p->tile_live = run;
            memcpy(p->tile_key, key, sizeof(key));
        }
        tile_prefix = p->tile_prefix.as<uint32_t>();
    }
#endif

    VMN_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
    for (uint64_t lp0 = 0; lp0 < n_local_total; lp0 += chunk) {
        RoundParams rp{};
        rp.seed = prm->seed; rp.lp0 = lp0; rp.n_local = (n_local_total - lp0 < chunk) ? n_local_total - lp0 : chunk;
        rp.per = spp / 4; rp.mode = mode; rp.max_segments = max_seg;
        if (mode == VMN_MODE_COMPAT) {
            if (rp.per == 0) continue; // spp < 4: the reference takes no samples at all
            VMN_CUDA(c, cudaMemsetAsync(p->cstate.p, 0, 4 * rp.n_local, c->stream));
            uint64_t live_px = rp.n_local;
            for (uint32_t round = 0; live_px; ++round) {
                rp.s0 = round; rp.batch = 1;
                clk.flush();
                VMN_CUDA(c, cudaMemsetAsync(p->ctl.p, 0, 64 + 8 * kMaxClaims, c->stream));
                clk.begin(ST_RAYGEN);
                k_raygen<<<cdiv64(rp.n_local, kWfBlock), kWfBlock, 0, c->stream>>>(cd, pm, rp, p->cstate.as<uint32_t>(), p->ray[0].as<float4>(), qc,
                                                                                   p->radiance.as<float4>(), p->meta.as<uint32_t>(), nullptr, 0u);
                clk.end();
                ++c->stats.kernel_launches;
                clk.flush();
                VMN_CUDA(c, cudaMemcpyAsync(p->h_count + 2, qc, 4, cudaMemcpyDeviceToHost, c->stream)); // samples started; read after the next sync
                int cur = 0;
                uint64_t live = 1;
                for (uint32_t it = 0; live;) {
                    // four bounces per host round trip: the queue counts live on the device and a kernel that finds its queue empty returns
                    // at once, so the only thing the host needs is to know when to stop
                    for (int k = 0; k < 4; ++k, ++it) {
                        if (it && it % kMaxClaims == 0) { clk.flush(); VMN_CUDA(c, cudaMemsetAsync(p->ctl.as<char>() + 64, 0, 8 * kMaxClaims, c->stream)); }
                        if ((s = bounce(c, p, pm, rp, cur, it % kMaxClaims, 0, &clk, grid)) != VMN_OK) return s;
                        cur ^= 1;
                    }
                    clk.flush();
                    VMN_CUDA(c, cudaMemcpyAsync(p->h_count, qc + cur, 4, cudaMemcpyDeviceToHost, c->stream));
                    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
                    live = p->h_count[0];
                }
                c->stats.paths_traced += p->h_count[2];
                clk.flush();
                VMN_CUDA(c, cudaMemsetAsync(qc + 4, 0, 4, c->stream));
                clk.begin(ST_RESOLVE);
                k_accumulate<<<cdiv64(rp.n_local, 256), 256, 0, c->stream>>>(pm, rp, spp, p->radiance.as<float4>(),
                                                                             c->film.as<float4>(), p->cstate.as<uint32_t>(), qc + 4);
                clk.end();
                ++c->stats.kernel_launches;
                clk.flush();
                VMN_CUDA(c, cudaMemcpyAsync(p->h_count, qc + 4, 4, cudaMemcpyDeviceToHost, c->stream));
                VMN_CUDA(c, cudaStreamSynchronize(c->stream));
                live_px = p->h_count[0];
            }
        } else {
            for (uint32_t s0 = s_begin; s0 < s_end; s0 += batch) {
                rp.s0 = s0; rp.batch = (s_end - s0 < batch) ? s_end - s0 : batch;
                clk.flush();
                VMN_CUDA(c, cudaMemsetAsync(p->ctl.p, 0, 64 + 8 * kMaxClaims, c->stream));
                clk.begin(ST_RAYGEN);
                k_raygen<<<cdiv64(rp.n_local * rp.batch, kWfBlock), kWfBlock, 0, c->stream>>>(cd, pm, rp, nullptr, p->ray[0].as<float4>(), qc,
                                                                                              p->radiance.as<float4>(),
                                                                                              p->meta.as<uint32_t>(), tile_prefix, p->tile_live);
                clk.end();
                ++c->stats.kernel_launches;
                int cur = 0;
                for (uint32_t it = 0; it < max_seg; ++it) {
                    // a claim counter serves one launch: recycle the block every kMaxClaims bounces (max_segments may exceed it)
                    if (it && it % kMaxClaims == 0) { clk.flush(); VMN_CUDA(c, cudaMemsetAsync(p->ctl.as<char>() + 64, 0, 8 * kMaxClaims, c->stream)); }
                    if ((s = bounce(c, p, pm, rp, cur, it % kMaxClaims, 0, &clk, grid)) != VMN_OK) return s;
                    cur ^= 1;
                }
                clk.begin(ST_RESOLVE);
                k_accumulate<<<cdiv64(rp.n_local, 256), 256, 0, c->stream>>>(pm, rp, spp, p->radiance.as<float4>(),
                                                                             c->film.as<float4>(), nullptr, nullptr);
                clk.end();
                ++c->stats.kernel_launches;
                c->stats.paths_traced += rp.n_local * rp.batch;
            }
        }
    }
    VMN_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    VMN_CUDA(c, cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]);
    c->stats.total_ms = ms;
    double st[4] = {0, 0, 0, 0};
    clk.collect(st);
    c->stats.raygen_ms = st[ST_RAYGEN]; c->stats.extend_ms = st[ST_EXTEND]; c->stats.shade_ms = st[ST_SHADE]; c->stats.resolve_ms = st[ST_RESOLVE];
    unsigned long long h[3] = {0, 0, 0};
    VMN_CUDA(c, cudaMemcpy(h, p->ctl.as<char>() + kOffFrame, 24, cudaMemcpyDeviceToHost));
    c->stats.nodes_visited = h[0]; c->stats.tris_tested = h[1]; c->stats.rays_traced = h[2];
    return VMN_OK;
}

// ---- BruteForceTracer::Render (integrators.cpp:9-186) ---------------------------------------------------------------------------
// The RenderEngine's default integrator (renderEngine.cpp:51) on the same wavefront: every live pixel takes one sample per round
//   k_bf_raygen  jittered primary ray (integrators.cpp:65-81)                      -> queue 0, slot = local pixel
//   k_extend     closest hit
//   k_bf_shade1  RayCast finish, point-light N.L, normal-map perturbation (:83-106); a miss finishes the sample here;
//                a hit emits the mirror-direction probe (:119-121) and, with VMN_RENDER_BF_SHADOW, the shadow ray -> queue 1
//   k_extend     closest hit of the probes (and shadow rays: the any-hit helper RayCastCollision is a closest hit with t > 1e-3)
//   k_bf_shade2  shadow (:108-110), probe miss lightening (:133-137), colour (:141-156), running mean and early-out (:158-173);
//                a pixel that stops writes its film entry (:176-183)
// RNG: Philox stream (seed, pixel, sample), draws 0 and 1 (the reference races one std::mt19937(time(0)) across threads, :30).
struct BfParams {
    uint64_t seed, n_local, lp0;
    uint32_t sample, spp, flags;
};

// integrators.cpp:65-81
__device__ __forceinline__ void bf_camera_ray(const CamDev& c, uint64_t p, float jx, float jy, V3* o, V3* d) {
    uint64_t py = p / c.width, px = p - py * c.width;
    if ((p >> 32) == 0) { const uint32_t y32 = (uint32_t)p / c.width; py = y32; px = (uint32_t)p - y32 * c.width; }
    float hx = (float)((((double)(float(px) + jx) - 0.25) / (double)c.width) * 2 - 1);
    float hy = (float)((((double)(float(py) + jy) - 0.25) / (double)c.height) * 2 - 1);
    float bx = (float)((double)(hx * c.sx) * 0.5), by = (float)((double)(hy * c.sy) * 0.5);
    float g0 = bx, g1 = -by, g2 = -c.dist, g3 = 1.0f;
    float rd[4];
#pragma unroll
    for (int r = 0; r < 4; ++r) rd[r] = (c.m[r] * g0 + c.m[4 + r] * g1) + (c.m[8 + r] * g2 + c.m[12 + r] * g3);
    V3 dir(rd[0] / rd[3], rd[1] / rd[3], rd[2] / rd[3]);
    *d = normalize(dir);
    *o = V3(c.px, c.py, c.pz);
}

__global__ void __launch_bounds__(kWfBlock) k_bf_raygen(CamDev cam, PixelMap pm, BfParams bp, const uint2* __restrict__ meta, float4* __restrict__ rays,
                                                        uint32_t* __restrict__ qcount, unsigned long long* __restrict__ samples_started) {
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool live = lp < bp.n_local;
    uint64_t pixel = 0;
    if (live) live = map_pixel(pm, bp.lp0 + lp, &pixel) && (meta[lp].y & 1u) == 0u;
    V3 o(0, 0, 0), d(0, 0, 0);
    if (live) {
        Rng rng(bp.seed, pixel, bp.sample, 0);
        const float jx = rng.jitter(), jy = rng.jitter();
        bf_camera_ray(cam, pixel, jx, jy, &o, &d);
    }
    const uint32_t ballot = __ballot_sync(0xffffffffu, live);
    if (ballot == 0u) return;
    const uint32_t lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (uint32_t)(__ffs(ballot) - 1)) { base = atomicAdd(qcount, (uint32_t)__popc(ballot)); atomicAdd(samples_started, (unsigned long long)__popc(ballot)); }
    base = __shfl_sync(0xffffffffu, base, __ffs(ballot) - 1);
    if (live) {
        const uint32_t q = base + __popc(ballot & ((1u << lane) - 1u));
        rays[2ull * q] = make_float4(o.x, o.y, o.z, __uint_as_float((uint32_t)lp));
        rays[2ull * q + 1] = make_float4(d.x, d.y, d.z, 0.f);
    }
}

// integrators.cpp:170: `abs` resolves to ::abs(int) in that translation unit, so the float sum is truncated to int first (x86
// cvttss2si: INT_MIN for NaN / out of range; abs(INT_MIN) stays INT_MIN)
__device__ __forceinline__ float abs_via_int(float x) {
    const int i = (fabsf(x) < 2147483648.f) ? __float2int_rz(x) : (int)0x80000000;
    const uint32_t a = i < 0 ? 0u - (uint32_t)i : (uint32_t)i;
    return (float)(int)a;
}

// the end of one sample of pixel `lp` (integrators.cpp:141-173) and, when the pixel stops, its film entry (:176-183)
// the end of one sample (integrators.cpp:141-173) on register state: adds the sample, runs the early-out; true = the pixel stops
__device__ __forceinline__ bool bf_sample_end(const SceneDev& sc, uint32_t spp, bool hit, float vNDL, float u, float v, float4& accum, float4& last,
                                              uint32_t n /* nTotalSamples after ++ (:63) */) {
    if (hit) {
        float4 cur;
        if (sc.tex1) { // :141-147
            float c4[4];
            texture_sample(sc.tex1, sc.tex1_w, sc.tex1_h, sc.tex1_c, u, v, c4);
            cur = make_float4(c4[0] * vNDL, c4[1] * vNDL, c4[2] * vNDL, 1.0f);
        } else {
            cur = make_float4(0.890196078f * vNDL, 0.258823529f * vNDL, 0.203921569f * vNDL, 1.0f); // :150-155
        }
        accum.x += cur.x; accum.y += cur.y; accum.z += cur.z; accum.w += cur.w; // :158
    }
    const float fn = (float)n;
    bool done = false;
    if (n > 2u) { // :167-172
        last.x -= accum.x / fn; last.y -= accum.y / fn; last.z -= accum.z / fn; last.w -= accum.w / fn;
        if (abs_via_int(last.x + last.y + last.z + last.w) < 0.001f) done = true;
    }
    if (!done) last = make_float4(accum.x / fn, accum.y / fn, accum.z / fn, accum.w / fn); // :173
    if (n >= spp) done = true;
    return done;
}

// the end of one sample of pixel `lp` and, when the pixel stops, its film entry (:176-183)
__device__ __forceinline__ void bf_finish_sample(const SceneDev& sc, const PixelMap& pm, const BfParams& bp, uint32_t lp, bool hit, float vNDL, float u,
                                                 float v, float dist, float4* __restrict__ accum_a, float4* __restrict__ last_a,
                                                 uint2* __restrict__ meta_a, float4* __restrict__ film, float2* __restrict__ aux) {
    float4 accum = accum_a[lp], last = last_a[lp];
    const uint32_t n = meta_a[lp].x + 1u; // ++nTotalSamples (:63)
    const bool done = bf_sample_end(sc, bp.spp, hit, vNDL, u, v, accum, last, n);
    accum_a[lp] = accum; last_a[lp] = last; meta_a[lp] = make_uint2(n, done ? 1u : 0u);
    if (done) {
        uint64_t pixel;
        if (map_pixel(pm, bp.lp0 + lp, &pixel)) {
            film[pixel] = make_float4(accum.x, accum.y, accum.z, (float)n);
            aux[pixel] = make_float2(accum.w, dist);
        }
    }
}

// integrators.cpp:83-119 for a primary hit: light direction, N.L (with the normal map of texture slot 0), mirror-probe direction
__device__ __forceinline__ void bf_primary_shade(const SceneDev& sc, const RayCastOut& h, float* vNDL_out, V3* sdir, V3* ldir_out) {
    const V3 lightLocation(500.f, 1100.f, 2000.f); // :16
    const V3 ldir = normalize(lightLocation - h.P); // :83-84
    V3 hn = h.N;
    float vNDL = dot(ldir, normalize(hn)); // :88
    if (sc.tex) { // :98-106
        float n4[4];
        texture_sample(sc, h.uv0, h.uv1, n4);
        hn = hn + V3(n4[0], n4[1], n4[2]);
        vNDL = dot(ldir, normalize(hn));
    }
    const V3 nl = -ldir;
    *sdir = nl - (hn * 2.f) * dot(hn, nl); // :119
    *vNDL_out = vNDL; *ldir_out = ldir;
}

__global__ void __launch_bounds__(kWfBlock) k_bf_shade1(SceneDev sc, const SphereDev* __restrict__ sph, PixelMap pm, BfParams bp,
                                                        const float4* __restrict__ rays, const float2* __restrict__ hits,
                                                        const uint32_t* __restrict__ qcount_in, float4* __restrict__ rays_out,
                                                        uint32_t* __restrict__ qcount_out, float4* __restrict__ accum, float4* __restrict__ last,
                                                        uint2* __restrict__ meta, float4* __restrict__ tmp, float4* __restrict__ film,
                                                        float2* __restrict__ aux) {
    const uint32_t n = *qcount_in;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t stride = (bp.flags & VMN_RENDER_BF_SHADOW) ? 2u : 1u;
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) & ~31ull; base < n; base += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = base + lane;
        bool emit = false;
        V3 P(0, 0, 0), sdir(0, 0, 0), ldir(0, 0, 0);
        uint32_t slot = 0, mt_out = 0;
        if (i < n) {
            const float4 ra = __ldg(&rays[2 * i]), rb = __ldg(&rays[2 * i + 1]);
            const float2 hr = __ldg(&hits[i]);
            slot = __float_as_uint(ra.w);
            const V3 rStart(ra.x, ra.y, ra.z), rDir(rb.x, rb.y, rb.z);
            const RayCastOut h = raycast_finish(sc, sph, rStart, rDir, hr.x, __float_as_uint(hr.y)); // :81
            if (!h.hit) {
                bf_finish_sample(sc, pm, bp, slot, false, 0.f, 0.f, 0.f, h.dist, accum, last, meta, film, aux); // :160-164
            } else {
                float vNDL;
                bf_primary_shade(sc, h, &vNDL, &sdir, &ldir);
                P = h.P;
                tmp[slot] = make_float4(vNDL, h.uv0, h.uv1, h.dist);
                emit = true;
            }
        }
        const uint32_t ballot = __ballot_sync(0xffffffffu, emit);
        if (ballot) {
            const int leader = __ffs(ballot) - 1;
            uint32_t qb = 0;
            if ((int)lane == leader) qb = atomicAdd(qcount_out, stride * (uint32_t)__popc(ballot));
            qb = __shfl_sync(0xffffffffu, qb, leader);
            if (emit) {
                const uint64_t q = qb + stride * __popc(ballot & ((1u << lane) - 1u));
                rays_out[2 * q] = make_float4(P.x, P.y, P.z, __uint_as_float(slot));
                rays_out[2 * q + 1] = make_float4(sdir.x, sdir.y, sdir.z, 0.f);
                if (stride == 2u) { // the shadow ray sits right behind its probe
                    rays_out[2 * q + 2] = make_float4(P.x, P.y, P.z, __uint_as_float(slot));
                    rays_out[2 * q + 3] = make_float4(ldir.x, ldir.y, ldir.z, 0.f);
                }
            }
        }
    }
}

__global__ void __launch_bounds__(kWfBlock) k_bf_shade2(SceneDev sc, const SphereDev* __restrict__ sph, PixelMap pm, BfParams bp,
                                                        const float4* __restrict__ rays, const float2* __restrict__ hits,
                                                        const uint32_t* __restrict__ qcount_in, float4* __restrict__ accum, float4* __restrict__ last,
                                                        uint2* __restrict__ meta, const float4* __restrict__ tmp, float4* __restrict__ film,
                                                        float2* __restrict__ aux) {
    const uint32_t stride = (bp.flags & VMN_RENDER_BF_SHADOW) ? 2u : 1u;
    const uint32_t n = *qcount_in / stride;
    for (uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = (uint64_t)stride * j;
        const float4 ra = __ldg(&rays[2 * i]), rb = __ldg(&rays[2 * i + 1]);
        const float2 hr = __ldg(&hits[i]);
        const uint32_t slot = __float_as_uint(ra.w);
        const RayCastOut h2 = raycast_finish(sc, sph, V3(ra.x, ra.y, ra.z), V3(rb.x, rb.y, rb.z), hr.x, __float_as_uint(hr.y)); // :121
        const float4 t4 = tmp[slot];
        float vNDL = t4.x;
        if (stride == 2u) { // :108-110 with RayCastCollision's test (meshEngine.cpp:196-207)
            const float2 hs = __ldg(&hits[i + 1]);
            if (__float_as_uint(hs.y) != VMN_MISS && (double)hs.x > 1e-3) vNDL *= 0.00f;
        }
        if (!h2.hit) { vNDL *= 0.9f; vNDL += 0.1f; } // :133-137
        bf_finish_sample(sc, pm, bp, slot, true, vNDL, t4.y, t4.z, t4.w, accum, last, meta, film, aux);
    }
}

// ---- one-launch frames -----------------------------------------------------------------------------------------------------------
// The adaptive estimators (COMPAT PathTracer, BruteForceTracer) take a data-dependent number of samples per pixel and of rays per
// sample, so their wavefront needs the host to read a counter after every few launches to know when to stop: C1 (512 x 512, 16 spp)
// was 240 launches and a dozen stream synchronisations for 4.8 ms of work. For frames of up to a few million pixels the whole pixel
// loop of pathtracer.cpp:242-324 / integrators.cpp:59-183 runs in ONE kernel instead: one thread per pixel, the sample loop, the early-out
// state machine and the path loop in registers, the single-ray form of the same BVH8 traversal (trace_wide), the same raycast_finish /
// radiance_bounce / bf_* code and the same Philox streams as the wavefront stages — so the film is bit-identical to the wavefront's
// (tests run both). Divergence costs lanes, but these frames are launch-bound, not throughput-bound.
struct MegaParams {
    uint64_t seed, n_local;
    uint32_t spp, per, max_segments, flags;
};

__device__ __forceinline__ uint32_t mega_trace(const SceneDev& sc, V3 o, V3 d, float* t, uint2* sstack) {
    TravCounters cnt{0ull, 0ull};
    return leaf_original_id(sc.tris, trace_wide<false>(sc, o, d, t, sstack, kWfBlock, &cnt));
}

__global__ void __launch_bounds__(kWfBlock) k_mega_compat(SceneDev sc, const SphereDev* __restrict__ sph, CamDev cam, PixelMap pm, MegaParams mp,
                                                          float4* __restrict__ film, unsigned long long* __restrict__ frame_ctr) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kWfBlock];
    uint2* sstack = sstack_all + threadIdx.x;
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t pixel;
    if (lp >= mp.n_local || !map_pixel(pm, lp, &pixel)) return;
    float4 f = film[pixel];
    uint32_t st = 0, sz = 0, rays = 0, samples = 0;
    for (uint32_t round = 0;; ++round) { // one sample per iteration: k_raygen + (k_extend, k_shade)* + k_accumulate of the wavefront
        Rng jr(mp.seed, pixel, round, 0);
        const float jx = jr.jitter(), jy = jr.jitter();
        V3 o, d;
        camera_ray(cam, pixel, (st >> 1) & 1u, st & 1u, jx, jy, &o, &d);
        float4 acc = make_float4(0.f, 0.f, 0.f, -100.f), tp = make_float4(1.f, 1.f, 1.f, 1.f);
        uint32_t depth = 0, segs = 0;
        bool thr_written = false;
        for (bool alive = true; alive;) {
            float t;
            const uint32_t id = mega_trace(sc, o, d, &t, sstack);
            ++rays;
            const RayCastOut h = raycast_finish(sc, sph, o, d, t, id);
            ++segs;
            alive = h.hit; // :37-41
            if (alive) {
                if (thr_written || depth == 0 || h.Le.x != 0.f || h.Le.y != 0.f || h.Le.z != 0.f) { // as k_shade (x + 1*0 = x otherwise)
                    float4 a = depth == 0 ? make_float4(0.f, 0.f, 0.f, -100.f) : acc;
                    a.x += tp.x * h.Le.x; a.y += tp.y * h.Le.y; a.z += tp.z * h.Le.z; a.w += tp.w * 0.f; // :43
                    if (depth == 0) a.w = h.dist; // :44-47
                    acc = a;
                }
                if (length(h.Le) > 1.f) alive = false; // :52
            } else if (depth == 0) acc = make_float4(0.f, 0.f, 0.f, -100.f);
            if (alive) {
                bool thr_dirty = false;
                V3 no, nd;
                alive = radiance_bounce(sc, VMN_MODE_COMPAT, mp.max_segments, mp.seed, h, d, depth, segs, tp, thr_dirty, &no, &nd,
                                        [&](uint64_t* px, uint32_t* smp) { *px = pixel; *smp = round; });
                thr_written = thr_written || thr_dirty;
                o = no; d = nd;
            }
        }
        ++samples;
        // k_accumulate, COMPAT branch (pathtracer.cpp:282-312)
        f.x += acc.x; f.y += acc.y; f.z += acc.z; f.w += 1.f;
        const uint32_t n = (uint32_t)f.w;
        bool brk = false;
        if ((double)n > sqrt((double)mp.spp)) { // :290
            const float fn = (float)n, fn1 = (float)(n + 1);
            V3 a(f.x / fn, f.y / fn, f.z / fn), b((f.x + acc.x) / fn1, (f.y + acc.y) / fn1, (f.z + acc.z) / fn1);
            if (fabsf(length(a - b)) < 0.00001f) brk = true;
        }
        ++sz;
        if (brk || sz >= mp.per) { sz = 0; if (st == 3) break; ++st; }
    }
    film[pixel] = f;
    atomicAdd(frame_ctr + 2, (unsigned long long)rays);
    atomicAdd(frame_ctr + 3, (unsigned long long)samples);
}

__global__ void __launch_bounds__(kWfBlock) k_mega_bf(SceneDev sc, const SphereDev* __restrict__ sph, CamDev cam, PixelMap pm, MegaParams mp,
                                                      float4* __restrict__ film, float2* __restrict__ aux, unsigned long long* __restrict__ frame_ctr) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kWfBlock];
    uint2* sstack = sstack_all + threadIdx.x;
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t pixel;
    if (lp >= mp.n_local || !map_pixel(pm, lp, &pixel)) return;
    float4 accum = make_float4(0.f, 0.f, 0.f, 0.f), last = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t n = 0, rays = 0;
    float dist = 0.f;
    for (bool done = false; !done;) {
        Rng jr(mp.seed, pixel, n, 0); // sample index = samples taken so far
        const float jx = jr.jitter(), jy = jr.jitter();
        V3 o, d;
        bf_camera_ray(cam, pixel, jx, jy, &o, &d);
        float t;
        uint32_t id = mega_trace(sc, o, d, &t, sstack);
        ++rays;
        const RayCastOut h = raycast_finish(sc, sph, o, d, t, id); // :81
        ++n;
        dist = h.dist;
        if (!h.hit) { done = bf_sample_end(sc, mp.spp, false, 0.f, 0.f, 0.f, accum, last, n); continue; } // :160-164
        float vNDL;
        V3 sdir, ldir;
        bf_primary_shade(sc, h, &vNDL, &sdir, &ldir);
        id = mega_trace(sc, h.P, sdir, &t, sstack); // the mirror-direction probe (:119-121)
        ++rays;
        const RayCastOut h2 = raycast_finish(sc, sph, h.P, sdir, t, id);
        if (mp.flags & VMN_RENDER_BF_SHADOW) { // :108-110 with RayCastCollision's test (meshEngine.cpp:196-207)
            float ts;
            const uint32_t ids = mega_trace(sc, h.P, ldir, &ts, sstack);
            ++rays;
            if (ids != VMN_MISS && (double)ts > 1e-3) vNDL *= 0.00f;
        }
        if (!h2.hit) { vNDL *= 0.9f; vNDL += 0.1f; } // :133-137
        done = bf_sample_end(sc, mp.spp, true, vNDL, h.uv0, h.uv1, accum, last, n);
    }
    film[pixel] = make_float4(accum.x, accum.y, accum.z, (float)n);
    aux[pixel] = make_float2(accum.w, dist);
    atomicAdd(frame_ctr + 2, (unsigned long long)rays);
    atomicAdd(frame_ctr + 3, (unsigned long long)n);
}

// VMN_MEGA=0 never, 1 always, unset: frames of up to 4 Mi pixels over scenes of up to 64 Ki triangles. Measured (profiles/r02c_*): C1
// (30 triangles, 512 x 512, COMPAT) 6.08 -> 1.40 ms end to end, 240 launches -> 1; but BruteForce at 1080p over C2's 983 040 triangles is
// throughput-bound, not launch-bound — 2.35 ms as a wavefront (5.3 Grays/s through the pooled persistent traversal) against 3.74 ms as one
// divergent per-thread kernel — so large scenes keep the wavefront. Read at every render (tests switch it).
static bool mega_wanted(const vmn_ctx* c, uint64_t n_local) {
    const char* e = getenv("VMN_MEGA");
    if (e && *e) return atoi(e) != 0;
    return n_local <= (4ull << 20) && c->n_tris <= (64ull << 10);
}

static vmn_status render_mega(vmn_ctx* c, const vmn_render_params* prm, const PixelMap& pm, uint64_t n_local) {
    vmn_status s = pool_ensure(c, 0);
    if (s != VMN_OK) return s;
    PathPool* p = c->pool;
    unsigned long long* ctr = reinterpret_cast<unsigned long long*>(p->ctl.as<char>() + kOffFrame);
    VMN_CUDA(c, cudaMemsetAsync(ctr, 0, 32, c->stream));
    MegaParams mp{};
    mp.seed = prm->seed; mp.n_local = n_local; mp.spp = c->cam.spp; mp.per = c->cam.spp / 4; mp.max_segments = prm->max_segments ? prm->max_segments : 5;
    mp.flags = prm->flags;
    const CamDev cd = cam_dev(c);
    VMN_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
    const uint32_t grid = cdiv64(n_local, kWfBlock);
    if (prm->mode == VMN_MODE_BRUTEFORCE)
        k_mega_bf<<<grid, kWfBlock, 0, c->stream>>>(c->scene, c->spheres.as<SphereDev>(), cd, pm, mp, c->film.as<float4>(), c->film_aux.as<float2>(), ctr);
    else
        k_mega_compat<<<grid, kWfBlock, 0, c->stream>>>(c->scene, c->spheres.as<SphereDev>(), cd, pm, mp, c->film.as<float4>(), ctr);
    VMN_CUDA(c, cudaGetLastError());
    VMN_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
    unsigned long long h[4] = {0, 0, 0, 0};
    VMN_CUDA(c, cudaMemcpyAsync(h, ctr, 32, cudaMemcpyDeviceToHost, c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]);
    c->stats.total_ms = ms; c->stats.extend_ms = ms; // one kernel: traversal and shading are not separable
    c->stats.kernel_launches = 1;
    c->stats.rays_traced = h[2]; c->stats.paths_traced = h[3];
    c->stats.nodes_visited = 0; c->stats.tris_tested = 0;
    return VMN_OK;
}

vmn_status render_bruteforce(vmn_ctx* c, const vmn_render_params* prm) {
    const uint64_t npix = c->film_pixels;
    const uint32_t spp = c->cam.spp;
    if (spp > 65535u) return fail(c, VMN_ERR_INVALID, "BRUTEFORCE: spp %u > 65535 (the reference's uint16_t sample loop never ends, integrators.cpp:59)", spp);
    if (prm->flags & VMN_RENDER_ACCUMULATE) return fail(c, VMN_ERR_INVALID, "VMN_RENDER_ACCUMULATE needs VMN_MODE_FIXED");
    PixelMap pm{};
    uint64_t n_local_total;
    vmn_status s = make_pixel_map(c, prm, &pm, &n_local_total);
    if (s != VMN_OK) return s;
    c->stats.kernel_launches = 0; c->stats.rays_traced = 0; c->stats.paths_traced = 0;
    c->stats.raygen_ms = c->stats.extend_ms = c->stats.shade_ms = c->stats.resolve_ms = c->stats.total_ms = 0;
    if ((s = dev_reserve(c, c->film_aux, 8 * npix)) != VMN_OK) return s;
    VMN_CUDA(c, cudaMemsetAsync(c->film.p, 0, 16 * npix, c->stream));
    VMN_CUDA(c, cudaMemsetAsync(c->film_aux.p, 0, 8 * npix, c->stream));
    c->film_bf = true;
    if (n_local_total == 0 || spp == 0) return VMN_OK; // spp == 0: the reference divides 0 by 0; here the film stays empty
    if (mega_wanted(c, n_local_total)) return render_mega(c, prm, pm, n_local_total);

    const uint32_t stride = (prm->flags & VMN_RENDER_BF_SHADOW) ? 2u : 1u;
    uint64_t chunk = c->max_paths / stride;
    if (chunk > n_local_total) chunk = n_local_total;
    if (chunk * stride >= 0xFFFFFFFFull) chunk = 0xFFFFFFF0ull / stride;
    if ((s = pool_ensure(c, chunk * stride)) != VMN_OK) return s;
    PathPool* p = c->pool;
    if ((s = dev_reserve(c, p->bf_last, 16 * chunk)) != VMN_OK) return s;
    p->ev_used = 0;
    StageClock clk{c, p, {}};
    uint32_t* qc = p->ctl.as<uint32_t>();
    const CamDev cd = cam_dev(c);
    const int grid = extend_grid(c, (c->flags & VMN_CFG_COUNTERS) != 0);
    const SphereDev* sph = c->spheres.as<SphereDev>();
    clk.flush();
    VMN_CUDA(c, cudaMemsetAsync(p->ctl.as<char>() + kOffFrame, 0, 32, c->stream));

    VMN_CUDA(c, cudaEventRecord(c->ev[4], c->stream));
    for (uint64_t lp0 = 0; lp0 < n_local_total; lp0 += chunk) {
        BfParams bp{};
        bp.seed = prm->seed; bp.lp0 = lp0; bp.n_local = (n_local_total - lp0 < chunk) ? n_local_total - lp0 : chunk;
        bp.spp = spp; bp.flags = prm->flags;
        VMN_CUDA(c, cudaMemsetAsync(p->radiance.p, 0, 16 * bp.n_local, c->stream)); // accum
        VMN_CUDA(c, cudaMemsetAsync(p->bf_last.p, 0, 16 * bp.n_local, c->stream));  // lastSampleColour
        VMN_CUDA(c, cudaMemsetAsync(p->meta.p, 0, 8 * bp.n_local, c->stream));      // (nTotalSamples, done)
        for (uint32_t sample = 0; sample < spp; ++sample) {
            bp.sample = sample;
            clk.flush();
            VMN_CUDA(c, cudaMemsetAsync(p->ctl.p, 0, 64 + 16, c->stream)); // queue counts + the two claim counters used below
            clk.begin(ST_RAYGEN);
            k_bf_raygen<<<cdiv64(bp.n_local, kWfBlock), kWfBlock, 0, c->stream>>>(cd, pm, bp, p->meta.as<uint2>(), p->ray[0].as<float4>(), qc,
                                                                                  reinterpret_cast<unsigned long long*>(p->ctl.as<char>() + kOffFrame) + 3);
            clk.end();
            ++c->stats.kernel_launches;
            // The host only needs the live count to know when to stop launching: it is read after every fourth sample (a stage that
            // finds its queue empty returns at once), the samples started are counted on the device (frame counter 3).
            if ((sample & 3u) == 3u || sample + 1 == spp) {
                clk.flush();
                VMN_CUDA(c, cudaMemcpyAsync(p->h_count, qc, 4, cudaMemcpyDeviceToHost, c->stream));
                VMN_CUDA(c, cudaStreamSynchronize(c->stream));
                if (p->h_count[0] == 0) break;
            }
            if ((s = launch_extend(c, p, 0, 0, &clk, grid)) != VMN_OK) return s;
            clk.begin(ST_SHADE);
            k_bf_shade1<<<c->sm_count * 8, kWfBlock, 0, c->stream>>>(c->scene, sph, pm, bp, p->ray[0].as<float4>(), p->hit.as<float2>(), qc,
                                                                     p->ray[1].as<float4>(), qc + 1, p->radiance.as<float4>(), p->bf_last.as<float4>(),
                                                                     p->meta.as<uint2>(), p->thr.as<float4>(), c->film.as<float4>(),
                                                                     c->film_aux.as<float2>());
            clk.end();
            if ((s = launch_extend(c, p, 1, 1, &clk, grid)) != VMN_OK) return s;
            clk.begin(ST_SHADE);
            k_bf_shade2<<<c->sm_count * 8, kWfBlock, 0, c->stream>>>(c->scene, sph, pm, bp, p->ray[1].as<float4>(), p->hit.as<float2>(), qc + 1,
                                                                     p->radiance.as<float4>(), p->bf_last.as<float4>(), p->meta.as<uint2>(),
                                                                     p->thr.as<float4>(), c->film.as<float4>(), c->film_aux.as<float2>());
            clk.end();
            VMN_CUDA(c, cudaGetLastError());
            c->stats.kernel_launches += 2;
        }
    }
    VMN_CUDA(c, cudaEventRecord(c->ev[5], c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    VMN_CUDA(c, cudaGetLastError());
    float ms = 0.f;
    cudaEventElapsedTime(&ms, c->ev[4], c->ev[5]);
    c->stats.total_ms = ms;
    double st[4] = {0, 0, 0, 0};
    clk.collect(st);
    c->stats.raygen_ms = st[ST_RAYGEN]; c->stats.extend_ms = st[ST_EXTEND]; c->stats.shade_ms = st[ST_SHADE]; c->stats.resolve_ms = st[ST_RESOLVE];
    unsigned long long h[4] = {0, 0, 0, 0};
    VMN_CUDA(c, cudaMemcpy(h, p->ctl.as<char>() + kOffFrame, 32, cudaMemcpyDeviceToHost));
    c->stats.nodes_visited = h[0]; c->stats.tris_tested = h[1]; c->stats.rays_traced = h[2]; c->stats.paths_traced = h[3];
    return VMN_OK;
}

} // namespace vmn
