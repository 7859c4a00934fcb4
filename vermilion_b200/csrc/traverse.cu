// traverse.cu — ray-query kernels behind vmn_intersect / vmn_occluded / vmn_raycast.
//
// Persistent warps: the grid is sized to (SM count x resident blocks per SM); each warp claims 32
// consecutive rays at a time with one atomicAdd (lane 0) + shuffle, so long rays do not leave a
// tail of idle blocks. The BVH8 traversal stack lives in shared memory, strided so that the 32
// lanes of a warp hit 32 different banks pairs (uint2 entries, conflict-free 64-bit accesses).
#include "engine.cuh"
#include "traverse.cuh"

#include <mutex>

namespace vmn {

constexpr int kTraceBlock = 128;
#ifndef VMN_TRACE_MINB
#define VMN_TRACE_MINB 8 // 64 registers: 8 resident CTAs per SM measured fastest (22.3 ms vs 26.7 ms at 96 registers on C2)
#endif

__device__ __forceinline__ bool claim_rays(unsigned long long* next, uint64_t n, uint64_t* i_out) {
    unsigned long long base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(next, 32ull);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (base >= n) return false;
    *i_out = base + (threadIdx.x & 31);
    return true;
}

__device__ __forceinline__ void load_ray(const vmn_ray* __restrict__ rays, uint64_t i, V3* o, V3* d) {
    const float4* p = reinterpret_cast<const float4*>(rays + i);
    const float4 a = ld_stream(p), b = ld_stream(p + 1);
    *o = V3(a.x, a.y, a.z); *d = V3(b.x, b.y, b.z);
}

#ifndef VMN_TRACE_MINB_OOC
#define VMN_TRACE_MINB_OOC 9 // 56 registers for scenes the L2 cannot hold: latency-bound, one more CTA per SM is +2 % on C3 (r02u); 10 spills (-7 %)
#endif

template <bool WIDE, bool COUNT, int MINB = VMN_TRACE_MINB>
__global__ void __launch_bounds__(kTraceBlock, MINB) k_intersect(SceneDev sc, const vmn_ray* __restrict__ rays, uint64_t n,
                                                           vmn_hit* __restrict__ out, unsigned long long* next, TravCounters* gcnt,
                                                           uint32_t refill_min) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kTraceBlock];
    __shared__ __align__(16) uint32_t wscratch_all[VMN_WSCRATCH * (kTraceBlock / 32)];
    uint2* sstack = sstack_all + threadIdx.x;
    TravCounters cnt{0ull, 0ull};
    auto finish = [&](uint64_t i, V3 o, V3 d, float t, uint32_t id) { // id: original triangle id or VMN_MISS
        float u = 0.f, v = 0.f;
        if (id != VMN_MISS) {
            const V3 hp = o + d * t; // bvh.cpp:140
            tri_bary(load_tri(sc.tris_orig + id), hp, &u, &v);
        }
        st_stream(&reinterpret_cast<float4*>(out)[i], make_float4(t, u, v, __uint_as_float(id)));
    };
    if (WIDE) {
        persistent_trace<COUNT>(sc, n, next, pick_chunk(n, gridDim.x * (kTraceBlock / 32)), refill_min, sstack, kTraceBlock, wscratch_all + VMN_WSCRATCH * (threadIdx.x >> 5), &cnt,
                                [&](uint64_t i, V3* o, V3* d) { load_ray(rays, i, o, d); },
                                [&](uint64_t i, const Trav& s) { finish(i, s.o, s.d, s.best_t, leaf_original_id(sc.tris, s.best)); });
    } else {
        uint64_t i;
        while (claim_rays(next, n, &i)) {
            if (i >= n) continue;
            V3 o, d;
            load_ray(rays, i, &o, &d);
            float t;
            const uint32_t k = trace_bvh2<COUNT>(sc, o, d, &t, &cnt);
            finish(i, o, d, t, k == VMN_MISS ? VMN_MISS : __float_as_uint(__ldg(&reinterpret_cast<const float4*>(sc.tris_sorted + k)->w)));
        }
    }
    if (COUNT) {
        atomicAdd(&gcnt->nodes, cnt.nodes);
        atomicAdd(&gcnt->tris, cnt.tris);
    }
}

__global__ void __launch_bounds__(kTraceBlock, VMN_TRACE_MINB) k_occluded(SceneDev sc, const vmn_ray* __restrict__ rays, uint64_t n, uint8_t* __restrict__ out,
                                                          unsigned long long* next, uint32_t refill_min) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kTraceBlock];
    __shared__ __align__(16) uint32_t wscratch_all[VMN_WSCRATCH * (kTraceBlock / 32)];
    uint2* sstack = sstack_all + threadIdx.x;
    TravCounters cnt{0ull, 0ull};
    persistent_trace<false>(sc, n, next, pick_chunk(n, gridDim.x * (kTraceBlock / 32)), refill_min, sstack, kTraceBlock, wscratch_all + VMN_WSCRATCH * (threadIdx.x >> 5), &cnt,
                            [&](uint64_t i, V3* o, V3* d) { load_ray(rays, i, o, d); },
                            [&](uint64_t i, const Trav& s) { out[i] = (s.best != VMN_MISS && (double)s.best_t > 1e-3) ? 1 : 0; }); // meshEngine.cpp:196-207
}

__global__ void __launch_bounds__(kTraceBlock, 6) k_raycast(SceneDev sc, const SphereDev* __restrict__ sph, const vmn_ray* __restrict__ rays, uint64_t n,
                                                         vmn_raycast_rec* __restrict__ out, unsigned long long* next, uint32_t refill_min) {
    __shared__ uint2 sstack_all[VMN_SH_STACK * kTraceBlock];
    __shared__ __align__(16) uint32_t wscratch_all[VMN_WSCRATCH * (kTraceBlock / 32)];
    uint2* sstack = sstack_all + threadIdx.x;
    TravCounters cnt{0ull, 0ull};
    persistent_trace<false>(sc, n, next, pick_chunk(n, gridDim.x * (kTraceBlock / 32)), refill_min, sstack, kTraceBlock, wscratch_all + VMN_WSCRATCH * (threadIdx.x >> 5), &cnt,
                            [&](uint64_t i, V3* o, V3* d) { load_ray(rays, i, o, d); },
                            [&](uint64_t i, const Trav& s) {
                                const RayCastOut r = raycast_finish(sc, sph, s.o, s.d, s.best_t, leaf_original_id(sc.tris, s.best));
                                float4* q = reinterpret_cast<float4*>(out + i);
                                q[0] = make_float4(r.P.x, r.P.y, r.P.z, r.dist);
                                q[1] = make_float4(r.N.x, r.N.y, r.N.z, __uint_as_float((r.hit ? VMN_RC_HIT : 0u) | (r.material ? VMN_RC_MATERIAL : 0u)));
                                q[2] = make_float4(r.uv0, r.uv1, r.Le.x, r.Le.y);
                                q[3] = make_float4(r.Le.z, __uint_as_float(r.tri), 0.f, 0.f);
                            });
}

template <class K>
static int resident_grid(vmn_ctx* c, K kernel) {
    int per_sm = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kTraceBlock, 0) != cudaSuccess || per_sm < 1) per_sm = 4;
    return c->sm_count * per_sm;
}

// Idle lanes of a persistent warp that trigger a refill from the ray queue. Measured (profiles/r02a_ab_c2.txt, r02a_ab_c3.txt): 8 on a
// scene the L2 holds (C2: 4 and 8 equal, 12 is 3.7 % slower); 4 when every node fetch goes to HBM (C3 soup: +1.9 % over 8, 16 is 6 % slower)
// — a lane that waits for seven neighbours to finish wastes more when each of their steps is a DRAM round trip. VMN_REFILL overrides.
uint32_t refill_min(const vmn_ctx* c) {
    static const int env = [] { const char* e = getenv("VMN_REFILL"); int x = e ? atoi(e) : 0; return x < 0 ? 0 : (x > 32 ? 32 : x); }();
    if (env) return (uint32_t)env;
    return c->trav_bytes > (96ull << 20) ? 4u : 8u;
}

static vmn_status reset_counters(vmn_ctx* c) {
    vmn_status s = dev_reserve(c, c->counters, 64);
    if (s != VMN_OK) return s;
    VMN_CUDA(c, cudaMemsetAsync(c->counters.p, 0, 64, c->stream));
    return VMN_OK;
}

vmn_status launch_intersect(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, vmn_hit* d_out, uint32_t flags) {
    vmn_status s = reset_counters(c);
    if (s != VMN_OK) return s;
    auto* next = c->counters.as<unsigned long long>();
    auto* gcnt = reinterpret_cast<TravCounters*>(c->counters.as<unsigned long long>() + 2);
    const bool count = (c->flags & VMN_CFG_COUNTERS) != 0;
    const bool wide = !(flags & VMN_ISECT_BVH2);
    c->stats.kernel_launches = 0;
    if (n) {
        VMN_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
#define LAUNCH(W, C)                                                                                                           \
    k_intersect<W, C><<<resident_grid(c, k_intersect<W, C>), kTraceBlock, 0, c->stream>>>(c->scene, d_rays, n, d_out, next, gcnt, refill_min(c))
        // out-of-cache scenes (same threshold as refill_min) run the 9-CTA build of the plain wide kernel; VMN_OOC_OCC=0 turns that off
        static const bool ooc_occ = [] { const char* e = getenv("VMN_OOC_OCC"); return !(e && e[0] == '0'); }();
        if (wide && !count && ooc_occ && c->trav_bytes > (96ull << 20))
            k_intersect<true, false, VMN_TRACE_MINB_OOC><<<resident_grid(c, k_intersect<true, false, VMN_TRACE_MINB_OOC>), kTraceBlock, 0, c->stream>>>(
                c->scene, d_rays, n, d_out, next, gcnt, refill_min(c));
        else if (wide) { if (count) LAUNCH(true, true); else LAUNCH(true, false); }
        else { if (count) LAUNCH(false, true); else LAUNCH(false, false); }
#undef LAUNCH
        VMN_CUDA(c, cudaGetLastError());
        VMN_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
        c->stats.kernel_launches = 1;
    }
    c->stats.rays_traced = n;
    return VMN_OK;
}

vmn_status launch_occluded(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, uint8_t* d_out) {
    vmn_status s = reset_counters(c);
    if (s != VMN_OK) return s;
    if (n) {
        k_occluded<<<resident_grid(c, k_occluded), kTraceBlock, 0, c->stream>>>(c->scene, d_rays, n, d_out, c->counters.as<unsigned long long>(), refill_min(c));
        VMN_CUDA(c, cudaGetLastError());
    }
    c->stats.rays_traced = n; c->stats.kernel_launches = n ? 1 : 0;
    return VMN_OK;
}

vmn_status launch_raycast(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, vmn_raycast_rec* d_out) {
    vmn_status s = reset_counters(c);
    if (s != VMN_OK) return s;
    if (n) {
        k_raycast<<<resident_grid(c, k_raycast), kTraceBlock, 0, c->stream>>>(c->scene, c->spheres.as<SphereDev>(), d_rays, n, d_out,
                                                                            c->counters.as<unsigned long long>(), refill_min(c));
        VMN_CUDA(c, cudaGetLastError());
    }
    c->stats.rays_traced = n; c->stats.kernel_launches = n ? 1 : 0;
    return VMN_OK;
}

} // namespace vmn

// ---- vmn_intersect_static: closest-hit queries for callers that may not allocate -------------------------------------------------
// No counterpart in the reference (declared in include/vermilion_b200.h). It is
// vmn_intersect — host rays in, host hits out, the same k_intersect instantiations — with ALL device scratch (the persistent kernel's
// claim cursor + visit counters, the staged rays, the hits) in one file-scope __device__ array instead of the context's grow-only
// cudaMalloc buffers. Every device that loads the library holds the array (3 MiB + 256 B) until exit. Calls share it: it keeps the last
// call's contents (the counter block is cleared on the stream at the start of each call; rays and hits are overwritten before they are
// read), and two calls that overlap on one GPU — two contexts, two host threads — race on it.
namespace {

constexpr uint64_t kStaticCounterBytes = 256; // claim cursor (u64), pad, TravCounters: the first 64 bytes in use, like vmn_ctx::counters
constexpr uint64_t kStaticRayOff = kStaticCounterBytes;
constexpr uint64_t kStaticHitOff = kStaticRayOff + sizeof(vmn_ray) * (uint64_t)VMN_STATIC_MAX_RAYS;
constexpr uint64_t kStaticBytes = kStaticHitOff + sizeof(vmn_hit) * (uint64_t)VMN_STATIC_MAX_RAYS;

static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address a kernel may use comes from
// cudaGetSymbolAddress, asked once per device with that device current.
constexpr int kStaticMaxDevices = 64;
unsigned char* g_static_ptr[kStaticMaxDevices] = {};
std::mutex g_static_mu;

vmn_status static_scratch(vmn_ctx* c, unsigned char** out) {
    if (c->device < 0 || c->device >= kStaticMaxDevices) return vmn::fail(c, VMN_ERR_INVALID, "device %d: the static scratch serves devices 0..%d", c->device, kStaticMaxDevices - 1);
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[c->device]) {
        VMN_CUDA(c, cudaSetDevice(c->device));
        void* p = nullptr;
        VMN_CUDA(c, cudaGetSymbolAddress(&p, g_static_scratch));
        if (!p) return vmn::fail(c, VMN_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_static_ptr[c->device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_ptr[c->device];
    return VMN_OK;
}

} // namespace

extern "C" vmn_status vmn_intersect_static(vmn_ctx* c, const vmn_ray* rays, uint64_t n, vmn_hit* out, uint32_t flags) {
    using namespace vmn;
    if (!c) return VMN_ERR_INVALID;
    if (!c->committed) return fail(c, VMN_ERR_STATE, "scene not committed (call vmn_scene_commit)");
    if (n && (!rays || !out)) return fail(c, VMN_ERR_INVALID, "null ray/hit pointer");
    if (n > (uint64_t)VMN_STATIC_MAX_RAYS)
        return fail(c, VMN_ERR_INVALID, "vmn_intersect_static: %llu rays exceed VMN_STATIC_MAX_RAYS = %u (the static device scratch holds no more; "
                    "split the batch or use vmn_intersect)", (unsigned long long)n, (unsigned)VMN_STATIC_MAX_RAYS);
    VMN_CUDA(c, cudaSetDevice(c->device));
    unsigned char* base = nullptr;
    vmn_status s = static_scratch(c, &base);
    if (s != VMN_OK) return s;
    auto* next = reinterpret_cast<unsigned long long*>(base);
    auto* gcnt = reinterpret_cast<TravCounters*>(next + 2);
    auto* d_rays = reinterpret_cast<vmn_ray*>(base + kStaticRayOff);
    auto* d_out = reinterpret_cast<vmn_hit*>(base + kStaticHitOff);
    const bool count = (c->flags & VMN_CFG_COUNTERS) != 0;
    const bool wide = !(flags & VMN_ISECT_BVH2);
    c->stats.kernel_launches = 0;
    c->stats.rays_traced = n;
    if (n) {
        VMN_CUDA(c, cudaMemsetAsync(base, 0, kStaticCounterBytes, c->stream)); // the cursor and the counters must start at zero
        VMN_CUDA(c, cudaMemcpyAsync(d_rays, rays, sizeof(vmn_ray) * n, cudaMemcpyHostToDevice, c->stream));
        VMN_CUDA(c, cudaEventRecord(c->ev[2], c->stream));
#define LAUNCH(W, C)                                                                                                           \
    k_intersect<W, C><<<resident_grid(c, k_intersect<W, C>), kTraceBlock, 0, c->stream>>>(c->scene, d_rays, n, d_out, next, gcnt, refill_min(c))
        if (wide) { if (count) LAUNCH(true, true); else LAUNCH(true, false); }
        else { if (count) LAUNCH(false, true); else LAUNCH(false, false); }
#undef LAUNCH
        VMN_CUDA(c, cudaGetLastError());
        VMN_CUDA(c, cudaEventRecord(c->ev[3], c->stream));
        c->stats.kernel_launches = 1;
        VMN_CUDA(c, cudaMemcpyAsync(out, d_out, sizeof(vmn_hit) * n, cudaMemcpyDeviceToHost, c->stream));
    }
    unsigned long long h[4] = {0, 0, 0, 0};
    if (n && count) VMN_CUDA(c, cudaMemcpyAsync(h, base, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
    if ((s = check_device_error(c)) != VMN_OK) return s; // synchronises the stream; VMN_ERR_STATE if a traversal stack overflowed
    if (n) { float ms = 0.f; cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]); c->stats.extend_ms = ms; c->stats.total_ms = ms; }
    if (count) { c->stats.nodes_visited = h[2]; c->stats.tris_tested = h[3]; }
    return VMN_OK;
}
