// dist.cu — several B200s behind ONE call: the vmn_dist_* entry points of include/vermilion_b200.h.
//
// The reference's host makes exactly one mIntegrator->Render(mCameras, mMeshEngine) per frame (core/engines/renderEngine.cpp:147-166)
// and parallelises inside it over pixels (core/integrators/pathtracer.cpp:226-227). This layer keeps that shape: one group object owns
// one engine context per GPU, splits the frame (sample slices or interleaved tiles), and assembles the film on the root.
//
//   local group  one process, N devices (the C++ host / vermilion_headless --gpus N): one context and one host thread per device,
//                communicators from ncclCommInitAll
//   rank group   one process per device (torchrun, bench.py): ncclCommInitRank with an id the caller hands from rank 0 to the others
//
// What crosses NVLink, and nothing else:
//   scene   the host arrays go up ONCE, on the root's PCIe link, and are ncclBroadcast to the other devices; every device then builds its
//           own BVH (1.4 ms for C2's 983 040 triangles; broadcasting the built arrays instead would move 2.6 x the bytes)
//   film    sample slices: ncclReduce(sum) of the (sum r, g, b, n) film to the root, in place
//           tiles: every device packs the pixels of ITS tiles (1/N of the film) and the root receives them (ncclSend / ncclRecv inside
//           one group call, the gather NCCL has no primitive for) and unpacks them into its film — not an all-reduce of the whole buffer
// Only the root resolves and reads the image back.
//
// NCCL is loaded at run time (dlopen libnccl.so.2: the copy the process already holds, e.g. torch's, else the system's), so the library
// itself has no link-time dependency on it and a single-GPU user never needs it. A local group whose device list repeats a device (a
// one-GPU test box) cannot use NCCL (it refuses duplicate devices); it falls back to peer copies + a sum kernel on the root's stream,
// which exercises the same split / pack / unpack code.
#include "engine.cuh"

#include <dlfcn.h>
#include <algorithm>
#include <thread>

namespace vmn {

// ---- the slice of NCCL's C API this file uses (declared here: no build-time dependency on nccl.h) -------------------------------
typedef struct ncclComm* ncclComm_t;
struct ncclUniqueId { char internal[128]; };
enum { ncclSuccess = 0 };
enum { ncclUint8 = 1, ncclFloat32 = 7 };
enum { ncclSum = 0 };
struct NcclApi {
    void* handle = nullptr;
    int (*GetUniqueId)(ncclUniqueId*) = nullptr;
    int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    int (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*Broadcast)(const void*, void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Reduce)(const void*, void*, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    int (*GetVersion)(int*) = nullptr;
    std::string error;
    bool load() {
        if (handle) return true;
        for (const char* name : {"libnccl.so.2", "libnccl.so"}) {
            handle = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (handle) break;
        }
        if (!handle) { error = std::string("cannot load libnccl.so.2: ") + (dlerror() ? dlerror() : "?"); return false; }
#define VMN_NCCL_SYM(field, sym) field = reinterpret_cast<decltype(field)>(dlsym(handle, sym)); if (!field) { error = std::string("libnccl lacks ") + sym; return false; }
        VMN_NCCL_SYM(GetUniqueId, "ncclGetUniqueId") VMN_NCCL_SYM(CommInitRank, "ncclCommInitRank") VMN_NCCL_SYM(CommInitAll, "ncclCommInitAll")
        VMN_NCCL_SYM(CommDestroy, "ncclCommDestroy") VMN_NCCL_SYM(Broadcast, "ncclBroadcast") VMN_NCCL_SYM(Reduce, "ncclReduce")
        VMN_NCCL_SYM(Send, "ncclSend") VMN_NCCL_SYM(Recv, "ncclRecv") VMN_NCCL_SYM(GroupStart, "ncclGroupStart") VMN_NCCL_SYM(GroupEnd, "ncclGroupEnd")
        VMN_NCCL_SYM(GetErrorString, "ncclGetErrorString") VMN_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef VMN_NCCL_SYM
        return true;
    }
};
static NcclApi g_nccl;

// ---- film pack / unpack / add (tile gather, copy transport) -----------------------------------------------------------------------
// lp-th local pixel of (tile, rank, world) <-> film pixel: the same map the render kernels use (map_pixel, common.cuh)
template <int F>
__global__ void k_film_pack(PixelMap pm, uint64_t n_local, const float* __restrict__ film, float* __restrict__ packed) {
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= n_local) return;
    uint64_t pixel;
    const bool ok = map_pixel(pm, lp, &pixel);
#pragma unroll
    for (int k = 0; k < F; ++k) packed[F * lp + k] = ok ? film[F * pixel + k] : 0.f;
}
template <int F>
__global__ void k_film_unpack(PixelMap pm, uint64_t n_local, const float* __restrict__ packed, float* __restrict__ film) {
    const uint64_t lp = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= n_local) return;
    uint64_t pixel;
    if (!map_pixel(pm, lp, &pixel)) return;
#pragma unroll
    for (int k = 0; k < F; ++k) film[F * pixel + k] = packed[F * lp + k];
}
__global__ void k_film_add(float4* __restrict__ dst, const float4* __restrict__ src, uint64_t n4) {
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n4) return;
    float4 a = dst[i];
    const float4 b = src[i];
    a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
    dst[i] = a;
}

} // namespace vmn

using namespace vmn;

struct vmn_dist {
    struct Member {
        vmn_ctx* ctx = nullptr;
        int device = 0;
        uint32_t rank = 0;      // rank of this member in the group
        ncclComm_t comm = nullptr;
        DevBuf packed;          // this member's packed tiles (send side) / scratch
        vmn_status status = VMN_OK;
    };
    std::vector<Member> m;      // local members: all of them in a local group, exactly one in a rank group
    uint32_t world = 1;
    bool use_nccl = false;
    std::string err;
    DevBuf stage;               // root: receive staging for the tile gather / the copy transport
    DevBuf hdr;                 // 32-byte device word for the scene header broadcast
    vmn_dist_stats stats{};
    cudaEvent_t ev[4] = {};
    Member* root() { return (!m.empty() && m[0].rank == 0) ? &m[0] : nullptr; }
};

namespace {

vmn_status dfail(vmn_dist* g, vmn_status s, const std::string& msg) { if (g) g->err = msg; return s; }
vmn_status member_fail(vmn_dist* g, vmn_dist::Member& mb, vmn_status s, const char* what) {
    return dfail(g, s, std::string(what) + " (rank " + std::to_string(mb.rank) + ", device " + std::to_string(mb.device) + "): " + vmn_last_error(mb.ctx));
}
#define DCUDA(g, expr)                                                                                                    \
    do {                                                                                                                  \
        cudaError_t e__ = (expr);                                                                                         \
        if (e__ != cudaSuccess) return dfail((g), VMN_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(e__));     \
    } while (0)
#define DNCCL(g, expr)                                                                                                    \
    do {                                                                                                                  \
        int r__ = (expr);                                                                                                 \
        if (r__ != ncclSuccess) return dfail((g), VMN_ERR_NCCL, std::string(#expr) + ": " + g_nccl.GetErrorString(r__)); \
    } while (0)

// run f(member) for every local member, one host thread each (a member's calls block on its own stream)
template <class F>
vmn_status for_members(vmn_dist* g, F f) {
    if (g->m.size() == 1) { g->m[0].status = f(g->m[0]); }
    else {
        std::vector<std::thread> th;
        for (auto& mb : g->m) th.emplace_back([&mb, &f] { cudaSetDevice(mb.device); mb.status = f(mb); });
        for (auto& t : th) t.join();
    }
    for (auto& mb : g->m) if (mb.status != VMN_OK) return member_fail(g, mb, mb.status, "member call failed");
    return VMN_OK;
}

vmn_status sync_all(vmn_dist* g) {
    for (auto& mb : g->m) { DCUDA(g, cudaSetDevice(mb.device)); DCUDA(g, cudaStreamSynchronize(mb.ctx->stream)); }
    return VMN_OK;
}

uint64_t tiles_local_pixels(uint32_t width, uint32_t height, uint32_t tile, uint32_t rank, uint32_t world) {
    const uint64_t tx = (width + tile - 1) / tile, ty = (height + tile - 1) / tile, tiles = tx * ty;
    const uint64_t mine = tiles > rank ? (tiles - rank + world - 1) / world : 0;
    return mine * tile * tile;
}
PixelMap tile_map(uint32_t width, uint32_t height, uint32_t tile, uint32_t rank, uint32_t world) {
    PixelMap pm{};
    pm.width = width; pm.height = height; pm.tile = tile; pm.rank = rank; pm.world = world; pm.tiles_x = (width + tile - 1) / tile; pm.tile_log2p1 = tile_log2p1(tile);
    return pm;
}
inline uint32_t cdiv64(uint64_t a, uint32_t b) { return (uint32_t)((a + b - 1) / b); }

// ---- collectives over the group (NCCL, or peer copies inside one process) ------------------------------------------------------
// broadcast `bytes` from the root member's buffer to the same buffer of every other member; ptr(member) gives the buffer
// (on_copy_stream: issue it on the members' copy streams instead of their main ones — the scene attributes, beside the geometry build)
template <class P>
vmn_status bcast(vmn_dist* g, P ptr, size_t bytes, bool on_copy_stream = false) {
    if (g->world == 1 || bytes == 0) return VMN_OK;
    auto stream_of = [&](vmn_dist::Member& mb) { return on_copy_stream ? mb.ctx->copy_stream : mb.ctx->stream; };
    if (g->use_nccl) {
        if (g->m.size() > 1) DNCCL(g, g_nccl.GroupStart());
        for (auto& mb : g->m) {
            DCUDA(g, cudaSetDevice(mb.device));
            DNCCL(g, g_nccl.Broadcast(ptr(mb), ptr(mb), bytes, ncclUint8, 0, mb.comm, stream_of(mb)));
        }
        if (g->m.size() > 1) DNCCL(g, g_nccl.GroupEnd());
        return VMN_OK;
    }
    auto* r = g->root();
    DCUDA(g, cudaSetDevice(r->device));
    DCUDA(g, cudaStreamSynchronize(stream_of(*r)));
    for (auto& mb : g->m) {
        if (mb.rank == 0) continue;
        DCUDA(g, cudaSetDevice(mb.device));
        DCUDA(g, cudaMemcpyPeerAsync(ptr(mb), mb.device, ptr(*r), r->device, bytes, stream_of(mb)));
    }
    return VMN_OK;
}

// film(root) = sum over members of film(member), n floats; the other members' films are left as they were
template <class P>
vmn_status reduce_to_root(vmn_dist* g, P ptr, uint64_t n_floats) {
    if (g->world == 1) return VMN_OK;
    if (g->use_nccl) {
        if (g->m.size() > 1) DNCCL(g, g_nccl.GroupStart());
        for (auto& mb : g->m) {
            DCUDA(g, cudaSetDevice(mb.device));
            DNCCL(g, g_nccl.Reduce(ptr(mb), ptr(mb), n_floats, ncclFloat32, ncclSum, 0, mb.comm, mb.ctx->stream));
        }
        if (g->m.size() > 1) DNCCL(g, g_nccl.GroupEnd());
        return VMN_OK;
    }
    auto* r = g->root();
    vmn_status s = dev_reserve(r->ctx, g->stage, 4 * n_floats);
    if (s != VMN_OK) return member_fail(g, *r, s, "staging");
    for (auto& mb : g->m) { // rank order: a deterministic sum
        if (mb.rank == 0) continue;
        DCUDA(g, cudaSetDevice(mb.device));
        DCUDA(g, cudaStreamSynchronize(mb.ctx->stream));
        DCUDA(g, cudaSetDevice(r->device));
        DCUDA(g, cudaMemcpyPeerAsync(g->stage.p, r->device, ptr(mb), mb.device, 4 * n_floats, r->ctx->stream));
        k_film_add<<<cdiv64(n_floats / 4, 256), 256, 0, r->ctx->stream>>>(reinterpret_cast<float4*>(ptr(*r)), g->stage.as<float4>(), n_floats / 4);
        DCUDA(g, cudaGetLastError());
    }
    return VMN_OK;
}

// tile gather of one film plane with F floats per pixel: every non-root member packs its tiles, the root receives and unpacks them
template <int F, class P>
vmn_status gather_tiles(vmn_dist* g, P plane, uint32_t width, uint32_t height, uint32_t tile) {
    if (g->world == 1) return VMN_OK;
    const uint32_t W = g->world;
    std::vector<uint64_t> nloc(W), off(W + 1, 0);
    for (uint32_t r = 0; r < W; ++r) { nloc[r] = tiles_local_pixels(width, height, tile, r, W); off[r + 1] = off[r] + (r ? nloc[r] : 0); }
    auto* root = g->root();
    if (root) {
        vmn_status s = dev_reserve(root->ctx, g->stage, 4ull * F * std::max<uint64_t>(off[W], 1));
        if (s != VMN_OK) return member_fail(g, *root, s, "staging");
    }
    for (auto& mb : g->m) { // pack
        if (mb.rank == 0 || nloc[mb.rank] == 0) continue;
        DCUDA(g, cudaSetDevice(mb.device));
        vmn_status s = dev_reserve(mb.ctx, mb.packed, 4ull * F * nloc[mb.rank]);
        if (s != VMN_OK) return member_fail(g, mb, s, "pack buffer");
        k_film_pack<F><<<cdiv64(nloc[mb.rank], 256), 256, 0, mb.ctx->stream>>>(tile_map(width, height, tile, mb.rank, W), nloc[mb.rank], plane(mb), mb.packed.as<float>());
        DCUDA(g, cudaGetLastError());
    }
    if (g->use_nccl) {
        DNCCL(g, g_nccl.GroupStart());
        for (auto& mb : g->m) {
            DCUDA(g, cudaSetDevice(mb.device));
            if (mb.rank == 0) {
                for (uint32_t r = 1; r < W; ++r)
                    if (nloc[r]) DNCCL(g, g_nccl.Recv(g->stage.as<float>() + F * off[r], F * nloc[r], ncclFloat32, (int)r, mb.comm, mb.ctx->stream));
            } else if (nloc[mb.rank]) DNCCL(g, g_nccl.Send(mb.packed.p, F * nloc[mb.rank], ncclFloat32, 0, mb.comm, mb.ctx->stream));
        }
        DNCCL(g, g_nccl.GroupEnd());
    } else {
        for (auto& mb : g->m) {
            if (mb.rank == 0 || nloc[mb.rank] == 0) continue;
            DCUDA(g, cudaSetDevice(mb.device));
            DCUDA(g, cudaStreamSynchronize(mb.ctx->stream));
            DCUDA(g, cudaSetDevice(root->device));
            DCUDA(g, cudaMemcpyPeerAsync(g->stage.as<float>() + F * off[mb.rank], root->device, mb.packed.p, mb.device, 4ull * F * nloc[mb.rank], root->ctx->stream));
        }
    }
    if (root) {
        DCUDA(g, cudaSetDevice(root->device));
        for (uint32_t r = 1; r < W; ++r) {
            if (!nloc[r]) continue;
            k_film_unpack<F><<<cdiv64(nloc[r], 256), 256, 0, root->ctx->stream>>>(tile_map(width, height, tile, r, W), nloc[r], g->stage.as<float>() + F * off[r], plane(*root));
            DCUDA(g, cudaGetLastError());
        }
    }
    return VMN_OK;
}

vmn_status create_member(vmn_dist* g, int device, uint32_t rank, uint32_t flags, uint64_t max_paths, void* stream) {
    vmn_dist::Member mb;
    mb.device = device; mb.rank = rank;
    vmn_config cfg{};
    cfg.device = device; cfg.flags = flags; cfg.max_paths = max_paths; cfg.stream = stream;
    vmn_status s = vmn_create(&cfg, &mb.ctx);
    if (s != VMN_OK) {
        g->err = std::string("vmn_create on device ") + std::to_string(device) + ": " + (mb.ctx ? vmn_last_error(mb.ctx) : "out of memory");
        if (mb.ctx) vmn_destroy(mb.ctx);
        return s;
    }
    g->m.push_back(mb);
    return VMN_OK;
}

} // namespace

extern "C" {

vmn_status vmn_dist_create_local(const int32_t* devices, uint32_t n, uint32_t flags, uint64_t max_paths, vmn_dist** out) {
    if (!out) return VMN_ERR_INVALID;
    *out = nullptr;
    if (!devices || n == 0 || n > VMN_DIST_MAX_RANKS) return VMN_ERR_INVALID;
    vmn_dist* g = new (std::nothrow) vmn_dist();
    if (!g) return VMN_ERR_OOM;
    *out = g; // returned even on failure so that vmn_dist_last_error can be read; the caller still destroys it
    g->world = n;
    for (uint32_t r = 0; r < n; ++r) {
        vmn_status s = create_member(g, devices[r], r, flags, max_paths, nullptr);
        if (s != VMN_OK) return s;
    }
    bool distinct = true;
    for (uint32_t a = 0; a < n; ++a) for (uint32_t b = a + 1; b < n; ++b) if (devices[a] == devices[b]) distinct = false;
    if (n > 1 && distinct) {
        if (!g_nccl.load()) return dfail(g, VMN_ERR_NCCL, g_nccl.error);
        std::vector<ncclComm_t> comms(n);
        std::vector<int> devs(devices, devices + n);
        DNCCL(g, g_nccl.CommInitAll(comms.data(), (int)n, devs.data()));
        for (uint32_t r = 0; r < n; ++r) g->m[r].comm = comms[r];
        g->use_nccl = true;
    } else if (n > 1) {
        // a device appears twice (one-GPU test box): NCCL refuses that; members exchange by peer copies on the root's stream
        for (uint32_t a = 0; a < n; ++a)
            for (uint32_t b = 0; b < n; ++b)
                if (devices[a] != devices[b]) { cudaSetDevice(devices[a]); cudaDeviceEnablePeerAccess(devices[b], 0); cudaGetLastError(); }
    }
    DCUDA(g, cudaSetDevice(g->m[0].device));
    for (auto& e : g->ev) DCUDA(g, cudaEventCreate(&e));
    return VMN_OK;
}

vmn_status vmn_dist_unique_id(uint8_t* id) {
    if (!id) return VMN_ERR_INVALID;
    if (!g_nccl.load()) return VMN_ERR_NCCL;
    ncclUniqueId u;
    if (g_nccl.GetUniqueId(&u) != ncclSuccess) return VMN_ERR_NCCL;
    static_assert(sizeof(u) == VMN_DIST_ID_BYTES, "ncclUniqueId is 128 bytes");
    memcpy(id, &u, sizeof(u));
    return VMN_OK;
}

vmn_status vmn_dist_create_rank(const uint8_t* id, uint32_t rank, uint32_t world, const vmn_config* cfg, vmn_dist** out) {
    if (!out) return VMN_ERR_INVALID;
    *out = nullptr;
    if (world == 0 || rank >= world || world > VMN_DIST_MAX_RANKS || (world > 1 && !id)) return VMN_ERR_INVALID;
    vmn_dist* g = new (std::nothrow) vmn_dist();
    if (!g) return VMN_ERR_OOM;
    *out = g;
    g->world = world;
    vmn_status s = create_member(g, cfg ? cfg->device : 0, rank, cfg ? cfg->flags : 0, cfg ? cfg->max_paths : 0, cfg ? cfg->stream : nullptr);
    if (s != VMN_OK) return s;
    if (world > 1) {
        if (!g_nccl.load()) return dfail(g, VMN_ERR_NCCL, g_nccl.error);
        ncclUniqueId u;
        memcpy(&u, id, sizeof(u));
        DCUDA(g, cudaSetDevice(g->m[0].device));
        DNCCL(g, g_nccl.CommInitRank(&g->m[0].comm, (int)world, u, (int)rank));
        g->use_nccl = true;
    }
    for (auto& e : g->ev) DCUDA(g, cudaEventCreate(&e));
    return VMN_OK;
}

void vmn_dist_destroy(vmn_dist* g) {
    if (!g) return;
    for (auto& mb : g->m) {
        cudaSetDevice(mb.device);
        if (mb.ctx && mb.ctx->stream) cudaStreamSynchronize(mb.ctx->stream);
        if (mb.comm) g_nccl.CommDestroy(mb.comm);
        dev_free(mb.packed);
    }
    if (!g->m.empty()) { cudaSetDevice(g->m[0].device); dev_free(g->stage); dev_free(g->hdr); for (auto& e : g->ev) if (e) cudaEventDestroy(e); }
    for (auto& mb : g->m) if (mb.ctx) vmn_destroy(mb.ctx);
    delete g;
}

const char* vmn_dist_last_error(const vmn_dist* g) { return g ? g->err.c_str() : "null group"; }
uint32_t vmn_dist_world(const vmn_dist* g) { return g ? g->world : 0; }
uint32_t vmn_dist_rank(const vmn_dist* g) { return (g && !g->m.empty()) ? g->m[0].rank : 0; }
uint32_t vmn_dist_local_count(const vmn_dist* g) { return g ? (uint32_t)g->m.size() : 0; }
vmn_ctx* vmn_dist_context(vmn_dist* g, uint32_t local_index) { return (g && local_index < g->m.size()) ? g->m[local_index].ctx : nullptr; }

vmn_status vmn_dist_scene_set_triangles(vmn_dist* g, const float* pos, const float* nrm, const float* uv, uint64_t n, uint32_t flags) {
    if (!g || g->m.empty()) return VMN_ERR_INVALID;
    const bool bc = (flags & VMN_DIST_SCENE_BROADCAST) != 0 && g->world > 1;
    g->stats.scene_ms = 0;
    if (!bc) { // every local member uploads the host arrays over its own PCIe link
        return for_members(g, [&](vmn_dist::Member& mb) { return vmn_scene_set_triangles(mb.ctx, pos, nrm, uv, n); });
    }
    // header (n, which arrays) from the root, so that the other ranks of a rank group need no host copy of the scene at all
    auto* root = g->root();
    unsigned long long hdr[4] = {n, nrm ? 1ull : 0ull, uv ? 1ull : 0ull, 0ull};
    if (root && n && !pos) return dfail(g, VMN_ERR_INVALID, "pos is NULL on the root");
    auto& first = g->m[0];
    DCUDA(g, cudaSetDevice(first.device));
    if (g->m.size() == 1) { // rank group: learn the header from the root
        vmn_status s = dev_reserve(first.ctx, g->hdr, sizeof(hdr));
        if (s != VMN_OK) return member_fail(g, first, s, "header");
        if (root) DCUDA(g, cudaMemcpyAsync(g->hdr.p, hdr, sizeof(hdr), cudaMemcpyHostToDevice, first.ctx->stream));
        DNCCL(g, g_nccl.Broadcast(g->hdr.p, g->hdr.p, sizeof(hdr), ncclUint8, 0, first.comm, first.ctx->stream));
        DCUDA(g, cudaMemcpyAsync(hdr, g->hdr.p, sizeof(hdr), cudaMemcpyDeviceToHost, first.ctx->stream));
        DCUDA(g, cudaStreamSynchronize(first.ctx->stream));
    }
    const uint64_t nn = hdr[0];
    const bool have_nrm = hdr[1] != 0, have_uv = hdr[2] != 0;
    for (auto& mb : g->m) { // staging buffers everywhere; host arrays go up on the root only
        DCUDA(g, cudaSetDevice(mb.device));
        vmn_ctx* c = mb.ctx;
        c->committed = false; c->geometry_built = false; c->n_tris = 0;
        vmn_status s;
        if ((s = dev_reserve(c, c->pos, 36ull * nn)) != VMN_OK || (have_nrm && (s = dev_reserve(c, c->nrm, 36ull * nn)) != VMN_OK) ||
            (have_uv && (s = dev_reserve(c, c->uv, 24ull * nn)) != VMN_OK))
            return member_fail(g, mb, s, "scene staging");
        c->have_nrm = have_nrm; c->have_uv = have_uv;
        if (!c->copy_stream) DCUDA(g, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        if (mb.rank == 0 && nn) {
            DCUDA(g, cudaEventRecord(g->ev[0], c->stream));
            DCUDA(g, cudaMemcpyAsync(c->pos.p, pos, 36ull * nn, cudaMemcpyHostToDevice, c->stream));
        }
    }
    // Positions first, on the main streams: as soon as they have arrived every member builds the position-only part of its BVH, while
    // the normals and uvs go up on the root's copy stream and cross NVLink on the members' copy streams (same order on every rank).
    vmn_status s;
    const bool overlap = load_overlap();
    if ((s = bcast(g, [](vmn_dist::Member& mb) { return mb.ctx->pos.p; }, 36ull * nn)) != VMN_OK) return s;
    for (auto& mb : g->m) {
        DCUDA(g, cudaSetDevice(mb.device));
        vmn_ctx* c = mb.ctx;
        if (overlap) {
            DCUDA(g, cudaEventRecord(c->ev[6], c->stream));
            DCUDA(g, cudaStreamWaitEvent(c->copy_stream, c->ev[6], 0));
        }
        if (mb.rank == 0 && nn) {
            cudaStream_t as = overlap ? c->copy_stream : c->stream;
            if (have_nrm) DCUDA(g, cudaMemcpyAsync(c->nrm.p, nrm, 36ull * nn, cudaMemcpyHostToDevice, as));
            if (have_uv) DCUDA(g, cudaMemcpyAsync(c->uv.p, uv, 24ull * nn, cudaMemcpyHostToDevice, as));
        }
    }
    if (have_nrm && (s = bcast(g, [](vmn_dist::Member& mb) { return mb.ctx->nrm.p; }, 36ull * nn, overlap)) != VMN_OK) return s;
    if (have_uv && (s = bcast(g, [](vmn_dist::Member& mb) { return mb.ctx->uv.p; }, 24ull * nn, overlap)) != VMN_OK) return s;
    for (auto& mb : g->m) mb.ctx->n_tris = nn;
    if (overlap) s = for_members(g, [&](vmn_dist::Member& mb) { // synchronises each main stream; a failed build is reported by vmn_dist_scene_commit
        vmn_ctx* c = mb.ctx;
        c->geometry_status = build_geometry(c);
        c->geometry_err = c->geometry_status == VMN_OK ? std::string() : c->err;
        c->geometry_built = true;
        return VMN_OK;
    });
    if (s != VMN_OK) return s;
    for (auto& mb : g->m) { // the main stream joins the copy stream: everything of this scene is ordered before what follows on it
        DCUDA(g, cudaSetDevice(mb.device));
        DCUDA(g, cudaEventRecord(mb.ctx->ev[7], mb.ctx->copy_stream));
        DCUDA(g, cudaStreamWaitEvent(mb.ctx->stream, mb.ctx->ev[7], 0));
    }
    if (root && nn) { DCUDA(g, cudaSetDevice(root->device)); DCUDA(g, cudaEventRecord(g->ev[1], root->ctx->stream)); }
    if ((s = sync_all(g)) != VMN_OK) return s; // host pointers are borrowed for the call only
    if (root && nn) { float ms = 0.f; cudaEventElapsedTime(&ms, g->ev[0], g->ev[1]); g->stats.scene_ms = ms; }
    for (auto& mb : g->m) mb.ctx->n_tris = nn;
    return VMN_OK;
}

vmn_status vmn_dist_scene_set_spheres(vmn_dist* g, const vmn_sphere* sp, uint32_t n) {
    if (!g) return VMN_ERR_INVALID;
    return for_members(g, [&](vmn_dist::Member& mb) { return vmn_scene_set_spheres(mb.ctx, sp, n); });
}

vmn_status vmn_dist_scene_set_texture(vmn_dist* g, uint32_t slot, const float* data, uint32_t w, uint32_t h, uint32_t ch) {
    if (!g) return VMN_ERR_INVALID;
    return for_members(g, [&](vmn_dist::Member& mb) { return vmn_scene_set_texture(mb.ctx, slot, data, w, h, ch); });
}

vmn_status vmn_dist_scene_commit(vmn_dist* g) {
    if (!g) return VMN_ERR_INVALID;
    return for_members(g, [&](vmn_dist::Member& mb) { return vmn_scene_commit(mb.ctx); });
}

vmn_status vmn_dist_camera_set(vmn_dist* g, const vmn_camera* cam) {
    if (!g) return VMN_ERR_INVALID;
    return for_members(g, [&](vmn_dist::Member& mb) { return vmn_camera_set(mb.ctx, cam); });
}

vmn_status vmn_dist_render(vmn_dist* g, const vmn_render_params* p, int32_t split, uint32_t tile_size) {
    if (!g || g->m.empty() || !p) return VMN_ERR_INVALID;
    const uint32_t W = g->world;
    if (split != VMN_DIST_SPLIT_SAMPLES && split != VMN_DIST_SPLIT_TILES) return dfail(g, VMN_ERR_INVALID, "bad split");
    if (split == VMN_DIST_SPLIT_SAMPLES && p->mode != VMN_MODE_FIXED && W > 1)
        return dfail(g, VMN_ERR_INVALID, "sample slices need VMN_MODE_FIXED (the adaptive early-out of the other estimators is sequential per pixel): use tiles");
    if (p->tile_size || p->pixel_begin || p->pixel_end) return dfail(g, VMN_ERR_INVALID, "the group chooses the shares: leave tile_size / pixel range zero");
    if (!g->m[0].ctx->have_cam) return dfail(g, VMN_ERR_STATE, "no camera set");
    const vmn_camera cam = g->m[0].ctx->cam;
    const uint32_t tile = tile_size ? tile_size : 64u;
    const uint32_t s_begin = p->sample_begin, s_end = p->sample_end ? p->sample_end : cam.spp;
    if (s_begin > s_end) return dfail(g, VMN_ERR_INVALID, "bad sample range");
    auto* root = g->root();
    if (root) { DCUDA(g, cudaSetDevice(root->device)); DCUDA(g, cudaEventRecord(g->ev[0], root->ctx->stream)); }
    vmn_status s = for_members(g, [&](vmn_dist::Member& mb) {
        vmn_render_params rp = *p;
        if (W > 1 && split == VMN_DIST_SPLIT_TILES) { rp.tile_size = tile; rp.tile_rank = mb.rank; rp.tile_world = W; }
        else if (W > 1) { // slices of [s_begin, s_end) that differ by at most one sample
            const uint32_t len = s_end - s_begin, base = len / W, rem = len % W;
            rp.sample_begin = s_begin + mb.rank * base + std::min(mb.rank, rem);
            rp.sample_end = rp.sample_begin + base + (mb.rank < rem ? 1u : 0u);
            if (rp.sample_end == rp.sample_begin) { // nothing to do for this rank: an empty share still clears its film
                if (!(rp.flags & VMN_RENDER_ACCUMULATE)) cudaMemsetAsync(mb.ctx->film.p, 0, 16 * mb.ctx->film_pixels, mb.ctx->stream);
                mb.ctx->stats.rays_traced = 0; mb.ctx->stats.total_ms = 0; mb.ctx->film_bf = false;
                return VMN_OK;
            }
            if (rp.sample_end == 0) rp.sample_end = rp.sample_begin; // (never: sample_end == 0 means "to spp" in vmn_render)
        }
        return vmn_render(mb.ctx, &rp);
    });
    if (s != VMN_OK) return s;
    if (root) { DCUDA(g, cudaSetDevice(root->device)); DCUDA(g, cudaEventRecord(g->ev[1], root->ctx->stream)); }
    // assemble the film on the root
    if (W > 1) {
        const uint64_t npix = (uint64_t)cam.width * cam.height;
        if (split == VMN_DIST_SPLIT_SAMPLES) {
            // Progressive passes (VMN_RENDER_ACCUMULATE) stay in each rank's own film. A pass that asks for the assembled frame
            // (VMN_DIST_RENDER_ASSEMBLE, or any non-accumulating pass) reduces the films into the root's and then clears the other
            // ranks' films, so that what they hold is never counted twice by a later assembly: the root's film is the running total.
            if (!(p->flags & VMN_RENDER_ACCUMULATE) || (p->flags & VMN_DIST_RENDER_ASSEMBLE)) {
                s = reduce_to_root(g, [](vmn_dist::Member& mb) { return mb.ctx->film.as<float>(); }, 4 * npix);
                if (s != VMN_OK) return s;
                if (!g->use_nccl) { auto* r = g->root(); DCUDA(g, cudaSetDevice(r->device)); DCUDA(g, cudaStreamSynchronize(r->ctx->stream)); } // peer copies read the films on the root's stream
                for (auto& mb : g->m)
                    if (mb.rank != 0) { DCUDA(g, cudaSetDevice(mb.device)); DCUDA(g, cudaMemsetAsync(mb.ctx->film.p, 0, 16 * npix, mb.ctx->stream)); }
            }
        } else {
            s = gather_tiles<4>(g, [](vmn_dist::Member& mb) { return mb.ctx->film.as<float>(); }, cam.width, cam.height, tile);
            if (s == VMN_OK && p->mode == VMN_MODE_BRUTEFORCE)
                s = gather_tiles<2>(g, [](vmn_dist::Member& mb) { return mb.ctx->film_aux.as<float>(); }, cam.width, cam.height, tile);
        }
        if (s != VMN_OK) return s;
    }
    if (root) DCUDA(g, cudaEventRecord(g->ev[2], root->ctx->stream));
    if ((s = sync_all(g)) != VMN_OK) return s;
    g->stats.world = W; g->stats.local = (uint32_t)g->m.size();
    g->stats.rays_traced = 0;
    for (uint32_t i = 0; i < g->m.size() && i < VMN_DIST_MAX_RANKS; ++i) {
        g->stats.render_ms[i] = g->m[i].ctx->stats.total_ms;
        g->stats.rays_traced += g->m[i].ctx->stats.rays_traced;
    }
    if (root) {
        float a = 0.f, b = 0.f;
        cudaEventElapsedTime(&a, g->ev[0], g->ev[1]); cudaEventElapsedTime(&b, g->ev[1], g->ev[2]);
        g->stats.frame_ms = a + b; g->stats.collective_ms = b;
    }
    return VMN_OK;
}

vmn_status vmn_dist_read_image(vmn_dist* g, float* dst, int32_t layout) {
    if (!g) return VMN_ERR_INVALID;
    auto* root = g->root();
    if (!root) return VMN_OK; // only the root holds the assembled film
    DCUDA(g, cudaSetDevice(root->device));
    vmn_status s = vmn_read_image(root->ctx, dst, layout);
    return s == VMN_OK ? s : member_fail(g, *root, s, "vmn_read_image");
}

vmn_status vmn_dist_get_stats(vmn_dist* g, vmn_dist_stats* out) {
    if (!g || !out) return VMN_ERR_INVALID;
    *out = g->stats;
    return VMN_OK;
}

} // extern "C"
