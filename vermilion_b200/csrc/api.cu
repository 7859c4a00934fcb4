// api.cu — the extern "C" surface of libvermilion_b200.so (include/vermilion_b200.h).
#include "engine.cuh"

#include <cstdarg>
#include <cmath>
#include <new>

namespace vmn {

vmn_status fail(vmn_ctx* c, vmn_status s, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    if (c) c->err = buf;
    return s;
}

void dev_free(DevBuf& b) {
    if (b.p) cudaFree(b.p);
    b.p = nullptr; b.bytes = 0;
}

vmn_status dev_alloc(vmn_ctx* c, DevBuf& b, size_t bytes) {
    dev_free(b);
    if (bytes == 0) return VMN_OK;
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) {
        b.p = nullptr;
        cudaGetLastError();
        return fail(c, e == cudaErrorMemoryAllocation ? VMN_ERR_OOM : VMN_ERR_CUDA, "cudaMalloc(%zu): %s", bytes, cudaGetErrorString(e));
    }
    b.bytes = bytes;
    return VMN_OK;
}

vmn_status check_device_error(vmn_ctx* c) {
    if (!c->errflag.p) return VMN_OK;
    uint32_t h = 0;
    VMN_CUDA(c, cudaMemcpyAsync(&h, c->errflag.p, 4, cudaMemcpyDeviceToHost, c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    if (h == 0u) return VMN_OK;
    VMN_CUDA(c, cudaMemsetAsync(c->errflag.p, 0, 4, c->stream));
    return fail(c, VMN_ERR_STATE, "device error word 0x%x:%s the results of this call are unreliable", h,
                (h & VMN_DEVERR_STACK) ? " a traversal stack overflowed (tree deeper than the stack);" : "");
}

vmn_status dev_reserve(vmn_ctx* c, DevBuf& b, size_t bytes) {
    if (b.bytes >= bytes && b.p) return VMN_OK;
    return dev_alloc(c, b, bytes);
}

// glm::rotate (extern/glm/glm/ext/matrix_transform.inl:18-46) composed as pathtracer.cpp:216-221, in host fp32.
// Compiled by nvcc's host compiler without fast-math; column-major m[col*4+row].
static void rotate(float m[16], float angle, float ax, float ay, float az) {
    const float c = std::cos(angle), s = std::sin(angle);
    const float inv = 1.0f / std::sqrt((ax * ax + ay * ay) + az * az);
    const float x = ax * inv, y = ay * inv, z = az * inv;
    const float tx = x * (1.f - c), ty = y * (1.f - c), tz = z * (1.f - c);
    float R[3][3];
    R[0][0] = c + tx * x; R[0][1] = tx * y + s * z; R[0][2] = tx * z - s * y;
    R[1][0] = ty * x - s * z; R[1][1] = c + ty * y; R[1][2] = ty * z + s * x;
    R[2][0] = tz * x + s * y; R[2][1] = tz * y - s * x; R[2][2] = c + tz * z;
    float out[16];
    for (int col = 0; col < 3; ++col)
        for (int r = 0; r < 4; ++r) out[col * 4 + r] = m[0 * 4 + r] * R[col][0] + m[1 * 4 + r] * R[col][1] + m[2 * 4 + r] * R[col][2];
    for (int r = 0; r < 4; ++r) out[12 + r] = m[12 + r];
    memcpy(m, out, sizeof(out));
}

void camera_matrix(const vmn_camera& cam, float m[16]) {
    for (int i = 0; i < 16; ++i) m[i] = (i % 5 == 0) ? 1.f : 0.f;
    rotate(m, cam.rotation[1], 0, 1, 0);
    rotate(m, cam.rotation[0], 1, 0, 0);
    rotate(m, cam.rotation[2], 0, 0, 1);
}

} // namespace vmn

using namespace vmn;

// Optional (env VMN_L2_PERSIST=1; OFF by default): a persisting L2 access-policy window over the arrays every ray walks (leaf triangles +
// wide nodes, one allocation) when they fit the device's persisting carve-out, so that the ray queues — gigabytes per frame, each byte
// touched once — cannot evict them. Measured on C2 (profiles/r02a_ab_c2.txt, 16 spp frame): k_extend 17.48 -> 17.38 ms, but k_raygen
// 0.76 -> 0.95 ms and k_shade 7.69 -> 7.98 ms, because the carve-out is taken from the L2 the other stages' read-modify-write traffic
// lives in: the frame gets 1.5 % SLOWER (26.22 -> 26.61 ms). The evict-first loads / stores on the queue traffic (VMN_STREAM_HINTS,
// traverse.cuh) are kept: k_extend 17.66 -> 17.48 ms at no cost elsewhere.
static void apply_l2_policy(vmn_ctx* c) {
    static const int enabled = [] { const char* e = getenv("VMN_L2_PERSIST"); return e ? atoi(e) : 0; }();
    int max_persist = 0, max_window = 0;
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, c->device);
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, c->device);
    cudaStreamAttrValue v{};
    // (a window over the wide nodes alone, with a hit ratio, for scenes that exceed the carve-out — C3: 107 MB of nodes — measured no
    // gain either: 910.7 against 909.1 Mrays/s, profiles/r02c_ab_c3.txt; removed)
    const bool want = enabled == 1 && c->trav_bytes > 0 && max_persist > 0 && c->trav_bytes <= (uint64_t)max_persist && c->trav_bytes <= (uint64_t)max_window;
    if (want) {
        cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)max_persist);
        v.accessPolicyWindow.base_ptr = c->tris.p;
        v.accessPolicyWindow.num_bytes = (size_t)c->trav_bytes;
        v.accessPolicyWindow.hitRatio = 1.0f;
        v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
        v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
        c->l2_window = cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &v) == cudaSuccess;
    } else if (c->l2_window) {
        v.accessPolicyWindow.num_bytes = 0; // disables the window
        cudaStreamSetAttribute(c->stream, cudaStreamAttributeAccessPolicyWindow, &v);
        cudaCtxResetPersistingL2Cache();
        c->l2_window = false;
    }
    cudaGetLastError();
    c->stats.l2_persist_bytes = c->l2_window ? c->trav_bytes : 0;
}


#define CHECK_CTX(c) do { if (!(c)) return VMN_ERR_INVALID; } while (0)
#define SET_DEVICE(c) VMN_CUDA(c, cudaSetDevice((c)->device))

extern "C" {

int vmn_abi_version(void) { return VMN_ABI_VERSION; }

vmn_status vmn_create(const vmn_config* cfg, vmn_ctx** out) {
    if (!out) return VMN_ERR_INVALID;
    *out = nullptr;
    vmn_ctx* c = new (std::nothrow) vmn_ctx();
    if (!c) return VMN_ERR_OOM;
    *out = c; // returned even on failure so that vmn_last_error can be read; caller still destroys it
    c->device = cfg ? cfg->device : 0;
    c->flags = cfg ? cfg->flags : 0;
    c->max_paths = (cfg && cfg->max_paths) ? cfg->max_paths : (16ull << 20);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(c, VMN_ERR_CUDA, "no CUDA device (%s); this engine has no CPU fallback", e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
    if (c->device < 0 || c->device >= ndev) return fail(c, VMN_ERR_INVALID, "device %d out of range [0,%d)", c->device, ndev);
    SET_DEVICE(c);
    cudaDeviceProp prop;
    VMN_CUDA(c, cudaGetDeviceProperties(&prop, c->device));
    c->sm_count = prop.multiProcessorCount;
    if (cfg && cfg->stream) { c->stream = (cudaStream_t)cfg->stream; c->own_stream = false; }
    else { VMN_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
    for (auto& ev : c->ev) VMN_CUDA(c, cudaEventCreate(&ev));
    vmn_status s = dev_alloc(c, c->errflag, 4);
    if (s != VMN_OK) return s;
    VMN_CUDA(c, cudaMemsetAsync(c->errflag.p, 0, 4, c->stream));
    return VMN_OK;
}

void vmn_destroy(vmn_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    if (c->l2_window) { c->trav_bytes = 0; apply_l2_policy(c); } // a borrowed stream must not keep a window over freed memory
    pool_destroy(c);
    for (DevBuf* b : {&c->pos, &c->nrm, &c->uv, &c->tris, &c->tris_sorted, &c->tris_orig, &c->shade, &c->b_min, &c->b_max,
                      &c->b_child, &c->spheres, &c->tex, &c->tex1, &c->film, &c->film_aux, &c->q_in, &c->q_out, &c->counters, &c->errflag})
        dev_free(*b);
    for (DevBuf& b : c->build_scratch) dev_free(b);
    for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

const char* vmn_last_error(const vmn_ctx* c) { return c ? c->err.c_str() : "null context"; }

vmn_status vmn_set_stream(vmn_ctx* c, void* s) {
    CHECK_CTX(c);
    SET_DEVICE(c);
    if (c->stream) VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    const uint64_t tb = c->trav_bytes;
    if (c->l2_window) { c->trav_bytes = 0; apply_l2_policy(c); c->trav_bytes = tb; }
    if (c->own_stream && c->stream) cudaStreamDestroy(c->stream);
    c->own_stream = false;
    if (s) c->stream = (cudaStream_t)s;
    else { VMN_CUDA(c, cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); c->own_stream = true; }
    if (c->committed) apply_l2_policy(c);
    return VMN_OK;
}

vmn_status vmn_synchronize(vmn_ctx* c) {
    CHECK_CTX(c);
    SET_DEVICE(c);
    return check_device_error(c); // synchronises; reports a traversal error raised by an earlier asynchronous (_device) call
}

static vmn_status upload(vmn_ctx* c, DevBuf& b, const void* src, size_t bytes) {
    vmn_status s = dev_reserve(c, b, bytes); // grow-only: re-uploading a same-sized scene does not reallocate
    if (s != VMN_OK) return s;
    if (bytes) VMN_CUDA(c, cudaMemcpyAsync(b.p, src, bytes, cudaMemcpyHostToDevice, c->stream));
    return VMN_OK;
}

vmn_status vmn_scene_set_triangles(vmn_ctx* c, const float* pos, const float* nrm, const float* uv, uint64_t n) {
    CHECK_CTX(c);
    if (n && !pos) return fail(c, VMN_ERR_INVALID, "pos is NULL");
    SET_DEVICE(c);
    c->committed = false;
    c->geometry_built = false;
    c->n_tris = 0;
    vmn_status s;
    if ((s = upload(c, c->pos, pos, 36ull * n)) != VMN_OK) return s;
    c->have_nrm = nrm != nullptr; c->have_uv = uv != nullptr;
    // The normals and uvs (60 of the 96 bytes per triangle) follow the positions on a second stream while the main stream already builds
    // the BVH from the positions (with pinned host arrays the two overlap: -1 ms per C2 scene load; pageable arrays are staged by the
    // driver and simply serialise as before). The event keeps the two uploads from sharing the link, which would only delay the build.
    const bool overlap = load_overlap();
    cudaStream_t attr_stream = c->stream;
    if ((nrm || uv) && overlap) {
        if (!c->copy_stream) VMN_CUDA(c, cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        VMN_CUDA(c, cudaEventRecord(c->ev[6], c->stream));
        VMN_CUDA(c, cudaStreamWaitEvent(c->copy_stream, c->ev[6], 0));
        attr_stream = c->copy_stream;
    }
    if (nrm && (s = dev_reserve(c, c->nrm, 36ull * n)) != VMN_OK) return s;
    if (uv && (s = dev_reserve(c, c->uv, 24ull * n)) != VMN_OK) return s;
    if (nrm && n) VMN_CUDA(c, cudaMemcpyAsync(c->nrm.p, nrm, 36ull * n, cudaMemcpyHostToDevice, attr_stream));
    if (uv && n) VMN_CUDA(c, cudaMemcpyAsync(c->uv.p, uv, 24ull * n, cudaMemcpyHostToDevice, attr_stream));
    c->n_tris = n;
    if (overlap) {
        c->geometry_status = build_geometry(c); // synchronises the main stream; a failure is reported by vmn_scene_commit
        c->geometry_err = c->geometry_status == VMN_OK ? std::string() : c->err;
        c->geometry_built = true;
    }
    if (c->copy_stream) VMN_CUDA(c, cudaStreamSynchronize(c->copy_stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream)); // host pointers are borrowed for the call only
    return VMN_OK;
}

vmn_status vmn_scene_set_spheres(vmn_ctx* c, const vmn_sphere* sp, uint32_t n) {
    CHECK_CTX(c);
    if (n > kMaxSpheres || (n && !sp)) return fail(c, VMN_ERR_INVALID, "at most 16 spheres");
    SET_DEVICE(c);
    SphereDev dev[kMaxSpheres];
    memset(dev, 0, sizeof(dev));
    for (uint32_t i = 0; i < n; ++i) {
        const vmn_sphere& a = sp[i];
        SphereDev& d = dev[i];
        d.cx = a.center[0]; d.cy = a.center[1]; d.cz = a.center[2];
        const bool emitter = a.emission[0] != 0.f || a.emission[1] != 0.f || a.emission[2] != 0.f;
        d.flags = (a.flags & VMN_SPH_FLIP_NORMAL) | (emitter ? VMN_SPHD_EMITTER : 0u);
        const volatile float r2f = a.radius * a.radius; // one float rounding, as `rad * rad` in sphereIntersect (meshEngine.cpp:187)
        d.r2 = (double)r2f;
        d.r = a.radius;
        d.ex = a.emission[0]; d.ey = a.emission[1]; d.ez = a.emission[2];
        d.nx = a.normal_center[0]; d.ny = a.normal_center[1]; d.nz = a.normal_center[2];
    }
    vmn_status s = upload(c, c->spheres, dev, sizeof(SphereDev) * (size_t)n);
    if (s != VMN_OK) return s;
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    c->n_spheres = n;
    c->scene.n_spheres = n;
    return VMN_OK;
}

void vmn_reference_spheres(vmn_sphere out[8]) {
    memset(out, 0, sizeof(vmn_sphere) * 8);
    auto set = [&](int i, float cx, float cy, float cz, float r, float ex, float ey, float ez, uint32_t flags, float nx, float ny, float nz) {
        out[i].center[0] = cx; out[i].center[1] = cy; out[i].center[2] = cz; out[i].radius = r;
        out[i].emission[0] = ex; out[i].emission[1] = ey; out[i].emission[2] = ez; out[i].flags = flags;
        out[i].normal_center[0] = nx; out[i].normal_center[1] = ny; out[i].normal_center[2] = nz;
    };
    // meshEngine.cpp:377-387 and :410-420 — emitters; normals are measured from other points than the centres
    set(0, 15.f, 140.f, 25.f, 3.5f, 0.f * 15.f, .5f * 15.f, 1.0f * 15.f, VMN_SPH_FLIP_NORMAL, -55.f, 350.f, -150.f);
    set(1, 0.f, 3300.f, 1300.f, 250.f, 1.0f * 15.2f, 1.0f * 15.2f, 1.0f * 15.2f, 0u, 500.f, 800.f, 1300.f);
    // meshEngine.cpp:424-500 — six radius-5e7 walls; centres computed in double, narrowed by the vec3 constructor
    const double S = 1e7 * 5;
    const float R = (float)S;
    set(2, 0.f, (float)-S, 0.f, R, 0, 0, 0, 0u, 0.f, (float)-S, 0.f);
    set(3, 0.f, (float)(S + 1000), 0.f, R, 0, 0, 0, 0u, 0.f, (float)(S + 1000), 0.f);
    set(4, (float)(-S + 2000), 0.f, 0.f, R, 0, 0, 0, VMN_SPH_FLIP_NORMAL, (float)(-S + 2000), 0.f, 0.f);
    set(5, (float)(S - 2000), 0.f, 0.f, R, 0, 0, 0, VMN_SPH_FLIP_NORMAL, (float)(S - 2000), 0.f, 0.f);
    set(6, 0.f, 0.f, (float)(-S + 2000), R, 0, 0, 0, VMN_SPH_FLIP_NORMAL, 0.f, 0.f, (float)(-S + 2000));
    set(7, 0.f, 0.f, (float)(S - 2000), R, 0, 0, 0, 0u, 0.f, 0.f, (float)(S - 2000));
}

static void scene_bind_textures(vmn_ctx* c) {
    c->scene.tex = c->have_tex ? c->tex.as<float>() : nullptr;
    c->scene.tex_w = c->tex_w; c->scene.tex_h = c->tex_h; c->scene.tex_c = c->tex_c;
    c->scene.tex1 = c->have_tex1 ? c->tex1.as<float>() : nullptr;
    c->scene.tex1_w = c->tex1_w; c->scene.tex1_h = c->tex1_h; c->scene.tex1_c = c->tex1_c;
}

vmn_status vmn_scene_set_texture(vmn_ctx* c, uint32_t slot, const float* data, uint32_t w, uint32_t h, uint32_t ch) {
    CHECK_CTX(c);
    if (slot > 1) return fail(c, VMN_ERR_INVALID, "texture slot %u: only slots 0 (pathtracer.cpp:62-80, integrators.cpp:98-106) and 1 (integrators.cpp:141-147) are sampled", slot);
    if (!data) { // unbind
        if (slot == 0) c->have_tex = false; else c->have_tex1 = false;
        scene_bind_textures(c);
        return VMN_OK;
    }
    if (!w || !h || ch < 1 || ch > 4) return fail(c, VMN_ERR_INVALID, "bad texture");
    SET_DEVICE(c);
    vmn_status s = upload(c, slot == 0 ? c->tex : c->tex1, data, sizeof(float) * (size_t)w * h * ch);
    if (s != VMN_OK) return s;
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    if (slot == 0) { c->tex_w = w; c->tex_h = h; c->tex_c = ch; c->have_tex = true; }
    else { c->tex1_w = w; c->tex1_h = h; c->tex1_c = ch; c->have_tex1 = true; }
    scene_bind_textures(c);
    return VMN_OK;
}

vmn_status vmn_scene_commit(vmn_ctx* c) {
    CHECK_CTX(c);
    SET_DEVICE(c);
    vmn_status s;
    if (c->geometry_built) { // vmn_scene_set_triangles built it beside the upload
        c->geometry_built = false;
        if (c->geometry_status != VMN_OK) return fail(c, c->geometry_status, "%s", c->geometry_err.c_str());
        s = build_records(c);
    } else {
        s = build_bvh(c);
    }
    if (s != VMN_OK) return s;
    apply_l2_policy(c);
    c->scene.n_spheres = c->n_spheres;
    scene_bind_textures(c);
    // the staged arrays stay allocated (reused by the next vmn_scene_set_triangles); they are 96 B/triangle
    c->committed = true;
    return VMN_OK;
}

} // extern "C"

// ---- ray queries ----------------------------------------------------------------------------------
#define NEED_COMMIT(c) do { if (!(c)->committed) return fail((c), VMN_ERR_STATE, "scene not committed (call vmn_scene_commit)"); } while (0)

namespace {
// stage host rays on the device in the context's scratch buffers
vmn_status stage_rays(vmn_ctx* c, const vmn_ray* rays, uint64_t n, size_t out_bytes_per_ray) {
    vmn_status s;
    if ((s = dev_reserve(c, c->q_in, sizeof(vmn_ray) * n)) != VMN_OK) return s;
    if ((s = dev_reserve(c, c->q_out, out_bytes_per_ray * n)) != VMN_OK) return s;
    if (n) VMN_CUDA(c, cudaMemcpyAsync(c->q_in.p, rays, sizeof(vmn_ray) * n, cudaMemcpyHostToDevice, c->stream));
    return VMN_OK;
}
vmn_status fetch(vmn_ctx* c, void* dst, size_t bytes) {
    if (bytes) VMN_CUDA(c, cudaMemcpyAsync(dst, c->q_out.p, bytes, cudaMemcpyDeviceToHost, c->stream));
    return check_device_error(c); // synchronises the stream
}
vmn_status read_counters(vmn_ctx* c) {
    if ((c->flags & VMN_CFG_COUNTERS) && c->counters.p) {
        unsigned long long h[4] = {0, 0, 0, 0};
        VMN_CUDA(c, cudaMemcpy(h, c->counters.p, sizeof(h), cudaMemcpyDeviceToHost));
        c->stats.nodes_visited = h[2]; c->stats.tris_tested = h[3];
    }
    return VMN_OK;
}
} // namespace

extern "C" {

vmn_status vmn_intersect_device(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, vmn_hit* d_out, uint32_t flags) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (n && (!d_rays || !d_out)) return fail(c, VMN_ERR_INVALID, "null ray/hit pointer");
    SET_DEVICE(c);
    return launch_intersect(c, d_rays, n, d_out, flags);
}

vmn_status vmn_intersect(vmn_ctx* c, const vmn_ray* rays, uint64_t n, vmn_hit* out, uint32_t flags) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (n && (!rays || !out)) return fail(c, VMN_ERR_INVALID, "null ray/hit pointer");
    SET_DEVICE(c);
    vmn_status s;
    if ((s = stage_rays(c, rays, n, sizeof(vmn_hit))) != VMN_OK) return s;
    if ((s = launch_intersect(c, c->q_in.as<vmn_ray>(), n, c->q_out.as<vmn_hit>(), flags)) != VMN_OK) return s;
    if ((s = fetch(c, out, sizeof(vmn_hit) * n)) != VMN_OK) return s;
    if (n) { float ms = 0.f; cudaEventElapsedTime(&ms, c->ev[2], c->ev[3]); c->stats.extend_ms = ms; c->stats.total_ms = ms; }
    return read_counters(c);
}

vmn_status vmn_occluded(vmn_ctx* c, const vmn_ray* rays, uint64_t n, uint8_t* out) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (n && (!rays || !out)) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    vmn_status s;
    if ((s = stage_rays(c, rays, n, 1)) != VMN_OK) return s;
    if ((s = launch_occluded(c, c->q_in.as<vmn_ray>(), n, c->q_out.as<uint8_t>())) != VMN_OK) return s;
    return fetch(c, out, n);
}

vmn_status vmn_raycast(vmn_ctx* c, const vmn_ray* rays, uint64_t n, vmn_raycast_rec* out) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (n && (!rays || !out)) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    vmn_status s;
    if ((s = stage_rays(c, rays, n, sizeof(vmn_raycast_rec))) != VMN_OK) return s;
    if ((s = launch_raycast(c, c->q_in.as<vmn_ray>(), n, c->q_out.as<vmn_raycast_rec>())) != VMN_OK) return s;
    return fetch(c, out, sizeof(vmn_raycast_rec) * n);
}

vmn_status vmn_radiance(vmn_ctx* c, const vmn_ray* rays, uint64_t n, uint64_t seed, int32_t mode, uint32_t max_segments, float* out,
                        uint32_t* rays_cast) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (n && (!rays || !out)) return fail(c, VMN_ERR_INVALID, "null pointer");
    if (mode != VMN_MODE_COMPAT && mode != VMN_MODE_FIXED) return fail(c, VMN_ERR_INVALID, "bad mode %d", mode);
    SET_DEVICE(c);
    vmn_status s;
    if ((s = stage_rays(c, rays, n, 16 + 4)) != VMN_OK) return s;
    float* d_out = c->q_out.as<float>();
    uint32_t* d_cnt = reinterpret_cast<uint32_t*>(d_out + 4 * n);
    if ((s = launch_radiance(c, c->q_in.as<vmn_ray>(), n, seed, mode, max_segments ? max_segments : 5, d_out, d_cnt)) != VMN_OK) return s;
    if (n) VMN_CUDA(c, cudaMemcpyAsync(out, d_out, 16 * n, cudaMemcpyDeviceToHost, c->stream));
    if (n && rays_cast) VMN_CUDA(c, cudaMemcpyAsync(rays_cast, d_cnt, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    return check_device_error(c); // synchronises the stream
}

// ---- camera, film, render -------------------------------------------------------------------------
vmn_status vmn_camera_set(vmn_ctx* c, const vmn_camera* cam) {
    CHECK_CTX(c);
    if (!cam || !cam->width || !cam->height) return fail(c, VMN_ERR_INVALID, "bad camera");
    SET_DEVICE(c);
    c->cam = *cam;
    camera_matrix(c->cam, c->cam_m);
    c->have_cam = true;
    const uint64_t npix = (uint64_t)cam->width * cam->height;
    if (npix != c->film_pixels) {
        vmn_status s = dev_alloc(c, c->film, 16 * npix);
        if (s != VMN_OK) return s;
        c->film_pixels = npix;
    }
    VMN_CUDA(c, cudaMemsetAsync(c->film.p, 0, 16 * npix, c->stream));
    c->film_bf = false;
    return VMN_OK;
}

vmn_status vmn_camera_rays(vmn_ctx* c, const uint64_t* pixel, const uint32_t* sx, const uint32_t* sy, const float* jx, const float* jy,
                           uint64_t n, vmn_ray* out) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (n && (!pixel || !sx || !sy || !jx || !jy || !out)) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    vmn_status s;
    if ((s = dev_reserve(c, c->q_in, 24 * n)) != VMN_OK) return s;
    if ((s = dev_reserve(c, c->q_out, sizeof(vmn_ray) * n)) != VMN_OK) return s;
    char* base = c->q_in.as<char>();
    uint64_t* d_px = reinterpret_cast<uint64_t*>(base);
    uint32_t* d_sx = reinterpret_cast<uint32_t*>(base + 8 * n);
    uint32_t* d_sy = d_sx + n;
    float* d_jx = reinterpret_cast<float*>(d_sy + n);
    float* d_jy = d_jx + n;
    if (n) {
        VMN_CUDA(c, cudaMemcpyAsync(d_px, pixel, 8 * n, cudaMemcpyHostToDevice, c->stream));
        VMN_CUDA(c, cudaMemcpyAsync(d_sx, sx, 4 * n, cudaMemcpyHostToDevice, c->stream));
        VMN_CUDA(c, cudaMemcpyAsync(d_sy, sy, 4 * n, cudaMemcpyHostToDevice, c->stream));
        VMN_CUDA(c, cudaMemcpyAsync(d_jx, jx, 4 * n, cudaMemcpyHostToDevice, c->stream));
        VMN_CUDA(c, cudaMemcpyAsync(d_jy, jy, 4 * n, cudaMemcpyHostToDevice, c->stream));
    }
    if ((s = launch_camera_rays(c, d_px, d_sx, d_sy, d_jx, d_jy, n, c->q_out.as<vmn_ray>())) != VMN_OK) return s;
    return fetch(c, out, sizeof(vmn_ray) * n);
}

vmn_status vmn_render(vmn_ctx* c, const vmn_render_params* p) {
    CHECK_CTX(c);
    NEED_COMMIT(c);
    if (!p) return fail(c, VMN_ERR_INVALID, "null params");
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (p->mode != VMN_MODE_COMPAT && p->mode != VMN_MODE_FIXED && p->mode != VMN_MODE_BRUTEFORCE) return fail(c, VMN_ERR_INVALID, "bad mode %d", p->mode);
    SET_DEVICE(c);
    vmn_status s;
    if (p->mode == VMN_MODE_BRUTEFORCE) s = render_bruteforce(c, p);
    else { c->film_bf = false; s = render_frame(c, p); }
    return s != VMN_OK ? s : check_device_error(c);
}

vmn_status vmn_film_device(vmn_ctx* c, float** d_sum, uint64_t* n_floats) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (d_sum) *d_sum = c->film.as<float>();
    if (n_floats) *n_floats = 4 * c->film_pixels;
    return VMN_OK;
}

vmn_status vmn_film_upload(vmn_ctx* c, const float* sum) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (!sum) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    VMN_CUDA(c, cudaMemcpyAsync(c->film.p, sum, 16 * c->film_pixels, cudaMemcpyHostToDevice, c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream)); // the host pointer is borrowed for the call only
    c->film_bf = false;
    return VMN_OK;
}

vmn_status vmn_film_aux_device(vmn_ctx* c, float** d_aux, uint64_t* n_floats) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (!c->film_bf) return fail(c, VMN_ERR_STATE, "the film does not hold a BRUTEFORCE render");
    if (d_aux) *d_aux = c->film_aux.as<float>();
    if (n_floats) *n_floats = 2 * c->film_pixels;
    return VMN_OK;
}

vmn_status vmn_film_resolve_device(vmn_ctx* c, const float* d_sum, float* dst) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (!d_sum || !dst) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    vmn_status s = dev_reserve(c, c->q_out, 20 * c->film_pixels);
    if (s != VMN_OK) return s;
    if ((s = film_resolve(c, d_sum, c->film_bf ? c->film_aux.as<float>() : nullptr, c->q_out.as<float>(), c->film_pixels)) != VMN_OK) return s;
    return fetch(c, dst, 20 * c->film_pixels);
}

vmn_status vmn_read_image(vmn_ctx* c, float* dst, int32_t layout) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (!dst) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    if (layout == VMN_FILM_SUM) {
        VMN_CUDA(c, cudaMemcpyAsync(dst, c->film.p, 16 * c->film_pixels, cudaMemcpyDeviceToHost, c->stream));
        VMN_CUDA(c, cudaStreamSynchronize(c->stream));
        return VMN_OK;
    }
    if (layout != VMN_FILM_RGBAZ) return fail(c, VMN_ERR_INVALID, "bad layout %d", layout);
    return vmn_film_resolve_device(c, c->film.as<float>(), dst);
}

vmn_status vmn_read_image_rgba8(vmn_ctx* c, uint8_t* rgba8, float* depth) {
    CHECK_CTX(c);
    if (!c->have_cam) return fail(c, VMN_ERR_STATE, "no camera set");
    if (!rgba8) return fail(c, VMN_ERR_INVALID, "null pointer");
    SET_DEVICE(c);
    const uint64_t npix = c->film_pixels;
    vmn_status s = dev_reserve(c, c->q_out, 8 * npix);
    if (s != VMN_OK) return s;
    uint8_t* d_rgba = c->q_out.as<uint8_t>();
    float* d_depth = reinterpret_cast<float*>(d_rgba + 4 * npix);
    if ((s = film_pack_rgba8(c, c->film.as<float>(), c->film_bf ? c->film_aux.as<float>() : nullptr, d_rgba, depth ? d_depth : nullptr, npix)) != VMN_OK) return s;
    VMN_CUDA(c, cudaMemcpyAsync(rgba8, d_rgba, 4 * npix, cudaMemcpyDeviceToHost, c->stream));
    if (depth) VMN_CUDA(c, cudaMemcpyAsync(depth, d_depth, 4 * npix, cudaMemcpyDeviceToHost, c->stream));
    VMN_CUDA(c, cudaStreamSynchronize(c->stream));
    return VMN_OK;
}

vmn_status vmn_get_stats(vmn_ctx* c, vmn_stats* out) {
    CHECK_CTX(c);
    if (!out) return VMN_ERR_INVALID;
    *out = c->stats;
    return VMN_OK;
}

} // extern "C"
