// engine.cuh — the context behind the C ABI (include/vermilion_b200.h) and the internal launch API.
#pragma once
#include "common.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

namespace vmn {

// device buffer owned by the context
struct DevBuf {
    void* p = nullptr;
    size_t bytes = 0;
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct PathPool; // wavefront state (wavefront.cu)

} // namespace vmn

struct vmn_ctx {
    int device = 0;
    uint32_t flags = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    int sm_count = VMN_SM_COUNT_DEFAULT;
    std::string err;

    // host-provided scene, staged on the device until commit
    vmn::DevBuf pos, nrm, uv;
    uint64_t n_tris = 0;
    bool have_nrm = false, have_uv = false;
    bool committed = false;
    // vmn_scene_set_triangles has already run build_geometry on these triangles (its status and message are kept for vmn_scene_commit,
    // which reports a failed build exactly as it did when it ran the whole build itself)
    bool geometry_built = false;
    vmn_status geometry_status = VMN_OK;
    std::string geometry_err;
    cudaStream_t copy_stream = nullptr; // normals / uvs upload, behind the positions, beside the geometry build

    // built scene
    vmn::DevBuf tris, tris_sorted, tris_orig, shade, b_min, b_max, b_child; // tris: leaf triangles + wide nodes (one allocation)
    uint64_t trav_bytes = 0;   // bytes of `tris` the traversal reads (leaf triangles + the wide nodes in use)
    bool l2_window = false;    // an L2 persisting access-policy window is set on `stream`
    vmn::DevBuf spheres, tex, tex1;
    vmn::DevBuf build_scratch[16]; // bvh_build.cu, grow-only
    uint32_t n_spheres = 0, tex_w = 0, tex_h = 0, tex_c = 0, tex1_w = 0, tex1_h = 0, tex1_c = 0;
    bool have_tex = false, have_tex1 = false;
    uint64_t n_wide = 0;
    vmn::SceneDev scene{};

    // camera / film
    vmn_camera cam{};
    bool have_cam = false;
    float cam_m[16];
    vmn::DevBuf film;      // float4 (sum r,g,b, n) per pixel
    vmn::DevBuf film_aux;  // BRUTEFORCE only: float2 (sum alpha, last primary hit distance) per pixel
    bool film_bf = false;  // the film holds a BRUTEFORCE render (resolve: alpha = sum alpha / n, depth = distance)
    uint64_t film_pixels = 0;

    // wavefront pool
    uint64_t max_paths = 0;
    vmn::PathPool* pool = nullptr;

    // scratch for the ray-query entry points
    vmn::DevBuf q_in, q_out, counters;
    vmn::DevBuf errflag;   // one uint32 of VMN_DEVERR_* bits set by kernels (SceneDev::err)

    vmn_stats stats{};
    cudaEvent_t ev[8] = {};
};

namespace vmn {

vmn_status fail(vmn_ctx* c, vmn_status s, const char* fmt, ...);
vmn_status dev_alloc(vmn_ctx* c, DevBuf& b, size_t bytes);       // (re)allocate exactly
vmn_status dev_reserve(vmn_ctx* c, DevBuf& b, size_t bytes);     // grow only
void dev_free(DevBuf& b);
vmn_status check_device_error(vmn_ctx* c); // synchronises the stream; VMN_ERR_STATE if a kernel raised a VMN_DEVERR_* bit (and clears it)

#define VMN_CUDA(c, expr)                                                                                   \
    do {                                                                                                    \
        cudaError_t e__ = (expr);                                                                           \
        if (e__ != cudaSuccess) return vmn::fail((c), VMN_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e__)); \
    } while (0)

// bvh_build.cu
// VMN_LOAD_OVERLAP=0 (read at every scene load): upload everything first and build in vmn_scene_commit, as before r02x (A/B switch)
inline bool load_overlap() { const char* e = getenv("VMN_LOAD_OVERLAP"); return !(e && e[0] == '0'); }
vmn_status build_bvh(vmn_ctx* c);      // = build_geometry + build_records
vmn_status build_geometry(vmn_ctx* c); // positions only
vmn_status build_records(vmn_ctx* c);  // + normals, uvs

// traverse.cu
uint32_t refill_min(const vmn_ctx* c); // idle lanes that trigger a refill in the persistent kernels (by scene size; env VMN_REFILL overrides)
vmn_status launch_intersect(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, vmn_hit* d_out, uint32_t flags);
vmn_status launch_occluded(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, uint8_t* d_out);
vmn_status launch_raycast(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, vmn_raycast_rec* d_out);

// wavefront.cu
vmn_status launch_radiance(vmn_ctx* c, const vmn_ray* d_rays, uint64_t n, uint64_t seed, int mode, uint32_t max_segments,
                           float* d_out, uint32_t* d_rays_cast);
vmn_status launch_camera_rays(vmn_ctx* c, const uint64_t* d_pixel, const uint32_t* d_sx, const uint32_t* d_sy, const float* d_jx,
                              const float* d_jy, uint64_t n, vmn_ray* d_out);
vmn_status render_frame(vmn_ctx* c, const vmn_render_params* p);
vmn_status render_bruteforce(vmn_ctx* c, const vmn_render_params* p);
// d_aux: BRUTEFORCE second plane or nullptr (PathTracer film: alpha 1, depth = samples)
vmn_status film_resolve(vmn_ctx* c, const float* d_sum, const float* d_aux, float* d_rgbaz, uint64_t npix);
vmn_status film_pack_rgba8(vmn_ctx* c, const float* d_sum, const float* d_aux, uint8_t* d_rgba, float* d_depth, uint64_t npix);
vmn_status make_pixel_map(vmn_ctx* c, const vmn_render_params* prm, PixelMap* out, uint64_t* n_local_total); // this context's share of the frame
void pool_destroy(vmn_ctx* c);
void camera_matrix(const vmn_camera& cam, float m[16]);

} // namespace vmn
