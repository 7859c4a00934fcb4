/*
 * vermilion_b200.h — C ABI of libvermilion_b200.so, the B200 (sm_100a) engine for
 * Vermilion's ray-tracing hot path.
 *
 * The reference (CasperTheCat/Vermilion) has no FFI or plugin registry; the seam the
 * path sits behind is the C++ virtual interface Vermilion::Integrator::Render
 * (core/integrators/integrators.h:11-16) plus the accelerator seam
 * BVH::getIntersection (core/accelerators/bvh.h:29-30) and MeshEngine::RayCast
 * (core/engines/meshEngine.h:113-114). Each entry point below names the reference
 * interface it replaces. The C++ adapter vermilion_b200/host/b200_pathtracer.{h,cpp}
 * (class B200PathTracer : Vermilion::Integrator) is the reference-side binding; see
 * INTEGRATION.md.
 *
 * Conventions: every call returns vmn_status (0 = OK); no exceptions or callbacks
 * cross the ABI; host pointers are borrowed for the duration of the call only;
 * `_device` variants take CUDA device pointers valid on the context's device; the
 * context owns all other device memory. One context drives ONE device (multi-GPU =
 * one process/context per GPU, the film buffer exchanged by the host with NCCL). A
 * context is not thread-safe. There is no CPU fallback: without a CUDA device
 * vmn_create fails with VMN_ERR_CUDA.
 *
 * All file:line citations are into the reference tree.
 */
#ifndef VERMILION_B200_H
#define VERMILION_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built -fvisibility=hidden; these are its only exports */
#endif

#define VMN_ABI_VERSION 2

typedef enum {
    VMN_OK = 0,
    VMN_ERR_INVALID = 1, /* bad argument */
    VMN_ERR_CUDA = 2,    /* CUDA runtime / launch failure, or no device */
    VMN_ERR_OOM = 3,     /* device or host allocation failed */
    VMN_ERR_NCCL = 5,    /* a collective failed, or libnccl.so.2 could not be loaded (vmn_dist_* only) */
    VMN_ERR_STATE = 4    /* call order violated (e.g. render before commit), or a device-side consistency check failed
                          * (a traversal stack overflowed: the results of that call are unreliable) */
} vmn_status;

typedef struct vmn_ctx vmn_ctx;

typedef struct {
    int32_t device;        /* CUDA device ordinal */
    uint32_t flags;        /* VMN_CFG_* */
    uint64_t max_paths;    /* wavefront pool size (paths in flight, 112 B each); 0 = default 16 Mi */
    void* stream;          /* cudaStream_t to launch on; NULL = a stream the context creates */
} vmn_config;
#define VMN_CFG_COUNTERS 1u /* keep per-ray node/triangle visit counters (slower; for B_ray) */

/* 32-byte ray = core/accelerators/Ray.h:4-11 without inv_d. pads are ignored. */
typedef struct { float o[3]; float pad0; float d[3]; float pad1; } vmn_ray;

/* 16-byte closest-hit record = IntersectionInfo (core/accelerators/IntersectionInfo.h:6-10)
 * with the Object* replaced by the canonical triangle id (running face index of
 * MeshEngine::createBVH, meshEngine.cpp:660-717) and (u,v) = the barycentric weights
 * (w1,w2) Triangle::getNormal derives from the hit point (triangle.cpp:67-78).
 * Miss: tri = VMN_MISS, t = 999999999.f (bvh.cpp:48), u = v = 0. */
typedef struct { float t; float u; float v; uint32_t tri; } vmn_hit;
#define VMN_MISS 0xFFFFFFFFu

/* 64-byte record of every output of MeshEngine::RayCast (meshEngine.cpp:239-509). */
#define VMN_RC_HIT 1u      /* RayCast's return value */
#define VMN_RC_MATERIAL 2u /* *ppImpactMaterial != nullptr (a triangle lies on the ray, :502-503) */
typedef struct {
    float P[3];
    float dist;
    float N[3];
    uint32_t flags;
    float uv[2];
    float Le[3];
    uint32_t tri;
    uint32_t pad[2];
} vmn_raycast_rec;

/* One analytic sphere of MeshEngine::RayCast (meshEngine.cpp:377-500). The reference
 * hard-codes eight; vmn_reference_spheres() returns them in its test order. */
#define VMN_SPH_FLIP_NORMAL 1u /* N = -normalize(P - normal_center) */
typedef struct {
    float center[3];
    float radius;
    float emission[3];
    uint32_t flags;
    float normal_center[3]; /* the reference measures emitter normals from other points (:385,:418) */
    float pad;
} vmn_sphere;

/* Camera as Vermilion::Camera holds it after its constructor (camera.cpp:34-56):
 * rotation in radians with x,y already negated. */
typedef struct {
    uint32_t width, height;
    float position[3];
    float rotation[3];
    float dist_to_film;
    float sensor_x, sensor_y;
    uint32_t spp; /* uSamplesPerPixel */
} vmn_camera;

enum { VMN_MODE_COMPAT = 0, /* PathTracer, the reference estimator exactly: adaptive early-out, r2 = 10*xi, roulette */
       VMN_MODE_FIXED = 1,  /* PathTracer, exact spp, r2 = xi, at most max_segments rays per path */
       VMN_MODE_BRUTEFORCE = 2 }; /* BruteForceTracer::Render (core/integrators/integrators.cpp:9-186), the RenderEngine default
                                   * integrator (renderEngine.cpp:51): point light at (500,1100,2000), normal map = texture slot 0,
                                   * albedo = texture slot 1, one mirror-direction probe, its own early-out. spp <= 65535 (the
                                   * reference's uint16_t sample counter never terminates above that). */

typedef struct {
    int32_t mode;           /* VMN_MODE_* */
    uint32_t max_segments;  /* FIXED only; 0 = 5 (primary + 4 bounces) */
    uint64_t seed;          /* Philox key */
    uint64_t pixel_begin, pixel_end; /* pixel range [begin,end) this context renders; end = 0 -> width*height */
    uint32_t sample_begin, sample_end; /* FIXED: sample slice [begin,end) of the spp-long sequence; end = 0 -> spp */
    uint32_t tile_size;     /* 0 = contiguous pixel range; else pixel_begin/end index interleaved tiles (see vmn_render) */
    uint32_t tile_rank, tile_world; /* tile t belongs to this context iff t % tile_world == tile_rank */
    uint32_t flags;         /* VMN_RENDER_* */
} vmn_render_params;
#define VMN_RENDER_ACCUMULATE 1u /* keep the film and add to it (progressive rendering: successive sample slices, FIXED mode) */
#define VMN_RENDER_BF_SHADOW 2u  /* BRUTEFORCE: enable the shadow ray the reference has commented out (integrators.cpp:108-110):
                                  * if RayCastCollision(hitLocation, lightDirection) (meshEngine.cpp:196-207) then vNDL *= 0 */

typedef struct {
    uint64_t n_triangles, n_nodes, n_wide_nodes;
    uint64_t node_bytes, tri_bytes;   /* bytes of one node / triangle record as laid out in HBM */
    double build_ms;                  /* last BVH build (vmn_scene_set_triangles + vmn_scene_commit), device time */
    uint64_t rays_traced;             /* extend-stage rays of the last render / intersect call */
    uint64_t paths_traced;            /* samples finished by the last render */
    uint64_t nodes_visited, tris_tested; /* only with VMN_CFG_COUNTERS */
    double extend_ms, shade_ms, raygen_ms, resolve_ms, total_ms; /* device time of the last render, per stage */
    uint64_t kernel_launches;         /* kernels launched by the last render / intersect call */
    uint64_t wide_depth;              /* levels of the BVH8 built by the last vmn_scene_commit (<= 128, the traversal stack) */
    uint64_t l2_persist_bytes;        /* bytes of nodes + leaf triangles under an L2 persisting access-policy window (0 = none: the
                                       * arrays exceed the device's persisting carve-out, or VMN_L2_PERSIST=0) */
} vmn_stats;

/* lifetime (replaces `new PathTracer()` + RenderEngine::assignIntegrator, main.cpp:60-63) */
int vmn_abi_version(void);
vmn_status vmn_create(const vmn_config* cfg, vmn_ctx** out);
void vmn_destroy(vmn_ctx* ctx);
const char* vmn_last_error(const vmn_ctx* ctx);
vmn_status vmn_set_stream(vmn_ctx* ctx, void* cuda_stream);
vmn_status vmn_synchronize(vmn_ctx* ctx);

/* scene (replaces MeshEngine::createBVH, meshEngine.cpp:649-724, and `new BVH(tris)`, bvh.cpp:155-279).
 * pos: 9 floats per triangle; nrm: 9 per triangle or NULL (zeros); uv: 6 per triangle or NULL.
 * The arrays are borrowed for the call only. The call already builds the position-only part of the BVH while the normals and uvs are
 * still in flight (pinned host arrays: the two overlap); vmn_scene_commit finishes the build and reports its errors. */
vmn_status vmn_scene_set_triangles(vmn_ctx* ctx, const float* pos, const float* nrm, const float* uv, uint64_t n);
vmn_status vmn_scene_set_spheres(vmn_ctx* ctx, const vmn_sphere* spheres, uint32_t n); /* n <= 16; 0 disables */
void vmn_reference_spheres(vmn_sphere out[8]); /* the eight spheres of meshEngine.cpp:377-500 */
/* replaces MeshEngine::bindTexture's upload (meshEngine.cpp:85-91); slot = index in MeshEngine::boundTextures. Slot 0 is the one
 * Radiance samples (pathtracer.cpp:62-80) and BruteForceTracer's normal map (integrators.cpp:98-106); slot 1 is BruteForceTracer's
 * albedo (integrators.cpp:141-147). data = NULL unbinds the slot. */
vmn_status vmn_scene_set_texture(vmn_ctx* ctx, uint32_t slot, const float* data, uint32_t w, uint32_t h, uint32_t c);
vmn_status vmn_scene_commit(vmn_ctx* ctx); /* finishes the GPU BVH build (all of it when the triangles did not come through this context's own upload) */

/* ray queries */
#define VMN_ISECT_BVH2 1u /* traverse the binary LBVH instead of the compressed BVH8 (baseline / cross-check) */
/* replaces BVH::getIntersection(ray, &info, false) (bvh.cpp:47-145) for n rays */
vmn_status vmn_intersect(vmn_ctx* ctx, const vmn_ray* rays, uint64_t n, vmn_hit* out, uint32_t flags);
vmn_status vmn_intersect_device(vmn_ctx* ctx, const vmn_ray* d_rays, uint64_t n, vmn_hit* d_out, uint32_t flags);
/* vmn_intersect for callers that may not allocate (code under CUDA graph capture nearby, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. Same arguments, same kernels and
 * bit-identical results as vmn_intersect (host rays in, host hits out), but every byte of device scratch — the persistent kernel's
 * ray-claim cursor and visit counters, the staged rays, the hits — lives in one file-scope `static __device__` array of the library
 * (3 MiB + 256 B, held by every GPU that loads the library until exit) instead of the context's grow-only cudaMalloc buffers: the
 * call makes no cudaMalloc / cudaFree and calls no allocating routine. n > VMN_STATIC_MAX_RAYS is rejected with VMN_ERR_INVALID (the
 * message names the limit). All calls on one GPU share the array; it keeps the previous call's contents (the call clears the counter
 * block itself), and calls that overlap on the GPU — two contexts or two host threads on one device — race on it: serialise them. */
#define VMN_STATIC_MAX_RAYS 65536u
vmn_status vmn_intersect_static(vmn_ctx* ctx, const vmn_ray* rays, uint64_t n, vmn_hit* out, uint32_t flags);
/* replaces MeshEngine::RayCastCollision (meshEngine.cpp:196-207): out[i] = 1 iff closest hit has t > 1e-3 */
vmn_status vmn_occluded(vmn_ctx* ctx, const vmn_ray* rays, uint64_t n, uint8_t* out);
/* replaces MeshEngine::RayCast (meshEngine.cpp:239-509) for n rays */
vmn_status vmn_raycast(vmn_ctx* ctx, const vmn_ray* rays, uint64_t n, vmn_raycast_rec* out);
/* replaces Radiance() (pathtracer.cpp:21-198) for n start rays: Philox stream (seed, pixel = i, sample 0),
 * iteration i (1-based) of the bounce loop draws from Philox block i. out: 4 floats per ray; rays_cast (optional): RayCast count per path. */
vmn_status vmn_radiance(vmn_ctx* ctx, const vmn_ray* rays, uint64_t n, uint64_t seed, int32_t mode, uint32_t max_segments,
                        float* out, uint32_t* rays_cast);

/* camera + film (replaces the Camera fields PathTracer::Render reads, camera.h:66-104) */
vmn_status vmn_camera_set(vmn_ctx* ctx, const vmn_camera* cam);
/* primary rays of pixel loop pathtracer.cpp:251-280 for given (pixel, stratum, jitter) tuples; parity helper */
vmn_status vmn_camera_rays(vmn_ctx* ctx, const uint64_t* pixel, const uint32_t* sx, const uint32_t* sy, const float* jx,
                           const float* jy, uint64_t n, vmn_ray* out);

/* replaces PathTracer::Render's pixel loop (pathtracer.cpp:226-324) over this context's share of the frame */
vmn_status vmn_render(vmn_ctx* ctx, const vmn_render_params* params);
/* film readback. VMN_FILM_RGBAZ = Camera::mImage layout, 5 floats per pixel (camera.cpp:106-113): clamped rgb, alpha 1,
 * depth = samples taken (after a BRUTEFORCE render: alpha = hits / samples, depth = distance of the last primary RayCast,
 * integrators.cpp:176-183). VMN_FILM_SUM = raw (sum r, sum g, sum b, n) 4 floats per pixel (for sample-slice all-reduce). */
enum { VMN_FILM_RGBAZ = 0, VMN_FILM_SUM = 1 };
vmn_status vmn_read_image(vmn_ctx* ctx, float* dst, int32_t layout);
/* device pointer to the raw sum buffer (4 floats per pixel, width*height pixels) for NCCL all-reduce / gather by the host */
vmn_status vmn_film_device(vmn_ctx* ctx, float** d_sum, uint64_t* n_floats);
/* resolve an (externally reduced) sum buffer already on the device into RGBAZ on the host */
vmn_status vmn_film_resolve_device(vmn_ctx* ctx, const float* d_sum, float* dst_rgbaz);

/* restores the raw sum buffer (VMN_FILM_SUM layout, width*height*4 floats) from the host, e.g. from a checkpoint written by
 * vermilion_b200/host/film_io.h, so that a progressive FIXED-mode render resumes with VMN_RENDER_ACCUMULATE at sample_begin =
 * the checkpoint's sample count. (The reference has no checkpointing: its image lives only in Camera::mImage, main.cpp:94-95.) */
vmn_status vmn_film_upload(vmn_ctx* ctx, const float* sum);

/* BRUTEFORCE film: device pointer to the second accumulation plane, 2 floats per pixel (sum alpha, depth of the last primary
 * RayCast), which a multi-GPU host reduces together with vmn_film_device's plane (tiles sum with zeros). */
vmn_status vmn_film_aux_device(vmn_ctx* ctx, float** d_aux, uint64_t* n_floats);

/* replaces the 32f -> 8-bit conversion loop of Camera::saveFrame (camera.cpp:138-175, RGBAZ case): rgba8[4*p+k] =
 * (unsigned char)floor(rgbaz[5*p+k] * 255), depth[p] = rgbaz[5*p+4]. Packed on the device: 8 bytes per pixel cross PCIe instead
 * of 20. depth may be NULL. (Writing the PNG/EXR files stays host code of the reference, camera.cpp:178-189.) */
vmn_status vmn_read_image_rgba8(vmn_ctx* ctx, uint8_t* rgba8, float* depth);

vmn_status vmn_get_stats(vmn_ctx* ctx, vmn_stats* out);

/* Diagnostic for the traversal roofline (no counterpart in the reference): streams a zero-filled device buffer of
 * `bytes` (>= 1 MiB) `passes` times with the 128-bit loads the traversal kernels use (L2-cached only) and returns the read
 * bandwidth in GB/s. A working set well below the L2 size (126 MB) measures L2 read bandwidth — what bounds k_extend on a scene
 * that stays L2-resident (C2) —, one well above it HBM read bandwidth. Not on the render path. */
vmn_status vmn_probe_read_bandwidth(vmn_ctx* ctx, uint64_t bytes, uint32_t passes, double* gb_per_s);

/* ---- several GPUs behind one call --------------------------------------------------------------------------------------------
 * The reference's host makes ONE mIntegrator->Render(mCameras, mMeshEngine) per frame (core/engines/renderEngine.cpp:147-166) and
 * parallelises inside it over pixels (core/integrators/pathtracer.cpp:226-227). A vmn_dist group keeps that shape for 1..16 B200s
 * of one node: it owns one context per device, splits the frame, and assembles the film on the root (rank 0) over NCCL / NVLink:
 *   sample slices (FIXED mode): ncclReduce(sum) of the (sum r, g, b, n) film to the root;
 *   interleaved tiles (any mode): every rank sends the pixels of ITS tiles (1/N of the film), the root receives and unpacks them.
 * The scene goes up once on the root's PCIe link and is ncclBroadcast to the other devices (VMN_DIST_SCENE_BROADCAST), each device
 * builds its own BVH; only the root resolves and reads the image back. NCCL is loaded at run time (dlopen): a single-GPU user of this
 * library never needs it.
 *   local group: one process, N devices, one host thread per device (ncclCommInitAll) — what B200PathTracer and vermilion_headless
 *                --gpus N use;
 *   rank group:  one process per device (torchrun): rank 0 calls vmn_dist_unique_id and hands the 128 bytes to the others by any
 *                means (bench.py: torch.distributed broadcast), every rank calls vmn_dist_create_rank. All vmn_dist_* calls are
 *                then collective: every rank makes them in the same order. */
typedef struct vmn_dist vmn_dist;
#define VMN_DIST_MAX_RANKS 16
#define VMN_DIST_ID_BYTES 128
enum { VMN_DIST_SPLIT_SAMPLES = 0, VMN_DIST_SPLIT_TILES = 1 };
#define VMN_DIST_SCENE_BROADCAST 1u /* vmn_dist_scene_set_triangles: read the host arrays on the root only and broadcast them over NVLink
                                     * (the other ranks of a rank group may pass NULL / 0); without it every member uploads the arrays */
#define VMN_DIST_RENDER_ASSEMBLE 0x100u /* vmn_render_params.flags, sample slices with VMN_RENDER_ACCUMULATE: progressive passes stay in
                                         * each rank's own film; the pass that carries this flag also reduces the films to the root */
typedef struct {
    uint32_t world, local;                     /* ranks in the group / members in this process */
    double render_ms[VMN_DIST_MAX_RANKS];      /* device time of the last vmn_render of each LOCAL member */
    double frame_ms, collective_ms, scene_ms;  /* root's stream: render + assembly, assembly alone; last broadcast scene load (upload, broadcast and the geometry build beside them) */
    uint64_t rays_traced;                      /* sum over the local members */
} vmn_dist_stats;

vmn_status vmn_dist_create_local(const int32_t* devices, uint32_t n, uint32_t cfg_flags, uint64_t max_paths, vmn_dist** out);
vmn_status vmn_dist_unique_id(uint8_t* id /* VMN_DIST_ID_BYTES */);
vmn_status vmn_dist_create_rank(const uint8_t* id, uint32_t rank, uint32_t world, const vmn_config* cfg, vmn_dist** out);
void vmn_dist_destroy(vmn_dist* g);
const char* vmn_dist_last_error(const vmn_dist* g);
uint32_t vmn_dist_world(const vmn_dist* g);
uint32_t vmn_dist_rank(const vmn_dist* g);           /* rank of this process's first member (0 in a local group) */
uint32_t vmn_dist_local_count(const vmn_dist* g);
vmn_ctx* vmn_dist_context(vmn_dist* g, uint32_t local_index); /* borrowed: a member's context (stats, ray queries) */
/* scene / camera: as the vmn_scene_* / vmn_camera_set calls, applied to every member */
vmn_status vmn_dist_scene_set_triangles(vmn_dist* g, const float* pos, const float* nrm, const float* uv, uint64_t n, uint32_t flags);
vmn_status vmn_dist_scene_set_spheres(vmn_dist* g, const vmn_sphere* spheres, uint32_t n);
vmn_status vmn_dist_scene_set_texture(vmn_dist* g, uint32_t slot, const float* data, uint32_t w, uint32_t h, uint32_t c);
vmn_status vmn_dist_scene_commit(vmn_dist* g);
vmn_status vmn_dist_camera_set(vmn_dist* g, const vmn_camera* cam);
/* replaces the one Render call of renderEngine.cpp:164 for the whole group: params as vmn_render (tile / pixel range fields zero: the
 * group chooses the shares; sample_begin / sample_end select the pass of a progressive render), split = VMN_DIST_SPLIT_*,
 * tile_size 0 = 64. On return the root's film holds the assembled frame. */
vmn_status vmn_dist_render(vmn_dist* g, const vmn_render_params* params, int32_t split, uint32_t tile_size);
vmn_status vmn_dist_read_image(vmn_dist* g, float* dst, int32_t layout); /* root: as vmn_read_image; other ranks: no-op */
vmn_status vmn_dist_get_stats(vmn_dist* g, vmn_dist_stats* out);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif
