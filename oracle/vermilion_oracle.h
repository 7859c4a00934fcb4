/*
 * vermilion_oracle.h — C API of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing under vermilion_b200/ may include, link or
 * load this. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference leg use it, and only as the checker or the timed CPU arm.
 *
 * The oracle is a plain C++ restatement (no glm, no Assimp) of the reference
 * path: core/accelerators/{bvh,bbox,triangle}.cpp, core/engines/meshEngine.cpp
 * (RayCast, sphereIntersect, VermiTexture::Sample) and
 * core/integrators/pathtracer.cpp (Radiance, PathTracer::Render) and
 * core/integrators/integrators.cpp (BruteForceTracer::Render). Each
 * function in vermilion_oracle.cpp cites the reference file:line it follows.
 *
 * Parity status: PINNED — against the reference's own code compiled from
 * /root/reference into oracle/_ref/libvermref.so (see oracle/Makefile,
 * oracle/ref_harness.cpp); tests/test_oracle_vs_ref.py checks bit-exact
 * agreement here, tests/golden/ holds vectors generated from that build for the
 * GPU box where /root/reference does not exist. The reference ships no tests,
 * golden vectors or fixtures of its own.
 *
 * Record layouts are shared with include/vermilion_b200.h so that the tests
 * can compare arrays directly.
 */
#ifndef VERMILION_ORACLE_H
#define VERMILION_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* 32-byte ray: o.xyz, pad, d.xyz, pad (pads ignored). */
typedef struct { float o[3]; float pad0; float d[3]; float pad1; } orc_ray;

/* 16-byte BVH closest-hit record. Miss: tri = 0xFFFFFFFF, t = 999999999.f
 * (bvh.cpp:48), u = v = 0. u,v are the barycentric weights w1,w2 that
 * Triangle::getNormal derives from the hit point (triangle.cpp:67-78). */
typedef struct { float t; float u; float v; uint32_t tri; } orc_hit;

/* 64-byte MeshEngine::RayCast record (meshEngine.cpp:239-509). */
#define ORC_RC_HIT 1u      /* return value of RayCast */
#define ORC_RC_MATERIAL 2u /* *ppImpactMaterial != nullptr */
typedef struct {
    float P[3];
    float dist;
    float N[3];
    uint32_t flags;
    float uv[2];
    float Le[3];
    uint32_t tri; /* BVH closest triangle or 0xFFFFFFFF (oracle-only extra) */
    uint32_t pad[2];
} orc_raycast;

/* Camera exactly as Vermilion::Camera holds it after its constructor ran
 * (camera.cpp:34-56): rotation already in radians with x,y negated. */
typedef struct {
    uint32_t width, height;
    float position[3];
    float rotation[3];
    float dist_to_film;
    float sensor_x, sensor_y;
    uint32_t spp;
} orc_camera;

enum { ORC_MODE_COMPAT = 0, ORC_MODE_FIXED = 1 };
enum { ORC_RNG_MT64 = 0, ORC_RNG_PHILOX = 1, ORC_RNG_MT32_SHARED = 2 };
#define ORC_BF_SHADOW 1u /* BruteForceTracer with the commented-out shadow test of integrators.cpp:109-110 enabled */

typedef struct orc_scene orc_scene;

/* pos: 9 floats per triangle (v0,v1,v2); nrm: 9 per triangle or NULL (zeros);
 * uv: 6 per triangle or NULL (zeros). Triangle id = index in these arrays =
 * running face index of MeshEngine::createBVH (meshEngine.cpp:660-717). */
orc_scene* orc_scene_create(const float* pos, const float* nrm, const float* uv, uint64_t ntri, uint32_t leaf_size);
void orc_scene_destroy(orc_scene*);
void orc_scene_set_texture(orc_scene*, const float* data, uint32_t w, uint32_t h, uint32_t c); /* replaces all bound textures by this one */
void orc_scene_add_texture(orc_scene*, const float* data, uint32_t w, uint32_t h, uint32_t c); /* appends, as MeshEngine::bindTexture does */
/* enable / disable the 8 hard-coded spheres of RayCast (default: enabled) */
void orc_scene_set_spheres(orc_scene*, int enabled);
void orc_scene_stats(const orc_scene*, uint32_t* n_nodes, uint32_t* n_leafs);

/* BVH::getIntersection (bvh.cpp:47-145) + barycentrics of getNormal. */
void orc_intersect(const orc_scene*, const orc_ray* rays, uint64_t n, orc_hit* out, int threads);
/* counts nodes popped / triangles tested, summed over rays (for B_ray of the CPU tree) */
void orc_intersect_stats(const orc_scene*, const orc_ray* rays, uint64_t n, uint64_t* nodes, uint64_t* tris);
/* MeshEngine::RayCastCollision (meshEngine.cpp:196-207) */
void orc_occluded(const orc_scene*, const orc_ray* rays, uint64_t n, uint8_t* out, int threads);
/* MeshEngine::RayCast */
void orc_raycast_many(const orc_scene*, const orc_ray* rays, uint64_t n, orc_raycast* out, int threads);

/* Radiance (pathtracer.cpp:21-198) for n start rays. rng_kind MT64: engine i is
 * std::mt19937_64(seed + i). PHILOX: stream (seed, pixel=i, sample=0), draws
 * of bounce-loop iteration i (1-based) come from Philox block i. mode FIXED = documented fixed estimator.
 * out: 4 floats per ray; rays_cast (optional): RayCast calls per ray. */
void orc_radiance_many(const orc_scene*, const orc_ray* rays, uint64_t n, uint64_t seed, int rng_kind, int mode,
                       float* out, uint32_t* rays_cast, int threads);

/* Camera ray of PathTracer::Render (pathtracer.cpp:216-221, 251-280) for
 * pixel p, stratum (sx, sy) and the two U[0,0.5) jitters jx, jy. */
void orc_camera_matrix(const orc_camera*, float* m16);
void orc_camera_ray(const orc_camera*, uint64_t p, uint32_t sx, uint32_t sy, float jx, float jy, orc_ray* out);

/* PathTracer::Render pixel loop (pathtracer.cpp:226-324) with a seeded
 * per-pixel RNG. image: 5 floats per pixel (RGBAZ, camera.cpp:106-113).
 * pixel range [p0, p1) lets a caller time or check a subset; sample range
 * [s0, s1) only applies to FIXED mode (sample-slice split); in FIXED mode
 * raw_sum (optional, 4 floats per pixel) receives the un-normalised sums.
 * Returns the number of RayCast calls made. */
uint64_t orc_render(const orc_scene*, const orc_camera*, int mode, int rng_kind, uint64_t seed,
                    uint64_t p0, uint64_t p1, uint32_t s0, uint32_t s1, uint32_t max_segments,
                    float* image, float* raw_sum, int threads);

/* BruteForceTracer::Render pixel loop (integrators.cpp:21-185). rng_kind MT32_SHARED: one std::mt19937((uint32)seed)
 * shared by all pixels, drawn in pixel order (what the reference does with one OpenMP thread; its seed is time(0),
 * integrators.cpp:30). PHILOX: stream (seed, pixel, sample), draws 0 and 1. image: 5 floats per pixel = clamped rgb,
 * alpha = hits / samples, depth = distance of the last primary RayCast (integrators.cpp:176-183). samples_taken
 * (optional): nTotalSamples per pixel. raw (optional): 6 floats per pixel = the un-normalised state (sum r, sum g, sum b, n,
 * sum alpha, last distance), i.e. the two film planes of the GPU engine. Returns the number of BVH queries made. */
uint64_t orc_render_bruteforce(const orc_scene*, const orc_camera*, int rng_kind, uint64_t seed, uint32_t flags, uint64_t p0,
                               uint64_t p1, float* image, uint32_t* samples_taken, float* raw, int threads);

/* Counter-based RNG shared by the oracle's PHILOX mode and the GPU path:
 * Philox4x32-10, key (seed_lo, seed_hi), counter (pixel_lo, pixel_hi, sample, draw/4). */
void orc_philox(uint64_t seed, uint64_t pixel, uint32_t sample, uint32_t block, uint32_t out[4]);

int orc_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
