/*
 * ref_harness.cpp — C entry points around the reference's OWN, UNMODIFIED sources.
 *
 * TEST INFRASTRUCTURE ONLY. This file is linked (oracle/Makefile, target `ref`)
 * with core/accelerators/{bvh,bbox,triangle}.cpp, core/engines/{meshEngine,
 * loggingEngine}.cpp, core/integrators/{pathtracer,integrators}.cpp and
 * core/camera/{camera.cpp,frameTile/frameTile.cpp} compiled where they lie under
 * /root/reference, into oracle/_ref/libvermref.so (git-ignored, travels to the
 * GPU box). No reference source is copied into this repository.
 *
 * What is NOT the reference here and why:
 *  - OpenImageIO is absent from the image: oracle/stubs/OpenImageIO/imagebuf.h
 *    declares the five names meshEngine.cpp:76-88 / camera.cpp:178-188 use (image
 *    file I/O only, never on the measured path); its ImageInput reads PFM files,
 *    so MeshEngine::bindTexture itself runs (ref_scene_bind_texture_file).
 *  - Assimp: the reference's vendored extern/assimp (4.1.0) is compiled from its own
 *    sources by oracle/Makefile (Importer core + OBJ importer + the post-process
 *    steps MeshEngine::load asks for; no cmake) and linked in, so
 *    ref_scene_load_file goes through the real MeshEngine::load -> ReadFile ->
 *    processScene (meshEngine.cpp:95-150). Procedural scenes are still injected
 *    as hand-filled aiScene/aiMesh structs (the exact structs
 *    MeshEngine::processScene and createBVH walk, meshEngine.cpp:533-724).
 *  - `private` is re-declared public for meshEngine.h ONLY in this translation
 *    unit, to set MeshEngine::pScene and call processScene() without
 *    Importer::ReadFile. Class layout is unaffected; the reference objects are
 *    compiled from the untouched sources.
 *  - The pixel loop of PathTracer::Render hard-wires std::random_device
 *    (pathtracer.cpp:231); ref_render_seeded restates that loop (and only that
 *    loop) with a per-pixel seeded std::mt19937_64 and calls the reference's own
 *    Radiance(). ref_render_pathtracer calls the real PathTracer::Render.
 *  - BruteForceTracer::Render seeds its shared std::mt19937 with time(0) (integrators.cpp:30). The library is
 *    linked with -Wl,--wrap=time (oracle/Makefile), so that the reference object's call lands in __wrap_time
 *    below, which returns a caller-chosen value while ref_render_bruteforce runs (and the real time otherwise).
 *    The source is untouched; with one OpenMP thread the real Render is then reproducible.
 */
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <ctime>
#include <cmath>
#include <random>
#include <string>
#include <unordered_map>
#include <vector>
#include <unistd.h>
#include <fcntl.h>
#include <omp.h>

#define private public
#include "engines/meshEngine.h"
#undef private
#include "accelerators/triangle.h"
#include "accelerators/bvh.h"
#include "camera/camera.h"
#include "integrators/integrators.h"
#include "glm/gtc/matrix_transform.hpp"

#include "vermilion_oracle.h" /* record layouts only */

// The reference's own bounce loop, external linkage (pathtracer.cpp:21)
glm::vec4 Radiance(Vermilion::MeshEngine* mEng, glm::vec3 rStart, glm::vec3 rDir, std::mt19937_64& mtRanEngine);

static long g_fixed_time = 0; // != 0: what the reference's time(0) returns (integrators.cpp:30)
extern "C" time_t __real_time(time_t*);
extern "C" time_t __wrap_time(time_t* t) {
    if (g_fixed_time) { if (t) *t = (time_t)g_fixed_time; return (time_t)g_fixed_time; }
    return __real_time(t);
}

namespace {

struct Quiet { // silence the reference's printf / per-pixel fprintf(stderr)
    int so, se, dn;
    Quiet() {
        fflush(stdout); fflush(stderr);
        so = dup(1); se = dup(2); dn = open("/dev/null", O_WRONLY);
        dup2(dn, 1); dup2(dn, 2);
    }
    ~Quiet() {
        fflush(stdout); fflush(stderr);
        dup2(so, 1); dup2(se, 2); close(so); close(se); close(dn);
    }
};

inline int clamp_threads(int t) { int m = omp_get_max_threads(); return (t <= 0 || t > m) ? m : t; }

} // namespace

struct ref_scene {
    // direct accelerator tier
    std::vector<Object*> tris;
    std::unordered_map<const Object*, uint32_t> ids;
    BVH* bvh = nullptr;
    // MeshEngine tier
    Vermilion::LogEngine* logger = nullptr;
    Vermilion::MeshEngine* mesh = nullptr;
    aiScene* scene = nullptr; // hand-filled; owned here (null for a file-loaded scene: the Importer owns that one)
    std::vector<float> dpos, dnrm, duv; // ref_scene_load_file: what createBVH's walk read, for ref_scene_dump_triangles
};

struct ref_camera_settings {
    uint32_t width, height;
    float position[3];
    float rotation_deg[3];
    float back_distance, back_size_x, back_size_y;
    uint32_t rays_per_pixel;
};

namespace {

Vermilion::Camera* make_camera(const ref_camera_settings* s) {
    Vermilion::cameraSettings vcs;
    vcs.imageResX = s->width; vcs.imageResY = s->height;
    vcs.position = Vermilion::float3(s->position[0], s->position[1], s->position[2]);
    vcs.rotation = Vermilion::float3(s->rotation_deg[0], s->rotation_deg[1], s->rotation_deg[2]);
    vcs.fBackDistance = s->back_distance;
    vcs.horAngleOfView = 90.f;
    vcs.raysPerPixel = s->rays_per_pixel;
    vcs.fBackSizeX = s->back_size_x; vcs.fBackSizeY = s->back_size_y;
    vcs.rayMaxBounces = 5; vcs.tileSize = 16;
    vcs.renderMode = Vermilion::vermRenderMode::RGBAZ;
    return new Vermilion::Camera(vcs);
}

// Fixed-mode estimator over the reference's own MeshEngine::RayCast. Mirrors
// oracle/vermilion_oracle.cpp radiance(mode=FIXED): same control flow and draw
// order as pathtracer.cpp:21-198 with r2 = xi and a max_segments cap.
glm::vec4 radiance_fixed(Vermilion::MeshEngine* mEng, glm::vec3 rStart, glm::vec3 rDir, std::mt19937_64& eng,
                         uint32_t max_segments, uint32_t* nrays) {
    std::uniform_real_distribution<double> distrib(0.0, 1.0);
    aiMaterial* hitMaterial; glm::vec3 hitLocation, hitNormal, hitColour; float hitDistance; glm::vec2 hitUV;
    glm::vec4 accumColour(0, 0, 0, -100), accumRadiance(1, 1, 1, 1);
    uint32_t segs = 0;
    while (1) {
        bool hit = mEng->RayCast(rStart, rDir, &hitMaterial, &hitLocation, &hitNormal, &hitDistance, &hitUV, &hitColour);
        ++segs; if (nrays) ++*nrays;
        if (!hit) return accumColour;
        accumColour += accumRadiance * glm::vec4(hitColour, 0.f);
        if (segs == 1) accumColour.w = hitDistance;
        if (glm::length(hitColour) > 1.f) return accumColour;
        if (segs >= max_segments) return accumColour;
        glm::vec4 sampleColour(1.f, 1.f, 1.f, 1.f);
        if (hitMaterial && mEng->boundTextures.size() > 0) { sampleColour = glm::vec4(); mEng->boundTextures[0].Sample(hitUV, &sampleColour); }
        if (hitMaterial && distrib(eng) >= 0.96) {
            distrib(eng); distrib(eng); distrib(eng);
            rStart = hitLocation - rDir * 0.001f;
            rDir = glm::normalize(rDir - hitNormal * 2.f * glm::dot(hitNormal, rDir));
        } else if (hitMaterial) {
            accumRadiance *= sampleColour;
            float r1 = 2 * M_PI * distrib(eng);
            float r2 = 1.f * distrib(eng);
            float r2s = sqrt(r2);
            glm::vec3 w = glm::dot(hitNormal, rDir) < 0.f ? hitNormal : hitNormal * -1.f;
            glm::vec3 u = glm::normalize(glm::cross(fabs(w.x) > .1 ? glm::vec3(0, 1, 0) : glm::vec3(1, 0, 0), w));
            glm::vec3 v = glm::cross(w, u);
            glm::vec3 d = glm::normalize(u * float(cos(r1)) * r2s + v * float(sin(r1)) * r2s + w * float(sqrt(1 - r2)));
            rStart = hitLocation - rDir * 0.001f;
            rDir = d;
        } else {
            double r1 = 2 * M_PI * distrib(eng);
            double r2 = 1.f * distrib(eng);
            float r2s = sqrt(r2);
            distrib(eng); distrib(eng); distrib(eng);
            glm::vec3 w = glm::dot(hitNormal, rDir) < 0.f ? hitNormal : hitNormal * -1.f;
            glm::vec3 u = glm::normalize(glm::cross(fabs(w.x) > .1 ? glm::vec3(0, 1, 0) : glm::vec3(1, 0, 0), w));
            glm::vec3 v = glm::cross(w, u);
            glm::vec3 d = glm::normalize(u * float(cos(r1)) * r2s + v * float(sin(r1)) * r2s + w * float(sqrt(1 - r2)));
            rStart = hitLocation - rDir * 0.001f;
            rDir = d;
        }
    }
}

// Radiance() has no ray counter; count by wrapping RayCast is impossible without touching
// the reference, so compat-mode ray counts come from the oracle port (bit-identical paths).

} // namespace

extern "C" {

int ref_max_threads(void) { return omp_get_max_threads(); }

/* Build (a) a BVH directly over reference Triangle objects and (b) a MeshEngine whose
 * pScene is a hand-filled aiScene with one mesh holding 3*ntri unshared vertices. */
ref_scene* ref_scene_create(const float* pos, const float* nrm, const float* uv, uint64_t ntri, int with_mesh_engine) {
    Quiet q;
    ref_scene* s = new ref_scene();
    s->tris.reserve(ntri);
    for (uint64_t i = 0; i < ntri; ++i) {
        const float* p = pos + 9 * i;
        glm::vec3 n0(0), n1(0), n2(0); glm::vec2 t0(0), t1(0), t2(0);
        if (nrm) { const float* n = nrm + 9 * i; n0 = glm::vec3(n[0], n[1], n[2]); n1 = glm::vec3(n[3], n[4], n[5]); n2 = glm::vec3(n[6], n[7], n[8]); }
        if (uv) { const float* t = uv + 6 * i; t0 = glm::vec2(t[0], t[1]); t1 = glm::vec2(t[2], t[3]); t2 = glm::vec2(t[4], t[5]); }
        Triangle* tr = new Triangle(glm::vec3(p[0], p[1], p[2]), glm::vec3(p[3], p[4], p[5]), glm::vec3(p[6], p[7], p[8]), n0, n1, n2, t0, t1, t2);
        s->ids[tr] = (uint32_t)i;
        s->tris.push_back(tr);
    }
    if (ntri) s->bvh = new BVH(s->tris); // bvh.cpp:155-162, leafSize = 4
    if (with_mesh_engine && ntri) {
        aiScene* sc = (aiScene*)calloc(1, sizeof(aiScene)); // aiScene ctor/dtor live in Assimp's Version.cpp (not linked)
        aiMesh* m = new aiMesh();
        m->mPrimitiveTypes = aiPrimitiveType_TRIANGLE;
        m->mNumVertices = (unsigned)(3 * ntri); m->mNumFaces = (unsigned)ntri;
        m->mVertices = new aiVector3D[3 * ntri];
        m->mNormals = new aiVector3D[3 * ntri];
        if (uv) { m->mTextureCoords[0] = new aiVector3D[3 * ntri]; m->mNumUVComponents[0] = 2; }
        m->mFaces = new aiFace[ntri];
        for (uint64_t i = 0; i < ntri; ++i) {
            for (int k = 0; k < 3; ++k) {
                const float* p = pos + 9 * i + 3 * k;
                m->mVertices[3 * i + k] = aiVector3D(p[0], p[1], p[2]);
                if (nrm) { const float* n = nrm + 9 * i + 3 * k; m->mNormals[3 * i + k] = aiVector3D(n[0], n[1], n[2]); }
                else m->mNormals[3 * i + k] = aiVector3D(0, 0, 0);
                if (uv) { const float* t = uv + 6 * i + 2 * k; m->mTextureCoords[0][3 * i + k] = aiVector3D(t[0], t[1], 0); }
            }
            m->mFaces[i].mNumIndices = 3;
            m->mFaces[i].mIndices = new unsigned int[3]{(unsigned)(3 * i), (unsigned)(3 * i + 1), (unsigned)(3 * i + 2)};
        }
        sc->mNumMeshes = 1; sc->mMeshes = new aiMesh*[1]{m};
        // a non-null material pointer is all RayCast needs (meshEngine.cpp:502-503); never dereferenced
        sc->mNumMaterials = 1; sc->mMaterials = new aiMaterial*[1]{reinterpret_cast<aiMaterial*>(sc)};
        s->scene = sc;
        s->logger = new Vermilion::LogEngine("/dev/null", Vermilion::VermiLogFile, Vermilion::VermiLogLevelNone);
        s->mesh = new Vermilion::MeshEngine(s->logger);
        s->mesh->pScene = sc;
        if (!s->mesh->processScene()) { /* leaves sceneAccelerator null */ }
    }
    return s;
}

/* The real loading path: MeshEngine::load(fName) with the reference's own flags (meshEngine.cpp:95-125) through the
 * vendored Assimp, then processScene -> createBVH. The harness-owned accelerator tier (tris / ids / bvh, used by
 * ref_intersect for triangle ids) is rebuilt from MeshEngine::sceneMeshes in createBVH's order (meshEngine.cpp:660-717),
 * which defines the canonical triangle ids. Returns null when load() fails. */
ref_scene* ref_scene_load_file(const char* path) {
    Quiet q;
    ref_scene* s = new ref_scene();
    s->logger = new Vermilion::LogEngine("/dev/null", Vermilion::VermiLogFile, Vermilion::VermiLogLevelNone);
    s->mesh = new Vermilion::MeshEngine(s->logger);
    std::string f(path);
    if (!s->mesh->load(f)) { delete s->mesh; delete s->logger; delete s; return nullptr; }
    for (aiMesh* mesh : s->mesh->sceneMeshes)
        for (uint32_t i = 0; i < mesh->mNumFaces; ++i) {
            glm::vec3 v[3], n[3]; glm::vec2 t[3] = {glm::vec2(0), glm::vec2(0), glm::vec2(0)};
            for (int k = 0; k < 3; ++k) {
                unsigned idx = mesh->mFaces[i].mIndices[k];
                v[k] = glm::vec3(mesh->mVertices[idx].x, mesh->mVertices[idx].y, mesh->mVertices[idx].z);
                n[k] = glm::vec3(mesh->mNormals[idx].x, mesh->mNormals[idx].y, mesh->mNormals[idx].z);
                if (mesh->HasTextureCoords(0)) t[k] = glm::vec2(mesh->mTextureCoords[0][idx].x, mesh->mTextureCoords[0][idx].y);
            }
            Triangle* tr = new Triangle(v[0], v[1], v[2], n[0], n[1], n[2], t[0], t[1], t[2]);
            s->ids[tr] = (uint32_t)s->tris.size();
            s->tris.push_back(tr);
            for (int k = 0; k < 3; ++k) {
                s->dpos.insert(s->dpos.end(), {v[k].x, v[k].y, v[k].z});
                s->dnrm.insert(s->dnrm.end(), {n[k].x, n[k].y, n[k].z});
                s->duv.insert(s->duv.end(), {t[k].x, t[k].y});
            }
        }
    if (!s->tris.empty()) s->bvh = new BVH(s->tris);
    return s;
}

uint64_t ref_scene_triangle_count(const ref_scene* s) { return s ? s->tris.size() : 0; }
uint32_t ref_scene_mesh_count(const ref_scene* s) { return s && s->mesh ? (uint32_t)s->mesh->sceneMeshes.size() : 0; }

/* The canonical triangle arrays (9 / 9 / 6 floats per triangle) of a scene made by ref_scene_load_file */
void ref_scene_dump_triangles(const ref_scene* s, float* pos, float* nrm, float* uv) {
    if (pos) memcpy(pos, s->dpos.data(), sizeof(float) * s->dpos.size());
    if (nrm) memcpy(nrm, s->dnrm.data(), sizeof(float) * s->dnrm.size());
    if (uv) memcpy(uv, s->duv.data(), sizeof(float) * s->duv.size());
}

void ref_scene_destroy(ref_scene* s) {
    if (!s) return;
    Quiet q;
    delete s->mesh; // deletes its BVH; Triangles made by createBVH leak in the reference (bvh.cpp:147-153)
    delete s->logger;
    delete s->bvh;
    for (auto t : s->tris) delete static_cast<Triangle*>(t);
    if (s->scene) {
        delete s->scene->mMeshes[0]; delete[] s->scene->mMeshes; delete[] s->scene->mMaterials; free(s->scene);
    }
    delete s;
}

void ref_scene_set_texture(ref_scene* s, const float* data, uint32_t w, uint32_t h, uint32_t c) {
    // what bindTexture does after the OIIO read (meshEngine.cpp:85-91); MeshEngine owns and delete[]s pData
    float* copy = new float[(size_t)w * h * c];
    memcpy(copy, data, sizeof(float) * (size_t)w * h * c);
    s->mesh->boundTextures.emplace_back((uint16_t)w, (uint16_t)h, (uint16_t)c, copy);
}

/* The reference's own MeshEngine::bindTexture (meshEngine.cpp:74-93) on a file (PFM, see the OpenImageIO stand-in). 1 = bound. */
int ref_scene_bind_texture_file(ref_scene* s, const char* path) {
    std::string f(path);
    return s && s->mesh && s->mesh->bindTexture(f) ? 1 : 0;
}

/* BVH::getIntersection + Triangle::getNormal(uv) over the harness-owned triangles */
void ref_intersect(const ref_scene* s, const orc_ray* rays, uint64_t n, orc_hit* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        Ray r(glm::vec3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), glm::vec3(rays[i].d[0], rays[i].d[1], rays[i].d[2]));
        IntersectionInfo ii{};
        bool hit = s->bvh && s->bvh->getIntersection(r, &ii, false);
        out[i].t = s->bvh ? ii.t : 999999999.f; out[i].u = 0; out[i].v = 0; out[i].tri = 0xFFFFFFFFu;
        if (hit) {
            out[i].tri = s->ids.at(ii.object);
            glm::vec2 uv(0);
            ii.object->getNormal(ii, &uv); // with uv corners (0,0),(1,0),(0,1) this returns (w1,w2)
            out[i].u = uv.x; out[i].v = uv.y;
        }
    }
}

void ref_occluded(const ref_scene* s, const orc_ray* rays, uint64_t n, uint8_t* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i)
        out[i] = s->mesh->RayCastCollision(glm::vec3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), glm::vec3(rays[i].d[0], rays[i].d[1], rays[i].d[2])) ? 1 : 0;
}

void ref_raycast_many(const ref_scene* s, const orc_ray* rays, uint64_t n, orc_raycast* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 4096) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        aiMaterial* mat; glm::vec3 P, N, Le; float dist; glm::vec2 uv;
        bool hit = s->mesh->RayCast(glm::vec3(rays[i].o[0], rays[i].o[1], rays[i].o[2]), glm::vec3(rays[i].d[0], rays[i].d[1], rays[i].d[2]),
                                    &mat, &P, &N, &dist, &uv, &Le);
        orc_raycast& o = out[i];
        o.P[0] = P.x; o.P[1] = P.y; o.P[2] = P.z; o.dist = dist;
        o.N[0] = N.x; o.N[1] = N.y; o.N[2] = N.z;
        o.flags = (hit ? ORC_RC_HIT : 0u) | (mat ? ORC_RC_MATERIAL : 0u);
        o.uv[0] = uv.x; o.uv[1] = uv.y; o.Le[0] = Le.x; o.Le[1] = Le.y; o.Le[2] = Le.z;
        o.tri = 0xFFFFFFFFu; o.pad[0] = o.pad[1] = 0;
    }
}

/* The reference's Radiance() with engine i = std::mt19937_64(seed + i); mode FIXED uses radiance_fixed */
void ref_radiance_many(const ref_scene* s, const orc_ray* rays, uint64_t n, uint64_t seed, int mode, float* out, int threads) {
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 256) num_threads(threads)
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        std::mt19937_64 eng(seed + (uint64_t)i);
        glm::vec3 o(rays[i].o[0], rays[i].o[1], rays[i].o[2]), d(rays[i].d[0], rays[i].d[1], rays[i].d[2]);
        glm::vec4 v = (mode == ORC_MODE_FIXED) ? radiance_fixed(s->mesh, o, d, eng, 5, nullptr) : Radiance(s->mesh, o, d, eng);
        out[4 * i + 0] = v.x; out[4 * i + 1] = v.y; out[4 * i + 2] = v.z; out[4 * i + 3] = v.w;
    }
}

/* Vermilion::Camera's constructor (camera.cpp:34-81) → the fields the integrator reads */
void ref_camera_fields(const ref_camera_settings* cs, orc_camera* out) {
    Vermilion::Camera* cam = make_camera(cs);
    out->width = cam->uImageU; out->height = cam->uImageV;
    for (int i = 0; i < 3; ++i) { out->position[i] = cam->mPosition[i]; out->rotation[i] = cam->mRotation[i]; }
    out->dist_to_film = cam->mDistToFilm; out->sensor_x = cam->sensorSizeX; out->sensor_y = cam->sensorSizeY;
    out->spp = cam->uSamplesPerPixel;
    delete cam;
}

/* Camera ray of the reference pixel loop, glm arithmetic (pathtracer.cpp:216-221, 251-280) */
void ref_camera_ray(const ref_camera_settings* cs, uint64_t p, uint32_t sampleX, uint32_t sampleY, float jx, float jy, orc_ray* out) {
    Vermilion::Camera* cam = make_camera(cs);
    glm::mat4 cameraTransform(1);
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.y, glm::vec3(0, 1, 0));
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.x, glm::vec3(1, 0, 0));
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.z, glm::vec3(0, 0, 1));
    float homogenousX = ((float(p % cam->uImageU) + (sampleX * 0.5f - 0.5f) + jx - 0.25) / cam->uImageU) - 0.5;
    float homogenousY = ((float(p / cam->uImageU) + (sampleY * 0.5f - 0.5f) + jy - 0.25) / cam->uImageV) - 0.5;
    float cameraBackXCm = homogenousX * cam->sensorSizeX;
    float cameraBackYCm = homogenousY * cam->sensorSizeY;
    auto gridPane = glm::vec4(cameraBackXCm, -cameraBackYCm, -cam->mDistToFilm, 1.0f);
    auto rayDirection = cameraTransform * gridPane;
    glm::vec3 d = glm::normalize(glm::vec3(rayDirection) / rayDirection.w);
    out->o[0] = cam->mPosition.x; out->o[1] = cam->mPosition.y; out->o[2] = cam->mPosition.z; out->pad0 = 0;
    out->d[0] = d.x; out->d[1] = d.y; out->d[2] = d.z; out->pad1 = 0;
    delete cam;
}

/* The real PathTracer::Render through the Integrator interface (integrators.h:11-26).
 * Non-reproducible (std::random_device). image: 5 floats per pixel. Returns seconds. */
double ref_render_pathtracer(const ref_scene* s, const ref_camera_settings* cs, float* image, int threads) {
    Vermilion::Camera* cam = make_camera(cs);
    std::vector<Vermilion::Camera*> cams{cam};
    Vermilion::Integrator* integ = new Vermilion::PathTracer();
    int saved = omp_get_max_threads();
    omp_set_num_threads(clamp_threads(threads));
    double t0 = omp_get_wtime();
    { Quiet q; integ->Render(cams, s->mesh); }
    double t1 = omp_get_wtime();
    omp_set_num_threads(saved);
    memcpy(image, cam->mImage, sizeof(float) * 5 * (size_t)cam->RenderTargetSize);
    delete integ; delete cam;
    return t1 - t0;
}

/* Seeded restatement of the pixel loop only (pathtracer.cpp:226-324) around the reference's own
 * Radiance / RayCast. Engine for pixel p = std::mt19937_64(seed + p). FIXED: samples [s0,s1). */
void ref_render_seeded(const ref_scene* s, const ref_camera_settings* cs, int mode, uint64_t seed, uint64_t p0, uint64_t p1,
                       uint32_t s0, uint32_t s1, uint32_t max_segments, float* image, float* raw_sum, int threads) {
    Vermilion::Camera* cam = make_camera(cs);
    glm::mat4 cameraTransform(1);
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.y, glm::vec3(0, 1, 0));
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.x, glm::vec3(1, 0, 0));
    cameraTransform = glm::rotate(cameraTransform, cam->mRotation.z, glm::vec3(0, 0, 1));
    if (max_segments == 0) max_segments = 5;
    threads = clamp_threads(threads);
#pragma omp parallel for schedule(dynamic, 16) num_threads(threads)
    for (uint64_t p = p0; p < p1; ++p) {
        std::uniform_real_distribution<float> distrib(0, 0.5);
        std::mt19937_64 mtRanEngine(seed + p);
        glm::vec4 accum(0, 0, 0, 0);
        uint32_t nTotalSamples = 0;
        auto one = [&](uint32_t sampleX, uint32_t sampleY) {
            float jx = distrib(mtRanEngine);
            float jy = distrib(mtRanEngine);
            float homogenousX = ((float(p % cam->uImageU) + (sampleX * 0.5f - 0.5f) + jx - 0.25) / cam->uImageU) - 0.5;
            float homogenousY = ((float(p / cam->uImageU) + (sampleY * 0.5f - 0.5f) + jy - 0.25) / cam->uImageV) - 0.5;
            float cameraBackXCm = homogenousX * cam->sensorSizeX;
            float cameraBackYCm = homogenousY * cam->sensorSizeY;
            auto gridPane = glm::vec4(cameraBackXCm, -cameraBackYCm, -cam->mDistToFilm, 1.0f);
            auto rayDirection = cameraTransform * gridPane;
            glm::vec3 d = glm::normalize(glm::vec3(rayDirection) / rayDirection.w);
            return (mode == ORC_MODE_FIXED) ? radiance_fixed(s->mesh, cam->mPosition, d, mtRanEngine, max_segments, nullptr)
                                            : Radiance(s->mesh, cam->mPosition, d, mtRanEngine);
        };
        if (mode == ORC_MODE_COMPAT) {
            for (uint16_t sampleX = 0; sampleX < 2; ++sampleX)
                for (uint16_t sampleY = 0; sampleY < 2; ++sampleY)
                    for (uint32_t sampleZ = 0; sampleZ < (cam->uSamplesPerPixel / 4); ++sampleZ) {
                        ++nTotalSamples;
                        auto sampleRadiance = one(sampleX, sampleY);
                        accum += sampleRadiance;
                        if (nTotalSamples > sqrt(cam->uSamplesPerPixel) &&
                            fabs(glm::length(glm::vec3(accum) / float(nTotalSamples) - glm::vec3(accum + sampleRadiance) / float(nTotalSamples + 1))) < 0.00001f)
                            break;
                    }
        } else {
            uint32_t per = cam->uSamplesPerPixel / 4;
            // the engine is sequential: skip nothing, so a slice [s0,s1) only equals the full run when s0 == 0
            for (uint32_t sidx = s0; sidx < s1; ++sidx) {
                uint32_t sampleX = per ? (sidx / (2 * per)) & 1u : 0, sampleY = per ? (sidx / per) & 1u : 0;
                ++nTotalSamples;
                accum += one(sampleX, sampleY);
            }
        }
        if (raw_sum) { float* r = raw_sum + 4 * p; r[0] = accum.x; r[1] = accum.y; r[2] = accum.z; r[3] = (float)nTotalSamples; }
        Vermilion::pixelValue px;
        px.pixel = p;
        px.red = std::max(std::min(accum.x / nTotalSamples, 1.f), 0.f);
        px.green = std::max(std::min(accum.y / nTotalSamples, 1.f), 0.f);
        px.blue = std::max(std::min(accum.z / nTotalSamples, 1.f), 0.f);
        px.alpha = 1.f;
        px.depth = nTotalSamples;
        cam->setPixelValue(px);
    }
    for (uint64_t p = p0; p < p1; ++p) memcpy(image + 5 * p, cam->mImage + 5 * p, 5 * sizeof(float));
    delete cam;
}

/* The real BruteForceTracer::Render through the Integrator interface (integrators.cpp:9-186) with its time(0) seed
 * pinned to `seed` (see __wrap_time). threads = 1 makes it deterministic (pixels draw from the shared engine in order);
 * more threads race on the engine exactly as the reference does. image: 5 floats per pixel. Returns seconds. */
double ref_render_bruteforce(const ref_scene* s, const ref_camera_settings* cs, uint64_t seed, float* image, int threads) {
    Vermilion::Camera* cam = make_camera(cs);
    std::vector<Vermilion::Camera*> cams{cam};
    Vermilion::Integrator* integ = new Vermilion::BruteForceTracer();
    int saved = omp_get_max_threads();
    omp_set_num_threads(clamp_threads(threads));
    g_fixed_time = (long)seed;
    double t0 = omp_get_wtime();
    { Quiet q; integ->Render(cams, s->mesh); }
    double t1 = omp_get_wtime();
    g_fixed_time = 0;
    omp_set_num_threads(saved);
    memcpy(image, cam->mImage, sizeof(float) * 5 * (size_t)cam->RenderTargetSize);
    delete integ; delete cam;
    return t1 - t0;
}

/* the reference MeshEngine of a scene, for oracle/adapter_harness.cpp (the Integrator adapter under test) */
void* ref_scene_mesh_engine(ref_scene* s) { return s ? s->mesh : nullptr; }

const char* ref_build_flags(void) {
#ifdef VERMREF_FLAGS
    return VERMREF_FLAGS;
#else
    return "unknown";
#endif
}

} // extern "C"
