/*
 * adapter_harness.cpp — TEST INFRASTRUCTURE. Drives the product's reference-side binding
 * (vermilion_b200/host/b200_pathtracer.cpp, class Vermilion::B200PathTracer : Vermilion::Integrator) the way the
 * reference's RenderEngine::draw() does (core/engines/renderEngine.cpp:147-166): a reference Camera built by the
 * reference constructor, a reference MeshEngine holding the scene, and a call through the Integrator vtable.
 * Linked into oracle/_ref/libvermadapter.so together with the adapter; MeshEngine / Camera come from libvermref.so
 * (the reference's unmodified sources), the GPU work from libvermilion_b200.so.
 */
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>

#include "b200_pathtracer.h"

struct ref_scene;
struct ref_camera_settings {
    uint32_t width, height;
    float position[3];
    float rotation_deg[3];
    float back_distance, back_size_x, back_size_y;
    uint32_t rays_per_pixel;
};
extern "C" void* ref_scene_mesh_engine(ref_scene*);

static const char* g_checkpoint_path = nullptr;
static uint32_t g_checkpoint_every = 0;
extern "C" void adapter_set_checkpoint(const char* path, uint32_t every_spp) { g_checkpoint_path = path; g_checkpoint_every = every_spp; }
/* B200Options::devices / splitSamples for the following adapter_render calls (n = 0: one device, the default) */
static std::vector<int> g_devices;
static bool g_split_samples = false;
extern "C" void adapter_set_devices(const int* devices, int n, int split_samples) { g_devices.assign(devices, devices + n); g_split_samples = split_samples != 0; }

extern "C" int adapter_render(ref_scene* scene, const ref_camera_settings* s, int render_mode, int mode, uint64_t seed, uint32_t max_segments,
                              float* image, uint64_t* rays_traced, char* err, int errlen) {
    Vermilion::cameraSettings vcs;
    vcs.imageResX = s->width; vcs.imageResY = s->height;
    vcs.position = Vermilion::float3(s->position[0], s->position[1], s->position[2]);
    vcs.rotation = Vermilion::float3(s->rotation_deg[0], s->rotation_deg[1], s->rotation_deg[2]);
    vcs.fBackDistance = s->back_distance;
    vcs.horAngleOfView = 90.f;
    vcs.raysPerPixel = s->rays_per_pixel;
    vcs.fBackSizeX = s->back_size_x; vcs.fBackSizeY = s->back_size_y;
    vcs.rayMaxBounces = 5; vcs.tileSize = 16;
    vcs.renderMode = (Vermilion::vermRenderMode)render_mode;
    Vermilion::Camera* cam = new Vermilion::Camera(vcs);
    std::vector<Vermilion::Camera*> cams{cam};

    Vermilion::B200Options opt;
    opt.mode = mode; opt.seed = seed; opt.maxSegments = max_segments;
    // mode 2 = VMN_MODE_BRUTEFORCE selects the BruteForceTracer drop-in; bit 8 of `mode` turns its shadow rays on
    if (g_checkpoint_path) { opt.checkpointPath = g_checkpoint_path; opt.checkpointEverySpp = g_checkpoint_every; }
    opt.shadowRays = (mode & 0x100) != 0;
    opt.mode = mode & 0xFF;
    opt.devices = g_devices; opt.splitSamples = g_split_samples;
    Vermilion::B200PathTracer* b200 = opt.mode == VMN_MODE_BRUTEFORCE ? new Vermilion::B200BruteForceTracer(opt) : new Vermilion::B200PathTracer(opt);
    Vermilion::Integrator* integrator = b200; // main.cpp:60-63
    integrator->Render(cams, static_cast<Vermilion::MeshEngine*>(ref_scene_mesh_engine(scene)));
    int rc = b200->ok() ? 0 : 1;
    if (err && errlen > 0) { strncpy(err, b200->lastError().c_str(), errlen - 1); err[errlen - 1] = 0; }
    if (rays_traced) *rays_traced = b200->stats().rays_traced;
    size_t per = render_mode == (int)Vermilion::vermRenderMode::RGBAZ ? 5 : render_mode == (int)Vermilion::vermRenderMode::RGBA ? 4
               : render_mode == (int)Vermilion::vermRenderMode::RGB ? 3 : 1;
    if (rc == 0) memcpy(image, cam->mImage, sizeof(float) * per * (size_t)cam->RenderTargetSize);
    delete integrator;
    delete cam;
    return rc;
}

/* Several cameras in one Render call (pathtracer.cpp:210 loops over cameraList): RGBAZ films, images[i] receives camera i's. */
extern "C" int adapter_render_cameras(ref_scene* scene, const ref_camera_settings* s, int ncam, int mode, uint64_t seed, float** images,
                                      char* err, int errlen) {
    std::vector<Vermilion::Camera*> cams;
    for (int i = 0; i < ncam; ++i) {
        Vermilion::cameraSettings vcs;
        vcs.imageResX = s[i].width; vcs.imageResY = s[i].height;
        vcs.position = Vermilion::float3(s[i].position[0], s[i].position[1], s[i].position[2]);
        vcs.rotation = Vermilion::float3(s[i].rotation_deg[0], s[i].rotation_deg[1], s[i].rotation_deg[2]);
        vcs.fBackDistance = s[i].back_distance; vcs.horAngleOfView = 90.f; vcs.raysPerPixel = s[i].rays_per_pixel;
        vcs.fBackSizeX = s[i].back_size_x; vcs.fBackSizeY = s[i].back_size_y; vcs.rayMaxBounces = 5; vcs.tileSize = 16;
        vcs.renderMode = Vermilion::vermRenderMode::RGBAZ;
        cams.push_back(new Vermilion::Camera(vcs));
    }
    Vermilion::B200Options opt;
    opt.mode = mode; opt.seed = seed;
    Vermilion::B200PathTracer* b200 = new Vermilion::B200PathTracer(opt);
    Vermilion::Integrator* integrator = b200;
    integrator->Render(cams, static_cast<Vermilion::MeshEngine*>(ref_scene_mesh_engine(scene)));
    int rc = b200->ok() ? 0 : 1;
    if (err && errlen > 0) { strncpy(err, b200->lastError().c_str(), errlen - 1); err[errlen - 1] = 0; }
    for (int i = 0; i < ncam; ++i) {
        if (rc == 0) memcpy(images[i], cams[i]->mImage, sizeof(float) * 5 * (size_t)cams[i]->RenderTargetSize);
        delete cams[i];
    }
    delete integrator;
    return rc;
}

/* The scene cache of the adapter is keyed on content: ONE B200PathTracer renders, the harness then edits the MeshEngine in place
 * (every vertex of every mesh moved by dy: same face count, same pointers — the case a (pointer, count) key misses) and renders
 * again; image1 / image2 receive the two RGBAZ films. The vertices are restored before returning. */
extern "C" int adapter_render_edit_render(ref_scene* scene, const ref_camera_settings* s, int mode, uint64_t seed, float dy, float* image1,
                                          float* image2, char* err, int errlen) {
    Vermilion::cameraSettings vcs;
    vcs.imageResX = s->width; vcs.imageResY = s->height;
    vcs.position = Vermilion::float3(s->position[0], s->position[1], s->position[2]);
    vcs.rotation = Vermilion::float3(s->rotation_deg[0], s->rotation_deg[1], s->rotation_deg[2]);
    vcs.fBackDistance = s->back_distance; vcs.horAngleOfView = 90.f; vcs.raysPerPixel = s->rays_per_pixel;
    vcs.fBackSizeX = s->back_size_x; vcs.fBackSizeY = s->back_size_y; vcs.rayMaxBounces = 5; vcs.tileSize = 16;
    vcs.renderMode = Vermilion::vermRenderMode::RGBAZ;
    Vermilion::Camera* cam = new Vermilion::Camera(vcs);
    std::vector<Vermilion::Camera*> cams{cam};
    Vermilion::B200Options opt;
    opt.mode = mode; opt.seed = seed;
    Vermilion::B200PathTracer* b200 = new Vermilion::B200PathTracer(opt);
    Vermilion::Integrator* integrator = b200;
    Vermilion::MeshEngine* mEng = static_cast<Vermilion::MeshEngine*>(ref_scene_mesh_engine(scene));
    const size_t bytes = sizeof(float) * 5 * (size_t)cam->RenderTargetSize;
    integrator->Render(cams, mEng);
    int rc = b200->ok() ? 0 : 1;
    if (rc == 0) memcpy(image1, cam->mImage, bytes);
    for (aiMesh* m : mEng->sceneMeshes) for (unsigned v = 0; v < m->mNumVertices; ++v) m->mVertices[v].y += dy;
    if (rc == 0) { integrator->Render(cams, mEng); rc = b200->ok() ? 0 : 1; }
    if (rc == 0) memcpy(image2, cam->mImage, bytes);
    for (aiMesh* m : mEng->sceneMeshes) for (unsigned v = 0; v < m->mNumVertices; ++v) m->mVertices[v].y -= dy;
    if (err && errlen > 0) { strncpy(err, b200->lastError().c_str(), errlen - 1); err[errlen - 1] = 0; }
    delete integrator;
    delete cam;
    return rc;
}
