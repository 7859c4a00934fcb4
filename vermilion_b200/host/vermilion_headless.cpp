// vermilion_headless.cpp — the reference's command line (main.cpp:58-104) without the Vulkan/GLFW/ImGui preview, with the
// integrator selectable: `--engine cpu` is the reference's own PathTracer (what main.cpp:61 constructs), `--engine b200` the
// drop-in Vermilion::B200PathTracer; `cpu-bf` / `b200-bf` are BruteForceTracer (the RenderEngine's default, renderEngine.cpp:51)
// and its drop-in. Everything else is the reference's code path: MeshEngine::load through Assimp (meshEngine.cpp:95-150),
// bindTexture for the optional normal map / texture arguments (main.cpp:71-92), a reference Camera, and
// Integrator::Render called through the base-class pointer exactly as RenderEngine::draw does (renderEngine.cpp:147-166).
//
// RenderEngine itself is not used because renderEngine.cpp includes the Vulkan preview (renderer.h -> GLFW + Vulkan, absent
// here); the three things it does for a headless run are restated below: the default camera (renderEngine.cpp:114-142), the
// Render call, and saveFrame (camera.cpp:126-194: 8-bit floor(x*255) RGBA + the float depth plane). The image is written with
// the host library's PNG / PFM writers because OpenImageIO is not in this image.
//
// Built by oracle/Makefile against the reference headers and objects into oracle/_ref/vermilion_headless. There is no
// fallback: with `--engine b200` and no usable GPU the program prints the engine's error and exits 2 without writing an image.
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>
#include <fcntl.h>

#include "b200_pathtracer.h"
#include "camera/camera.h"
#include "film_io.h"

namespace {

void usage() {
    fprintf(stderr,
            "usage: vermilion_headless <scene file> [normal map] [texture] [options]\n"
            "  --engine cpu|b200|cpu-bf|b200-bf   integrator (default cpu = the reference's PathTracer)\n"
            "  --out NAME                         NAME.png + NAME_depth.pfm (default output, as main.cpp:95)\n"
            "  --res WxH  --spp N                 image size and raysPerPixel (default 648x432, 128: renderEngine.cpp:119-134)\n"
            "  --camera x,y,z,rx,ry,rz            position and rotation in degrees (default -4000,1600,8000,0,25,0)\n"
            "  --film dist,sx,sy                  film distance and size (default 6,3.6,2.4)\n"
            "  --mode compat|fixed  --seed S  --segments K  --shadow   (b200 engines only)\n"
            "  --gpus N | --devices a,b,...       (b200 engines) render the frame on several GPUs of this node: one Render call, the film\n"
            "                                     assembled on the first device over NCCL / NVLink; --split tiles|samples (default tiles)\n"
            "  --quiet                            drop the reference's per-pixel stderr chatter (pathtracer.cpp:229)\n");
}

bool parse_floats(const char* s, float* out, int n) {
    for (int i = 0; i < n; ++i) {
        char* end = nullptr;
        out[i] = strtof(s, &end);
        if (end == s) return false;
        s = end;
        if (i + 1 < n) { if (*s != ',') return false; ++s; }
    }
    return *s == 0;
}

} // namespace

int main(int argc, char** argv) {
    std::vector<std::string> files;
    std::string engine = "cpu", out = "output", mode = "compat";
    Vermilion::cameraSettings vcs; // defaults: RenderEngine::CreateInternalDefaultCamera (renderEngine.cpp:114-142)
    vcs.imageResX = 2592 / 4; vcs.imageResY = 1728 / 4;
    vcs.position = Vermilion::float3(-4000, 1600, 8000);
    vcs.rotation = Vermilion::float3(0, 25, 0);
    vcs.fBackDistance = 3.f * 2; vcs.horAngleOfView = 90.f; vcs.raysPerPixel = 128;
    vcs.fBackSizeX = 3.6f; vcs.fBackSizeY = 2.4f; vcs.rayMaxBounces = 5; vcs.tileSize = 16;
    vcs.renderMode = Vermilion::vermRenderMode::RGBAZ;
    Vermilion::B200Options opt;
    bool quiet = false;

    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto value = [&]() -> const char* { if (i + 1 >= argc) { usage(); exit(1); } return argv[++i]; };
        if (a == "--engine") engine = value();
        else if (a == "--out") out = value();
        else if (a == "--mode") mode = value();
        else if (a == "--seed") opt.seed = strtoull(value(), nullptr, 10);
        else if (a == "--segments") opt.maxSegments = (uint32_t)strtoul(value(), nullptr, 10);
        else if (a == "--spp") vcs.raysPerPixel = (uint32_t)strtoul(value(), nullptr, 10);
        else if (a == "--shadow") opt.shadowRays = true;
        else if (a == "--gpus") { int n = atoi(value()); if (n < 1 || n > 16) { usage(); return 1; } opt.devices.clear(); for (int d = 0; d < n; ++d) opt.devices.push_back(d); }
        else if (a == "--devices") { opt.devices.clear(); const char* p = value(); while (*p) { char* e = nullptr; long d = strtol(p, &e, 10); if (e == p) { usage(); return 1; } opt.devices.push_back((int)d); p = (*e == ',') ? e + 1 : e; } }
        else if (a == "--split") { std::string v = value(); if (v == "samples") opt.splitSamples = true; else if (v == "tiles") opt.splitSamples = false; else { usage(); return 1; } }
        else if (a == "--quiet") quiet = true;
        else if (a == "--res") { unsigned w = 0, h = 0; if (sscanf(value(), "%ux%u", &w, &h) != 2 || !w || !h) { usage(); return 1; } vcs.imageResX = w; vcs.imageResY = h; }
        else if (a == "--camera") { float f[6]; if (!parse_floats(value(), f, 6)) { usage(); return 1; }
            vcs.position = Vermilion::float3(f[0], f[1], f[2]); vcs.rotation = Vermilion::float3(f[3], f[4], f[5]); }
        else if (a == "--film") { float f[3]; if (!parse_floats(value(), f, 3)) { usage(); return 1; }
            vcs.fBackDistance = f[0]; vcs.fBackSizeX = f[1]; vcs.fBackSizeY = f[2]; }
        else if (a.rfind("--", 0) == 0) { usage(); return 1; }
        else files.push_back(a);
    }
    if (files.empty()) { usage(); return 1; } // the reference opens its Vulkan preview here (main.cpp:65-68): unavailable
    if (mode != "compat" && mode != "fixed") { usage(); return 1; }
    opt.mode = mode == "fixed" ? VMN_MODE_FIXED : VMN_MODE_COMPAT;

    // main.cpp:60-63
    auto mEng = new Vermilion::MeshEngine();
    Vermilion::Integrator* integrator = nullptr;
    Vermilion::B200PathTracer* b200 = nullptr;
    if (engine == "cpu") integrator = new Vermilion::PathTracer();
    else if (engine == "cpu-bf") integrator = new Vermilion::BruteForceTracer();
    else if (engine == "b200") integrator = b200 = new Vermilion::B200PathTracer(opt);
    else if (engine == "b200-bf") integrator = b200 = new Vermilion::B200BruteForceTracer(opt);
    else { usage(); return 1; }

    // main.cpp:84-92
    if (!mEng->load(files[0])) { fprintf(stderr, "vermilion_headless: cannot load %s\n", files[0].c_str()); return 3; }
    std::string normalName = files.size() > 1 ? files[1] : std::string(), textureName = files.size() > 2 ? files[2] : std::string();
    if (mEng->bindTexture(normalName)) printf("Using a texture\n");
    mEng->bindTexture(textureName);
    uint64_t nFaces = 0;
    for (aiMesh* m : mEng->sceneMeshes) nFaces += m->mNumFaces;

    // RenderEngine::draw (renderEngine.cpp:147-166)
    std::vector<Vermilion::Camera*> cameras{new Vermilion::Camera(vcs)};
    int saved_err = -1;
    if (quiet) { fflush(stderr); saved_err = dup(2); int dn = open("/dev/null", O_WRONLY); dup2(dn, 2); close(dn); }
    auto t0 = std::chrono::steady_clock::now();
    integrator->Render(cameras, mEng);
    double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    if (quiet) { fflush(stderr); dup2(saved_err, 2); close(saved_err); }
    if (b200 && !b200->ok()) {
        fprintf(stderr, "vermilion_headless: B200 engine failed: %s\n", b200->lastError().c_str());
        return 2;
    }

    // RenderEngine::saveFrame -> Camera::saveFrame (camera.cpp:126-194), RGBAZ branch
    Vermilion::Camera* cam = cameras[0];
    const uint64_t np = cam->RenderTargetSize;
    std::vector<uint8_t> rgba(np * 4);
    std::vector<float> depth(np);
    double mean_depth = 0;
    for (uint64_t p = 0; p < np; ++p) {
        for (int c = 0; c < 4; ++c) rgba[p * 4 + c] = static_cast<unsigned char>(std::floor(cam->mImage[p * 5 + c] * 255));
        depth[p] = cam->mImage[p * 5 + 4];
        mean_depth += depth[p];
    }
    if (vmnh_save_frame(out.c_str(), rgba.data(), depth.data(), cam->uImageU, cam->uImageV) != 0) {
        fprintf(stderr, "vermilion_headless: cannot write %s.png\n", out.c_str());
        return 4;
    }
    printf("{\"engine\": \"%s\", \"triangles\": %llu, \"width\": %u, \"height\": %u, \"spp\": %u, \"render_s\": %.6f, \"mean_depth_plane\": %.4f",
           engine.c_str(), (unsigned long long)nFaces, cam->uImageU, cam->uImageV, cam->uSamplesPerPixel, seconds, mean_depth / (double)np);
    if (b200) {
        printf(", \"rays_traced\": %llu, \"gpus\": %u", (unsigned long long)b200->stats().rays_traced, b200->groupStats().world ? b200->groupStats().world : 1u);
        if (b200->groupStats().world > 1) {
            printf(", \"frame_ms\": %.3f, \"collective_ms\": %.3f, \"render_ms\": [", b200->groupStats().frame_ms, b200->groupStats().collective_ms);
            for (uint32_t i = 0; i < b200->groupStats().local; ++i) printf("%s%.3f", i ? ", " : "", b200->groupStats().render_ms[i]);
            printf("]");
        }
    }
    printf(", \"out\": \"%s.png\"}\n", out.c_str());

    // main.cpp:98-100
    delete mEng;
    delete integrator;
    delete cam;
    return 0;
}
