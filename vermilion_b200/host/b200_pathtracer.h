// b200_pathtracer.h — the reference-side binding of libvermilion_b200.so: a Vermilion::Integrator.
//
// Drop-in for `new PathTracer()` in the reference's main.cpp:60-63:
//
//     Vermilion::Integrator* integrator = new Vermilion::B200PathTracer();   // was: new Vermilion::PathTracer()
//     rEng->assignIntegrator(integrator);                                    // renderEngine.cpp:70-78
//     ...
//     rEng->draw();                                                          // -> integrator->Render(mCameras, mMeshEngine)
//
// and Vermilion::B200BruteForceTracer for `new BruteForceTracer()`, the RenderEngine's default integrator
// (renderEngine.cpp:51, core/integrators/integrators.cpp:9-186).
//
// It is compiled against the reference's own headers (core/integrators/integrators.h, core/camera/camera.h,
// core/engines/meshEngine.h) and links only the C ABI in include/vermilion_b200.h — no CUDA headers, no CPU fallback:
// when the GPU engine cannot be created Render() logs the error and leaves the film untouched (the reference's
// Render returns void and reports failures through its logs, renderEngine.cpp:150-154).
#ifndef VERMILION_B200_PATHTRACER_H
#define VERMILION_B200_PATHTRACER_H
#include <cstdint>
#include <string>
#include <vector>

#include "integrators/integrators.h"
#include "engines/meshEngine.h"
#include "vermilion_b200.h"

namespace Vermilion
{
    struct B200Options
    {
        int device = 0;
        // several B200s of this node behind the ONE Render call the host makes (renderEngine.cpp:164): empty = {device}. The frame is
        // split by interleaved 64x64 tiles (any estimator) or, with splitSamples, by sample slices (FIXED mode); the film is assembled
        // on devices[0] over NCCL / NVLink (vmn_dist_* in vermilion_b200.h) and written to Camera::mImage as for one GPU.
        std::vector<int> devices;
        bool splitSamples = false;
        int mode = VMN_MODE_COMPAT;      // the reference estimator; VMN_MODE_FIXED = exact spp, r2 = xi, <= maxSegments rays per path;
                                         // VMN_MODE_BRUTEFORCE = BruteForceTracer::Render (what B200BruteForceTracer sets)
        bool shadowRays = false;         // BRUTEFORCE: enable the shadow test the reference has commented out (integrators.cpp:108-110)
        uint32_t maxSegments = 5;
        uint64_t seed = 1;               // the reference seeds from std::random_device (pathtracer.cpp:231); here it is explicit
        uint64_t maxPathsInFlight = 0;   // 0 = library default
        // one share of the frame rendered by this process alone (one process per GPU; the caller exchanges the film, e.g. with
        // vmn_dist_create_rank — the in-process alternative is `devices` above)
        uint32_t tileSize = 0, tileRank = 0, tileWorld = 1;
        uint32_t sampleBegin = 0, sampleEnd = 0;
        // progressive rendering with checkpoints (FIXED mode, whole sample range): the film sums are written to checkpointPath
        // (film_io.h) every checkpointEverySpp samples and a matching checkpoint found there is resumed instead of starting over
        std::string checkpointPath;
        uint32_t checkpointEverySpp = 0;
    };

    class B200PathTracer : public Integrator
    {
    public:
        explicit B200PathTracer(const B200Options& options = B200Options());
        ~B200PathTracer() override;
        void Render(std::vector<Vermilion::Camera*>& cameraList, MeshEngine* mEng) override;

        bool ok() const { return mGroup != nullptr && mLastStatus == VMN_OK; }
        const std::string& lastError() const { return mLastError; }
        const vmn_stats& stats() const { return mStats; }            // of devices[0]; rays_traced summed over all devices
        const vmn_dist_stats& groupStats() const { return mGroupStats; }
        vmn_ctx* context() { return mGroup ? vmn_dist_context(mGroup, 0) : nullptr; }
        // The GPU copy of the scene is keyed on the CONTENT of what Render uploads (vertex, normal, uv and texture data), so an edit of
        // the MeshEngine is picked up by the next Render without any call; invalidate() only forces the next Render to upload again.
        void invalidate() { mSceneHash = 0; mHaveScene = false; }

    private:
        bool uploadScene(MeshEngine* mEng);
        bool renderProgressive(const vmn_camera& vc);
        bool fail(const char* what);
        uint64_t frameFingerprint(const vmn_camera& vc) const;

        B200Options mOpt;
        vmn_dist* mGroup = nullptr;
        vmn_status mLastStatus = VMN_OK;
        std::string mLastError;
        vmn_stats mStats{};
        vmn_dist_stats mGroupStats{};
        uint64_t mSceneHash = 0; // content hash of the uploaded scene
        bool mHaveScene = false;
        uint64_t mProgressiveRays = 0;
    };

    // Drop-in for `new BruteForceTracer()` (integrators.h:18-21): same adapter, VMN_MODE_BRUTEFORCE. boundTextures[0] is the
    // normal map and boundTextures[1] the albedo, as in integrators.cpp:98-106 / :141-147.
    class B200BruteForceTracer : public B200PathTracer
    {
    public:
        explicit B200BruteForceTracer(B200Options options = B200Options()) : B200PathTracer((options.mode = VMN_MODE_BRUTEFORCE, options)) {}
    };
}
#endif
