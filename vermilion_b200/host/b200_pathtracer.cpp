// b200_pathtracer.cpp — Vermilion::B200PathTracer, see b200_pathtracer.h.
//
// Replaces PathTracer::Render (reference core/integrators/pathtracer.cpp:200-328) with calls into the C ABI:
//   geometry  : MeshEngine::sceneMeshes walked exactly as MeshEngine::createBVH does (core/engines/meshEngine.cpp:660-717),
//               which defines the canonical triangle order; uv only when the mesh HasTextureCoords(0)
//   spheres   : the eight analytic spheres hard-coded in MeshEngine::RayCast (meshEngine.cpp:377-500) via vmn_reference_spheres
//   texture   : MeshEngine::boundTextures[0], the only one Radiance samples (pathtracer.cpp:62-80), and boundTextures[1],
//               BruteForceTracer's albedo (integrators.cpp:141-147)
//   camera    : the public Camera fields the pixel loop reads (core/camera/camera.h:66-104)
//   film      : written back into Camera::mImage in the layout of Camera::renderMode (camera.cpp:88-124)
#include "b200_pathtracer.h"
#include "film_io.h"

#include <cstdio>
#include <cstring>

namespace Vermilion
{

B200PathTracer::B200PathTracer(const B200Options& options) : mOpt(options)
{
    std::vector<int32_t> devices(mOpt.devices.begin(), mOpt.devices.end());
    if (devices.empty()) devices.push_back(mOpt.device);
    mLastStatus = vmn_dist_create_local(devices.data(), (uint32_t)devices.size(), 0, mOpt.maxPathsInFlight, &mGroup);
    if (mLastStatus != VMN_OK)
    {
        mLastError = mGroup ? vmn_dist_last_error(mGroup) : "vmn_dist_create_local: out of memory";
        fprintf(stderr, "[B200PathTracer] cannot create the GPU engine: %s\n", mLastError.c_str());
        if (mGroup) vmn_dist_destroy(mGroup);
        mGroup = nullptr;
    }
}

B200PathTracer::~B200PathTracer()
{
    if (mGroup) vmn_dist_destroy(mGroup);
}

bool B200PathTracer::fail(const char* what)
{
    mLastError = std::string(what) + ": " + (mGroup ? vmn_dist_last_error(mGroup) : "no engine");
    fprintf(stderr, "[B200PathTracer] %s\n", mLastError.c_str());
    return false;
}

bool B200PathTracer::uploadScene(MeshEngine* mEng)
{
    uint64_t nFaces = 0;
    for (aiMesh* mesh : mEng->sceneMeshes) nFaces += mesh->mNumFaces;
    std::vector<float> pos(9 * nFaces), nrm(9 * nFaces), uv(6 * nFaces, 0.f);
    uint64_t f = 0;
    for (aiMesh* mesh : mEng->sceneMeshes)
    {
        // createBVH dereferences mNormals unconditionally (meshEngine.cpp:690-704); a mesh without normals crashes the reference —
        // here it is an error the caller can read
        if (mesh->mNumFaces && (!mesh->mVertices || !mesh->mNormals || !mesh->mFaces))
        {
            mLastStatus = VMN_ERR_INVALID;
            mLastError = "a mesh has no vertex normals (MeshEngine::load asks Assimp for aiProcess_GenNormals; createBVH needs them)";
            fprintf(stderr, "[B200PathTracer] %s\n", mLastError.c_str());
            return false;
        }
        const bool hasUV = mesh->HasTextureCoords(0);
        for (uint32_t i = 0; i < mesh->mNumFaces; ++i, ++f)
        {
            for (int k = 0; k < 3; ++k)
            {
                const unsigned idx = mesh->mFaces[i].mIndices[k];
                pos[9 * f + 3 * k + 0] = mesh->mVertices[idx].x;
                pos[9 * f + 3 * k + 1] = mesh->mVertices[idx].y;
                pos[9 * f + 3 * k + 2] = mesh->mVertices[idx].z;
                nrm[9 * f + 3 * k + 0] = mesh->mNormals[idx].x;
                nrm[9 * f + 3 * k + 1] = mesh->mNormals[idx].y;
                nrm[9 * f + 3 * k + 2] = mesh->mNormals[idx].z;
                if (hasUV)
                {
                    uv[6 * f + 2 * k + 0] = mesh->mTextureCoords[0][idx].x;
                    uv[6 * f + 2 * k + 1] = mesh->mTextureCoords[0][idx].y;
                }
            }
        }
    }
    // cache key = the content that would be uploaded: same-count geometry edits and textures replaced in place change it
    uint64_t h = vmnh_hash64(pos.data(), pos.size() * 4, nFaces + 1);
    h = vmnh_hash64(nrm.data(), nrm.size() * 4, h);
    h = vmnh_hash64(uv.data(), uv.size() * 4, h);
    for (uint32_t slot = 0; slot < 2 && slot < mEng->boundTextures.size(); ++slot)
    {
        const VermiTexture& t = mEng->boundTextures[slot];
        const uint32_t dims[4] = {slot, t.nWidth, t.nHeight, t.nChannels};
        h = vmnh_hash64(dims, sizeof(dims), h);
        if (t.pData) h = vmnh_hash64(t.pData, (uint64_t)t.nWidth * t.nHeight * t.nChannels * sizeof(float), h);
    }
    if (h == 0) h = 1;
    if (mHaveScene && h == mSceneHash) return true;
    mHaveScene = false;

    if ((mLastStatus = vmn_dist_scene_set_triangles(mGroup, pos.data(), nrm.data(), uv.data(), nFaces, VMN_DIST_SCENE_BROADCAST)) != VMN_OK) return fail("vmn_dist_scene_set_triangles");
    vmn_sphere spheres[8];
    vmn_reference_spheres(spheres);
    if ((mLastStatus = vmn_dist_scene_set_spheres(mGroup, spheres, 8)) != VMN_OK) return fail("vmn_dist_scene_set_spheres");
    for (uint32_t slot = 0; slot < 2; ++slot)
    {
        if (slot < mEng->boundTextures.size())
        {
            const VermiTexture& t = mEng->boundTextures[slot];
            if ((mLastStatus = vmn_dist_scene_set_texture(mGroup, slot, t.pData, t.nWidth, t.nHeight, t.nChannels)) != VMN_OK) return fail("vmn_dist_scene_set_texture");
        }
        else vmn_dist_scene_set_texture(mGroup, slot, nullptr, 0, 0, 0); // unbind
    }
    if ((mLastStatus = vmn_dist_scene_commit(mGroup)) != VMN_OK) return fail("vmn_dist_scene_commit");
    mSceneHash = h; mHaveScene = true;
    return true;
}

// what a checkpointed film is a picture of: camera, estimator parameters and scene content
uint64_t B200PathTracer::frameFingerprint(const vmn_camera& vc) const
{
    uint64_t h = vmnh_hash64(&vc, sizeof(vc), mSceneHash);
    const uint64_t est[3] = {(uint64_t)mOpt.mode, mOpt.maxSegments, mOpt.seed};
    return vmnh_hash64(est, sizeof(est), h);
}

// FIXED mode in slices of checkpointEverySpp samples, the film sums saved after every slice; a checkpoint of the same frame (size, spp,
// seed, segment cap AND fingerprint: camera pose, film geometry, scene and texture content) is picked up where it stopped; any other
// file at that path is ignored and overwritten. Slices accumulate in the per-pixel sample order, so on one GPU the result is
// bit-identical to the one-shot render.
bool B200PathTracer::renderProgressive(const vmn_camera& vc)
{
    const uint64_t npix = (uint64_t)vc.width * vc.height;
    std::vector<float> film(4 * npix);
    uint32_t done = 0;
    const uint64_t fp = frameFingerprint(vc);
    vmnh_checkpoint_header h{};
    vmn_ctx* root = vmn_dist_context(mGroup, 0);
    if (vmnh_checkpoint_info(mOpt.checkpointPath.c_str(), &h) == 0)
    {
        const bool same = h.width == vc.width && h.height == vc.height && h.spp_total == vc.spp && h.seed == mOpt.seed &&
                          h.max_segments == mOpt.maxSegments && h.spp_done <= vc.spp && h.frame_fingerprint == fp;
        if (!same) fprintf(stderr, "[B200PathTracer] %s is a checkpoint of another frame (size, spp, seed, camera or scene differ): ignored\n", mOpt.checkpointPath.c_str());
        else if (vmnh_checkpoint_load(mOpt.checkpointPath.c_str(), film.data(), film.size(), &h) == 0)
        {
            if ((mLastStatus = vmn_film_upload(root, film.data())) != VMN_OK) return fail("vmn_film_upload");
            done = h.spp_done;
            fprintf(stderr, "[B200PathTracer] resuming %s at %u of %u spp\n", mOpt.checkpointPath.c_str(), done, vc.spp);
        }
    }
    const uint32_t every = mOpt.checkpointEverySpp ? mOpt.checkpointEverySpp : vc.spp;
    bool first = done == 0;
    mProgressiveRays = 0;
    while (done < vc.spp)
    {
        vmn_render_params rp{};
        rp.mode = VMN_MODE_FIXED; rp.max_segments = mOpt.maxSegments; rp.seed = mOpt.seed;
        rp.sample_begin = done; rp.sample_end = done + every < vc.spp ? done + every : vc.spp;
        // the root's film is the running total (a resumed checkpoint lives there); every pass is assembled before it is saved
        rp.flags = (first ? 0u : VMN_RENDER_ACCUMULATE) | VMN_DIST_RENDER_ASSEMBLE;
        if (first && vmn_dist_world(mGroup) > 1) rp.flags |= 0u; // first pass clears every film
        if ((mLastStatus = vmn_dist_render(mGroup, &rp, VMN_DIST_SPLIT_SAMPLES, 0)) != VMN_OK) return fail("vmn_dist_render");
        first = false;
        done = rp.sample_end;
        { vmn_dist_stats gs{}; vmn_dist_get_stats(mGroup, &gs); mProgressiveRays += gs.rays_traced; }
        if ((mLastStatus = vmn_dist_read_image(mGroup, film.data(), VMN_FILM_SUM)) != VMN_OK) return fail("vmn_dist_read_image");
        memcpy(h.magic, "VMNFILM2", 8);
        h.width = vc.width; h.height = vc.height; h.spp_done = done; h.spp_total = vc.spp; h.seed = mOpt.seed; h.max_segments = mOpt.maxSegments;
        h.frame_fingerprint = fp;
        if (vmnh_checkpoint_save(mOpt.checkpointPath.c_str(), film.data(), &h) != 0)
            fprintf(stderr, "[B200PathTracer] cannot write checkpoint %s\n", mOpt.checkpointPath.c_str());
    }
    return true;
}

void B200PathTracer::Render(std::vector<Vermilion::Camera*>& cameraList, MeshEngine* mEng)
{
    if (!mGroup) { fprintf(stderr, "[B200PathTracer] no GPU engine (there is no CPU fallback); frame not rendered\n"); return; }
    if (!mEng) { mLastStatus = VMN_ERR_INVALID; mLastError = "null MeshEngine"; return; }
    if (!uploadScene(mEng)) return;
    vmn_ctx* root = vmn_dist_context(mGroup, 0);

    for (Vermilion::Camera* cam : cameraList) // pathtracer.cpp:210
    {
        vmn_camera vc{};
        vc.width = cam->uImageU; vc.height = cam->uImageV;
        for (int i = 0; i < 3; ++i) { vc.position[i] = cam->mPosition[i]; vc.rotation[i] = cam->mRotation[i]; }
        vc.dist_to_film = cam->mDistToFilm;
        vc.sensor_x = cam->sensorSizeX; vc.sensor_y = cam->sensorSizeY;
        vc.spp = cam->uSamplesPerPixel;
        if ((mLastStatus = vmn_dist_camera_set(mGroup, &vc)) != VMN_OK) { fail("vmn_dist_camera_set"); return; }

        vmn_render_params rp{};
        rp.mode = mOpt.mode; rp.max_segments = mOpt.maxSegments; rp.seed = mOpt.seed;
        rp.sample_begin = mOpt.sampleBegin; rp.sample_end = mOpt.sampleEnd;
        if (mOpt.mode == VMN_MODE_BRUTEFORCE && mOpt.shadowRays) rp.flags |= VMN_RENDER_BF_SHADOW;
        const bool progressive = !mOpt.checkpointPath.empty() && mOpt.mode == VMN_MODE_FIXED && mOpt.tileSize == 0 && mOpt.sampleEnd == 0;
        if (progressive) { if (!renderProgressive(vc)) return; }
        else if (mOpt.tileSize)
        {
            // one share of the frame on this process's (single) device; the caller assembles the film
            if (vmn_dist_world(mGroup) != 1) { mLastStatus = VMN_ERR_INVALID; mLastError = "tileSize/tileRank describe a single-device share: leave `devices` empty"; return; }
            rp.tile_size = mOpt.tileSize; rp.tile_rank = mOpt.tileRank; rp.tile_world = mOpt.tileWorld;
            if ((mLastStatus = vmn_render(root, &rp)) != VMN_OK) { mLastError = std::string("vmn_render: ") + vmn_last_error(root); fprintf(stderr, "[B200PathTracer] %s\n", mLastError.c_str()); return; }
        }
        else
        {
            const int split = (mOpt.splitSamples && mOpt.mode == VMN_MODE_FIXED) ? VMN_DIST_SPLIT_SAMPLES : VMN_DIST_SPLIT_TILES;
            if ((mLastStatus = vmn_dist_render(mGroup, &rp, split, 0)) != VMN_OK) { fail("vmn_dist_render"); return; }
        }
        vmn_get_stats(root, &mStats);
        vmn_dist_get_stats(mGroup, &mGroupStats);
        if (progressive) mStats.rays_traced = mProgressiveRays; // rays of the passes this call rendered (a resumed frame: fewer)
        else if (vmn_dist_world(mGroup) > 1 && !mOpt.tileSize) mStats.rays_traced = mGroupStats.rays_traced;

        const uint64_t npix = (uint64_t)vc.width * vc.height;
        if (cam->renderMode == vermRenderMode::RGBAZ)
        {
            // Camera::mImage already has the RGBAZ layout (camera.cpp:106-113): bulk write
            if ((mLastStatus = vmn_dist_read_image(mGroup, cam->mImage, VMN_FILM_RGBAZ)) != VMN_OK) { fail("vmn_dist_read_image"); return; }
        }
        else
        {
            std::vector<float> img(5 * npix);
            if ((mLastStatus = vmn_dist_read_image(mGroup, img.data(), VMN_FILM_RGBAZ)) != VMN_OK) { fail("vmn_dist_read_image"); return; }
            for (uint64_t p = 0; p < npix; ++p) // pathtracer.cpp:318-324
            {
                pixelValue px;
                px.pixel = p;
                px.red = img[5 * p]; px.green = img[5 * p + 1]; px.blue = img[5 * p + 2]; px.alpha = img[5 * p + 3]; px.depth = img[5 * p + 4];
                px.light = 0.f;
                cam->setPixelValue(px); // throws std::exception for unsupported modes, as the reference does (camera.cpp:122)
            }
        }
    }
}

}
