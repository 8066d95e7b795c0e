// scene.h — host-side scene object behind the opaque NovumScene handle.
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include "../../include/novum_b200.h"
#include "bvh8_build.h"
#include "scene_dev.cuh"

struct NovumScene {
    // device geometry in the reference's layouts (owned unless borrowed from the caller)
    float4* d_positions = nullptr;
    float4* d_normals = nullptr;
    float2* d_uvs = nullptr;
    NovumTriangle* d_tris = nullptr;
    float4* d_shadeTris = nullptr;        // nv::packShadeTri records, built by sceneFinalize (owned by the scene in every creation path)
    NovumTriangle* d_lights = nullptr;
    NovumMaterial* d_matsRef = nullptr;   // 176 B reference-layout table
    nv::DMat* d_mats = nullptr;           // compact table
    float4* d_textures = nullptr;
    bool borrowed = false;                // geometry/material arrays belong to the caller (from_device)
    int nPos = 0, nNorm = 0, nUV = 0, nTri = 0, nLights = 0, nMat = 0, nTexels = 0;
    nv::Bvh8Device bvh;
    float sceneMin[3] = {0, 0, 0}, sceneMax[3] = {0, 0, 0}, sceneCenter[3] = {0, 0, 0};
    float sceneRadius = 0.f;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float lastTraceMs = 0.f;
    float parseMs = 0.f;                  // host time of the config / OBJ front end (novum_scene_load_config only)
    // traversal counters
    bool countersOn = false;
    int profileFlags = 0;                 // novum_set_profile
    unsigned long long* d_counters = nullptr;   // [0] node visits, [1] tri tests
    int* d_queueCounter = nullptr;              // persistent-kernel work counter
    // staging for the host-pointer trace API (grown on demand)
    void* d_stage = nullptr; size_t stageBytes = 0;
    // host-side config meshes (when loaded from a .rendertron)
    std::vector<std::string> meshPaths;
    // lazily created integrator engines (owned; freed through the stored deleter)
    void* wfEngine = nullptr;  void (*wfEngineFree)(void*) = nullptr;
    void* vcmEngine = nullptr; void (*vcmEngineFree)(void*) = nullptr;
    void* d_frame = nullptr; size_t frameBytes = 0;     // device framebuffer of novum_render, kept across calls
    int device = 0;                                     // the CUDA device the scene lives on
    int features = 0; float envRadiance[3] = {0.f, 0.f, 0.f};   // novum_scene_set_features / _set_environment
    void* multiGpu = nullptr; void (*multiGpuFree)(void*) = nullptr;   // replicas + per-GPU engines of the in-process multi-GPU launch (multigpu.cu)

    nv::SceneView view() const;
};

namespace nv {
int sceneFinalize(NovumScene* s);     // compact materials, BVH8 build, bounds, stream/events
// an owned copy of `src` (which lives on device srcDevice) on the CURRENT device: arrays copied peer to peer, BVH8 built here
int sceneReplicate(const NovumScene* src, NovumScene** out);
int ensureStage(NovumScene* s, size_t bytes);
int smCount();
int renderDeviceMis(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample, NovumShard shard,
                    uint64_t seed, NovumFloat4* d_colors, NovumRenderStats* stats, int useMIS);      // wavefront.cu
void bdRebindScene(void* bdEngine, NovumScene* s);                                                    // vcm.cu: engine buffers outlive a scene rebuild
}
