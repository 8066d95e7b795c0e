// launch_api.cu — shims with the reference's launch signatures (deviceCode.cuh:8-18). They take the DEVICE buffers
// main.cu hands to launch_* unchanged, wrap them in a borrowed scene (which builds our own BVH8 from them; the
// incoming BVH2 is ignored) and run the new engines. Like the reference they are blocking and add the SUM over
// samples into `colors`; unlike it they return an error code instead of printing.
#include "scene.h"
#include <cstring>

namespace nv { const NovumConstants& currentConstants(); }   // wavefront.cu: what novum_update_constants stored

namespace {

int nMatFromTriangles(const NovumTriangle* d_scene, int triNum) {
    // the reference passes no material count; the table built by initRender has 24 entries (main.cu:443-467).
    (void)d_scene; (void)triNum;
    return 24;
}

int runShim(int integrator, int maxDepth, int eyeDepth, int lightDepth, const NovumCamera& cam, NovumMaterial* materials,
            NovumFloat4* textures, NovumVertices* vertices, int vertNum, NovumTriangle* scene, int triNum, NovumTriangle* lights,
            int lightNum, int numSample, int w, int h, NovumFloat4* colors, float mergePower, float radiusMult) {
    NovumScene* s = nullptr;
    int rc = novum_scene_from_device(vertices, vertNum, scene, triNum, lights, lightNum, materials, nMatFromTriangles(scene, triNum),
                                     textures, &s);
    if (rc) return rc;
    NovumRenderConfig cfg; memset(&cfg, 0, sizeof(cfg));
    cfg.width = w; cfg.height = h; cfg.integrator = integrator; cfg.sampleCount = numSample; cfg.maxDepth = maxDepth;
    cfg.bdptEyeDepth = eyeDepth; cfg.bdptLightDepth = lightDepth;
    cfg.vcmMergeConst = mergePower; cfg.vcmInitialMergeRadiusMultiplier = radiusMult;
    // the reference's kernels read the strategy toggles from __constant__ memory set by updateConstants (deviceCode.cu:9-51)
    const NovumConstants& k = nv::currentConstants();
    cfg.sampleEnvironment = k.sampleEnvironment; cfg.bdptLightTrace = k.bdptLightTrace; cfg.bdptNee = k.bdptNee; cfg.bdptNaive = k.bdptNaive;
    cfg.bdptConnection = k.bdptConnection; cfg.vcmDoMerge = k.vcmDoMerge; cfg.doSPPM = k.doSPPM; cfg.bdptDrawPath = k.bdptDrawPath;
    cfg.bdptDoMis = k.bdptDoMis; cfg.bdptPaintWeight = k.bdptPaintWeight;
    NovumShard shard = {0, 1};
    rc = novum_render_device(s, &cfg, &cam, 0, numSample, shard, 103033ull /* deviceCode.cu:215 */, colors, nullptr);
    novum_scene_destroy(s);
    return rc;
}

}  // namespace

extern "C" {

int novum_launch_unidirectional(int maxDepth, NovumCamera camera, NovumMaterial* materials, NovumFloat4* textures,
                                NovumBVHnode* BVH, int* BVHindices, NovumVertices* vertices, int vertNum, NovumTriangle* scene,
                                int triNum, NovumTriangle* lights, int lightNum, int numSample, int useMIS, int w, int h,
                                NovumFloat4* colors) {
    (void)BVH; (void)BVHindices; (void)useMIS;   // main.cu:565 always passes useMIS = true
    return runShim(NOVUM_UNIDIRECTIONAL, maxDepth, 0, 0, camera, materials, textures, vertices, vertNum, scene, triNum, lights, lightNum,
                   numSample, w, h, colors, 0.f, 0.f);
}

int novum_launch_naive_unidirectional(int maxDepth, NovumCamera camera, NovumMaterial* materials, NovumFloat4* textures,
                                      NovumBVHnode* BVH, int* BVHindices, NovumVertices* vertices, int vertNum, NovumTriangle* scene,
                                      int triNum, NovumTriangle* lights, int lightNum, int numSample, int useMIS, int w, int h,
                                      NovumFloat4* colors) {
    (void)BVH; (void)BVHindices; (void)useMIS;
    return runShim(NOVUM_NAIVE_UNIDIRECTIONAL, maxDepth, 0, 0, camera, materials, textures, vertices, vertNum, scene, triNum, lights,
                   lightNum, numSample, w, h, colors, 0.f, 0.f);
}

int novum_launch_bidirectional(int eyeDepth, int lightDepth, NovumCamera camera, void* eyePath, void* lightPath,
                               NovumMaterial* materials, NovumFloat4* textures, NovumBVHnode* BVH, int* BVHindices,
                               NovumVertices* vertices, int vertNum, NovumTriangle* scene, int triNum, NovumTriangle* lights,
                               int lightNum, int numSample, int w, int h, NovumFloat4 sceneCenter, float sceneRadius,
                               NovumFloat4* colors, NovumFloat4* overlay, int postProcess) {
    (void)eyePath; (void)lightPath; (void)BVH; (void)BVHindices; (void)sceneCenter; (void)sceneRadius; (void)overlay; (void)postProcess;
    return runShim(NOVUM_BIDIRECTIONAL, 0, eyeDepth, lightDepth, camera, materials, textures, vertices, vertNum, scene, triNum, lights,
                   lightNum, numSample, w, h, colors, 0.f, 0.f);
}

int novum_launch_vcm(int eyeDepth, int lightDepth, NovumCamera camera, void* lightPath, void* photons, void* photons_sorted,
                     NovumMaterial* materials, NovumFloat4* textures, NovumBVHnode* BVH, int* BVHindices, NovumVertices* vertices,
                     int vertNum, NovumTriangle* scene, int triNum, NovumTriangle* lights, int lightNum, int numSample, int w, int h,
                     NovumFloat4 sceneCenter, float sceneRadius, NovumFloat4 sceneMin, NovumFloat4* colors, NovumFloat4* overlay,
                     int postProcess, float mergeRadiusPower, float initialRadiusMultiplier) {
    (void)lightPath; (void)photons; (void)photons_sorted; (void)BVH; (void)BVHindices; (void)sceneCenter; (void)sceneRadius;
    (void)sceneMin; (void)overlay; (void)postProcess;
    // VCM vs SPPM is chosen through the DO_SPPM constant in the reference (main.cu:325-332, deviceCode.cu:50)
    return runShim(NOVUM_VCM, 0, eyeDepth, lightDepth, camera, materials, textures, vertices, vertNum, scene, triNum, lights, lightNum,
                   numSample, w, h, colors, mergeRadiusPower, initialRadiusMultiplier);
}

}  // extern "C"
