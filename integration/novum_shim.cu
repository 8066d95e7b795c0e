// novum_shim.cu — the reference-side binding of libnovum_b200.so: drop this file into the reference's build IN PLACE OF deviceCode.cu
// and link -lnovum_b200. It defines the five host functions declared in the reference's deviceCode.cuh:6-18 (C++ linkage, the
// reference's own types) and forwards each to the extern "C" entry point of include/novum_b200.h with the same argument list.
// main.cu / initRender (main.cu:235-923) stays untouched.
//
//   nvcc -arch=sm_100a -O3 main.cu novum_shim.cu imageUtil.cu -I<repo>/include -L<repo>/cudapathtracer_b200 -lnovum_b200 \
//        -Xlinker -rpath=<repo>/cudapathtracer_b200 -o novum
//
// The repository builds exactly this (the reference's unmodified main.cu + imageUtil.cu + this file) as oracle/_ref/novum_dropin and
// tests it against the reference's own renders (tests/test_gpu_dropin.py).
#include "deviceCode.cuh"            // the reference's declarations
#include "novum_b200.h"
#include <cstdio>
#include <cstdlib>

static_assert(sizeof(Camera) == sizeof(NovumCamera) && sizeof(Material) == sizeof(NovumMaterial) && sizeof(Triangle) == sizeof(NovumTriangle) &&
              sizeof(BVHnode) == sizeof(NovumBVHnode) && sizeof(Vertices) == sizeof(NovumVertices), "reference PODs and the C ABI structs must match");

// NOVUM_GPUS=N in the environment spreads a render over N GPUs of this process; the reference itself is single-GPU
static int novumGpus() { const char* e = getenv("NOVUM_GPUS"); int n = e ? atoi(e) : 1; return n > 0 ? n : 1; }

// the reference prints cudaGetLastError at the end of each launcher and carries on (deviceCode.cu:274-282, 1950-1955, 3488-3496)
static void novumReport(const char* what, int rc) {
    if (rc) fprintf(stderr, "RENDER ERROR in %s: %s (code %d)\n", what, novum_error_string(rc), rc);
    else printf("Render executed with no error (%s)\n", what);
}

static NovumFloat4 nf4(const float4& v) { NovumFloat4 r = {v.x, v.y, v.z, v.w}; return r; }
static NovumCamera ncam(const Camera& c) { NovumCamera r; memcpy(&r, &c, sizeof(r)); return r; }

__host__ void updateConstants(RenderConfig& c) {            // deviceCode.cu:38-51
    NovumConstants k = {c.sampleEnvironment, c.bdptLightTrace, c.bdptNee, c.bdptNaive, c.bdptConnection,
                        c.vcmDoMerge, c.doSPPM, c.bdptDrawPath, c.bdptDoMis, c.bdptPaintWeight};
    int rc = novum_update_constants(&k);
    if (rc) novumReport("updateConstants", rc);
}

__host__ void launch_unidirectional(int maxDepth, Camera camera, Material* materials, float4* textures, BVHnode* BVH, int* BVHindices,
                                    Vertices* vertices, int vertNum, Triangle* scene, int triNum, Triangle* lights, int lightNum, int numSample,
                                    bool useMIS, int w, int h, float4* colors) {
    novumReport("launch_unidirectional",
                novum_launch_unidirectional(maxDepth, ncam(camera), (NovumMaterial*)materials, (NovumFloat4*)textures, (NovumBVHnode*)BVH, BVHindices,
                                            (NovumVertices*)vertices, vertNum, (NovumTriangle*)scene, triNum, (NovumTriangle*)lights, lightNum,
                                            numSample, useMIS ? 1 : 0, w, h, (NovumFloat4*)colors, novumGpus()));
}

__host__ void launch_naive_unidirectional(int maxDepth, Camera camera, Material* materials, float4* textures, BVHnode* BVH, int* BVHindices,
                                          Vertices* vertices, int vertNum, Triangle* scene, int triNum, Triangle* lights, int lightNum,
                                          int numSample, bool useMIS, int w, int h, float4* colors) {
    novumReport("launch_naive_unidirectional",
                novum_launch_naive_unidirectional(maxDepth, ncam(camera), (NovumMaterial*)materials, (NovumFloat4*)textures, (NovumBVHnode*)BVH,
                                                  BVHindices, (NovumVertices*)vertices, vertNum, (NovumTriangle*)scene, triNum,
                                                  (NovumTriangle*)lights, lightNum, numSample, useMIS ? 1 : 0, w, h, (NovumFloat4*)colors, novumGpus()));
}

__host__ void launch_bidirectional(int eyeDepth, int lightDepth, Camera camera, PathVertices* eyePath, PathVertices* lightPath, Material* materials,
                                   float4* textures, BVHnode* BVH, int* BVHindices, Vertices* vertices, int vertNum, Triangle* scene, int triNum,
                                   Triangle* lights, int lightNum, int numSample, int w, int h, float4 sceneCenter, float sceneRadius,
                                   float4* colors, float4* overlay, bool postProcess) {
    novumReport("launch_bidirectional",
                novum_launch_bidirectional(eyeDepth, lightDepth, ncam(camera), eyePath, lightPath, (NovumMaterial*)materials, (NovumFloat4*)textures,
                                           (NovumBVHnode*)BVH, BVHindices, (NovumVertices*)vertices, vertNum, (NovumTriangle*)scene, triNum,
                                           (NovumTriangle*)lights, lightNum, numSample, w, h, nf4(sceneCenter), sceneRadius,
                                           (NovumFloat4*)colors, (NovumFloat4*)overlay, postProcess ? 1 : 0, novumGpus()));
}

__host__ void launch_VCM(int eyeDepth, int lightDepth, Camera camera, VCMPathVertices* lightPath, Photons* photons, Photons* photons_sorted,
                         Material* materials, float4* textures, BVHnode* BVH, int* BVHindices, Vertices* vertices, int vertNum, Triangle* scene,
                         int triNum, Triangle* lights, int lightNum, int numSample, int w, int h, float4 sceneCenter, float sceneRadius,
                         float4 sceneMin, float4* colors, float4* overlay, bool postProcess, float mergeRadiusPower, float initialRadiusMultiplier) {
    novumReport("launch_VCM",
                novum_launch_vcm(eyeDepth, lightDepth, ncam(camera), lightPath, photons, photons_sorted, (NovumMaterial*)materials,
                                 (NovumFloat4*)textures, (NovumBVHnode*)BVH, BVHindices, (NovumVertices*)vertices, vertNum, (NovumTriangle*)scene,
                                 triNum, (NovumTriangle*)lights, lightNum, numSample, w, h, nf4(sceneCenter), sceneRadius, nf4(sceneMin),
                                 (NovumFloat4*)colors, (NovumFloat4*)overlay, postProcess ? 1 : 0, mergeRadiusPower, initialRadiusMultiplier,
                                 novumGpus()));
}
