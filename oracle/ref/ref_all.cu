// ref_all.cu — builds the REFERENCE ITSELF (unmodified sources, included where they lie under /root/reference)
// into one binary with a small driver around its own public functions. TEST / BASELINE INFRASTRUCTURE ONLY.
//
// Nothing from the reference is copied into this repository: this file only #includes the reference's
// deviceCode.cu and main.cu at build time (oracle/build_ref.sh, output oracle/_ref/novum_ref, git-ignored).
// Two MSVC-isms stop the shipped sources compiling with g++; both are handled from outside:
//   * `inline __host__ __device__ __forceinline__` is a duplicate `inline` under g++: every system header is
//     included FIRST (include guards make the reference's own includes no-ops), then __forceinline__ is redefined
//     without the `inline` keyword for the reference's headers only;
//   * <cstdint>/<cfloat>/<cmath> are missing from util.cuh: included here.
// The reference's main() is renamed by macro; its initRender() is left unused. The driver below restates
// initRender's flow (main.cu:235-557) with the reference's own loadConfig / readObjSimple / computeInfoForBVH /
// buildBVH / Camera / Material factories, calls the reference's launch_* and dumps the raw float accumulator.
// Material fields the reference leaves indeterminate are zeroed — the only behavioural edit.
#include <cstdint>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include <cuda.h>
#include <curand_kernel.h>
#include <cuda_fp16.h>
#include <cub/cub.cuh>
#include <algorithm>
#include <sstream>
#include <numeric>
#include <iostream>
#include <fstream>
#include <string>
#include <vector>
#include <random>
#include <ctime>
#include <chrono>
#include <set>
#include <iomanip>
#include <unistd.h>
#include <utility>
#undef __forceinline__
#define __forceinline__ __attribute__((always_inline))
#ifdef NOVUM_REF_COUNT_RAYS
// Counting build (novum_ref_count, never the timed one): the reference's two intersection routines are compiled under other
// names and re-exported through wrappers that bump device counters, so every call site in deviceCode.cu is counted without
// touching a line of the reference. Gives the reference's OWN rays per path for the Mrays/s conversion in bench.py.
__device__ unsigned long long g_refRayCounters[2];
#define BVHSceneIntersect BVHSceneIntersect_refimpl
#define BVHShadowRay BVHShadowRay_refimpl
#include "integratorUtilities.cuh"
#undef BVHSceneIntersect
#undef BVHShadowRay
__device__ inline void BVHSceneIntersect(const Ray& r, BVHnode* BVH, int* BVHindices, Vertices* verts, Triangle* scene, Intersection& intersect,
                                         float max_t = 999999.0f, int skipTri = -1) {
    atomicAdd(&g_refRayCounters[0], 1ull);
    BVHSceneIntersect_refimpl(r, BVH, BVHindices, verts, scene, intersect, max_t, skipTri);
}
__device__ inline void BVHShadowRay(const Ray& r, BVHnode* BVH, int* BVHindices, Vertices* verts, Triangle* scene, Material* materials,
                                    float4& throughputScale, float max_t, int skip_tri) {
    atomicAdd(&g_refRayCounters[1], 1ull);
    BVHShadowRay_refimpl(r, BVH, BVHindices, verts, scene, materials, throughputScale, max_t, skip_tri);
}
#endif
#ifdef NOVUM_REF_NOSAVE
// novum_ref_nosave: util.cuh (already included through the headers above? no - include it now, under its own guard) declares
// `constexpr bool DO_PROGRESSIVERENDER = true` (util.cuh:33); inside deviceCode.cu the name is then read as `false`, so the BDPT / VCM
// launchers never enter their 5-second save (deviceCode.cu:1893, 3447). The unidirectional launcher has no such switch (:236-267).
#include "util.cuh"
#define DO_PROGRESSIVERENDER false
#endif
#include "deviceCode.cu"
#ifdef NOVUM_REF_NOSAVE
#undef DO_PROGRESSIVERENDER
#endif
#define main novum_reference_main_unused
#include "main.cu"
#undef main
#include "imageUtil.cu"   // same TU: the redefined __forceinline__ drops `inline`, so one definition per program

// ---- harness kernels: the reference's own intersection routines on a ray batch --------------------------------
__global__ void refClosestKernel(const Ray* rays, int n, BVHnode* BVH, int* BVHindices, Vertices* verts, Triangle* scene, float4* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r = rays[i];
    Intersection it;
    BVHSceneIntersect(r, BVH, BVHindices, verts, scene, it);
    float4 o = make_float4(__int_as_float(-1), 0.f, 0.f, 0.f);
    if (it.valid) {
        float4 bary; float t;
        triangleIntersect(verts, &scene[it.triIDX], r, bary, t);
        o = make_float4(__int_as_float(it.triIDX), it.dist, bary.x, bary.y);
    }
    out[i] = o;
}

__global__ void refShadowKernel(const Ray* rays, const float* maxT, const int* skip, int n, BVHnode* BVH, int* BVHindices, Vertices* verts,
                                Triangle* scene, Material* mats, float4* out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 thr;
    BVHShadowRay(rays[i], BVH, BVHindices, verts, scene, mats, thr, maxT[i], skip[i]);
    out[i] = thr;
}

#include "ref_scene.inl"
#include "ref_render.inl"
#include "ref_driver.inl"
