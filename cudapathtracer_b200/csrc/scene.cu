// scene.cu — scene object lifetime, material compaction, and the ray-batch intersection entry points
// (novum_trace_closest / novum_trace_shadow) that mirror BVHSceneIntersect / BVHShadowRay
// (integratorUtilities.cuh:84, :188) on whole batches.
#include "scene.h"
#include "trav_stream.cuh"
#include <cstdio>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <vector>

using namespace nv;

// layouts the reference's main.cu hands over unchanged (sizes printed from its own headers)
static_assert(sizeof(NovumTriangle) == 80, "Triangle");
static_assert(sizeof(NovumMaterial) == 176, "Material");
static_assert(sizeof(NovumCamera) == 112, "Camera");
static_assert(sizeof(NovumBVHnode) == 48, "BVHnode");
static_assert(sizeof(NovumVertices) == 32, "Vertices");
static_assert(sizeof(NovumRay) == 32 && sizeof(NovumHit) == 16, "Ray/Hit");
static_assert(sizeof(nv::DMat) == 96, "DMat");

namespace nv {

int smCount() {
    static int n = 0;
    if (!n) {
        int dev = 0; cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        if (n <= 0) n = 148;
    }
    return n;
}

__global__ void compactMaterialsKernel(const NovumMaterial* __restrict__ in, DMat* __restrict__ out, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const NovumMaterial& m = in[i];
    DMat d;
    d.albedo[0] = m.albedo.x; d.albedo[1] = m.albedo.y; d.albedo[2] = m.albedo.z; d.type = m.type;
    d.eta[0] = m.eta.x; d.eta[1] = m.eta.y; d.eta[2] = m.eta.z; d.roughness = m.roughness;
    d.k[0] = m.k.x; d.k[1] = m.k.y; d.k[2] = m.k.z; d.ior = m.ior;
    d.absorption[0] = m.absorption.x; d.absorption[1] = m.absorption.y; d.absorption[2] = m.absorption.z;
    d.transmission = m.transmission;
    d.texStart = m.startInd; d.texW = m.width; d.texH = m.height;
    // The reference's Material() constructor initialises none of hasTexture / hasTransMap / isSpecular / boundary / thinWalled /
    // absorption / priority, and each factory sets only some of them (objects.cuh:640-790): a table built by the reference's own
    // initRender arrives with indeterminate bytes in the rest. They are read here as the factories' evident intent - zero:
    //   hasTransMap is defined by Leaf and Mirror only; priority by SmoothDielectric only; Mirror leaves boundary / thinWalled /
    //   absorption unset; a texture of zero size (a missing file, imageUtil.cu:146-149) cannot be sampled (modulo by zero, reflectors.cuh:393).
    const bool leafOrMirror = m.type == NOVUM_MAT_LEAF || m.type == NOVUM_MAT_DELTAMIRROR;
    const bool mirror = m.type == NOVUM_MAT_DELTAMIRROR;
    const bool texOk = m.hasTexture && m.width > 0 && m.height > 0;
    if (!texOk) { d.texStart = 0; d.texW = 0; d.texH = 0; }
    if (mirror || m.type == NOVUM_MAT_LEAF) { d.absorption[0] = d.absorption[1] = d.absorption[2] = 0.0f; }
    d.flags = (texOk ? MF_TEXTURE : 0u) | ((leafOrMirror && m.hasTransMap && texOk) ? MF_TRANSMAP : 0u) | (m.isSpecular ? MF_SPECULAR : 0u) |
              ((!mirror && m.boundary) ? MF_BOUNDARY : 0u) | ((!mirror && m.thinWalled) ? MF_THINWALLED : 0u);
    d.priority = m.type == NOVUM_MAT_SMOOTHDIELECTRIC ? m.priority : 0; d.pad[0] = d.pad[1] = d.pad[2] = 0;
    out[i] = d;
}

__global__ void shadeTriKernel(const NovumTriangle* __restrict__ tris, const float4* __restrict__ normals, const float2* __restrict__ uvs,
                               int n, float4* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    TriRecord t = loadTri(tris, i);
    packShadeTri(out + (size_t)i * kShadeTriWords, V3(__ldg(normals + t.na)), V3(__ldg(normals + t.nb)), V3(__ldg(normals + t.nc)),
                 __ldg(uvs + t.uva), __ldg(uvs + t.uvb), __ldg(uvs + t.uvc), t.mat, t.emission, t.lightInd);
}

int ensureStage(NovumScene* s, size_t bytes) {
    if (bytes <= s->stageBytes) return 0;
    if (s->d_stage) cudaFree(s->d_stage);
    s->d_stage = nullptr; s->stageBytes = 0;
    size_t want = bytes + bytes / 4;
    cudaError_t e = cudaMalloc(&s->d_stage, want);
    if (e != cudaSuccess) return (int)e;
    s->stageBytes = want;
    return 0;
}

// largest normal / uv index a triangle names (+1): the lengths of those arrays do not cross the launch boundary
__global__ void attrCountKernel(const NovumTriangle* __restrict__ tris, int n, int* __restrict__ out) {
    int mn = -1, mu = -1;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const NovumTriangle& t = tris[i];
        mn = max(mn, max(t.naInd, max(t.nbInd, t.ncInd)));
        mu = max(mu, max(t.uvaInd, max(t.uvbInd, t.uvcInd)));
    }
    for (int o = 16; o > 0; o >>= 1) { mn = max(mn, __shfl_xor_sync(0xffffffffu, mn, o)); mu = max(mu, __shfl_xor_sync(0xffffffffu, mu, o)); }
    if ((threadIdx.x & 31) == 0) { atomicMax(out, mn); atomicMax(out + 1, mu); }
}

int sceneFinalize(NovumScene* s) {
    cudaError_t e;
    cudaGetDevice(&s->device);
    if (!s->stream && (e = cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
    if (!s->ev0) { cudaEventCreate(&s->ev0); cudaEventCreate(&s->ev1); }
    if ((e = cudaMalloc(&s->d_mats, sizeof(DMat) * (s->nMat > 0 ? s->nMat : 1))) != cudaSuccess) return (int)e;
    if (s->nMat > 0) compactMaterialsKernel<<<(s->nMat + 63) / 64, 64, 0, s->stream>>>(s->d_matsRef, s->d_mats, s->nMat);
    if ((e = cudaMalloc(&s->d_counters, 2 * sizeof(unsigned long long))) != cudaSuccess) return (int)e;
    cudaMemsetAsync(s->d_counters, 0, 2 * sizeof(unsigned long long), s->stream);
    if ((e = cudaMalloc(&s->d_queueCounter, 4 * sizeof(int))) != cudaSuccess) return (int)e;
    e = buildBvh8(s->d_positions, s->d_tris, s->nTri, s->d_mats, s->stream, s->bvh);
    if (e != cudaSuccess) return (int)e;
    // A traversal pushes at most one entry (the rest of a node's hit group) per BVH8 level it descends, so the tree's depth bounds the
    // stack: a tree deeper than the stack (12 shared + 40 local entries) would silently lose subtrees - refuse it here instead.
    if (s->bvh.levels > kSmemStack + kLocalStack) return NOVUM_E_TOODEEP;
    if ((e = cudaMalloc(&s->d_shadeTris, sizeof(float4) * kShadeTriWords * (size_t)s->nTri)) != cudaSuccess) return (int)e;
    shadeTriKernel<<<(s->nTri + 127) / 128, 128, 0, s->stream>>>(s->d_tris, s->d_normals, s->d_uvs, s->nTri, s->d_shadeTris);
    // main.cu:532-535: centre / radius / min of the root box
    for (int k = 0; k < 3; ++k) { s->sceneMin[k] = s->bvh.boundsMin[k]; s->sceneMax[k] = s->bvh.boundsMax[k]; }
    float c[3], d2 = 0.f;
    for (int k = 0; k < 3; ++k) { c[k] = (s->sceneMax[k] + s->sceneMin[k]) * 0.5f; float d = s->sceneMax[k] - c[k]; d2 += d * d; s->sceneCenter[k] = c[k]; }
    s->sceneRadius = sqrtf(d2) + 0.01f;
    return 0;
}

// ---- ray-batch kernels -----------------------------------------------------------------------------
constexpr int kTraceThreads = 128;

// Persistent warps stream the batch through their lanes with per-lane refill (trav_stream.cuh); stack in shared memory.
template <bool COUNT>
__global__ void __launch_bounds__(kTraceThreads) closestBatchKernel(SceneView sc, const NovumRay* __restrict__ rays, int n,
                                                                     NovumHit* __restrict__ hits, int* __restrict__ counter,
                                                                     TravCounters cnt) {
    __shared__ uint2 stackMem[kSmemStack * kTraceThreads];
    TravStack st; st.init(stackMem + threadIdx.x, kTraceThreads);
    streamClosest<COUNT>(sc.bvh, st, n, counter, cnt,
        [&](int i, V3& o, V3& d) {
            const float4* rp = reinterpret_cast<const float4*>(rays + i);
            float4 ro = __ldg(rp), rd = __ldg(rp + 1);
            o = V3(ro.x, ro.y, ro.z); d = V3(rd.x, rd.y, rd.z);
        },
        [&](int i, const HitRec& h) {
            reinterpret_cast<float4*>(hits)[i] = make_float4(__int_as_float(h.tri), h.tri >= 0 ? h.t : 0.f, h.u, h.v);
        },
        [](bool) {});
}

template <bool COUNT>
__global__ void __launch_bounds__(kTraceThreads) shadowBatchKernel(SceneView sc, const NovumRay* __restrict__ rays,
                                                                    const float* __restrict__ maxT, const int* __restrict__ skip, int n,
                                                                    float4* __restrict__ out, int* __restrict__ counter, TravCounters cnt) {
    __shared__ uint2 stackMem[kSmemStack * kTraceThreads];
    TravStack st; st.init(stackMem + threadIdx.x, kTraceThreads);
    streamShadow<COUNT>(sc, st, n, counter, cnt,
        [&](int i, V3& o, V3& d, float& mt, int& sk) {
            const float4* rp = reinterpret_cast<const float4*>(rays + i);
            float4 ro = __ldg(rp), rd = __ldg(rp + 1);
            o = V3(ro.x, ro.y, ro.z); d = V3(rd.x, rd.y, rd.z); mt = maxT[i]; sk = skip ? skip[i] : -1;
        },
        [&](int i, V3 t) { out[i] = make_float4(t.x, t.y, t.z, 0.f); });
}

}  // namespace nv

SceneView NovumScene::view() const {
    SceneView v;
    v.bvh.nodes = bvh.nodes; v.bvh.tris = bvh.tris; v.bvh.numNodes = bvh.numNodes; v.bvh.numTris = bvh.numTris;
    v.positions = d_positions; v.normals = d_normals; v.uvs = d_uvs;
    v.tris = d_tris; v.shadeTris = d_shadeTris; v.lights = d_lights; v.mats = d_mats; v.textures = d_textures;
    v.numTris = nTri; v.numLights = nLights; v.numMats = nMat;
    for (int k = 0; k < 3; ++k) { v.sceneMin[k] = sceneMin[k]; v.sceneCenter[k] = sceneCenter[k]; }
    v.sceneRadius = sceneRadius;
    v.features = features;
    for (int k = 0; k < 3; ++k) v.env[k] = envRadiance[k];
    return v;
}

extern "C" {

int novum_scene_create(const NovumFloat4* positions, int nPos, const NovumFloat4* normals, int nNorm,
                       const NovumFloat2* uvs, int nUV, const NovumTriangle* tris, int nTri,
                       const NovumMaterial* mats, int nMat, const NovumFloat4* textures, int nTexels, NovumScene** out) {
    if (!positions || !tris || !out || nPos <= 0 || nTri <= 0 || nMat <= 0 || !mats) return NOVUM_E_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return NOVUM_E_NOGPU;
    for (int i = 0; i < nTri; ++i) {   // the reference indexes normals[-1]/uvs[-1] when vn/vt are missing (main.cu:1044-1050): refuse
        const NovumTriangle& t = tris[i];
        if (t.aInd < 0 || t.bInd < 0 || t.cInd < 0 || t.aInd >= nPos || t.bInd >= nPos || t.cInd >= nPos) return NOVUM_E_ARG;
        if (t.naInd < 0 || t.nbInd < 0 || t.ncInd < 0 || t.naInd >= nNorm || t.nbInd >= nNorm || t.ncInd >= nNorm) return NOVUM_E_ARG;
        if (t.uvaInd < 0 || t.uvbInd < 0 || t.uvcInd < 0 || t.uvaInd >= nUV || t.uvbInd >= nUV || t.uvcInd >= nUV) return NOVUM_E_ARG;
        if (t.materialID < 0 || t.materialID >= nMat) return NOVUM_E_ARG;
    }
    NovumScene* s = new NovumScene();
    s->nPos = nPos; s->nNorm = nNorm; s->nUV = nUV; s->nTri = nTri; s->nMat = nMat; s->nTexels = nTexels;
    std::vector<NovumTriangle> lights;
    for (int i = 0; i < nTri; ++i) {   // main.cu:1018, 1060-1063: emissive triangles form the light list
        const NovumFloat4& e = tris[i].emission;
        if (e.x * e.x + e.y * e.y + e.z * e.z > 0.f) lights.push_back(tris[i]);
    }
    s->nLights = (int)lights.size();
    cudaError_t e = cudaSuccess;
    auto up = [&](void** dst, const void* src, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cudaMalloc(dst, bytes ? bytes : 16);
        if (e == cudaSuccess && bytes) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    up((void**)&s->d_positions, positions, sizeof(float4) * nPos);
    up((void**)&s->d_normals, normals, sizeof(float4) * nNorm);
    up((void**)&s->d_uvs, uvs, sizeof(float2) * nUV);
    up((void**)&s->d_tris, tris, sizeof(NovumTriangle) * nTri);
    up((void**)&s->d_lights, lights.data(), sizeof(NovumTriangle) * lights.size());
    up((void**)&s->d_matsRef, mats, sizeof(NovumMaterial) * nMat);
    up((void**)&s->d_textures, textures, sizeof(float4) * (nTexels > 0 ? nTexels : 0));
    // the uploads above are pageable copies on the legacy stream, what follows runs on the scene's own non-blocking stream
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { novum_scene_destroy(s); return (int)e; }
    int rc = sceneFinalize(s);
    if (rc) { novum_scene_destroy(s); return rc; }
    *out = s;
    return 0;
}

int novum_scene_from_device(const NovumVertices* d_vertices, int vertNum, const NovumTriangle* d_scene, int triNum,
                            const NovumTriangle* d_lights, int lightNum, const NovumMaterial* d_materials, int nMat,
                            const NovumFloat4* d_textures, NovumScene** out) {
    if (!d_vertices || !d_scene || !d_materials || !out || triNum <= 0 || nMat <= 0) return NOVUM_E_ARG;
    NovumVertices v;
    cudaError_t e = cudaMemcpy(&v, d_vertices, sizeof(v), cudaMemcpyDeviceToHost);   // the 32 B struct of device pointers
    if (e != cudaSuccess) return (int)e;
    NovumScene* s = new NovumScene();
    s->borrowed = true;
    s->d_positions = (float4*)v.positions; s->d_normals = (float4*)v.normals; s->d_uvs = (float2*)v.uvs;
    s->d_tris = (NovumTriangle*)d_scene; s->d_lights = (NovumTriangle*)d_lights;
    s->d_matsRef = (NovumMaterial*)d_materials; s->d_textures = (float4*)d_textures;
    s->nPos = vertNum; s->nNorm = 0; s->nUV = 0; s->nTri = triNum; s->nLights = lightNum; s->nMat = nMat;
    int rc = sceneFinalize(s);
    if (rc) { novum_scene_destroy(s); return rc; }
    {   // lengths of the normal / uv / texel arrays, as far as the scene uses them (needed to replicate the scene on another GPU)
        int h[2] = {-1, -1};
        cudaMemcpyAsync(s->d_queueCounter, h, sizeof(h), cudaMemcpyHostToDevice, s->stream);
        attrCountKernel<<<smCount() * 4, 256, 0, s->stream>>>(s->d_tris, s->nTri, s->d_queueCounter);
        cudaMemcpyAsync(h, s->d_queueCounter, sizeof(h), cudaMemcpyDeviceToHost, s->stream);
        std::vector<NovumMaterial> hm((size_t)nMat);
        cudaMemcpyAsync(hm.data(), d_materials, sizeof(NovumMaterial) * (size_t)nMat, cudaMemcpyDeviceToHost, s->stream);
        if ((e = cudaStreamSynchronize(s->stream)) != cudaSuccess) { novum_scene_destroy(s); return (int)e; }
        s->nNorm = h[0] + 1; s->nUV = h[1] + 1;
        long long texels = 0;
        for (const NovumMaterial& m : hm)
            if (m.hasTexture && m.width > 0 && m.height > 0 && m.startInd >= 0) texels = std::max(texels, (long long)m.startInd + (long long)m.width * m.height);
        s->nTexels = d_textures ? (int)std::min(texels, (long long)0x7fffffff) : 0;
    }
    *out = s;
    return 0;
}

}  // extern "C"

int nv::sceneReplicate(const NovumScene* src, NovumScene** out) {
    if (!src || !out) return NOVUM_E_ARG;
    int dev = 0;
    cudaGetDevice(&dev);
    NovumScene* s = new NovumScene();
    s->nPos = src->nPos; s->nNorm = src->nNorm; s->nUV = src->nUV; s->nTri = src->nTri; s->nLights = src->nLights; s->nMat = src->nMat;
    s->nTexels = src->nTexels;
    cudaError_t e = cudaSuccess;
    auto cp = [&](void** dst, const void* from, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cudaMalloc(dst, bytes ? bytes : 16);
        if (e == cudaSuccess && bytes) e = cudaMemcpyPeer(*dst, dev, from, src->device, bytes);
    };
    cp((void**)&s->d_positions, src->d_positions, sizeof(float4) * (size_t)s->nPos);
    cp((void**)&s->d_normals, src->d_normals, sizeof(float4) * (size_t)s->nNorm);
    cp((void**)&s->d_uvs, src->d_uvs, sizeof(float2) * (size_t)s->nUV);
    cp((void**)&s->d_tris, src->d_tris, sizeof(NovumTriangle) * (size_t)s->nTri);
    cp((void**)&s->d_lights, src->d_lights, sizeof(NovumTriangle) * (size_t)s->nLights);
    cp((void**)&s->d_matsRef, src->d_matsRef, sizeof(NovumMaterial) * (size_t)s->nMat);
    cp((void**)&s->d_textures, src->d_textures, sizeof(float4) * (size_t)s->nTexels);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { novum_scene_destroy(s); return (int)e; }
    int rc = sceneFinalize(s);
    if (rc) { novum_scene_destroy(s); return rc; }
    // the caller's root-box values, if the launch boundary put them there, travel with the scene
    for (int k = 0; k < 3; ++k) { s->sceneMin[k] = src->sceneMin[k]; s->sceneMax[k] = src->sceneMax[k]; s->sceneCenter[k] = src->sceneCenter[k]; }
    s->sceneRadius = src->sceneRadius;
    *out = s;
    return 0;
}

extern "C" {

void novum_scene_destroy(NovumScene* s) {
    if (!s) return;
    if (s->multiGpu && s->multiGpuFree) s->multiGpuFree(s->multiGpu);
    int prevDev = 0;
    cudaGetDevice(&prevDev);
    if (s->stream) cudaSetDevice(s->device);
    if (s->wfEngine && s->wfEngineFree) s->wfEngineFree(s->wfEngine);
    if (s->vcmEngine && s->vcmEngineFree) s->vcmEngineFree(s->vcmEngine);
    if (!s->borrowed) {
        cudaFree(s->d_positions); cudaFree(s->d_normals); cudaFree(s->d_uvs); cudaFree(s->d_tris);
        cudaFree(s->d_lights); cudaFree(s->d_matsRef); cudaFree(s->d_textures);
    }
    cudaFree(s->d_frame); cudaFree(s->d_shadeTris);
    cudaFree(s->d_mats); cudaFree(s->d_counters); cudaFree(s->d_queueCounter); cudaFree(s->d_stage);
    freeBvh8(s->bvh);
    if (s->ev0) { cudaEventDestroy(s->ev0); cudaEventDestroy(s->ev1); }
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaSetDevice(prevDev);
    delete s;
}

int novum_scene_info(const NovumScene* s, int* nTri, int* nLights, int* nBvh8Nodes, float sceneMin[3], float sceneMax[3],
                     float* sceneRadius) {
    if (!s) return NOVUM_E_NOSCENE;
    if (nTri) *nTri = s->nTri;
    if (nLights) *nLights = s->nLights;
    if (nBvh8Nodes) *nBvh8Nodes = s->bvh.numNodes;
    for (int k = 0; k < 3; ++k) { if (sceneMin) sceneMin[k] = s->sceneMin[k]; if (sceneMax) sceneMax[k] = s->sceneMax[k]; }
    if (sceneRadius) *sceneRadius = s->sceneRadius;
    return 0;
}

int novum_scene_timings(const NovumScene* s, float* parseMs, float* bvhBuildMs, int* bvhLaunches) {
    if (!s) return NOVUM_E_NOSCENE;
    if (parseMs) *parseMs = s->parseMs;
    if (bvhBuildMs) *bvhBuildMs = s->bvh.buildMs;
    if (bvhLaunches) *bvhLaunches = s->bvh.kernelLaunches;
    return 0;
}

int novum_scene_set_features(NovumScene* s, int features) {
    if (!s) return NOVUM_E_NOSCENE;
    if (features & ~(NOVUM_FEATURE_NESTED_BIDIR | NOVUM_FEATURE_ENVIRONMENT)) return NOVUM_E_ARG;
    s->features = features;
    return 0;
}

int novum_scene_set_environment(NovumScene* s, const float rgb[3]) {
    if (!s) return NOVUM_E_NOSCENE;
    if (!rgb || !(rgb[0] >= 0.0f && rgb[1] >= 0.0f && rgb[2] >= 0.0f) || !std::isfinite(rgb[0] + rgb[1] + rgb[2])) return NOVUM_E_ARG;
    for (int k = 0; k < 3; ++k) s->envRadiance[k] = rgb[k];
    return 0;
}

int novum_scene_bvh_depth(const NovumScene* s) { return s ? s->bvh.levels : NOVUM_E_NOSCENE; }

int novum_scene_counts(const NovumScene* s, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights) {
    if (!s) return NOVUM_E_NOSCENE;
    if (nPos) *nPos = s->nPos; if (nNorm) *nNorm = s->nNorm; if (nUV) *nUV = s->nUV;
    if (nTri) *nTri = s->nTri; if (nLights) *nLights = s->nLights;
    return 0;
}

int novum_scene_download(const NovumScene* s, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs,
                         NovumTriangle* tris, NovumTriangle* lights) {
    if (!s) return NOVUM_E_NOSCENE;
    cudaError_t e = cudaSuccess;
    auto dl = [&](void* dst, const void* src, size_t bytes) {
        if (dst && bytes && e == cudaSuccess) e = cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost);
    };
    dl(positions, s->d_positions, sizeof(float4) * s->nPos);
    dl(normals, s->d_normals, sizeof(float4) * s->nNorm);
    dl(uvs, s->d_uvs, sizeof(float2) * s->nUV);
    dl(tris, s->d_tris, sizeof(NovumTriangle) * s->nTri);
    dl(lights, s->d_lights, sizeof(NovumTriangle) * s->nLights);
    return (int)e;
}

static int traceGrid() { return smCount() * 8; }

int novum_trace_closest_device(NovumScene* s, const NovumRay* d_rays, int n, NovumHit* d_hits, int sync) {
    if (!s) return NOVUM_E_NOSCENE;
    if (n < 0 || (n > 0 && (!d_rays || !d_hits))) return NOVUM_E_ARG;
    if (n == 0) return 0;
    cudaMemsetAsync(s->d_queueCounter, 0, sizeof(int), s->stream);
    TravCounters cnt{s->d_counters, s->d_counters + 1};
    cudaEventRecord(s->ev0, s->stream);
    int grid = min(traceGrid(), (n + kTraceThreads - 1) / kTraceThreads);
    if (s->countersOn) closestBatchKernel<true><<<grid, kTraceThreads, 0, s->stream>>>(s->view(), d_rays, n, d_hits, s->d_queueCounter, cnt);
    else closestBatchKernel<false><<<grid, kTraceThreads, 0, s->stream>>>(s->view(), d_rays, n, d_hits, s->d_queueCounter, cnt);
    cudaEventRecord(s->ev1, s->stream);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    if (sync) { e = cudaStreamSynchronize(s->stream); if (e == cudaSuccess) cudaEventElapsedTime(&s->lastTraceMs, s->ev0, s->ev1); }
    return (int)e;
}

int novum_trace_shadow_device(NovumScene* s, const NovumRay* d_rays, const float* d_maxT, const int32_t* d_skipTri, int n,
                              NovumFloat4* d_throughput, int sync) {
    if (!s) return NOVUM_E_NOSCENE;
    if (n < 0 || (n > 0 && (!d_rays || !d_maxT || !d_throughput))) return NOVUM_E_ARG;
    if (n == 0) return 0;
    cudaMemsetAsync(s->d_queueCounter, 0, sizeof(int), s->stream);
    TravCounters cnt{s->d_counters, s->d_counters + 1};
    cudaEventRecord(s->ev0, s->stream);
    int grid = min(traceGrid(), (n + kTraceThreads - 1) / kTraceThreads);
    if (s->countersOn) shadowBatchKernel<true><<<grid, kTraceThreads, 0, s->stream>>>(s->view(), d_rays, d_maxT, d_skipTri, n, (float4*)d_throughput, s->d_queueCounter, cnt);
    else shadowBatchKernel<false><<<grid, kTraceThreads, 0, s->stream>>>(s->view(), d_rays, d_maxT, d_skipTri, n, (float4*)d_throughput, s->d_queueCounter, cnt);
    cudaEventRecord(s->ev1, s->stream);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    if (sync) { e = cudaStreamSynchronize(s->stream); if (e == cudaSuccess) cudaEventElapsedTime(&s->lastTraceMs, s->ev0, s->ev1); }
    return (int)e;
}

int novum_trace_closest(NovumScene* s, const NovumRay* rays, int n, NovumHit* hits) {
    if (!s) return NOVUM_E_NOSCENE;
    if (n < 0 || (n > 0 && (!rays || !hits))) return NOVUM_E_ARG;
    if (n == 0) return 0;
    size_t rb = sizeof(NovumRay) * (size_t)n, hb = sizeof(NovumHit) * (size_t)n;
    int rc = ensureStage(s, rb + hb);
    if (rc) return rc;
    NovumRay* dr = (NovumRay*)s->d_stage; NovumHit* dh = (NovumHit*)((char*)s->d_stage + rb);
    cudaMemcpyAsync(dr, rays, rb, cudaMemcpyHostToDevice, s->stream);
    rc = novum_trace_closest_device(s, dr, n, dh, 0);
    if (rc) return rc;
    cudaMemcpyAsync(hits, dh, hb, cudaMemcpyDeviceToHost, s->stream);
    cudaError_t e = cudaStreamSynchronize(s->stream);
    if (e == cudaSuccess) cudaEventElapsedTime(&s->lastTraceMs, s->ev0, s->ev1);
    return (int)e;
}

int novum_trace_shadow(NovumScene* s, const NovumRay* rays, const float* maxT, const int32_t* skipTri, int n,
                       NovumFloat4* throughput) {
    if (!s) return NOVUM_E_NOSCENE;
    if (n < 0 || (n > 0 && (!rays || !maxT || !throughput))) return NOVUM_E_ARG;
    if (n == 0) return 0;
    size_t rb = sizeof(NovumRay) * (size_t)n, tb = sizeof(float) * (size_t)n, ob = sizeof(float4) * (size_t)n;
    int rc = ensureStage(s, rb + 2 * tb + ob);
    if (rc) return rc;
    char* base = (char*)s->d_stage;
    NovumRay* dr = (NovumRay*)base; float4* dout = (float4*)(base + rb); float* dt = (float*)(base + rb + ob);
    int* dk = (int*)(base + rb + ob + tb);
    cudaMemcpyAsync(dr, rays, rb, cudaMemcpyHostToDevice, s->stream);
    cudaMemcpyAsync(dt, maxT, tb, cudaMemcpyHostToDevice, s->stream);
    if (skipTri) cudaMemcpyAsync(dk, skipTri, tb, cudaMemcpyHostToDevice, s->stream);
    rc = novum_trace_shadow_device(s, dr, dt, skipTri ? dk : nullptr, n, (NovumFloat4*)dout, 0);
    if (rc) return rc;
    cudaMemcpyAsync(throughput, dout, ob, cudaMemcpyDeviceToHost, s->stream);
    cudaError_t e = cudaStreamSynchronize(s->stream);
    if (e == cudaSuccess) cudaEventElapsedTime(&s->lastTraceMs, s->ev0, s->ev1);
    return (int)e;
}

int novum_set_profile(NovumScene* s, int flags) {
    if (!s) return NOVUM_E_NOSCENE;
    s->profileFlags = flags;
    return 0;
}

int novum_trace_last_ms(NovumScene* s, float* ms) {
    if (!s || !ms) return NOVUM_E_ARG;
    *ms = s->lastTraceMs;
    return 0;
}

int novum_trace_counters(NovumScene* s, int enable, uint64_t* nodeVisits, uint64_t* triTests) {
    if (!s) return NOVUM_E_NOSCENE;
    unsigned long long h[2] = {0, 0};
    cudaStreamSynchronize(s->stream);
    cudaError_t e = cudaMemcpy(h, s->d_counters, sizeof(h), cudaMemcpyDeviceToHost);
    if (nodeVisits) *nodeVisits = h[0];
    if (triTests) *triTests = h[1];
    cudaMemset(s->d_counters, 0, sizeof(h));
    s->countersOn = enable != 0;
    return (int)e;
}

}  // extern "C"
