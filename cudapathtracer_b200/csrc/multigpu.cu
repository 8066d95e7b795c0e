// multigpu.cu — in-process multi-GPU render behind the launch boundary (`numGpus` of novum_launch_*): the reference's
// single-threaded initRender (main.cu:235-923) can use N GPUs of one box without changing a line - one host thread per GPU here.
//   * scene: the caller's buffers live on the current device; every other GPU gets an owned replica (peer-to-peer copies of the
//     arrays, its own BVH8 build), cached with the primary scene;
//   * unidirectional / naive: sample passes striped over the GPUs (pass i on GPU i % N), no exchange;
//   * BDPT / VCM / SPPM: the tile renderer (vcm_host.inl): GPU g owns pixel rows y % N == g, the light pass stores its photons into every
//     peer's slab over NVLink (peer access enabled between all pairs), device-side rendezvous, several iterations in flight per GPU;
//   * frames: every GPU accumulates into its own full frame; GPU 0 adds the peers' frames into `colors` with one kernel that reads
//     them over NVLink (no NCCL: nothing here needs a collective library).
// The reference has no multi-GPU path at all.
#include "scene.h"
#include <thread>
#include <vector>

struct NovumTiles;
extern "C" {
int  novum_tiles_create(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, NovumShard shard, uint64_t seed, int lanes, NovumTiles** out);
void novum_tiles_destroy(NovumTiles* t);
int  novum_tiles_bases(NovumTiles* t, void** bases);
int  novum_tiles_set_view(NovumTiles* t, const NovumRenderConfig* cfg, const NovumCamera* cam, uint64_t seed);
int  novum_tiles_attach_pointers(NovumTiles* t, int peerRank, void* const* bases);
int  novum_tiles_render(NovumTiles* t, int firstIter, int numIter, NovumFloat4* d_colors, NovumRenderStats* stats);
}

namespace nv {

namespace {

constexpr int kMultiLanes = 4;

struct MultiGpu {
    int n = 0;
    std::vector<int> dev;                    // dev[0] = the primary scene's device
    std::vector<NovumScene*> scene;          // scene[0] = the primary (not owned)
    std::vector<float4*> frame;              // frame[g], g > 0: full-frame accumulator on GPU g
    size_t frameBytes = 0;
    std::vector<NovumTiles*> tiles;          // bidirectional engines, one per GPU
    NovumRenderConfig tilesCfg{};
    bool haveTiles = false;
};

__global__ void addFramesKernel(float4* __restrict__ dst, const float4* __restrict__ a, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        float4 d = dst[i]; const float4 s = a[i];
        d.x += s.x; d.y += s.y; d.z += s.z; d.w += s.w;
        dst[i] = d;
    }
}

void freeTiles(MultiGpu& m) {
    for (size_t g = 0; g < m.tiles.size(); ++g)
        if (m.tiles[g]) { cudaSetDevice(m.dev[g]); novum_tiles_destroy(m.tiles[g]); }
    m.tiles.clear(); m.haveTiles = false;
}

void freeMulti(void* p) {
    MultiGpu* m = (MultiGpu*)p;
    if (!m) return;
    int prev = 0;
    cudaGetDevice(&prev);
    freeTiles(*m);
    for (int g = 1; g < (int)m->scene.size(); ++g) {
        cudaSetDevice(m->dev[g]);
        if (g < (int)m->frame.size()) cudaFree(m->frame[g]);
        if (m->scene[g]) novum_scene_destroy(m->scene[g]);
    }
    cudaSetDevice(prev);
    delete m;
}

bool sameTilesCfg(const NovumRenderConfig& a, const NovumRenderConfig& b) {
    return a.width == b.width && a.height == b.height && a.bdptEyeDepth == b.bdptEyeDepth && a.bdptLightDepth == b.bdptLightDepth &&
           (a.integrator == NOVUM_BIDIRECTIONAL) == (b.integrator == NOVUM_BIDIRECTIONAL);
}

int ensureMulti(NovumScene* primary, int numGpus, MultiGpu** out) {
    MultiGpu* m = (MultiGpu*)primary->multiGpu;
    if (m && m->n == numGpus) { *out = m; return 0; }
    if (m) { freeMulti(m); primary->multiGpu = nullptr; }
    int ndev = 0;
    cudaGetDeviceCount(&ndev);
    if (numGpus > ndev || numGpus > 16) return NOVUM_E_ARG;
    m = new MultiGpu();
    m->n = numGpus;
    m->dev.push_back(primary->device);
    for (int d = 0; d < ndev && (int)m->dev.size() < numGpus; ++d) if (d != primary->device) m->dev.push_back(d);
    m->scene.assign(numGpus, nullptr); m->frame.assign(numGpus, nullptr);
    m->scene[0] = primary;
    primary->multiGpu = m; primary->multiGpuFree = freeMulti;
    for (int a = 0; a < numGpus; ++a) {                       // every GPU stores into every other GPU's slabs; GPU 0 reads the peers' frames
        cudaSetDevice(m->dev[a]);
        for (int b = 0; b < numGpus; ++b) {
            if (a == b) continue;
            int can = 0;
            cudaDeviceCanAccessPeer(&can, m->dev[a], m->dev[b]);
            if (!can) { cudaSetDevice(primary->device); return NOVUM_E_UNSUPPORTED; }
            cudaError_t e = cudaDeviceEnablePeerAccess(m->dev[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaSetDevice(primary->device); return (int)e; }
            cudaGetLastError();
        }
    }
    cudaSetDevice(primary->device);
    cudaStreamSynchronize(primary->stream);
    std::vector<int> rcs(numGpus, 0);
    std::vector<std::thread> th;
    for (int g = 1; g < numGpus; ++g)
        th.emplace_back([&, g]() { cudaSetDevice(m->dev[g]); rcs[g] = sceneReplicate(primary, &m->scene[g]); });
    for (std::thread& t : th) t.join();
    cudaSetDevice(primary->device);
    for (int g = 1; g < numGpus; ++g) if (rcs[g]) return rcs[g];
    *out = m;
    return 0;
}

}  // namespace

int launchMultiGpu(int numGpus, int integrator, const NovumRenderConfig& cfg, const NovumCamera& cam, NovumScene* primary, int numSample,
                   unsigned long long seed, float4* colors, int useMIS) {
    MultiGpu* m = nullptr;
    int rc = ensureMulti(primary, numGpus, &m);
    if (rc) return rc;
    const size_t px = (size_t)cfg.width * cfg.height, bytes = px * sizeof(float4);
    if (m->frameBytes < bytes) {
        for (int g = 1; g < numGpus; ++g) {
            cudaSetDevice(m->dev[g]);
            cudaFree(m->frame[g]); m->frame[g] = nullptr;
            cudaError_t e = cudaMalloc(&m->frame[g], bytes);
            if (e != cudaSuccess) { cudaSetDevice(primary->device); return (int)e; }
        }
        m->frameBytes = bytes;
    }
    for (int g = 1; g < numGpus; ++g) {
        cudaSetDevice(m->dev[g]);
        cudaMemsetAsync(m->frame[g], 0, bytes, m->scene[g]->stream);
        cudaStreamSynchronize(m->scene[g]->stream);
        // the caller's root box (merge radius schedule, grid origin) on every replica
        for (int k = 0; k < 3; ++k) { m->scene[g]->sceneMin[k] = primary->sceneMin[k]; m->scene[g]->sceneCenter[k] = primary->sceneCenter[k]; }
        m->scene[g]->sceneRadius = primary->sceneRadius;
    }
    m->frame[0] = colors;
    const bool bidir = integrator == NOVUM_BIDIRECTIONAL || integrator == NOVUM_VCM || integrator == NOVUM_SPPM;
    if (bidir) {
        if (m->haveTiles && !sameTilesCfg(m->tilesCfg, cfg)) freeTiles(*m);
        if (!m->haveTiles) {
            m->tiles.assign(numGpus, nullptr);
            for (int g = 0; g < numGpus && !rc; ++g) {
                cudaSetDevice(m->dev[g]);
                rc = novum_tiles_create(m->scene[g], &cfg, &cam, NovumShard{g, numGpus}, seed, kMultiLanes, &m->tiles[g]);
            }
            std::vector<std::vector<void*>> bases(numGpus, std::vector<void*>(kMultiLanes, nullptr));
            for (int g = 0; g < numGpus && !rc; ++g) rc = novum_tiles_bases(m->tiles[g], bases[g].data());
            for (int g = 0; g < numGpus && !rc; ++g)
                for (int q = 0; q < numGpus && !rc; ++q)
                    if (q != g) rc = novum_tiles_attach_pointers(m->tiles[g], q, bases[q].data());
            if (rc) { freeTiles(*m); cudaSetDevice(primary->device); return rc; }
            m->tilesCfg = cfg; m->haveTiles = true;
        }
        for (int g = 0; g < numGpus && !rc; ++g) rc = novum_tiles_set_view(m->tiles[g], &cfg, &cam, seed);      // this call's toggles (updateConstants), camera
        if (rc) { cudaSetDevice(primary->device); return rc; }
    }
    std::vector<int> rcs(numGpus, 0);
    auto work = [&](int g) {
        cudaSetDevice(m->dev[g]);
        if (bidir) {
            rcs[g] = novum_tiles_render(m->tiles[g], 0, numSample, (NovumFloat4*)m->frame[g], nullptr);
        } else {
            rcs[g] = renderDeviceMis(m->scene[g], &cfg, &cam, 0, numSample, NovumShard{g, numGpus}, seed, (NovumFloat4*)m->frame[g], nullptr, useMIS);
        }
    };
    std::vector<std::thread> th;
    for (int g = 1; g < numGpus; ++g) th.emplace_back(work, g);
    work(0);
    for (std::thread& t : th) t.join();
    cudaSetDevice(primary->device);
    for (int g = 0; g < numGpus; ++g) if (rcs[g]) return rcs[g];
    for (int g = 1; g < numGpus; ++g) addFramesKernel<<<smCount() * 4, 256, 0, primary->stream>>>(colors, m->frame[g], px);
    cudaError_t e = cudaStreamSynchronize(primary->stream);
    return e != cudaSuccess ? (int)e : (int)cudaGetLastError();
}

}  // namespace nv
