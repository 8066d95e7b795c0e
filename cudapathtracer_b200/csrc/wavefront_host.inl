// wavefront_host.inl — pool allocation and the host loop of the unidirectional wavefront engine.

namespace nv {

struct WfEngine {
    WfState st{};
    int* pinnedCounters = nullptr;      // host-pinned mirrors (2 slots) for the periodic "anything left?" read
    cudaEvent_t evRead[2] = {nullptr, nullptr};
    std::vector<cudaEvent_t> evPool;    // stage-boundary events (profile bit 0), grown on demand and reused
    size_t bytes = 0;
};

static int wfAlloc(WfEngine& e, int capacity) {
    WfState& s = e.st;
    s.capacity = capacity;
    cudaError_t err = cudaSuccess;
    auto A = [&](void** p, size_t b) { if (err == cudaSuccess) { err = cudaMalloc(p, b); e.bytes += b; } };
    size_t n = (size_t)capacity;
    A((void**)&s.slots, n * kSlotWords * sizeof(float4));
    A((void**)&s.extendQ[0], n * sizeof(int)); A((void**)&s.extendQ[1], n * sizeof(int));
    A((void**)&s.classQ, n * NUM_CLS * sizeof(int)); A((void**)&s.shadowQ, n * sizeof(ShadowRec));
    if (kRayBins) {
        A((void**)&s.binTmp, n * sizeof(int2)); A((void**)&s.binPos, n * sizeof(int));
        A((void**)&s.binCount, kNumRayBins * sizeof(int)); A((void**)&s.binStart, kNumRayBins * sizeof(int));
        if (err == cudaSuccess) err = cudaMemset(s.binCount, 0, kNumRayBins * sizeof(int));
    }
    A((void**)&s.counters, C_TOTAL * sizeof(int)); A((void**)&s.rayCounters, 4 * sizeof(unsigned long long));
    if (err != cudaSuccess) return (int)err;
    if ((err = cudaHostAlloc((void**)&e.pinnedCounters, 2 * C_TOTAL * sizeof(int), cudaHostAllocDefault)) != cudaSuccess) return (int)err;
    cudaEventCreateWithFlags(&e.evRead[0], cudaEventDisableTiming);
    cudaEventCreateWithFlags(&e.evRead[1], cudaEventDisableTiming);
    return 0;
}

static void wfFree(WfEngine& e) {
    WfState& s = e.st;
    cudaFree(s.slots); cudaFree(s.extendQ[0]); cudaFree(s.extendQ[1]); cudaFree(s.classQ); cudaFree(s.shadowQ);
    cudaFree(s.counters); cudaFree(s.rayCounters);
    cudaFree(s.binTmp); cudaFree(s.binPos); cudaFree(s.binCount); cudaFree(s.binStart);
    if (e.pinnedCounters) cudaFreeHost(e.pinnedCounters);
    for (int i = 0; i < 2; ++i) if (e.evRead[i]) cudaEventDestroy(e.evRead[i]);
    for (cudaEvent_t ev : e.evPool) cudaEventDestroy(ev);
    e = WfEngine();
}

// Renders `numSample` sample passes of this shard into d_colors (sum accumulator). Blocking.
static int wfRender(NovumScene* scene, WfEngine& eng, const NovumCamera& cam, int w, int h, int maxDepth, int useMIS, int naive,
                    int firstSample, int numSample, NovumShard shard, unsigned long long seed, float4* d_colors,
                    NovumRenderStats* stats) {
    if (shard.world < 1) shard.world = 1;
    int local = 0;
    for (int s = shard.rank; s < numSample; s += shard.world) ++local;
    WfParams P;
    P.cam = cam; P.w = w; P.h = h; P.maxDepth = maxDepth; P.useMIS = useMIS; P.naive = naive; P.seed = seed;
    P.firstSample = firstSample; P.localSamples = local; P.rank = shard.rank; P.world = shard.world;
    P.tilesX = (w + 7) / 8; P.tilesY = (h + 3) / 4;
    P.totalIds = (unsigned long long)local * P.tilesX * P.tilesY * 32ull;
    P.colors = d_colors;
    cudaStream_t stream = scene->stream;
    SceneView sc = scene->view();
    WfState& st = eng.st;
    unsigned long long poolWanted = P.totalIds < (unsigned long long)st.capacity ? P.totalIds : (unsigned long long)st.capacity;
    int pool = (int)poolWanted;
    unsigned long long launches = 0, stageLaunches[4] = {0, 0, 0, 0};
    const bool profTime = (scene->profileFlags & 1) != 0, profCount = (scene->profileFlags & 2) != 0;
    TravCounters cnt{scene->d_counters, scene->d_counters + 1};
    if (profCount) cudaMemsetAsync(scene->d_counters, 0, 2 * sizeof(unsigned long long), stream);
    size_t evUsed = 0;
    auto mark = [&]() {             // stage boundary event (profile bit 0)
        if (!profTime) return;
        if (evUsed == eng.evPool.size()) { cudaEvent_t e; cudaEventCreate(&e); eng.evPool.push_back(e); }
        cudaEventRecord(eng.evPool[evUsed++], stream);
    };
    cudaEventRecord(scene->ev0, stream);
    if (pool > 0) {
        cudaMemsetAsync(st.rayCounters, 0, 4 * sizeof(unsigned long long), stream);
        if (kRayBins) cudaMemsetAsync(st.binCount, 0, kNumRayBins * sizeof(int), stream);
        wfInit<<<(pool + 255) / 256, 256, 0, stream>>>(st, pool);
        ++launches;
        const int grid = smCount() * 8;
        const int tgrid = smCount() * NOVUM_TRAV_MINB;      // traversal kernels: one wave of resident CTAs
        int cur = 0;
        // Batches of kBatch iterations are enqueued back to back; the "anything left?" counter of batch b is read while
        // batch b+1 is already queued, so the GPU never waits for the host (one surplus batch of no-op launches at the end).
        const int kBatch = 4;
        int pending = -1;           // pinned slot whose copy is in flight
        for (long iter = 0;; ++iter) {
            mark();
            if (iter > 0) {
                wfAdvance<<<1, kAdvanceThreads, 0, stream>>>(st);
                if (kRayBins) { wfBinScatter<<<grid, 256, 0, stream>>>(st, st.extendQ[cur]); ++launches; }
                if (profCount) wfExtend<true><<<tgrid, kWfThreads, 0, stream>>>(sc, st, st.extendQ[cur], cnt);
                else wfExtend<false><<<tgrid, kWfThreads, 0, stream>>>(sc, st, st.extendQ[cur], cnt);
                launches += 2; stageLaunches[0] += 1; stageLaunches[3] += 1;
            }
            mark();
            wfShade<<<grid, kWfThreads, 0, stream>>>(sc, st, P, st.extendQ[cur ^ 1]);
            mark();
            wfShadow<<<tgrid, kWfThreads, 0, stream>>>(sc, st, d_colors);
            mark();
            launches += 2; stageLaunches[1] += 1; stageLaunches[2] += 1;
            cur ^= 1;
            if ((iter % kBatch) == kBatch - 1) {
                int slot = (int)((iter / kBatch) & 1);
                if (pending >= 0) {
                    cudaError_t e = cudaEventSynchronize(eng.evRead[pending]);
                    if (e != cudaSuccess) return (int)e;
                    if (eng.pinnedCounters[pending * C_TOTAL + C_NEXT_COUNT] == 0) break;
                }
                cudaMemcpyAsync(eng.pinnedCounters + slot * C_TOTAL, st.counters, C_TOTAL * sizeof(int), cudaMemcpyDeviceToHost, stream);
                cudaEventRecord(eng.evRead[slot], stream);
                pending = slot;
            }
            if (iter > 4000000) return (int)cudaErrorLaunchTimeout;
        }
    }
    cudaEventRecord(scene->ev1, stream);
    cudaError_t e = cudaStreamSynchronize(stream);
    if (e != cudaSuccess) return (int)e;
    if ((e = cudaGetLastError()) != cudaSuccess) return (int)e;
    if (stats) {
        unsigned long long rc[4] = {0, 0, 0, 0};
        if (pool > 0) cudaMemcpy(rc, st.rayCounters, sizeof(rc), cudaMemcpyDeviceToHost);
        memset(stats, 0, sizeof(*stats));
        stats->closestRays = rc[0]; stats->shadowRays = rc[1]; stats->paths = rc[3];
        stats->kernelLaunches = launches;
        cudaEventElapsedTime(&stats->deviceMs, scene->ev0, scene->ev1);
        for (int k = 0; k < 4; ++k) stats->stageLaunches[k] = stageLaunches[k];
        if (profTime) {
            for (size_t i = 0; i + 3 < evUsed; i += 4)
                for (int k = 0; k < 3; ++k) { float ms = 0; cudaEventElapsedTime(&ms, eng.evPool[i + k], eng.evPool[i + k + 1]); stats->stageMs[k] += ms; }
            stats->stageMs[3] = stats->deviceMs - stats->stageMs[0] - stats->stageMs[1] - stats->stageMs[2];
        }
        if (profCount) {
            unsigned long long c2[2] = {0, 0};
            cudaMemcpy(c2, scene->d_counters, sizeof(c2), cudaMemcpyDeviceToHost);
            stats->bvhNodeVisits = c2[0]; stats->triTests = c2[1];
        }
    }
    return 0;
}

}  // namespace nv

// ---- C ABI: render entry points ------------------------------------------------------------------------------
namespace nv {
int vcmRenderDevice(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                    NovumShard shard, unsigned long long seed, float4* d_colors, NovumRenderStats* stats);   // vcm.cu
static void wfEngineFree(void* p) { WfEngine* e = (WfEngine*)p; wfFree(*e); delete e; }
static NovumConstants g_constants = {0, 1, 1, 1, 1, 1, 0, 0, 1, 0};
const NovumConstants& currentConstants() { return g_constants; }
}

extern "C" {

int novum_update_constants(const NovumConstants* c) {
    if (!c) return NOVUM_E_ARG;
    nv::g_constants = *c;
    return 0;
}

int novum_render_device(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                        NovumShard shard, uint64_t seed, NovumFloat4* d_colors, NovumRenderStats* stats) {
    return nv::renderDeviceMis(s, cfg, cam, firstSample, numSample, shard, seed, d_colors, stats, 1);
}

}  // extern "C"

// novum_render_device with the unidirectional launcher's `useMIS` argument (deviceCode.cu:285-286, :442, :456) exposed: main.cu:565
// always passes true, the launch shims forward what they are given
int nv::renderDeviceMis(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                        NovumShard shard, uint64_t seed, NovumFloat4* d_colors, NovumRenderStats* stats, int useMIS) {
    if (!s) return NOVUM_E_NOSCENE;
    if (!cfg || !cam || !d_colors || numSample < 0 || cfg->width <= 0 || cfg->height <= 0) return NOVUM_E_ARG;
    if (shard.world < 1 || shard.rank < 0 || shard.rank >= shard.world) return NOVUM_E_ARG;
    int integ = cfg->integrator;
    if (integ == NOVUM_UNIDIRECTIONAL || integ == NOVUM_NAIVE_UNIDIRECTIONAL) {
        if (!s->wfEngine) {
            nv::WfEngine* e = new nv::WfEngine();
            long long px = (long long)cfg->width * cfg->height;
            long long cap = px < (1 << 20) ? (1 << 20) : (px > (1 << 22) ? (1 << 22) : px);   // 1M .. 4M path slots
            cap = (cap + 31) & ~31LL;
            int rc = nv::wfAlloc(*e, (int)cap);
            if (rc) { nv::wfFree(*e); delete e; return rc; }
            s->wfEngine = e; s->wfEngineFree = nv::wfEngineFree;
        }
        return nv::wfRender(s, *(nv::WfEngine*)s->wfEngine, *cam, cfg->width, cfg->height, cfg->maxDepth, useMIS != 0,
                            integ == NOVUM_NAIVE_UNIDIRECTIONAL, firstSample, numSample, shard, seed, (float4*)d_colors, stats);
    }
    if (integ == NOVUM_VCM || integ == NOVUM_SPPM || integ == NOVUM_BIDIRECTIONAL)
        return nv::vcmRenderDevice(s, cfg, cam, firstSample, numSample, shard, seed, (float4*)d_colors, stats);
    return NOVUM_E_UNSUPPORTED;
}

extern "C" {

int novum_render(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                 NovumShard shard, uint64_t seed, NovumFloat4* colors, NovumRenderStats* stats) {
    if (!s) return NOVUM_E_NOSCENE;
    if (!cfg || !colors) return NOVUM_E_ARG;
    size_t bytes = sizeof(float4) * (size_t)cfg->width * cfg->height;
    if (s->frameBytes < bytes) {          // the device framebuffer lives with the scene: no allocation per call
        cudaFree(s->d_frame); s->d_frame = nullptr; s->frameBytes = 0;
        cudaError_t e = cudaMalloc(&s->d_frame, bytes);
        if (e != cudaSuccess) return (int)e;
        s->frameBytes = bytes;
    }
    float4* d = (float4*)s->d_frame;
    cudaMemsetAsync(d, 0, bytes, s->stream);
    int rc = novum_render_device(s, cfg, cam, firstSample, numSample, shard, seed, (NovumFloat4*)d, stats);
    // `colors` may be pageable or page-locked host memory; page-locked (cudaHostAlloc / cudaHostRegister) copies at full PCIe rate
    if (rc == 0) rc = (int)cudaMemcpyAsync(colors, d, bytes, cudaMemcpyDeviceToHost, s->stream);
    if (rc == 0) rc = (int)cudaStreamSynchronize(s->stream);
    return rc;
}

}  // extern "C"
