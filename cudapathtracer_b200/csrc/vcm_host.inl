// vcm_host.inl — host side of the bidirectional engine: buffers, the per-iteration pass sequence, the multi-GPU phase API
// (light pass -> caller all-gathers the photon slabs -> grid build -> eye pass) and the single-call render entry.
// launch_VCM's host loop (deviceCode.cu:3256-3497) is the reference for what one iteration does: radius schedule,
// photon counter reset, light pass, hash grid, eye pass.

namespace nv {

struct BdEngine {
    BdState st{};
    NovumScene* scene = nullptr;
    NovumCamera cam{};
    int w = 0, h = 0, eyeDepth = 0, lightDepth = 0, mode = BD_MODE_VCM;
    int rank = 0, world = 1, nLocal = 0;
    unsigned long long seed = 0;
    NovumRenderConfig cfg{};
    // photons: `world` slabs (this rank appends into slab `rank`), compacted copy, cell-ordered copy
    uint4* slabs = nullptr; size_t slabWords = 0; int slabCap = 0;
    int slabBuffers = 1, slabBuf = 0;     // lane engines double-buffer their slabs: buffer (seq & 1) holds iteration seq's photons
    unsigned seq = 0;                     // light passes run by a lane engine (the same number on every rank): stamp = seq + 1
    cudaStream_t stream = nullptr;        // the scene's stream, or the lane's own
    bool ownStream = false;
    int device = 0;
    void* peerBase[16] = {nullptr};       // other ranks' slab buffers (index = rank), null = not attached
    bool peerIpc[16] = {false};           // opened through CUDA IPC here (to be closed), else a pointer the caller owns
    bool peerWrite = false;               // light pass stores photons into the peers as it appends them
    uint4* photonsAll = nullptr; uint4* photonsSorted = nullptr;
    int maxPhotons = 0, photonCount = 0;
    HashGridDevice grid;
    bool gridAllocated = false;
    RadixSortPlan mergePlan; bool mergePlanMade = false;      // orders a queue's merge records by photon-grid cell
    int* d_counts = nullptr; int* h_counts = nullptr; int* h_countsDev = nullptr;      // h_counts: mapped page-locked, h_countsDev its device address
    float mergeRadius = 0.f;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    unsigned long long launches = 0, iterations = 0;
    float deviceMs = 0.f;
    float stageMs[4] = {0, 0, 0, 0};      // profile bit 0: [0] closest-hit extends [1] light/eye step kernels + photon grid build + memsets [2] shadow resolves [3] merge kernels
    unsigned long long stageLaunches[4] = {0, 0, 0, 0};
    bool prof = false, profCount = false;
    std::vector<cudaEvent_t> evs; std::vector<int> evCls; size_t evUsed = 0;
    size_t bytes = 0;
};

// profile bit 0: an event after every launch, tagged with the launch's class; the gap to the previous event is charged to it
static void bdTick(BdEngine& e, int cls) {
    e.stageLaunches[cls]++;
    if (!e.prof) return;
    if (e.evUsed == e.evs.size()) { cudaEvent_t ev; cudaEventCreate(&ev); e.evs.push_back(ev); e.evCls.push_back(0); }
    e.evCls[e.evUsed] = cls;
    cudaEventRecord(e.evs[e.evUsed++], e.stream);
}

static void bdFree(BdEngine& e) {
    BdState& s = e.st;
    cudaFree(s.rayO); cudaFree(s.rayD); cudaFree(s.beta); cudaFree(s.prevP); cudaFree(s.dv); cudaFree(s.rng); cudaFree(s.hit);
    cudaFree(s.queue[0]); cudaFree(s.queue[1]); cudaFree(s.pathLen); cudaFree(s.medium); cudaFree(s.lvA); cudaFree(s.lvB); cudaFree(s.lvC);
    cudaFree(s.shadowQ); cudaFree(s.mergeQ); cudaFree(s.mergeOrder); cudaFree(s.mergeKeys); cudaFree(s.counters); cudaFree(s.stats);
    for (int r = 0; r < 16; ++r) if (e.peerBase[r] && e.peerIpc[r]) cudaIpcCloseMemHandle(e.peerBase[r]);
    cudaFree(e.slabs); cudaFree(e.photonsAll); cudaFree(e.photonsSorted); cudaFree(e.d_counts);
    if (e.h_counts) cudaFreeHost(e.h_counts);
    if (e.gridAllocated) hashGridFree(e.grid);
    if (e.mergePlanMade) radixSortPlanDestroy(e.mergePlan);
    if (e.ev0) { cudaEventDestroy(e.ev0); cudaEventDestroy(e.ev1); }
    for (cudaEvent_t ev : e.evs) cudaEventDestroy(ev);
    if (e.ownStream && e.stream) cudaStreamDestroy(e.stream);
    e = BdEngine();
}

static int bdModeOf(int integrator) { return integrator == NOVUM_BIDIRECTIONAL ? BD_MODE_BDPT : BD_MODE_VCM; }

static int bdAlloc(BdEngine& e, NovumScene* scene, const NovumRenderConfig& cfg, const NovumCamera& cam, NovumShard shard,
                   unsigned long long seed, bool lane = false) {
    e.scene = scene; e.cam = cam; e.cfg = cfg; e.seed = seed;
    cudaGetDevice(&e.device);
    if (lane) {          // a lane of a tile renderer: own stream, slabs double-buffered
        cudaError_t se = cudaStreamCreateWithFlags(&e.stream, cudaStreamNonBlocking);
        if (se != cudaSuccess) return (int)se;
        e.ownStream = true; e.slabBuffers = 2;
    } else e.stream = scene->stream;
    e.w = cfg.width; e.h = cfg.height; e.eyeDepth = cfg.bdptEyeDepth; e.lightDepth = cfg.bdptLightDepth;
    e.mode = bdModeOf(cfg.integrator);
    e.rank = shard.rank; e.world = shard.world;
    if (e.eyeDepth < 1 || e.lightDepth < 1 || e.eyeDepth > kMaxBdDepth || e.lightDepth > kMaxBdDepth) return NOVUM_E_ARG;
    if ((long long)e.w * e.h >= (1ll << 30)) return NOVUM_E_ARG;
    int rows = 0;
    for (int y = e.rank; y < e.h; y += e.world) ++rows;
    e.nLocal = rows * e.w;
    BdState& s = e.st;
    s.nLocal = e.nLocal;
    size_t n = (size_t)(e.nLocal > 0 ? e.nLocal : 1);
    cudaError_t err = cudaSuccess;
    auto A = [&](void** p, size_t b) { if (err == cudaSuccess) { err = cudaMalloc(p, b ? b : 16); e.bytes += b; } };
    A((void**)&s.rayO, n * 16); A((void**)&s.rayD, n * 16); A((void**)&s.beta, n * 16); A((void**)&s.prevP, n * 16); A((void**)&s.dv, n * 16);
    A((void**)&s.rng, n * 8); A((void**)&s.hit, n * 16);
    A((void**)&s.queue[0], n * 4); A((void**)&s.queue[1], n * 4); A((void**)&s.pathLen, n * 4);
    size_t nv = n * (size_t)e.lightDepth;
    A((void**)&s.lvA, nv * 16); A((void**)&s.lvB, nv * 16); A((void**)&s.lvC, nv * 16);
    size_t cap = n * (size_t)(e.lightDepth + 1);
    if (cap < 2 * n) cap = 2 * n;
    if (cap > 0x7fffffffull) cap = 0x7fffffffull;
    s.shadowCap = (int)cap;
    A((void**)&s.shadowQ, cap * sizeof(BdShadowRec));
    if (e.mode == BD_MODE_VCM) {
        A((void**)&s.mergeQ, n * (size_t)(1 + kMergeGroup) * sizeof(BdMergeRec));
        A((void**)&s.mergeOrder, n * (size_t)kMergeGroup * sizeof(int));
        A((void**)&s.mergeKeys, n * (size_t)kMergeGroup * sizeof(uint32_t));
    }
    A((void**)&s.counters, (kMaxBdDepth + 1) * BC_STRIDE * sizeof(int));
    A((void**)&s.stats, 8 * sizeof(unsigned long long));
    // photon storage: one slab per rank, sized for that rank's worst case (every vertex of every light path)
    int rowsMax = (e.h + e.world - 1) / e.world;
    e.slabCap = rowsMax * e.w * e.lightDepth;
    e.slabWords = (size_t)kSlabHeaderBytes / 16 + (size_t)e.slabCap * 2;
    e.maxPhotons = e.w * e.h * e.lightDepth;                 // deviceCode.cu:3276
    A((void**)&e.slabs, e.slabWords * 16 * e.world * e.slabBuffers);
    if (e.mode == BD_MODE_VCM) {
        if (e.world > 1) A((void**)&e.photonsAll, (size_t)e.maxPhotons * 32);
        A((void**)&e.photonsSorted, (((size_t)e.maxPhotons + 63) & ~(size_t)63) * 32);
    }
    A((void**)&e.d_counts, 32 * sizeof(int));        // [0, 16): photons per slab, [16]: bdWaitSlabs gave up (1 + rank it waited for)
    if (err != cudaSuccess) return (int)err;
    if ((err = cudaHostAlloc((void**)&e.h_counts, 32 * sizeof(int), cudaHostAllocMapped)) != cudaSuccess) return (int)err;
    if ((err = cudaHostGetDevicePointer((void**)&e.h_countsDev, e.h_counts, 0)) != cudaSuccess) return (int)err;
    memset(e.h_counts, 0, 32 * sizeof(int));               // page-locked memory arrives uninitialised; [16] is read as the time-out flag
    cudaMemsetAsync(e.d_counts, 0, 32 * sizeof(int), e.stream);
    if (e.mode == BD_MODE_VCM) {
        int rc = hashGridAlloc(e.grid, e.maxPhotons);
        if (rc) return rc;
        e.gridAllocated = true;
        if ((err = radixSortPlanCreate(e.mergePlan, (int)std::min<size_t>(n * (size_t)kMergeGroup, 0x7fffffffull))) != cudaSuccess) return (int)err;
        e.mergePlanMade = true;
    }
    cudaMemsetAsync(s.stats, 0, 8 * sizeof(unsigned long long), e.stream);
    cudaMemsetAsync(e.slabs, 0, e.slabWords * 16 * e.world * e.slabBuffers, e.stream);
    cudaEventCreate(&e.ev0); cudaEventCreate(&e.ev1);
    return 0;
}

static BdParams bdParams(const BdEngine& e, int iter, float4* colors) {
    BdParams P;
    memset(&P, 0, sizeof(P));
    const NovumRenderConfig& c = e.cfg;
    P.cam = e.cam; P.w = e.w; P.h = e.h; P.eyeDepth = e.eyeDepth; P.lightDepth = e.lightDepth; P.mode = e.mode;
    P.lightTrace = c.bdptLightTrace; P.nee = c.bdptNee; P.naive = c.bdptNaive; P.connection = c.bdptConnection;
    P.doMis = c.bdptDoMis; P.paintWeight = c.bdptPaintWeight; P.doMerge = c.vcmDoMerge; P.doSPPM = c.doSPPM; P.sampleEnv = c.sampleEnvironment;
    if (c.integrator == NOVUM_SPPM) {           // main.cu:314-333
        P.connection = 0; P.naive = 0; P.nee = 0; P.lightTrace = 0; P.doMis = 0; P.doMerge = 1; P.doSPPM = 1;
    }
    if (e.mode == BD_MODE_BDPT) { P.doMerge = 0; P.doSPPM = 0; }
    float r = e.mode == BD_MODE_VCM ? mergeRadiusAt(e.scene->sceneRadius * c.vcmInitialMergeRadiusMultiplier, c.vcmMergeConst, iter) : 0.0f;
    P.mergeRadius = r;
    P.etaVcm = e.mode == BD_MODE_VCM ? (float)(e.w * e.h) * kPi * r * r : 0.0f;      // deviceCode.cu:1962
    P.seed = e.seed; P.iter = iter; P.rank = e.rank; P.world = e.world; P.nLocal = e.nLocal;
    P.colors = colors;
    const size_t mySlab = ((size_t)e.slabBuf * e.world + e.rank) * e.slabWords;
    P.slab = e.slabs + mySlab; P.slabCap = e.slabCap;
    P.numPeers = 0; P.stamp = e.seq + 1u;
    if (e.peerWrite)
        for (int r = 0; r < e.world; ++r)
            if (r != e.rank && e.peerBase[r]) P.peerSlab[P.numPeers++] = (uint4*)e.peerBase[r] + mySlab;
    {   // PhotonSoA over the one allocation: a | b, each maxPhotons (rounded to 64) 16 B entries
        size_t np = ((size_t)e.maxPhotons + 63) & ~(size_t)63;
        P.photons.a = reinterpret_cast<const float4*>(e.photonsSorted);
        P.photons.b = e.photonsSorted + np;
    }
    P.cellStart = e.grid.cellStart; P.cellEnd = e.grid.cellEnd; P.grid = e.grid.params;
    P.gridMagic = e.grid.tableSize ? 0xFFFFFFFFFFFFFFFFull / e.grid.tableSize + 1ull : 0ull;
    return P;
}

}  // namespace nv

namespace nv {

static inline int bdGrid() { return smCount() * 8; }
static inline bool bdNested(const BdEngine& e) { return (e.scene->features & NOVUM_FEATURE_NESTED_BIDIR) != 0; }

// bdMergeLambert<MIS, MINB>: 4 eye vertices per warp; MINB = resident CTAs of 128 threads per SM the register allocation is held to
// (8 -> 64 registers, 7 -> 72, 6 -> 80); NOVUM_MERGE_MINB in the environment picks the build
static int mergeMinB() {
    static int v = 0;
    if (!v) { const char* e = getenv("NOVUM_MERGE_MINB"); const int k = e ? atoi(e) : kMergeCtasPerSm; v = (k == 6 || k == 7) ? k : 8; }
    return v;
}
static void launchMergeLambert(bool mis, cudaStream_t s, const SceneView& sc, const BdState& st, const BdParams& P, const MergeQueue& mq) {
    const int b = mergeMinB(), grid = smCount() * b;
    if (b == 8) { if (mis) bdMergeLambert<true, 8><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); else bdMergeLambert<false, 8><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); }
    else if (b == 7) { if (mis) bdMergeLambert<true, 7><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); else bdMergeLambert<false, 7><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); }
    else { if (mis) bdMergeLambert<true, 6><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); else bdMergeLambert<false, 6><<<grid, kMergeThreads, 0, s>>>(sc, st, P, mq); }
}

// light pass of iteration `iter` for this rank's rows: fills the light-vertex cache, appends photons to this rank's slab,
// splats the t = 1 connections into `colors`
static int bdLightPass(BdEngine& e, int iter, float4* colors) {
    cudaStream_t s = e.stream;
    BdState& st = e.st;
    BdParams P = bdParams(e, iter, colors);
    SceneView sc = e.scene->view();
    e.mergeRadius = P.mergeRadius;
    const bool nested = bdNested(e);
    if (nested && !st.medium) {          // the walks' medium stacks exist only once the feature has been asked for
        cudaError_t me = cudaMalloc((void**)&st.medium, sizeof(uint4) * (size_t)(e.nLocal > 0 ? e.nLocal : 1));
        if (me != cudaSuccess) return (int)me;
    }
    bdFillWords<<<2, 256, 0, s>>>(reinterpret_cast<uint4*>(st.counters), (size_t)(kMaxBdDepth + 1) * BC_STRIDE * sizeof(int) / 16, 0u);
    bdFillWords<<<1, 32, 0, s>>>(P.slab, kSlabHeaderBytes / 16, 0u);          // photon counter of this rank (deviceCode.cu:3330)
    const bool publish = P.numPeers > 0 || e.slabBuffers > 1;       // lane engines stamp their own header too (bdWaitSlabs reads all of them)
    if (e.nLocal == 0 || sc.numLights == 0) {
        if (e.nLocal > 0) { bdZeroInts<<<bdGrid(), 256, 0, s>>>(st.pathLen, e.nLocal); ++e.launches; }      // no light: no light vertices to connect to
        if (publish) { bdPushHeader<<<1, 32, 0, s>>>(P); ++e.launches; }
        return (int)cudaGetLastError();
    }
    const int numSteps = e.mode == BD_MODE_BDPT ? e.lightDepth - 1 : e.lightDepth;      // :1116 (depth 1..max-1) vs :2029 (0..max-1)
    bdTick(e, 1);
    bdLightStart<<<(e.nLocal + kBdThreads - 1) / kBdThreads, kBdThreads, 0, s>>>(sc, st, P);
    ++e.launches; bdTick(e, 1);
    const int grid = bdGrid();
    TravCounters tc{e.scene->d_counters, e.scene->d_counters + 1};
    for (int step = 0; step < numSteps; ++step) {
        int* cnt = st.counters + step * BC_STRIDE;
        if (e.profCount) bdExtend<true><<<grid, kBdThreads, 0, s>>>(sc, st, st.queue[step & 1], cnt, tc);
        else bdExtend<false><<<grid, kBdThreads, 0, s>>>(sc, st, st.queue[step & 1], cnt, tc);
        bdTick(e, 0);
        if (nested) bdLightStep<true><<<grid, kBdThreads, 0, s>>>(sc, st, P, step, numSteps, st.queue[step & 1], st.queue[(step + 1) & 1]);
        else bdLightStep<false><<<grid, kBdThreads, 0, s>>>(sc, st, P, step, numSteps, st.queue[step & 1], st.queue[(step + 1) & 1]);
        bdTick(e, 1);
        if (P.lightTrace) { bdResolve<<<grid, kBdThreads, 0, s>>>(sc, st, colors, cnt); ++e.launches; bdTick(e, 2); }
        e.launches += 2;
    }
    if (numSteps == 0 && P.lightTrace) { bdResolve<<<grid, kBdThreads, 0, s>>>(sc, st, colors, st.counters); ++e.launches; bdTick(e, 2); }
    if (publish) { bdPushHeader<<<1, 32, 0, s>>>(P); ++e.launches; bdTick(e, 1); }
    return (int)cudaGetLastError();
}

// after the slabs of all ranks are in place: count, compact, hash, sort, cell table, reorder into cell order
static int bdBuildGrid(BdEngine& e, int iter) {
    if (e.mode != BD_MODE_VCM) return 0;
    cudaStream_t s = e.stream;
    BdParams P = bdParams(e, iter, nullptr);
    if (!P.doMerge || P.mergeRadius <= 0.0f) { e.photonCount = 0; return 0; }
    const uint4* slabs = e.slabs + (size_t)e.slabBuf * e.world * e.slabWords;
    bdSlabCounts<<<1, 32, 0, s>>>(slabs, e.slabWords, e.world, e.slabCap, e.d_counts, e.h_countsDev);
    ++e.launches;
    cudaError_t err = cudaStreamSynchronize(s);           // the reference blocks on the same counter (deviceCode.cu:3352)
    if (err != cudaSuccess) return (int)err;
    if (e.h_counts[16]) return NOVUM_E_TIMEOUT;          // a peer never published its light pass
    long long total = 0;
    for (int r = 0; r < e.world; ++r) total += e.h_counts[r];
    if (total > e.maxPhotons) total = e.maxPhotons;
    e.photonCount = (int)total;
    const uint4* src = slabs + kSlabHeaderBytes / 16;
    if (e.world > 1) {
        dim3 g(256, e.world);
        bdCompactSlabs<<<g, 256, 0, s>>>(slabs, e.slabWords, e.world, e.d_counts, e.photonsAll);
        ++e.launches;
        src = e.photonsAll;
    }
    // cells of r / kMergeFine (vcm.cu): the merge looks at the cells within reach of a vertex, not at a fixed 27
    int rc = hashGridBuildAoS(e.grid, reinterpret_cast<const float*>(src), 8, e.photonCount, e.scene->sceneMin, P.mergeRadius / (float)kMergeFine, s);
    if (rc) return rc;
    e.launches += e.grid.launches;
    if (e.photonCount > 0) {
        bdReorderPhotons<<<(e.photonCount + 255) / 256, 256, 0, s>>>(src, e.grid.index, e.photonCount, 1.0f / P.etaVcm,
            const_cast<float4*>(P.photons.a), const_cast<uint4*>(P.photons.b));
        ++e.launches;
    }
    bdTick(e, 1);
    return (int)cudaGetLastError();
}

static int bdEyePass(BdEngine& e, int iter, float4* colors) {
    cudaStream_t s = e.stream;
    BdState& st = e.st;
    BdParams P = bdParams(e, iter, colors);
    SceneView sc = e.scene->view();
    if (e.nLocal == 0) return 0;
    const bool nested = bdNested(e) && st.medium != nullptr;
    bdFillWords<<<2, 256, 0, s>>>(reinterpret_cast<uint4*>(st.counters), (size_t)(kMaxBdDepth + 1) * BC_STRIDE * sizeof(int) / 16, 0u);
    const int numSteps = e.mode == BD_MODE_BDPT ? e.eyeDepth - 1 : e.eyeDepth;          // :794 vs :2723
    bdTick(e, 1);
    bdEyeStart<<<(e.nLocal + kBdThreads - 1) / kBdThreads, kBdThreads, 0, s>>>(st, P);
    ++e.launches; bdTick(e, 1);
    const int grid = bdGrid();
    const int mgrid = smCount() * kMergeCtasPerSm;   // merge kernels: 128 threads, 8 resident CTAs per SM
    const bool shadows = P.nee || P.connection;
    const bool merging = e.mode == BD_MODE_VCM && P.doMerge && P.mergeRadius > 0.0f;
    TravCounters tc{e.scene->d_counters, e.scene->d_counters + 1};
    for (int step = 0; step < numSteps; ++step) {
        int* cnt = st.counters + step * BC_STRIDE;
        if (e.profCount) bdExtend<true><<<grid, kBdThreads, 0, s>>>(sc, st, st.queue[step & 1], cnt, tc);
        else bdExtend<false><<<grid, kBdThreads, 0, s>>>(sc, st, st.queue[step & 1], cnt, tc);
        bdTick(e, 0);
        // merge records: step 0 has its own queue and is gathered at once; deeper steps append to the queue of their group
        const bool grouped = step > 0;
        const int groupFirst = grouped ? 1 + ((step - 1) / kMergeGroup) * kMergeGroup : 0;
        MergeQueue mq;
        mq.recs = grouped ? st.mergeQ + e.nLocal : st.mergeQ;
        mq.cap = grouped ? kMergeGroup * e.nLocal : e.nLocal;
        mq.cnt = st.counters + groupFirst * BC_STRIDE;
        mq.order = nullptr;
        if (nested) bdEyeStep<true><<<grid, kBdThreads, 0, s>>>(sc, st, P, step, numSteps, st.queue[step & 1], st.queue[(step + 1) & 1], mq);
        else bdEyeStep<false><<<grid, kBdThreads, 0, s>>>(sc, st, P, step, numSteps, st.queue[step & 1], st.queue[(step + 1) & 1], mq);
        bdTick(e, 1);
        if (shadows) { bdResolve<<<grid, kBdThreads, 0, s>>>(sc, st, colors, cnt); ++e.launches; bdTick(e, 2); }
        const bool flush = !grouped || step == groupFirst + kMergeGroup - 1 || step == numSteps - 1;
        if (merging && flush) {
            if (!P.paintWeight) {      // every Lambert queue is put in photon-cell order (bdMergeKeys): radix sort with the count on the device
                int cellBits = 1;
                for (int k = 0; k < 3; ++k) {
                    const float cells = (e.scene->sceneMax[k] - e.scene->sceneMin[k]) / P.grid.cell;
                    while (cellBits < 30 && (float)(1 << cellBits) < cells + 2.0f) ++cellBits;
                }
                const int shift = cellBits > 10 ? cellBits - 10 : 0;
                bdMergeKeys<<<grid, 256, 0, s>>>(mq, P.grid, shift, st.mergeKeys, st.mergeOrder);
                int sortLaunches = 0;
                radixSortPairsDeviceCount(e.mergePlan, st.mergeKeys, reinterpret_cast<uint32_t*>(st.mergeOrder), mq.cap, mq.cnt + BC_MERGE, 30, s, &sortLaunches);
                e.launches += 1 + sortLaunches;
                mq.order = st.mergeOrder;
            }
            // Lambert vertices (front of the queue) and all other materials (back): a kernel whose queue is empty exits at once
            if (P.paintWeight) {}
            else launchMergeLambert(P.doMis != 0, s, sc, st, P, mq);
            bdMerge<<<mgrid, kMergeThreads, 0, s>>>(sc, st, P, mq);
            e.launches += P.paintWeight ? 1 : 2; bdTick(e, 3);
        }
        e.launches += 2;
    }
    ++e.iterations;
    return (int)cudaGetLastError();
}

static void bdFillStats(BdEngine& e, NovumRenderStats* stats) {
    if (!stats) return;
    unsigned long long h[8] = {0};
    cudaMemcpy(h, e.st.stats, sizeof(h), cudaMemcpyDeviceToHost);
    memset(stats, 0, sizeof(*stats));
    stats->closestRays = h[0]; stats->shadowRays = h[1]; stats->photons = h[2]; stats->lightVertices = h[3];
    stats->mergeAccepted = h[4]; stats->mergeCandidates = h[5]; stats->mergeQueries = h[6];
    stats->paths = 2ull * (unsigned long long)e.nLocal * e.iterations;      // one light + one eye path per pixel per iteration
    stats->kernelLaunches = e.launches;
    stats->deviceMs = e.deviceMs;
    for (int k = 0; k < 4; ++k) { stats->stageMs[k] = e.stageMs[k]; stats->stageLaunches[k] = e.stageLaunches[k]; }
    if (e.profCount) {
        unsigned long long c2[2] = {0, 0};
        cudaMemcpy(c2, e.scene->d_counters, sizeof(c2), cudaMemcpyDeviceToHost);
        stats->bvhNodeVisits = c2[0];
        stats->triTests = c2[1];
    }
}

static void bdResetStats(BdEngine& e) {
    cudaMemsetAsync(e.st.stats, 0, 8 * sizeof(unsigned long long), e.stream);
    e.launches = 0; e.iterations = 0; e.deviceMs = 0.f; e.evUsed = 0;
    for (int k = 0; k < 4; ++k) { e.stageMs[k] = 0.f; e.stageLaunches[k] = 0; }
    e.prof = (e.scene->profileFlags & 1) != 0; e.profCount = (e.scene->profileFlags & 2) != 0;
    if (e.profCount) cudaMemsetAsync(e.scene->d_counters, 0, 2 * sizeof(unsigned long long), e.stream);
}

static bool bdMatches(const BdEngine& e, const NovumRenderConfig& cfg, NovumShard shard) {
    return e.w == cfg.width && e.h == cfg.height && e.eyeDepth == cfg.bdptEyeDepth && e.lightDepth == cfg.bdptLightDepth &&
           e.mode == bdModeOf(cfg.integrator) && e.rank == shard.rank && e.world == shard.world;
}

// ---- tile renderer: one rank of a row-sharded bidirectional render, `lanes` iterations in flight ----------------------------------
// Why lanes: a depth-synchronous iteration is ~130 launches whose later ones (deep bounces, a few long rays inside the glass) are
// latency chains that leave the GPU almost empty, and their length does not shrink when a rank owns 1/8 of the pixels - measured on
// one B200 emulating rank 0 of 8 (tools/tile_phase_timing.py): 14.4 ms per 4K VCM iteration where 53.5 / 8 = 6.7 ms was the ideal.
// Iterations are independent (the frame is a sum), so every rank runs `lanes` engines, lane k taking iterations first + k, first + k +
// lanes, ..., each on its own stream with its own host thread: the tails of one lane's kernels overlap the bulk of another's, and the
// one blocking read per iteration (the photon count) stalls only its own lane.
// Exchange: every light pass stores its photons into the same lane's slab on every peer as it appends them (bdLightStep), then stamps
// the headers (bdPushHeader); bdWaitSlabs spins ON THE DEVICE until all ranks' stamps for this light pass are in: no collective, no host.
// Slabs are double-buffered by the lane's light-pass count: a rank can start light pass q + 1 while a peer still compacts the slabs
// of q; it cannot reach q + 2 before that peer has stamped q + 1, which the peer does after its compaction of q (stream order).
struct TileRank {
    NovumScene* scene = nullptr;
    int rank = 0, world = 1, lanes = 1, device = 0;
    std::vector<BdEngine*> eng;
    cudaEvent_t evStart = nullptr; std::vector<cudaEvent_t> evEnd;
};

// iterations firstIter + j, j = offset, offset + stride, ... < numIter (stride > 1: whole iterations striped over ranks, as
// novum_render_device does); the lanes take them round-robin
static int tileLaneRun(TileRank& t, int lane, int firstIter, int numIter, int offset, int stride, float4* d_colors) {
    cudaSetDevice(t.device);
    BdEngine& e = *t.eng[lane];
    const bool exchange = e.world > 1 && e.mode == BD_MODE_VCM;
    for (int j = offset + stride * lane; j < numIter; j += stride * t.lanes) {
        const int iter = firstIter + j;
        e.slabBuf = (int)(e.seq & 1u);
        int rc = bdLightPass(e, iter, d_colors);
        if (rc) return rc;
        if (exchange) {
            const uint4* slabs = e.slabs + (size_t)e.slabBuf * e.world * e.slabWords;
            bdWaitSlabs<<<1, 32, 0, e.stream>>>(slabs, e.slabWords, e.world, e.seq + 1u, e.d_counts + 16);
            ++e.launches;
        }
        if ((rc = bdBuildGrid(e, iter)) != 0) return rc;
        if ((rc = bdEyePass(e, iter, d_colors)) != 0) return rc;
        ++e.seq;
    }
    return (int)cudaGetLastError();
}


int tileRender(TileRank& t, int firstIter, int numIter, int offset, int stride, float4* d_colors, NovumRenderStats* stats) {
    cudaSetDevice(t.device);
    if (t.world > 1)
        for (BdEngine* e : t.eng) {
            for (int r = 0; r < t.world; ++r)
                if (r != t.rank && !e->peerBase[r]) return NOVUM_E_ARG;          // every peer must be attached first
            e->peerWrite = true;
        }
    for (BdEngine* e : t.eng) bdResetStats(*e);
    cudaEventRecord(t.evStart, t.eng[0]->stream);
    for (int k = 1; k < t.lanes; ++k) cudaStreamWaitEvent(t.eng[k]->stream, t.evStart, 0);
    std::vector<int> rcs(t.lanes, 0);
    std::vector<std::thread> th;
    for (int k = 1; k < t.lanes; ++k) th.emplace_back([&, k]() { rcs[k] = tileLaneRun(t, k, firstIter, numIter, offset, stride, d_colors); });
    rcs[0] = tileLaneRun(t, 0, firstIter, numIter, offset, stride, d_colors);
    for (std::thread& x : th) x.join();
    int rc = 0;
    for (int k = 0; k < t.lanes; ++k) {
        cudaEventRecord(t.evEnd[k], t.eng[k]->stream);
        cudaError_t err = cudaStreamSynchronize(t.eng[k]->stream);
        if (!rc && rcs[k]) rc = rcs[k];
        if (!rc && err != cudaSuccess) rc = (int)err;
        if (!rc && t.eng[k]->h_counts[16]) rc = NOVUM_E_TIMEOUT;
    }
    if (rc) return rc;
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        for (int k = 0; k < t.lanes; ++k) {
            BdEngine& e = *t.eng[k];
            float ms = 0; cudaEventElapsedTime(&ms, t.evStart, t.evEnd[k]);
            for (size_t i = 1; i < e.evUsed; ++i) { float g = 0; cudaEventElapsedTime(&g, e.evs[i - 1], e.evs[i]); e.stageMs[e.evCls[i]] += g; }
            NovumRenderStats one;
            bdFillStats(e, &one);
            stats->closestRays += one.closestRays; stats->shadowRays += one.shadowRays; stats->paths += one.paths;
            stats->lightVertices += one.lightVertices; stats->kernelLaunches += one.kernelLaunches;
            stats->mergeQueries += one.mergeQueries; stats->mergeCandidates += one.mergeCandidates; stats->mergeAccepted += one.mergeAccepted;
            stats->bvhNodeVisits = one.bvhNodeVisits; stats->triTests = one.triTests;       // the lanes count into the scene's one pair of counters
            for (int c = 0; c < 4; ++c) { stats->stageMs[c] += one.stageMs[c]; stats->stageLaunches[c] += one.stageLaunches[c]; }
            stats->photons += one.photons;
            if (ms > stats->deviceMs) stats->deviceMs = ms;
        }
    }
    return 0;
}


static void tileFree(TileRank* t) {
    if (!t) return;
    cudaSetDevice(t->device);
    for (BdEngine* e : t->eng) { bdFree(*e); delete e; }
    for (cudaEvent_t ev : t->evEnd) cudaEventDestroy(ev);
    if (t->evStart) cudaEventDestroy(t->evStart);
    delete t;
}
static void tileEngineFree(void* p) { tileFree((TileRank*)p); }

// CUDA loads a kernel's code on its first launch, and that load may need every running kernel of the context to finish (lazy module
// loading). A lane that spins in bdWaitSlabs while ANOTHER lane of the same GPU launches a kernel for the first time would then never
// see that launch start: the lane behind it never publishes its light pass and the ranks waiting for it time out (seen in the first,
// warm-up step at 4 x 3 and 8 x 4 lanes). So every kernel of an iteration is loaded before the first device-side wait can exist.
static int bdPreloadKernels() {
    cudaFuncAttributes a;
    cudaError_t e = cudaSuccess;
    auto L = [&](const void* f) { if (e == cudaSuccess) e = cudaFuncGetAttributes(&a, f); };
    L((const void*)bdFillWords); L((const void*)bdZeroInts); L((const void*)bdLightStart); L((const void*)bdExtend<false>); L((const void*)bdExtend<true>);
    L((const void*)bdLightStep<false>); L((const void*)bdLightStep<true>); L((const void*)bdResolve); L((const void*)bdPushHeader); L((const void*)bdWaitSlabs); L((const void*)bdSlabCounts);
    L((const void*)bdCompactSlabs); L((const void*)bdReorderPhotons); L((const void*)bdEyeStart); L((const void*)bdEyeStep<false>); L((const void*)bdEyeStep<true>);
    L((const void*)bdMergeKeys); L((const void*)bdMergeLambert<true, 8>); L((const void*)bdMergeLambert<false, 8>); L((const void*)bdMergeLambert<true, 7>);
    L((const void*)bdMergeLambert<false, 7>); L((const void*)bdMergeLambert<true, 6>); L((const void*)bdMergeLambert<false, 6>);
    L((const void*)bdMerge);
    if (e != cudaSuccess) return (int)e;
    return hashGridPreload();
}

static int tileCreate(NovumScene* s, const NovumRenderConfig& cfg, const NovumCamera& cam, NovumShard shard, unsigned long long seed, int lanes,
                      TileRank** out) {
    int rc = bdPreloadKernels();
    if (rc) return rc;
    TileRank* t = new TileRank();
    t->scene = s; t->rank = shard.rank; t->world = shard.world;
    cudaGetDevice(&t->device);
    for (int k = 0; k < lanes; ++k) {
        BdEngine* e = new BdEngine();
        rc = bdAlloc(*e, s, cfg, cam, shard, seed, /*lane=*/true);
        if (rc) {
            bdFree(*e); delete e;
            // out of memory for this lane: run with the lanes that fit (world = 1 only: the ranks of a tile render must agree on the count)
            if (rc == (int)cudaErrorMemoryAllocation && k > 0 && shard.world == 1) { cudaGetLastError(); rc = 0; break; }
            tileFree(t);
            return rc;
        }
        t->eng.push_back(e);
        cudaEvent_t ev; cudaEventCreate(&ev); t->evEnd.push_back(ev);
    }
    t->lanes = (int)t->eng.size();
    cudaEventCreate(&t->evStart);
    for (BdEngine* e : t->eng) cudaStreamSynchronize(e->stream);          // slabs zeroed before a peer may store into them
    *out = t;
    return 0;
}

// iterations in flight of the single-call render (novum_render_device): NOVUM_LANES in the environment, default 4
// (4K VCM on one B200, 20 iterations per call: 49.6 / 45.1 / 44.1 / 43.8 ms per iteration with 1 / 2 / 3 / 4 lanes; a lane costs 16 GB
// there; a call of k iterations keeps min(k, lanes) busy, so callers should hand over several iterations at a time)
static int defaultLanes() {
    const char* e = getenv("NOVUM_LANES");
    int n = e ? atoi(e) : 4;
    return n < 1 ? 1 : (n > 8 ? 8 : n);
}

// the engines' buffers depend on (w, h, depths, mode) only: when the launch boundary rebuilds its scene (new geometry behind the same
// call, main.cu:925-933 re-renders 75 times) they move to the new scene instead of being freed and re-allocated (16 GB per lane at 4K)
void bdRebindScene(void* p, NovumScene* s) {
    TileRank* t = (TileRank*)p;
    if (!t) return;
    t->scene = s;
    for (BdEngine* e : t->eng) e->scene = s;
}

// Single-call form (novum_render_device): whole iterations are striped over the ranks of `shard` - iteration i runs all light and
// eye paths on rank i % world, so no exchange is needed and the caller only sums the framebuffers at the end. On the rank, the
// iterations go through the tile renderer's lanes (world = 1: the whole frame).
int vcmRenderDevice(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                    NovumShard shard, unsigned long long seed, float4* d_colors, NovumRenderStats* stats) {
    NovumShard whole = {0, 1};
    TileRank* t = (TileRank*)s->vcmEngine;
    if (t && !bdMatches(*t->eng[0], *cfg, whole)) { tileFree(t); s->vcmEngine = nullptr; t = nullptr; }
    if (!t) {
        int rc = tileCreate(s, *cfg, *cam, whole, seed, defaultLanes(), &t);
        if (rc) return rc;
        s->vcmEngine = t; s->vcmEngineFree = tileEngineFree;
    }
    for (BdEngine* e : t->eng) { e->cam = *cam; e->cfg = *cfg; e->seed = seed; }
    return tileRender(*t, firstSample, numSample, shard.rank, shard.world, d_colors, stats);
}

}  // namespace nv

// the opaque NovumTiles handle of the C ABI is a nv::TileRank
static inline nv::TileRank* tr(NovumTiles* T) { return reinterpret_cast<nv::TileRank*>(T); }
static inline const nv::TileRank* tr(const NovumTiles* T) { return reinterpret_cast<const nv::TileRank*>(T); }

extern "C" {

int novum_tiles_create(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, NovumShard shard, uint64_t seed, int lanes,
                       NovumTiles** out) {
    if (!s) return NOVUM_E_NOSCENE;
    if (!cfg || !cam || !out || shard.world < 1 || shard.world > 16 || shard.rank < 0 || shard.rank >= shard.world) return NOVUM_E_ARG;
    if (lanes < 1 || lanes > 8) return NOVUM_E_ARG;
    if (cfg->integrator != NOVUM_VCM && cfg->integrator != NOVUM_SPPM && cfg->integrator != NOVUM_BIDIRECTIONAL) return NOVUM_E_ARG;
    nv::TileRank* t = nullptr;
    int rc = nv::tileCreate(s, *cfg, *cam, shard, seed, lanes, &t);
    if (rc) return rc;
    *out = reinterpret_cast<NovumTiles*>(t);
    return 0;
}

void novum_tiles_destroy(NovumTiles* T) { nv::tileFree(tr(T)); }

/* camera, strategy toggles, merge parameters and seed of the following renders; resolution, depths and integrator family are fixed at creation */
int novum_tiles_set_view(NovumTiles* T, const NovumRenderConfig* cfg, const NovumCamera* cam, uint64_t seed) {
    if (!T || !cfg || !cam) return NOVUM_E_ARG;
    nv::TileRank& t = *tr(T);
    if (!nv::bdMatches(*t.eng[0], *cfg, NovumShard{t.rank, t.world})) return NOVUM_E_ARG;
    for (nv::BdEngine* e : t.eng) { e->cam = *cam; e->cfg = *cfg; e->seed = seed; }
    return 0;
}

int novum_tiles_lanes(const NovumTiles* T) { return T ? tr(T)->lanes : NOVUM_E_ARG; }

/* lanes x 64 bytes: the CUDA IPC handle of every lane's slab buffer, for the caller to carry to the other processes */
int novum_tiles_ipc_export(NovumTiles* T, void* handles) {
    if (!T || !handles) return NOVUM_E_ARG;
    for (int k = 0; k < tr(T)->lanes; ++k) {
        cudaIpcMemHandle_t h;
        cudaError_t err = cudaIpcGetMemHandle(&h, tr(T)->eng[k]->slabs);
        if (err != cudaSuccess) return (int)err;
        memcpy((char*)handles + 64 * k, &h, 64);
    }
    return 0;
}

int novum_tiles_ipc_attach(NovumTiles* T, int peerRank, const void* handles) {
    if (!T || !handles || peerRank < 0 || peerRank >= tr(T)->world || peerRank == tr(T)->rank) return NOVUM_E_ARG;
    for (int k = 0; k < tr(T)->lanes; ++k) {
        nv::BdEngine& e = *tr(T)->eng[k];
        if (e.peerBase[peerRank]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, (const char*)handles + 64 * k, 64);
        void* p = nullptr;
        cudaError_t err = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (err != cudaSuccess) { cudaGetLastError(); return (int)err; }
        e.peerBase[peerRank] = p; e.peerIpc[peerRank] = true;
    }
    return 0;
}

/* same-process ranks (one per GPU with peer access enabled, or several on one GPU): lane k's slab buffer of the peer, from novum_tiles_bases */
int novum_tiles_bases(NovumTiles* T, void** bases) {
    if (!T || !bases) return NOVUM_E_ARG;
    for (int k = 0; k < tr(T)->lanes; ++k) bases[k] = tr(T)->eng[k]->slabs;
    return 0;
}

int novum_tiles_attach_pointers(NovumTiles* T, int peerRank, void* const* bases) {
    if (!T || !bases || peerRank < 0 || peerRank >= tr(T)->world || peerRank == tr(T)->rank) return NOVUM_E_ARG;
    for (int k = 0; k < tr(T)->lanes; ++k) {
        nv::BdEngine& e = *tr(T)->eng[k];
        if (!bases[k]) return NOVUM_E_ARG;
        e.peerBase[peerRank] = bases[k]; e.peerIpc[peerRank] = false;
    }
    return 0;
}

/* Renders iterations [firstIter, firstIter + numIter) of this rank's rows into d_colors (full-frame SUM accumulator of this rank;
 * t = 1 splats land anywhere). Every rank of the render must make the same call. Blocking. */
int novum_tiles_render(NovumTiles* T, int firstIter, int numIter, NovumFloat4* d_colors, NovumRenderStats* stats) {
    if (!T || !d_colors || numIter < 0) return NOVUM_E_ARG;
    return nv::tileRender(*tr(T), firstIter, numIter, 0, 1, (float4*)d_colors, stats);
}

}  // extern "C"

namespace nv {
}  // namespace nv

// ---- C ABI: multi-GPU phase API (tile-sharded: rank r owns pixel rows y % world == r for BOTH walks) -----------------------
struct NovumVcmState { nv::BdEngine e; };

extern "C" {

int novum_vcm_create(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, NovumShard shard, uint64_t seed,
                     NovumVcmState** out) {
    if (!s) return NOVUM_E_NOSCENE;
    if (!cfg || !cam || !out || shard.world < 1 || shard.world > 16 || shard.rank < 0 || shard.rank >= shard.world) return NOVUM_E_ARG;
    if (cfg->integrator != NOVUM_VCM && cfg->integrator != NOVUM_SPPM && cfg->integrator != NOVUM_BIDIRECTIONAL) return NOVUM_E_ARG;
    NovumVcmState* v = new NovumVcmState();
    int rc = nv::bdAlloc(v->e, s, *cfg, *cam, shard, seed);
    if (rc) { nv::bdFree(v->e); delete v; return rc; }
    *out = v;
    return 0;
}

void novum_vcm_destroy(NovumVcmState* v) {
    if (!v) return;
    nv::bdFree(v->e);
    delete v;
}

int novum_vcm_light_pass(NovumVcmState* v, int iter, NovumFloat4* d_splat) {
    if (!v || !d_splat) return NOVUM_E_ARG;
    return nv::bdLightPass(v->e, iter, (float4*)d_splat);
}

int novum_vcm_cache_buffer(NovumVcmState* v, void** d_cacheBase, uint64_t* slabBytes, uint64_t* totalBytes) {
    if (!v) return NOVUM_E_ARG;
    if (d_cacheBase) *d_cacheBase = v->e.slabs;
    if (slabBytes) *slabBytes = v->e.slabWords * 16;
    if (totalBytes) *totalBytes = v->e.slabWords * 16 * v->e.world;
    return 0;
}

int novum_vcm_build_grid(NovumVcmState* v, int iter) {
    if (!v) return NOVUM_E_ARG;
    return nv::bdBuildGrid(v->e, iter);
}

int novum_vcm_eye_pass(NovumVcmState* v, int iter, NovumFloat4* d_colors) {
    if (!v || !d_colors) return NOVUM_E_ARG;
    return nv::bdEyePass(v->e, iter, (float4*)d_colors);
}

int novum_vcm_stats(NovumVcmState* v, NovumRenderStats* stats) {
    if (!v || !stats) return NOVUM_E_ARG;
    cudaError_t err = cudaStreamSynchronize(v->e.stream);
    if (err != cudaSuccess) return (int)err;
    nv::BdEngine& e = v->e;
    for (size_t i = 1; i < e.evUsed; ++i) { float ms = 0; cudaEventElapsedTime(&ms, e.evs[i - 1], e.evs[i]); e.stageMs[e.evCls[i]] += ms; }
    e.evUsed = 0;
    nv::bdFillStats(v->e, stats);
    stats->photons = (uint64_t)v->e.photonCount;      // photons in the last grid (all ranks)
    return 0;
}

/* Peer-write exchange: every rank exports its slab buffer (64-byte CUDA IPC handle), the caller moves the handles between the
   processes (any transport: torch.distributed all_gather_object, MPI, a pipe) and attaches them; novum_vcm_peer_write(1) then
   makes each light pass store its photons into all attached peers as it appends them. The caller must order the ranks between
   light pass and grid build, and between grid build and the next light pass (a tiny collective on novum_vcm_stream). */
int novum_vcm_ipc_export(NovumVcmState* v, void* handle64) {
    if (!v || !handle64) return NOVUM_E_ARG;
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    cudaIpcMemHandle_t h;
    cudaError_t err = cudaIpcGetMemHandle(&h, v->e.slabs);
    if (err != cudaSuccess) return (int)err;
    memcpy(handle64, &h, 64);
    return 0;
}

int novum_vcm_ipc_attach(NovumVcmState* v, int peerRank, const void* handle64) {
    if (!v || !handle64 || peerRank < 0 || peerRank >= v->e.world || peerRank == v->e.rank) return NOVUM_E_ARG;
    if (v->e.peerBase[peerRank]) return 0;
    cudaIpcMemHandle_t h;
    memcpy(&h, handle64, 64);
    void* p = nullptr;
    cudaError_t err = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (err != cudaSuccess) { cudaGetLastError(); return (int)err; }
    v->e.peerBase[peerRank] = p; v->e.peerIpc[peerRank] = true;
    return 0;
}

/* same-process form (ranks that are sessions of one process, on one GPU or on peer-enabled GPUs): attach by device pointer */
int novum_vcm_attach_pointer(NovumVcmState* v, int peerRank, void* d_peerCacheBase) {
    if (!v || !d_peerCacheBase || peerRank < 0 || peerRank >= v->e.world || peerRank == v->e.rank) return NOVUM_E_ARG;
    if (v->e.peerBase[peerRank] && v->e.peerIpc[peerRank]) cudaIpcCloseMemHandle(v->e.peerBase[peerRank]);
    v->e.peerBase[peerRank] = d_peerCacheBase; v->e.peerIpc[peerRank] = false;
    return 0;
}

int novum_vcm_peer_write(NovumVcmState* v, int enable) {
    if (!v) return NOVUM_E_ARG;
    if (enable)
        for (int r = 0; r < v->e.world; ++r)
            if (r != v->e.rank && !v->e.peerBase[r]) return NOVUM_E_ARG;      // every peer must be attached first
    v->e.peerWrite = enable != 0;
    return 0;
}

/* zeroes the session's counters and re-reads the scene's profile flags (novum_set_profile): stage times in novum_vcm_stats */
int novum_vcm_reset_stats(NovumVcmState* v) {
    if (!v) return NOVUM_E_ARG;
    nv::bdResetStats(v->e);
    return 0;
}

int novum_vcm_stream(NovumVcmState* v, void** stream) {
    if (!v || !stream) return NOVUM_E_ARG;
    *stream = (void*)v->e.stream;
    return 0;
}

}  // extern "C"
