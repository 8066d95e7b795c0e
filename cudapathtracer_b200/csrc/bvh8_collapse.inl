// bvh8_collapse.inl — second half of bvh8_build.cu: BVH2 -> compressed BVH8 and the host driver.

namespace nv {
namespace {

struct Bvh2View {
    const float4* lo; const float4* hi;
    const int* left; const int* right;
    const int* rangeFirst; const int* rangeLast;
    int n;   // number of leaves (triangles)
    __device__ __forceinline__ bool isLeaf(int node) const { return node >= n - 1; }
    __device__ __forceinline__ int triCount(int node) const { return isLeaf(node) ? 1 : rangeLast[node] - rangeFirst[node] + 1; }
    __device__ __forceinline__ int firstTri(int node) const { return isLeaf(node) ? node - (n - 1) : rangeFirst[node]; }
    __device__ __forceinline__ float area(int node) const {
        float4 a = lo[node], b = hi[node];
        float dx = b.x - a.x, dy = b.y - a.y, dz = b.z - a.z;
        return dx * dy + dy * dz + dz * dx;
    }
};

// smallest e with 2^e * 255 >= ext, as a biased float exponent byte
__device__ __forceinline__ uint32_t scaleExponent(float ext) {
    float s = ext / 255.0f;
    uint32_t b = __float_as_uint(s);
    uint32_t e = (b >> 23) & 0xffu;
    if (b & 0x007fffffu) e += 1;
    // guard against the division rounding down
    while (e < 254u && __uint_as_float(e << 23) * 255.0f < ext) e += 1;
    if (e < 1u) e = 1u;
    if (e > 254u) e = 254u;
    return e;
}

constexpr int kMaxLeafTris = 3;

// One thread per task (bvh2 node, node8 index). Writes the node, its triangles, and the child tasks.
__global__ void collapseKernel(Bvh2View b2, const uint2* __restrict__ tasksIn, int numTasks, uint2* __restrict__ tasksOut,
                               int* __restrict__ counters /* [0]=next task count, [1]=node count, [2]=tri count */,
                               float4* __restrict__ nodes, float4* __restrict__ trisOut, const uint32_t* __restrict__ sortedIdx,
                               const float4* __restrict__ pos, const NovumTriangle* __restrict__ tris,
                               const DMat* __restrict__ mats) {
    int tid = blockIdx.x * blockDim.x + threadIdx.x;
    if (tid >= numTasks) return;
    uint2 task = tasksIn[tid];
    int root = (int)task.x;
    int ch[8];
    int nc = 0;
    if (b2.isLeaf(root)) { ch[nc++] = root; }
    else { ch[nc++] = b2.left[root]; ch[nc++] = b2.right[root]; }
    // phase 1: open the largest-area child that is too big for a leaf slot
    while (nc < 8) {
        int best = -1; float bestA = -1.f;
        for (int i = 0; i < nc; ++i) {
            if (!b2.isLeaf(ch[i]) && b2.triCount(ch[i]) > kMaxLeafTris) {
                float a = b2.area(ch[i]);
                if (a > bestA) { bestA = a; best = i; }
            }
        }
        if (best < 0) break;
        int node = ch[best];
        ch[best] = b2.left[node];
        ch[nc++] = b2.right[node];
    }
    // phase 2: spare slots split the remaining multi-triangle leaf subtrees (tighter boxes, no extra nodes)
    while (nc < 8) {
        int best = -1; float bestA = -1.f;
        for (int i = 0; i < nc; ++i) {
            if (!b2.isLeaf(ch[i]) && b2.triCount(ch[i]) <= kMaxLeafTris) {
                float a = b2.area(ch[i]);
                if (a > bestA) { bestA = a; best = i; }
            }
        }
        if (best < 0) break;
        int node = ch[best];
        ch[best] = b2.left[node];
        ch[nc++] = b2.right[node];
    }
    // node box = union of the chosen children
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = 0; i < nc; ++i) {
        float4 a = b2.lo[ch[i]], c = b2.hi[ch[i]];
        lo[0] = fminf(lo[0], a.x); lo[1] = fminf(lo[1], a.y); lo[2] = fminf(lo[2], a.z);
        hi[0] = fmaxf(hi[0], c.x); hi[1] = fmaxf(hi[1], c.y); hi[2] = fmaxf(hi[2], c.z);
    }
    float cen[3] = {(lo[0] + hi[0]) * 0.5f, (lo[1] + hi[1]) * 0.5f, (lo[2] + hi[2]) * 0.5f};
    // greedy child -> slot assignment: slot s prefers direction ((s&4)?+:-, (s&2)?+:-, (s&1)?+:-)
    int slotOf[8]; int childAt[8];
    for (int i = 0; i < 8; ++i) { slotOf[i] = -1; childAt[i] = -1; }
    float ccx[8], ccy[8], ccz[8];
    for (int i = 0; i < nc; ++i) {
        float4 a = b2.lo[ch[i]], c = b2.hi[ch[i]];
        ccx[i] = (a.x + c.x) * 0.5f - cen[0]; ccy[i] = (a.y + c.y) * 0.5f - cen[1]; ccz[i] = (a.z + c.z) * 0.5f - cen[2];
    }
    for (int round = 0; round < nc; ++round) {
        float bestC = -FLT_MAX; int bi = -1, bs = -1;
        for (int i = 0; i < nc; ++i) {
            if (slotOf[i] >= 0) continue;
            for (int s = 0; s < 8; ++s) {
                if (childAt[s] >= 0) continue;
                float c = ((s & 4) ? ccx[i] : -ccx[i]) + ((s & 2) ? ccy[i] : -ccy[i]) + ((s & 1) ? ccz[i] : -ccz[i]);
                if (c > bestC) { bestC = c; bi = i; bs = s; }
            }
        }
        slotOf[bi] = bs; childAt[bs] = bi;
    }
    // classify, allocate
    uint32_t imask = 0; int numInner = 0, numLeafTris = 0;
    for (int s = 0; s < 8; ++s) {
        int i = childAt[s];
        if (i < 0) continue;
        int node = ch[i];
        bool inner = !b2.isLeaf(node) && b2.triCount(node) > kMaxLeafTris;
        if (inner) { imask |= 1u << s; ++numInner; }
        else numLeafTris += b2.triCount(node);
    }
    int childBase = numInner ? atomicAdd(&counters[1], numInner) : 0;
    int triBase = numLeafTris ? atomicAdd(&counters[2], numLeafTris) : 0;
    int taskBase = numInner ? atomicAdd(&counters[0], numInner) : 0;
    uint32_t ex = scaleExponent(hi[0] - lo[0]), ey = scaleExponent(hi[1] - lo[1]), ez = scaleExponent(hi[2] - lo[2]);
    float sx = __uint_as_float(ex << 23), sy = __uint_as_float(ey << 23), sz = __uint_as_float(ez << 23);
    uint32_t meta[8], q[6][8];
    int triOff = 0, innerRank = 0;
    for (int s = 0; s < 8; ++s) {
        int i = childAt[s];
        meta[s] = 0;
        for (int k = 0; k < 6; ++k) q[k][s] = 0;
        if (i < 0) continue;
        int node = ch[i];
        float4 a = b2.lo[node], c = b2.hi[node];
        // outward rounding, then verified against the dequantised value
        float clo[3] = {a.x, a.y, a.z}, chi[3] = {c.x, c.y, c.z};
        float sc[3] = {sx, sy, sz};
        for (int k = 0; k < 3; ++k) {
            int ql = (int)floorf((clo[k] - lo[k]) / sc[k]);
            int qh = (int)ceilf((chi[k] - lo[k]) / sc[k]);
            ql = max(0, min(255, ql)); qh = max(0, min(255, qh));
            while (ql > 0 && lo[k] + (float)ql * sc[k] > clo[k]) --ql;
            while (qh < 255 && lo[k] + (float)qh * sc[k] < chi[k]) ++qh;
            q[k][s] = (uint32_t)ql; q[3 + k][s] = (uint32_t)qh;
        }
        if (imask & (1u << s)) {
            meta[s] = (1u << 5) | (24u + (uint32_t)s);
            tasksOut[taskBase + innerRank] = make_uint2((uint32_t)node, (uint32_t)(childBase + innerRank));
            ++innerRank;
        } else {
            int cnt = b2.triCount(node), first = b2.firstTri(node);
            uint32_t unary = cnt == 1 ? 1u : (cnt == 2 ? 3u : 7u);
            meta[s] = (unary << 5) | (uint32_t)triOff;
            for (int t = 0; t < cnt; ++t) {
                uint32_t id = sortedIdx[first + t];
                const NovumTriangle* tr = tris + id;
                int4 idx = __ldg(reinterpret_cast<const int4*>(tr));
                int matId = __ldg(&tr->materialID);
                float4 pa = __ldg(pos + idx.x), pb = __ldg(pos + idx.y), pc = __ldg(pos + idx.z);
                int flags = matId & 0xffff;
                if (mats && mats[matId].type == MAT_LEAF) flags |= TF_LEAFMAT;
                float4* o = trisOut + (size_t)(triBase + triOff + t) * 3;
                o[0] = make_float4(pa.x, pa.y, pa.z, __int_as_float((int)id));
                o[1] = make_float4(pb.x - pa.x, pb.y - pa.y, pb.z - pa.z, __int_as_float(flags));
                o[2] = make_float4(pc.x - pa.x, pc.y - pa.y, pc.z - pa.z, 0.f);
            }
            triOff += cnt;
        }
    }
    auto pack4 = [](const uint32_t* v) { return v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24); };
    float4* np = nodes + (size_t)task.y * 5;
    np[0] = make_float4(lo[0], lo[1], lo[2], __uint_as_float(ex | (ey << 8) | (ez << 16) | (imask << 24)));
    np[1] = make_float4(__uint_as_float((uint32_t)childBase), __uint_as_float((uint32_t)triBase),
                        __uint_as_float(pack4(meta)), __uint_as_float(pack4(meta + 4)));
    np[2] = make_float4(__uint_as_float(pack4(q[0])), __uint_as_float(pack4(q[0] + 4)),
                        __uint_as_float(pack4(q[1])), __uint_as_float(pack4(q[1] + 4)));
    np[3] = make_float4(__uint_as_float(pack4(q[2])), __uint_as_float(pack4(q[2] + 4)),
                        __uint_as_float(pack4(q[3])), __uint_as_float(pack4(q[3] + 4)));
    np[4] = make_float4(__uint_as_float(pack4(q[4])), __uint_as_float(pack4(q[4] + 4)),
                        __uint_as_float(pack4(q[5])), __uint_as_float(pack4(q[5] + 4)));
}

#define NV_CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) return e_; } while (0)

}  // namespace

void freeBvh8(Bvh8Device& b) {
    cudaFree(b.nodes);          // nodes and triangles share one allocation (tris points into it)
    b = Bvh8Device();
}

cudaError_t buildBvh8(const float4* d_positions, const NovumTriangle* d_tris, int n, const DMat* d_mats,
                      cudaStream_t stream, Bvh8Device& out) {
    if (n <= 0) return cudaErrorInvalidValue;
    // everything the build owns for its duration, released on every way out (an early error return used to leak all of it)
    struct Owned {
        cudaEvent_t ev0 = nullptr, ev1 = nullptr;
        char* arena = nullptr;
        RadixSortPlan plan; bool planMade = false;
        ~Owned() {
            if (ev0) cudaEventDestroy(ev0);
            if (ev1) cudaEventDestroy(ev1);
            if (planMade) radixSortPlanDestroy(plan);
            cudaFree(arena);
        }
    } own;
    NV_CK(cudaEventCreate(&own.ev0)); NV_CK(cudaEventCreate(&own.ev1));
    cudaEvent_t ev0 = own.ev0, ev1 = own.ev1;
    NV_CK(cudaEventRecord(ev0, stream));
    int launches = 0;
    float4 *triLo, *triHi, *nodeLo, *nodeHi;
    uint32_t *bounds, *keys, *vals;
    int *left, *right, *parent, *rfirst, *rlast, *flags, *counters;
    uint2 *tasksA, *tasksB;
    float4* nodesTmp; float4* trisOut;
    const int numB2 = 2 * n - 1;
    const int ni = n > 1 ? n - 1 : 1;
    // ONE allocation for the scratch arrays (18 separate cudaMalloc calls before): carved in 256 B steps
    size_t arenaBytes = 0;
    auto carve = [&](size_t bytes) { const size_t at = arenaBytes; arenaBytes += (bytes + 255) & ~(size_t)255; return at; };
    const size_t oTriLo = carve(sizeof(float4) * n), oTriHi = carve(sizeof(float4) * n);
    const size_t oNodeLo = carve(sizeof(float4) * numB2), oNodeHi = carve(sizeof(float4) * numB2);
    const size_t oBounds = carve(18 * sizeof(uint32_t));
    const size_t oKeys = carve(sizeof(uint32_t) * n), oVals = carve(sizeof(uint32_t) * n);
    const size_t oLeft = carve(sizeof(int) * ni), oRight = carve(sizeof(int) * ni), oRfirst = carve(sizeof(int) * ni), oRlast = carve(sizeof(int) * ni);
    const size_t oFlags = carve(sizeof(int) * ni), oParent = carve(sizeof(int) * numB2), oCounters = carve(4 * sizeof(int));
    const size_t oTasksA = carve(sizeof(uint2) * n), oTasksB = carve(sizeof(uint2) * n);
    const size_t oNodesTmp = carve(sizeof(float4) * 5 * (size_t)n), oTrisOut = carve(sizeof(float4) * 3 * (size_t)n);
    NV_CK(cudaMalloc((void**)&own.arena, arenaBytes));
    char* const A = own.arena;
    triLo = (float4*)(A + oTriLo); triHi = (float4*)(A + oTriHi); nodeLo = (float4*)(A + oNodeLo); nodeHi = (float4*)(A + oNodeHi);
    bounds = (uint32_t*)(A + oBounds); keys = (uint32_t*)(A + oKeys); vals = (uint32_t*)(A + oVals);
    left = (int*)(A + oLeft); right = (int*)(A + oRight); rfirst = (int*)(A + oRfirst); rlast = (int*)(A + oRlast);
    flags = (int*)(A + oFlags); parent = (int*)(A + oParent); counters = (int*)(A + oCounters);
    tasksA = (uint2*)(A + oTasksA); tasksB = (uint2*)(A + oTasksB);
    nodesTmp = (float4*)(A + oNodesTmp); trisOut = (float4*)(A + oTrisOut);
    NV_CK(cudaMemsetAsync(parent, 0xff, sizeof(int) * numB2, stream));
    NV_CK(cudaMemsetAsync(flags, 0, sizeof(int) * ni, stream));

    const int T = 256, G = (n + T - 1) / T;
    initBounds<<<1, 32, 0, stream>>>(bounds);
    triBoundsKernel<<<G, T, 0, stream>>>(d_positions, d_tris, n, triLo, triHi, bounds);
    mortonKernel<<<G, T, 0, stream>>>(triLo, triHi, n, bounds, keys, vals);
    launches += 3;
    NV_CK(radixSortPlanCreate(own.plan, n));
    own.planMade = true;
    NV_CK(radixSortPairs(own.plan, keys, vals, n, 30, stream, &launches));
    if (n > 1) {
        radixTreeKernel<<<(n - 1 + T - 1) / T, T, 0, stream>>>(keys, n, left, right, parent, rfirst, rlast);
        ++launches;
    }
    refitKernel<<<G, T, 0, stream>>>(triLo, triHi, vals, n, left, right, parent, nodeLo, nodeHi, flags);
    ++launches;

    Bvh2View b2{nodeLo, nodeHi, left, right, rfirst, rlast, n};
    int hostCounters[4] = {0, 1, 0, 0};   // node 0 = root is pre-allocated
    uint2 rootTask = make_uint2(0u, 0u);  // bvh2 node 0 is the root (for n == 1 it is the only leaf)
    NV_CK(cudaMemcpyAsync(tasksA, &rootTask, sizeof(uint2), cudaMemcpyHostToDevice, stream));
    NV_CK(cudaMemcpyAsync(counters, hostCounters, sizeof(hostCounters), cudaMemcpyHostToDevice, stream));
    int numTasks = 1, levels = 0;
    uint2 *tin = tasksA, *tout = tasksB;
    while (numTasks > 0) {
        collapseKernel<<<(numTasks + 127) / 128, 128, 0, stream>>>(b2, tin, numTasks, tout, counters, nodesTmp, trisOut, vals,
                                                                   d_positions, d_tris, d_mats);
        ++launches; ++levels;
        NV_CK(cudaMemcpyAsync(hostCounters, counters, sizeof(hostCounters), cudaMemcpyDeviceToHost, stream));
        NV_CK(cudaStreamSynchronize(stream));
        numTasks = hostCounters[0];
        int zero = 0;
        NV_CK(cudaMemcpyAsync(counters, &zero, sizeof(int), cudaMemcpyHostToDevice, stream));
        uint2* t = tin; tin = tout; tout = t;
        if (levels > 4096) return cudaErrorUnknown;
    }
    out.numNodes = hostCounters[1];
    out.numTris = n;
    if (hostCounters[2] != n) { /* every triangle must land in exactly one leaf slot */ return cudaErrorAssert; }
    // nodes and triangles in ONE allocation, so that a single L2 access-policy window can cover the whole structure
    {
        size_t nodeWords = ((size_t)5 * out.numNodes + 7) & ~(size_t)7;          // keep the triangle array 128 B aligned
        out.bytes = sizeof(float4) * (nodeWords + 3 * (size_t)n);
        NV_CK(cudaMalloc(&out.nodes, out.bytes));
        out.tris = out.nodes + nodeWords;
        NV_CK(cudaMemcpyAsync(out.nodes, nodesTmp, sizeof(float4) * 5 * (size_t)out.numNodes, cudaMemcpyDeviceToDevice, stream));
        NV_CK(cudaMemcpyAsync(out.tris, trisOut, sizeof(float4) * 3 * (size_t)n, cudaMemcpyDeviceToDevice, stream));
    }
    uint32_t hb[18];
    NV_CK(cudaMemcpyAsync(hb, bounds, sizeof(hb), cudaMemcpyDeviceToHost, stream));
    NV_CK(cudaEventRecord(ev1, stream));
    NV_CK(cudaStreamSynchronize(stream));
    for (int k = 0; k < 3; ++k) { out.boundsMin[k] = decF(hb[12 + k]) - 0.000001f; out.boundsMax[k] = decF(hb[15 + k]) + 0.000001f; }
    cudaEventElapsedTime(&out.buildMs, ev0, ev1);
    out.kernelLaunches = launches; out.levels = levels;
    return cudaGetLastError();
}

}  // namespace nv
