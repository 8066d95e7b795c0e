// hashgrid.cu — radix-sorted photon hash grid (replaces buildHashGrid and its five synchronised launches,
// deviceCode.cu:3067-3198). Per build: hash kernel -> LSD radix sort on the ceil(log2(tableSize)) live key
// bits -> table kernel. The 0xFF memset of both whole tables per iteration (deviceCode.cu:3183-3184, 1.3 GB at
// 4K) is replaced by clearing only the slots the previous build touched; cellEnd is only read where cellStart
// is set, so it is never cleared.
#include "hashgrid.cuh"
#include "../../include/novum_b200.h"
#include <cmath>
#include <mutex>
#include <vector>

namespace nv {

int nextPrime(int n) {      // util.cuh:393-413
    auto isPrime = [](int v) {
        if (v <= 1) return false;
        if (v == 2 || v == 3) return true;
        if (v % 2 == 0 || v % 3 == 0) return false;
        for (long long i = 5; i * i <= v; i += 6) if (v % i == 0 || v % (i + 2) == 0) return false;
        return true;
    };
    if (n <= 2) return 2;
    if (n % 2 == 0) n++;
    while (!isPrime(n)) n += 2;
    return n;
}

float mergeRadiusAt(float initialRadius, float alpha, int currSample) {   // util.cuh:415-418
    return initialRadius * sqrtf(1.0f / powf((float)(currSample + 1), alpha));
}

namespace {

// floor((p - min) / cell) exactly as gridIndex computes it (an IEEE division), at the cost of a multiply: the product with the
// rounded reciprocal is within 2 ulp of the quotient, so the two floors can differ only when the product lies within a few ulp
// of an integer — then, and only then, the division is done (the literal form made this kernel instruction-bound: 184 warp
// instructions per 32 photons, three divisions + a 32-bit modulo by a run-time prime).
__device__ __forceinline__ int cellCoord(float v, float cell, float rcell) {
    const float q = v * rcell;
    if (q == 0.0f) return v < 0.0f ? -1 : 0;      // photons ON the scene's lower faces (v = 0) are common: no division for them
    float f = floorf(q);
    if (fminf(q - f, (f + 1.0f) - q) <= fabsf(q) * 6e-7f) f = floorf(__fdiv_rn(v, cell));
    return (int)f;
}

__global__ void hashKernel(const float* __restrict__ px, const float* __restrict__ py, const float* __restrict__ pz, int stride,
                           int n, GridParams g, float rcell, unsigned long long magic, uint32_t* __restrict__ keys, uint32_t* __restrict__ index) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    size_t o = (size_t)i * stride;
    const int cx = cellCoord(__ldg(px + o) - g.minX, g.cell, rcell), cy = cellCoord(__ldg(py + o) - g.minY, g.cell, rcell),
              cz = cellCoord(__ldg(pz + o) - g.minZ, g.cell, rcell);
    const unsigned h = (73856093u * (unsigned)cx) ^ (19349663u * (unsigned)cy) ^ (83492791u * (unsigned)cz);      // integratorUtilities.cuh:405-412
    keys[i] = fastMod(h, g.tableSize, magic);
    index[i] = (uint32_t)i;
}

// four consecutive keys per thread (one 16 B load) + the key before them
__global__ void clearCellsKernel(const uint32_t* __restrict__ prevKeys, int n, uint32_t* __restrict__ cellStart) {
    const int i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n) return;
    uint32_t k[4];
    if (i0 + 4 <= n) { const uint4 v = __ldg(reinterpret_cast<const uint4*>(prevKeys + i0)); k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w; }
    else for (int j = 0; j < 4; ++j) k[j] = i0 + j < n ? __ldg(prevKeys + i0 + j) : 0u;
    uint32_t before = i0 > 0 ? __ldg(prevKeys + i0 - 1) : ~k[0];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (i0 + j < n && k[j] != before) cellStart[k[j]] = 0xFFFFFFFFu;
        before = k[j];
    }
}

__global__ void tableKernel(const uint32_t* __restrict__ keys, int n, uint32_t* __restrict__ cellStart, uint32_t* __restrict__ cellEnd,
                            uint32_t* __restrict__ prevKeys) {     // deviceCode.cu:3113-3138
    const int i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
    if (i0 >= n) return;
    uint32_t k[4];
    const bool full = i0 + 4 <= n;
    if (full) {
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(keys + i0));
        k[0] = v.x; k[1] = v.y; k[2] = v.z; k[3] = v.w;
        *reinterpret_cast<uint4*>(prevKeys + i0) = v;
    } else {
        for (int j = 0; j < 4; ++j) { k[j] = i0 + j < n ? __ldg(keys + i0 + j) : 0u; if (i0 + j < n) prevKeys[i0 + j] = k[j]; }
    }
    const int last = min(i0 + 4, n) - 1;                      // last valid index of this thread
    const bool haveBefore = i0 > 0, haveAfter = last + 1 < n;
    const uint32_t before = haveBefore ? __ldg(keys + i0 - 1) : 0u, after = haveAfter ? __ldg(keys + last + 1) : 0u;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int i = i0 + j;
        if (i > last) break;
        const bool first = j == 0 ? (!haveBefore || before != k[0]) : k[j - 1] != k[j];
        const bool lastOfRun = (i == last || j == 3) ? (!haveAfter || after != k[j]) : k[j + 1] != k[j];
        if (first) cellStart[k[j]] = (uint32_t)i;
        if (lastOfRun) cellEnd[k[j]] = (uint32_t)(i + 1);
    }
}

// 27-cell radius count exactly as doEyePass walks the grid (deviceCode.cu:2968-2995)
__global__ void queryCountKernel(const float* __restrict__ spx, const float* __restrict__ spy, const float* __restrict__ spz,
                                 const uint32_t* __restrict__ cellStart, const uint32_t* __restrict__ cellEnd, GridParams g,
                                 const float* __restrict__ qx, const float* __restrict__ qy, const float* __restrict__ qz, int nq,
                                 uint32_t* __restrict__ counts) {
    int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= nq) return;
    float x = qx[q], y = qy[q], z = qz[q];
    int3 c = gridIndex(x, y, z, g);
    float r2 = g.cell * g.cell;
    uint32_t cnt = 0;
    for (int dz = -1; dz <= 1; ++dz) for (int dy = -1; dy <= 1; ++dy) for (int dx = -1; dx <= 1; ++dx) {
        unsigned h = hashCell(make_int3(c.x + dx, c.y + dy, c.z + dz), g.tableSize);
        uint32_t s = cellStart[h];
        if (s == 0xFFFFFFFFu) continue;
        uint32_t e = cellEnd[h];
        for (uint32_t i = s; i < e; ++i) {
            float ddx = x - spx[i], ddy = y - spy[i], ddz = z - spz[i];
            float d2 = ddx * ddx + ddy * ddy + ddz * ddz;
            if (d2 <= r2) ++cnt;
        }
    }
    counts[q] = cnt;
}

__global__ void gatherKernel(const float* __restrict__ in, const uint32_t* __restrict__ index, int n, float* __restrict__ out) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = in[index[i]];
}

}  // namespace

int hashGridPreload() {
    cudaFuncAttributes a;
    cudaError_t e;
    if ((e = cudaFuncGetAttributes(&a, hashKernel)) != cudaSuccess) return (int)e;
    if ((e = cudaFuncGetAttributes(&a, clearCellsKernel)) != cudaSuccess) return (int)e;
    if ((e = cudaFuncGetAttributes(&a, tableKernel)) != cudaSuccess) return (int)e;
    return (int)radixSortPreload();
}

int hashGridAlloc(HashGridDevice& g, int maxPhotons) {
    g.maxPhotons = maxPhotons;
    g.tableSize = (unsigned)nextPrime(maxPhotons * 2);     // deviceCode.cu:3283
    size_t nb = sizeof(uint32_t) * (size_t)(maxPhotons > 0 ? maxPhotons : 1), tb = sizeof(uint32_t) * (size_t)g.tableSize;
    cudaError_t e;
    if ((e = cudaMalloc(&g.keys, nb)) != cudaSuccess) return (int)e;
    if ((e = cudaMalloc(&g.index, nb)) != cudaSuccess) return (int)e;
    if ((e = cudaMalloc(&g.prevKeys, nb)) != cudaSuccess) return (int)e;
    if ((e = cudaMalloc(&g.cellStart, tb)) != cudaSuccess) return (int)e;
    if ((e = cudaMalloc(&g.cellEnd, tb)) != cudaSuccess) return (int)e;
    if ((e = radixSortPlanCreate(g.plan, maxPhotons)) != cudaSuccess) return (int)e;
    g.prevN = -1;
    return 0;
}

void hashGridFree(HashGridDevice& g) {
    cudaFree(g.keys); cudaFree(g.index); cudaFree(g.prevKeys); cudaFree(g.cellStart); cudaFree(g.cellEnd);
    radixSortPlanDestroy(g.plan);
    g = HashGridDevice();
}

static int hashGridBuildStrided(HashGridDevice& g, const float* px, const float* py, const float* pz, int stride, int n,
                                const float sceneMin[3], float mergeRadius, cudaStream_t stream) {
    if (n < 0 || n > g.maxPhotons) return NOVUM_E_ARG;
    if (!(mergeRadius > 0.0f) || !std::isfinite(mergeRadius)) return NOVUM_E_ARG;      // a cell size of 0 / inf / NaN would put every photon in one key
    g.params.minX = sceneMin[0]; g.params.minY = sceneMin[1]; g.params.minZ = sceneMin[2];
    g.params.cell = mergeRadius; g.params.tableSize = g.tableSize;
    g.launches = 0;
    const int T = 256;
    if (g.prevN < 0) {
        cudaMemsetAsync(g.cellStart, 0xFF, sizeof(uint32_t) * (size_t)g.tableSize, stream);
    } else if (g.prevN > 0) {
        clearCellsKernel<<<(g.prevN + 4 * T - 1) / (4 * T), T, 0, stream>>>(g.prevKeys, g.prevN, g.cellStart);
        ++g.launches;
    }
    g.n = n; g.prevN = n;
    if (n == 0) return 0;
    hashKernel<<<(n + T - 1) / T, T, 0, stream>>>(px, py, pz, stride, n, g.params, 1.0f / mergeRadius, 0xFFFFFFFFFFFFFFFFull / g.tableSize + 1ull, g.keys,
                                                  g.index);
    ++g.launches;
    int bits = 1;
    while (bits < 32 && (1ull << bits) < (unsigned long long)g.tableSize) ++bits;
    cudaError_t e = radixSortPairs(g.plan, g.keys, g.index, n, bits, stream, &g.launches);
    if (e != cudaSuccess) return (int)e;
    tableKernel<<<(n + 4 * T - 1) / (4 * T), T, 0, stream>>>(g.keys, n, g.cellStart, g.cellEnd, g.prevKeys);
    ++g.launches;
    return (int)cudaGetLastError();
}

int hashGridBuild(HashGridDevice& g, const float* d_px, const float* d_py, const float* d_pz, int n, const float sceneMin[3],
                  float mergeRadius, cudaStream_t stream) {
    return hashGridBuildStrided(g, d_px, d_py, d_pz, 1, n, sceneMin, mergeRadius, stream);
}

int hashGridBuildAoS(HashGridDevice& g, const float* d_records, int strideFloats, int n, const float sceneMin[3], float mergeRadius,
                     cudaStream_t stream) {
    return hashGridBuildStrided(g, d_records, d_records + 1, d_records + 2, strideFloats, n, sceneMin, mergeRadius, stream);
}

}  // namespace nv

// ---- C ABI ---------------------------------------------------------------------------------------------------------
struct NovumHashGrid {
    nv::HashGridDevice g;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    float lastMs = 0.f;
    float *px = nullptr, *py = nullptr, *pz = nullptr;       // staging for the host-pointer API
    float *spx = nullptr, *spy = nullptr, *spz = nullptr;    // positions in sorted order (for queries)
};

extern "C" {

int novum_hashgrid_create(int maxPhotons, NovumHashGrid** out) {
    if (!out || maxPhotons <= 0) return NOVUM_E_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return NOVUM_E_NOGPU;
    NovumHashGrid* h = new NovumHashGrid();
    int rc = nv::hashGridAlloc(h->g, maxPhotons);
    if (rc) { novum_hashgrid_destroy(h); return rc; }
    cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    cudaEventCreate(&h->ev0); cudaEventCreate(&h->ev1);
    size_t nb = sizeof(float) * (size_t)maxPhotons;
    float** bufs[6] = {&h->px, &h->py, &h->pz, &h->spx, &h->spy, &h->spz};
    for (auto b : bufs) { cudaError_t e = cudaMalloc(b, nb); if (e != cudaSuccess) { novum_hashgrid_destroy(h); return (int)e; } }
    *out = h;
    return 0;
}

void novum_hashgrid_destroy(NovumHashGrid* h) {
    if (!h) return;
    nv::hashGridFree(h->g);
    cudaFree(h->px); cudaFree(h->py); cudaFree(h->pz); cudaFree(h->spx); cudaFree(h->spy); cudaFree(h->spz);
    if (h->ev0) { cudaEventDestroy(h->ev0); cudaEventDestroy(h->ev1); }
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
}

int novum_hashgrid_table_size(const NovumHashGrid* h) { return h ? (int)h->g.tableSize : NOVUM_E_ARG; }

int novum_hashgrid_build_device(NovumHashGrid* h, const float* d_px, const float* d_py, const float* d_pz, int n,
                                const float sceneMin[3], float mergeRadius, int sync) {
    if (!h || !sceneMin || n < 0 || (n > 0 && (!d_px || !d_py || !d_pz))) return NOVUM_E_ARG;
    cudaEventRecord(h->ev0, h->stream);
    int rc = nv::hashGridBuild(h->g, d_px, d_py, d_pz, n, sceneMin, mergeRadius, h->stream);
    cudaEventRecord(h->ev1, h->stream);
    if (rc) return rc;
    if (n > 0) {   // keep sorted positions for novum_hashgrid_query_count
        const int T = 256, G = (n + T - 1) / T;
        nv::gatherKernel<<<G, T, 0, h->stream>>>(d_px, h->g.index, n, h->spx);
        nv::gatherKernel<<<G, T, 0, h->stream>>>(d_py, h->g.index, n, h->spy);
        nv::gatherKernel<<<G, T, 0, h->stream>>>(d_pz, h->g.index, n, h->spz);
    }
    if (sync) {
        cudaError_t e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess) return (int)e;
        cudaEventElapsedTime(&h->lastMs, h->ev0, h->ev1);
    }
    return (int)cudaGetLastError();
}

int novum_hashgrid_build(NovumHashGrid* h, const float* px, const float* py, const float* pz, int n, const float sceneMin[3],
                         float mergeRadius, uint32_t* sortedIndex, uint32_t* sortedHash, uint32_t* cellStart, uint32_t* cellEnd) {
    if (!h || !sceneMin || n < 0 || n > h->g.maxPhotons || (n > 0 && (!px || !py || !pz))) return NOVUM_E_ARG;
    size_t nb = sizeof(float) * (size_t)n;
    cudaMemcpyAsync(h->px, px, nb, cudaMemcpyHostToDevice, h->stream);
    cudaMemcpyAsync(h->py, py, nb, cudaMemcpyHostToDevice, h->stream);
    cudaMemcpyAsync(h->pz, pz, nb, cudaMemcpyHostToDevice, h->stream);
    int rc = novum_hashgrid_build_device(h, h->px, h->py, h->pz, n, sceneMin, mergeRadius, 0);
    if (rc) return rc;
    size_t ib = sizeof(uint32_t) * (size_t)n, tb = sizeof(uint32_t) * (size_t)h->g.tableSize;
    if (sortedIndex) cudaMemcpyAsync(sortedIndex, h->g.index, ib, cudaMemcpyDeviceToHost, h->stream);
    if (sortedHash) cudaMemcpyAsync(sortedHash, h->g.keys, ib, cudaMemcpyDeviceToHost, h->stream);
    if (cellStart) cudaMemcpyAsync(cellStart, h->g.cellStart, tb, cudaMemcpyDeviceToHost, h->stream);
    if (cellEnd) cudaMemcpyAsync(cellEnd, h->g.cellEnd, tb, cudaMemcpyDeviceToHost, h->stream);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (e == cudaSuccess) cudaEventElapsedTime(&h->lastMs, h->ev0, h->ev1);
    return (int)e;
}

int novum_hashgrid_query_count(NovumHashGrid* h, const float* qx, const float* qy, const float* qz, int nq, uint32_t* counts) {
    if (!h || nq < 0 || (nq > 0 && (!qx || !qy || !qz || !counts))) return NOVUM_E_ARG;
    if (nq == 0) return 0;
    float* d = nullptr; uint32_t* dc = nullptr;
    size_t nb = sizeof(float) * (size_t)nq;
    cudaError_t e = cudaMalloc(&d, 3 * nb);
    if (e != cudaSuccess) return (int)e;
    if ((e = cudaMalloc(&dc, sizeof(uint32_t) * (size_t)nq)) != cudaSuccess) { cudaFree(d); return (int)e; }
    cudaMemcpyAsync(d, qx, nb, cudaMemcpyHostToDevice, h->stream);
    cudaMemcpyAsync(d + nq, qy, nb, cudaMemcpyHostToDevice, h->stream);
    cudaMemcpyAsync(d + 2 * (size_t)nq, qz, nb, cudaMemcpyHostToDevice, h->stream);
    nv::queryCountKernel<<<(nq + 127) / 128, 128, 0, h->stream>>>(h->spx, h->spy, h->spz, h->g.cellStart, h->g.cellEnd, h->g.params,
                                                                   d, d + nq, d + 2 * (size_t)nq, nq, dc);
    cudaMemcpyAsync(counts, dc, sizeof(uint32_t) * (size_t)nq, cudaMemcpyDeviceToHost, h->stream);
    e = cudaStreamSynchronize(h->stream);
    cudaFree(d); cudaFree(dc);
    return (int)e;
}

int novum_hashgrid_last_ms(NovumHashGrid* h, float* ms) {
    if (!h || !ms) return NOVUM_E_ARG;
    *ms = h->lastMs;
    return 0;
}

}  // extern "C"

// ---- allocation-free grid build -------------------------------------------------------------------------------------
// novum_hashgrid_build_static: the same hash -> radix sort -> cell table pipeline as novum_hashgrid_build_device for callers that
// may not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). No counterpart in the reference.
// Inputs and outputs are the caller's device buffers; the sort's ping-pong buffers,
// per-tile digit histograms and digit totals live in ONE file-scope device array that every GPU loading this library holds until
// exit (8.25 MiB), which bounds the size: n <= maxPhotons <= NOVUM_STATIC_GRID_MAX_PHOTONS (2^20, the SPPM configuration's 1 M
// photons per iteration). Every call shares the array and finds the previous call's contents in it. Nothing in it has to start
// at zero — the histogram rows, digit totals and alternates are written in full for the tiles a call uses before they are read —
// and the caller's cellStart table is reset to 0xFFFFFFFF inside the call. Two calls that overlap on the GPU (different streams,
// or different host threads on one device) race on the array: serialise them.
namespace {

constexpr int kStaticGridMaxPhotons = NOVUM_STATIC_GRID_MAX_PHOTONS;
constexpr size_t kStaticGridScratchBytes = 2 * sizeof(uint32_t) * (size_t)kStaticGridMaxPhotons      // keysAlt, valsAlt
                                           + sizeof(uint32_t) * 256 * (size_t)(kStaticGridMaxPhotons / nv::kSortTile)   // hist
                                           + sizeof(uint32_t) * 256;                                   // digit totals
static __device__ __align__(256) unsigned char g_staticGridScratch[kStaticGridScratchBytes];

constexpr int kMaxDevices = 64;
struct StaticGridSlot { bool ready = false; nv::RadixSortPlan plan; };
StaticGridSlot g_staticGridSlots[kMaxDevices];      // one per device: the symbol has a different address on every GPU
std::mutex g_staticGridMutex;

// In host code `g_staticGridScratch` is only a handle for the runtime, not a device address: the address valid on the CURRENT
// device is asked for once per device and kept.
int staticGridPlan(nv::RadixSortPlan** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return NOVUM_E_UNSUPPORTED;
    StaticGridSlot& s = g_staticGridSlots[dev];
    std::lock_guard<std::mutex> lock(g_staticGridMutex);
    if (!s.ready) {
        void* ptr = nullptr;
        if ((e = cudaGetSymbolAddress(&ptr, g_staticGridScratch)) != cudaSuccess) return (int)e;
        if ((e = nv::radixSortPlanBind(s.plan, ptr, kStaticGridScratchBytes, kStaticGridMaxPhotons)) != cudaSuccess) return (int)e;
        s.ready = true;
    }
    *out = &s.plan;
    return 0;
}

}  // namespace

extern "C" int novum_hashgrid_build_static(const float* d_px, const float* d_py, const float* d_pz, int n, int maxPhotons,
                                           const float sceneMin[3], float mergeRadius, uint32_t* d_sortedHash, uint32_t* d_sortedIndex,
                                           uint32_t* d_cellStart, uint32_t* d_cellEnd, void* stream) {
    if (!sceneMin || !d_cellStart || !d_cellEnd || n < 0 || maxPhotons <= 0 || n > maxPhotons) return NOVUM_E_ARG;
    if (maxPhotons > kStaticGridMaxPhotons) return NOVUM_E_TOOBIG;
    if (n > 0 && (!d_px || !d_py || !d_pz || !d_sortedHash || !d_sortedIndex)) return NOVUM_E_ARG;
    if (!(mergeRadius > 0.0f) || !std::isfinite(mergeRadius)) return NOVUM_E_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return NOVUM_E_NOGPU;
    nv::RadixSortPlan* plan = nullptr;
    int rc = staticGridPlan(&plan);
    if (rc) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned tableSize = (unsigned)nv::nextPrime(maxPhotons * 2);          // as novum_hashgrid_create (deviceCode.cu:3283)
    cudaError_t e = cudaMemsetAsync(d_cellStart, 0xFF, sizeof(uint32_t) * (size_t)tableSize, st);
    if (e != cudaSuccess) return (int)e;
    if (n == 0) return 0;
    nv::GridParams gp;
    gp.minX = sceneMin[0]; gp.minY = sceneMin[1]; gp.minZ = sceneMin[2]; gp.cell = mergeRadius; gp.tableSize = tableSize;
    const int T = 256;
    nv::hashKernel<<<(n + T - 1) / T, T, 0, st>>>(d_px, d_py, d_pz, 1, n, gp, 1.0f / mergeRadius, 0xFFFFFFFFFFFFFFFFull / tableSize + 1ull,
                                                  d_sortedHash, d_sortedIndex);
    int bits = 1;
    while (bits < 32 && (1ull << bits) < (unsigned long long)tableSize) ++bits;
    if ((e = nv::radixSortPairs(*plan, d_sortedHash, d_sortedIndex, n, bits, st, nullptr)) != cudaSuccess) return (int)e;
    // tableKernel also records the keys for the next build's selective clear; here that copy goes to the (now free) alternate
    nv::tableKernel<<<(n + 4 * T - 1) / (4 * T), T, 0, st>>>(d_sortedHash, n, d_cellStart, d_cellEnd, plan->keysAlt);
    return (int)cudaGetLastError();
}
