// radix_sort.cu — LSD radix sort, 8-bit digits, three kernels per pass:
//   digitHistogram  : per-tile digit counts            (reads 4 B/key, 16 loads per thread in flight)
//   digitRowScan    : per-digit scan over the tiles    (tiny; the scan across digits happens inside every scatter CTA)
//   scatterPass     : stable warp-ranked scatter       (reads 8 B/key, writes 8 B/key)
// Algorithmic traffic: 20 B per key per pass. Ranking inside a tile is done per warp with __match_any_sync on the digit (no
// shared-memory atomics), warps are chained by a 256-thread column scan, so the sort is stable — required for LSD correctness.
// The scatter stages the tile's pairs in shared memory in digit order and writes each digit's run from consecutive threads.
#include "radix_sort.h"

namespace nv {

namespace {

constexpr int kThreads = 512;                          // scatter: 16 warps x 8 keys per thread (64 registers -> 2 CTAs = 32 warps per SM;
constexpr int kWarps = kThreads / 32;                  // the 256 x 16 form needed 106 registers and ran at 22 % occupancy, long-scoreboard bound)
constexpr int kItems = kSortTile / kThreads;           // 8 keys per thread
constexpr int kWarpSpan = kSortTile / kWarps;          // 256 contiguous keys per warp
constexpr int kHistThreads = 256;
constexpr int kHistItems = kSortTile / kHistThreads;   // 16 keys per thread

__global__ void __launch_bounds__(kHistThreads) digitHistogram(const uint32_t* __restrict__ keys, int n, int shift,
                                                               uint32_t* __restrict__ hist, int numTiles, const int* __restrict__ nDev) {
    __shared__ uint32_t cnt[256];
    if (nDev) n = min(max(__ldg(nDev), 0), n);              // the pair count lives on the device: the grid covers the capacity, tiles past the count report zeros
    cnt[threadIdx.x] = 0;
    const int base = blockIdx.x * kSortTile;
    uint32_t k[kHistItems];
#pragma unroll
    for (int i = 0; i < kHistItems; ++i) {                 // all loads in flight before the first shared-memory atomic
        const int idx = base + i * kHistThreads + threadIdx.x;
        k[i] = idx < n ? __ldg(keys + idx) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kHistItems; ++i)
        if (base + i * kHistThreads + threadIdx.x < n) atomicAdd(&cnt[(k[i] >> shift) & 0xffu], 1u);
    __syncthreads();
    hist[(size_t)threadIdx.x * numTiles + blockIdx.x] = cnt[threadIdx.x];
}

// One CTA per digit: exclusive scan of that digit's per-tile counts in place + the digit's total. Together with a 256-wide scan of
// the totals inside every scatter CTA this replaces the three-kernel scan of the whole [256][numTiles] array.
__global__ void __launch_bounds__(256) digitRowScan(uint32_t* __restrict__ hist, int numTiles, uint32_t* __restrict__ rowTotal) {
    __shared__ uint32_t warpSum[8];
    __shared__ uint32_t carry;
    uint32_t* row = hist + (size_t)blockIdx.x * numTiles;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < numTiles; base += 256) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < numTiles ? row[i] : 0u;
        uint32_t x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warpSum[warp] = x;
        __syncthreads();
        uint32_t before = carry;
#pragma unroll
        for (int w = 0; w < 8; ++w) before += w < warp ? warpSum[w] : 0u;
        if (i < numTiles) row[i] = before + x - v;
        __syncthreads();
        if (threadIdx.x == 255) carry = before + x;
        __syncthreads();
    }
    if (threadIdx.x == 0) rowTotal[blockIdx.x] = carry;
}

// Shared memory of one scatter CTA: per-warp digit counters, the tile's pairs in digit order (staged so that the global writes
// of a digit's run come from consecutive threads: written key by key from the ranking threads, every 4-byte store was its own
// L2 transaction — 21 M per pass at 10.5 M pairs, the first two passes took 170 / 136 us against 51 us for already-clustered
// input), and per digit the tile-local start and the offset from a tile-local to a global position.
struct ScatterSmem {
    uint32_t warpCnt[kWarps][256];
    uint32_t keys[kSortTile];
    uint32_t vals[kSortTile];
    uint32_t digitStart[256];
    uint32_t toGlobal[256];
    uint32_t scanWarp[16];
};

__global__ void __launch_bounds__(kThreads, 2) scatterPass(const uint32_t* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                           uint32_t* __restrict__ keysOut, uint32_t* __restrict__ valsOut, int n,
                                                           int shift, const uint32_t* __restrict__ histScanned, const uint32_t* __restrict__ rowTotal,
                                                           int numTiles, const int* __restrict__ nDev) {
    extern __shared__ __align__(16) unsigned char scatterSmemRaw[];
    if (nDev) n = min(max(__ldg(nDev), 0), n);
    if (blockIdx.x * kSortTile >= n) return;                 // (uniform per CTA)
    ScatterSmem& S = *reinterpret_cast<ScatterSmem*>(scatterSmemRaw);
    for (int i = threadIdx.x; i < kWarps * 256; i += kThreads) (&S.warpCnt[0][0])[i] = 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x;
    const int tileBase = tile * kSortTile;
    const int wbase = tileBase + warp * kWarpSpan;
    uint32_t key[kItems], val[kItems];
    uint32_t rank[kItems];
    const unsigned ltMask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kItems; ++r) {                     // keys and values of the whole tile requested up front
        const int idx = wbase + r * 32 + lane;
        key[r] = idx < n ? __ldg(keysIn + idx) : 0xffffffffu;
        val[r] = idx < n ? __ldg(valsIn + idx) : 0u;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) {                     // stable rank of every key among the warp's keys with the same digit
        const bool valid = wbase + r * 32 + lane < n;
        const unsigned digit = valid ? ((key[r] >> shift) & 0xffu) : 0x100u;
        const unsigned peers = __match_any_sync(0xffffffffu, digit);
        const int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) {
            prev = S.warpCnt[warp][digit];
            S.warpCnt[warp][digit] = prev + __popc(peers);
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[r] = prev + __popc(peers & ltMask);
        __syncwarp();
    }
    __syncthreads();
    if (threadIdx.x < 256) {   // thread d: chain the warps for digit d, then an exclusive scan over the digits of the tile
        const unsigned d = threadIdx.x;
        // global position of the tile's first key with digit d = all keys with smaller digits (scan of the digit totals, done here)
        //                                                       + this digit's keys in earlier tiles (digitRowScan)
        const uint32_t tot = __ldg(rowTotal + d);
        const uint32_t inRow = __ldg(histScanned + (size_t)d * numTiles + tile);
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const uint32_t c = S.warpCnt[w][d];
            S.warpCnt[w][d] = run;
            run += c;
        }
        uint32_t x = run, y = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t a = __shfl_up_sync(0xffffffffu, x, o);
            const uint32_t b = __shfl_up_sync(0xffffffffu, y, o);
            if (lane >= o) { x += a; y += b; }
        }
        if (lane == 31) { S.scanWarp[warp] = x; S.scanWarp[8 + warp] = y; }
        asm volatile("bar.sync 1, 256;");               // the 8 warps that hold digits
        uint32_t before = 0, smaller = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) {
            before += w < warp ? S.scanWarp[w] : 0u;
            smaller += w < warp ? S.scanWarp[8 + w] : 0u;
        }
        const uint32_t start = before + x - run;
        S.digitStart[d] = start;
        S.toGlobal[d] = smaller + (y - tot) + inRow - start;
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        if (wbase + r * 32 + lane < n) {
            const unsigned d = (key[r] >> shift) & 0xffu;
            const uint32_t lp = S.digitStart[d] + S.warpCnt[warp][d] + rank[r];
            S.keys[lp] = key[r];
            S.vals[lp] = val[r];
        }
    }
    __syncthreads();
    const int tileN = min(kSortTile, n - tileBase);
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        const int i = r * kThreads + threadIdx.x;
        if (i < tileN) {
            const uint32_t k = S.keys[i];
            const uint32_t pos = (uint32_t)i + S.toGlobal[(k >> shift) & 0xffu];
            keysOut[pos] = k;
            valsOut[pos] = S.vals[i];
        }
    }
}

__global__ void copyPairs(const uint32_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint32_t* __restrict__ kout, uint32_t* __restrict__ vout, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { kout[i] = kin[i]; vout[i] = vin[i]; }
}

// ---- exclusive scan, 2048 elements per block -------------------------------------------------
constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;

__device__ __forceinline__ uint32_t blockExclusiveScan(uint32_t v, uint32_t* total) {
    __shared__ uint32_t warpSums[kScanThreads / 32];
    __shared__ uint32_t blockTotal;
    int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) warpSums[warp] = x;
    __syncthreads();
    if (warp == 0) {
        uint32_t s = lane < kScanThreads / 32 ? warpSums[lane] : 0;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, s, o);
            if (lane >= o) s += y;
        }
        if (lane < kScanThreads / 32) warpSums[lane] = s;
        if (lane == 31) blockTotal = s;
    }
    __syncthreads();
    uint32_t prefix = (warp > 0 ? warpSums[warp - 1] : 0) + x - v;
    *total = blockTotal;
    __syncthreads();
    return prefix;
}

__global__ void __launch_bounds__(kScanThreads) scanReduce(const uint32_t* __restrict__ data, int n, uint32_t* __restrict__ sums) {
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) if (base + i < n) s += data[base + i];
    uint32_t total;
    blockExclusiveScan(s, &total);
    if (threadIdx.x == 0) sums[blockIdx.x] = total;
}

__global__ void __launch_bounds__(kScanThreads) scanSums(uint32_t* __restrict__ sums, int m) {
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < m; base += kScanThreads) {
        int i = base + threadIdx.x;
        uint32_t v = i < m ? sums[i] : 0;
        uint32_t total;
        uint32_t p = blockExclusiveScan(v, &total);
        uint32_t c = carry;
        if (i < m) sums[i] = c + p;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + total;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(kScanThreads) scanApply(uint32_t* __restrict__ data, int n, const uint32_t* __restrict__ sums) {
    int base = blockIdx.x * kScanTile + threadIdx.x * kScanItems;
    uint32_t v[kScanItems];
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) { v[i] = (base + i < n) ? data[base + i] : 0; s += v[i]; }
    uint32_t total;
    uint32_t p = blockExclusiveScan(s, &total) + sums[blockIdx.x];
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        if (base + i < n) data[base + i] = p;
        p += v[i];
    }
}

}  // namespace

cudaError_t radixSortPreload() {
    cudaFuncAttributes a;
    cudaError_t e;
    if ((e = cudaFuncGetAttributes(&a, digitHistogram)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, digitRowScan)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, scatterPass)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, copyPairs)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, scanReduce)) != cudaSuccess) return e;
    if ((e = cudaFuncGetAttributes(&a, scanSums)) != cudaSuccess) return e;
    return cudaFuncGetAttributes(&a, scanApply);
}

cudaError_t exclusiveScanU32(uint32_t* data, int n, uint32_t* scratch, cudaStream_t stream, int* launches) {
    if (n <= 0) return cudaSuccess;
    int blocks = (n + kScanTile - 1) / kScanTile;
    scanReduce<<<blocks, kScanThreads, 0, stream>>>(data, n, scratch);
    scanSums<<<1, kScanThreads, 0, stream>>>(scratch, blocks);
    scanApply<<<blocks, kScanThreads, 0, stream>>>(data, n, scratch);
    if (launches) *launches += 3;
    return cudaGetLastError();
}

cudaError_t radixSortPlanCreate(RadixSortPlan& plan, int maxN) {
    {   // 51 KB of dynamic shared memory per scatter CTA (once per device)
        cudaError_t ea = cudaFuncSetAttribute(scatterPass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScatterSmem));
        if (ea != cudaSuccess) return ea;
    }
    plan.maxN = maxN;
    plan.numTilesMax = (maxN + kSortTile - 1) / kSortTile;
    if (plan.numTilesMax < 1) plan.numTilesMax = 1;
    size_t histEntries = (size_t)256 * plan.numTilesMax;
    size_t scanBlocks = (histEntries + kScanTile - 1) / kScanTile + 1;
    if (scanBlocks < 256) scanBlocks = 256;              // digitRowScan keeps the 256 digit totals here
    cudaError_t e;
    size_t nb = (size_t)(maxN > 0 ? maxN : 1) * sizeof(uint32_t);
    if ((e = cudaMalloc(&plan.keysAlt, nb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.valsAlt, nb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.hist, histEntries * sizeof(uint32_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.blockSums, scanBlocks * sizeof(uint32_t))) != cudaSuccess) return e;
    plan.bytes = 2 * nb + (histEntries + scanBlocks) * sizeof(uint32_t);
    return cudaSuccess;
}

static void planLayout(int maxN, int& numTilesMax, size_t& nb, size_t& histBytes, size_t& sumsBytes) {
    numTilesMax = (maxN + kSortTile - 1) / kSortTile;
    if (numTilesMax < 1) numTilesMax = 1;
    const size_t histEntries = (size_t)256 * numTilesMax;
    size_t scanBlocks = (histEntries + kScanTile - 1) / kScanTile + 1;
    if (scanBlocks < 256) scanBlocks = 256;
    auto up = [](size_t b) { return (b + 255) & ~(size_t)255; };
    nb = up((size_t)(maxN > 0 ? maxN : 1) * sizeof(uint32_t));
    histBytes = up(histEntries * sizeof(uint32_t));
    sumsBytes = up(scanBlocks * sizeof(uint32_t));
}

size_t radixSortScratchBytes(int maxN) {
    int t; size_t nb, hb, sb;
    planLayout(maxN, t, nb, hb, sb);
    return 2 * nb + hb + sb;
}

cudaError_t radixSortPlanBind(RadixSortPlan& plan, void* d_scratch, size_t scratchBytes, int maxN) {
    if (!d_scratch || ((uintptr_t)d_scratch & 255) || scratchBytes < radixSortScratchBytes(maxN)) return cudaErrorInvalidValue;
    cudaError_t ea = cudaFuncSetAttribute(scatterPass, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ScatterSmem));
    if (ea != cudaSuccess) return ea;
    size_t nb, hb, sb;
    planLayout(maxN, plan.numTilesMax, nb, hb, sb);
    unsigned char* p = static_cast<unsigned char*>(d_scratch);
    plan.maxN = maxN;
    plan.keysAlt = reinterpret_cast<uint32_t*>(p);
    plan.valsAlt = reinterpret_cast<uint32_t*>(p + nb);
    plan.hist = reinterpret_cast<uint32_t*>(p + 2 * nb);
    plan.blockSums = reinterpret_cast<uint32_t*>(p + 2 * nb + hb);
    plan.bytes = 2 * nb + hb + sb;
    return cudaSuccess;
}

void radixSortPlanDestroy(RadixSortPlan& plan) {
    cudaFree(plan.keysAlt); cudaFree(plan.valsAlt); cudaFree(plan.hist); cudaFree(plan.blockSums);
    plan = RadixSortPlan();
}

cudaError_t radixSortPairs(RadixSortPlan& plan, uint32_t* keys, uint32_t* vals, int n, int numBits,
                           cudaStream_t stream, int* launches) {
    return radixSortPairsDeviceCount(plan, keys, vals, n, nullptr, numBits, stream, launches);
}

cudaError_t radixSortPairsDeviceCount(RadixSortPlan& plan, uint32_t* keys, uint32_t* vals, int n, const int* d_n, int numBits,
                                      cudaStream_t stream, int* launches) {
    if (n <= 1 && !d_n) return cudaSuccess;
    if (n < 1) return cudaSuccess;
    if (n > plan.maxN) return cudaErrorInvalidValue;
    int passes = (numBits + 7) / 8;
    if (passes < 1) passes = 1;
    if (passes > 4) passes = 4;
    int numTiles = (n + kSortTile - 1) / kSortTile;
    uint32_t *kin = keys, *vin = vals, *kout = plan.keysAlt, *vout = plan.valsAlt;
    for (int p = 0; p < passes; ++p) {
        int shift = p * 8;
        digitHistogram<<<numTiles, kHistThreads, 0, stream>>>(kin, n, shift, plan.hist, numTiles, d_n);
        if (launches) *launches += 1;
        digitRowScan<<<256, 256, 0, stream>>>(plan.hist, numTiles, plan.blockSums);
        scatterPass<<<numTiles, kThreads, sizeof(ScatterSmem), stream>>>(kin, vin, kout, vout, n, shift, plan.hist, plan.blockSums, numTiles, d_n);
        if (launches) *launches += 2;
        uint32_t* t;
        t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    if (kin != keys) {   // odd number of passes: result sits in the alternates (a kernel, not a copy-engine operation: see bdFillWords in vcm.cu)
        copyPairs<<<(n + 255) / 256, 256, 0, stream>>>(kin, vin, keys, vals, n);        // (with a device-side count: copies the capacity, harmless)
        if (launches) *launches += 1;
    }
    return cudaGetLastError();
}

}  // namespace nv
