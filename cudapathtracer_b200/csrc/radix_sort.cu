// radix_sort.cu — LSD radix sort, 8-bit digits, three kernels per pass:
//   digitHistogram  : per-tile digit counts            (reads 4 B/key)
//   exclusiveScan   : digit-major scan of the counts   (tiny)
//   scatterPass     : stable warp-ranked scatter       (reads 8 B/key, writes 8 B/key)
// Algorithmic traffic: 20 B per key per pass. Ranking inside a tile is done per warp with
// __match_any_sync on the digit (no shared-memory atomics), warps are then chained by a
// 256-thread column scan, so the sort is stable — required for LSD correctness.
#include "radix_sort.h"

namespace nv {

namespace {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kItems = kSortTile / kThreads;          // 16 keys per thread
constexpr int kWarpSpan = kSortTile / kWarps;         // 512 contiguous keys per warp

__global__ void __launch_bounds__(kThreads) digitHistogram(const uint32_t* __restrict__ keys, int n, int shift,
                                                           uint32_t* __restrict__ hist, int numTiles) {
    __shared__ uint32_t cnt[256];
    cnt[threadIdx.x] = 0;
    __syncthreads();
    int tile = blockIdx.x;
    int base = tile * kSortTile;
#pragma unroll 4
    for (int i = 0; i < kItems; ++i) {
        int idx = base + i * kThreads + threadIdx.x;
        if (idx < n) atomicAdd(&cnt[(keys[idx] >> shift) & 0xffu], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * numTiles + tile] = cnt[threadIdx.x];
}

__global__ void __launch_bounds__(kThreads) scatterPass(const uint32_t* __restrict__ keysIn, const uint32_t* __restrict__ valsIn,
                                                        uint32_t* __restrict__ keysOut, uint32_t* __restrict__ valsOut, int n,
                                                        int shift, const uint32_t* __restrict__ histScanned, int numTiles) {
    __shared__ uint32_t warpCnt[kWarps][256];
    for (int i = threadIdx.x; i < kWarps * 256; i += kThreads) (&warpCnt[0][0])[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tile = blockIdx.x;
    const int wbase = tile * kSortTile + warp * kWarpSpan;
    uint32_t key[kItems];
    uint32_t rank[kItems];
    const unsigned ltMask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        int idx = wbase + r * 32 + lane;
        bool valid = idx < n;
        uint32_t k = valid ? keysIn[idx] : 0xffffffffu;
        key[r] = k;
        unsigned digit = valid ? ((k >> shift) & 0xffu) : 0x100u;
        unsigned peers = __match_any_sync(0xffffffffu, digit);
        int leader = __ffs(peers) - 1;
        uint32_t prev = 0;
        if (valid && lane == leader) {
            prev = warpCnt[warp][digit];
            warpCnt[warp][digit] = prev + __popc(peers);
        }
        prev = __shfl_sync(0xffffffffu, prev, leader);
        rank[r] = prev + __popc(peers & ltMask);
        __syncwarp();
    }
    __syncthreads();
    {   // thread d chains the warps for digit d and adds the global offset of (digit, tile)
        unsigned d = threadIdx.x;
        uint32_t base = histScanned[(size_t)d * numTiles + tile];
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            uint32_t c = warpCnt[w][d];
            warpCnt[w][d] = base;
            base += c;
        }
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < kItems; ++r) {
        int idx = wbase + r * 32 + lane;
        if (idx < n) {
            uint32_t k = key[r];
            uint32_t pos = warpCnt[warp][(k >> shift) & 0xffu] + rank[r];
            keysOut[pos] = k;
            valsOut[pos] = valsIn[idx];
        }
    }
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
    plan.maxN = maxN;
    plan.numTilesMax = (maxN + kSortTile - 1) / kSortTile;
    if (plan.numTilesMax < 1) plan.numTilesMax = 1;
    size_t histEntries = (size_t)256 * plan.numTilesMax;
    size_t scanBlocks = (histEntries + kScanTile - 1) / kScanTile + 1;
    cudaError_t e;
    size_t nb = (size_t)(maxN > 0 ? maxN : 1) * sizeof(uint32_t);
    if ((e = cudaMalloc(&plan.keysAlt, nb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.valsAlt, nb)) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.hist, histEntries * sizeof(uint32_t))) != cudaSuccess) return e;
    if ((e = cudaMalloc(&plan.blockSums, scanBlocks * sizeof(uint32_t))) != cudaSuccess) return e;
    plan.bytes = 2 * nb + (histEntries + scanBlocks) * sizeof(uint32_t);
    return cudaSuccess;
}

void radixSortPlanDestroy(RadixSortPlan& plan) {
    cudaFree(plan.keysAlt); cudaFree(plan.valsAlt); cudaFree(plan.hist); cudaFree(plan.blockSums);
    plan = RadixSortPlan();
}

cudaError_t radixSortPairs(RadixSortPlan& plan, uint32_t* keys, uint32_t* vals, int n, int numBits,
                           cudaStream_t stream, int* launches) {
    if (n <= 1) return cudaSuccess;
    if (n > plan.maxN) return cudaErrorInvalidValue;
    int passes = (numBits + 7) / 8;
    if (passes < 1) passes = 1;
    if (passes > 4) passes = 4;
    int numTiles = (n + kSortTile - 1) / kSortTile;
    uint32_t *kin = keys, *vin = vals, *kout = plan.keysAlt, *vout = plan.valsAlt;
    for (int p = 0; p < passes; ++p) {
        int shift = p * 8;
        digitHistogram<<<numTiles, kThreads, 0, stream>>>(kin, n, shift, plan.hist, numTiles);
        if (launches) *launches += 1;
        cudaError_t e = exclusiveScanU32(plan.hist, 256 * numTiles, plan.blockSums, stream, launches);
        if (e != cudaSuccess) return e;
        scatterPass<<<numTiles, kThreads, 0, stream>>>(kin, vin, kout, vout, n, shift, plan.hist, numTiles);
        if (launches) *launches += 1;
        uint32_t* t;
        t = kin; kin = kout; kout = t;
        t = vin; vin = vout; vout = t;
    }
    if (kin != keys) {   // odd number of passes: result sits in the alternates
        cudaMemcpyAsync(keys, kin, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream);
        cudaMemcpyAsync(vals, vin, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream);
    }
    return cudaGetLastError();
}

}  // namespace nv
