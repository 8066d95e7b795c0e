// radix_sort.h — hand-written LSD radix sort of (u32 key, u32 value) pairs on the device.
// Replaces the reference's cub::DeviceRadixSort::SortPairs call (deviceCode.cu:3172, 3298) and is
// also the sorter of the LBVH builder. Only the key bits that can be non-zero are sorted
// (the reference always sorts all 32 bits of a hash that is < tableSize).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace nv {

struct RadixSortPlan {
    int maxN = 0;
    int numTilesMax = 0;
    uint32_t* keysAlt = nullptr;    // ping-pong buffers
    uint32_t* valsAlt = nullptr;
    uint32_t* hist = nullptr;       // [256][numTiles] per pass
    uint32_t* blockSums = nullptr;  // scan scratch
    size_t bytes = 0;
};

constexpr int kSortTile = 4096;     // keys per block per pass

cudaError_t radixSortPlanCreate(RadixSortPlan& plan, int maxN);
void radixSortPlanDestroy(RadixSortPlan& plan);
// Bytes of scratch a plan for maxN pairs needs, and a plan laid over caller-owned device scratch of at least that many bytes
// (256-byte aligned). Nothing is allocated; such a plan must NOT be passed to radixSortPlanDestroy.
size_t radixSortScratchBytes(int maxN);
cudaError_t radixSortPlanBind(RadixSortPlan& plan, void* d_scratch, size_t scratchBytes, int maxN);

// Sorts keys[0..n) ascending on bits [0, numBits) (rounded up to whole 8-bit digits), stable.
// On return the sorted data is in `keys`/`vals` (the plan's alternates are scratch).
// Returns the number of kernels launched through *launches (may be null).
cudaError_t radixSortPairs(RadixSortPlan& plan, uint32_t* keys, uint32_t* vals, int n, int numBits,
                           cudaStream_t stream, int* launches);

// forces the module load of every kernel of this file (see nv::bdPreloadKernels in vcm_host.inl)
cudaError_t radixSortPreload();

// The same with the pair count on the device: `n` is the capacity the launches are sized for, *d_n (clamped to [0, n]) the number of pairs
// actually sorted - no host round trip to learn it. d_n == nullptr sorts n pairs.
cudaError_t radixSortPairsDeviceCount(RadixSortPlan& plan, uint32_t* keys, uint32_t* vals, int n, const int* d_n, int numBits,
                                      cudaStream_t stream, int* launches);

// exclusive prefix sum over u32, in place; scratch must hold >= ceil(n/2048)+1 entries
cudaError_t exclusiveScanU32(uint32_t* data, int n, uint32_t* scratch, cudaStream_t stream, int* launches);

}  // namespace nv
