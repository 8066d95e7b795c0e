// hashgrid.cuh — photon / light-vertex hash grid: device functions shared by the grid builder and the
// merging kernels. Same cell function, primes and modulus as the reference
// (integratorUtilities.cuh:382-412), so a photon lands in the same table slot as there.
#pragma once
#include "vecmath.cuh"
#include "radix_sort.h"

namespace nv {

struct GridParams {
    float minX, minY, minZ;
    float cell;             // = merge radius
    unsigned tableSize;     // prime >= 2 * maxPhotons (util.cuh:404-413, deviceCode.cu:3283)
};

NV_HD int3 gridIndex(float px, float py, float pz, const GridParams& g) {      // :382-388
    return make_int3((int)floorf((px - g.minX) / g.cell), (int)floorf((py - g.minY) / g.cell), (int)floorf((pz - g.minZ) / g.cell));
}
NV_HD unsigned hashCell(int3 c, unsigned tableSize) {                           // :405-412
    unsigned n = (73856093u * (unsigned)c.x) ^ (19349663u * (unsigned)c.y) ^ (83492791u * (unsigned)c.z);
    return n % tableSize;
}

#ifdef __CUDACC__
// n % p for 32-bit n and a fixed p with magic = floor((2^64 - 1) / p) + 1 (Lemire, Kaser, Kurz 2019: exact for every 32-bit n)
__device__ __forceinline__ unsigned fastMod(unsigned n, unsigned p, unsigned long long magic) {
    return (unsigned)__umul64hi(magic * (unsigned long long)n, (unsigned long long)p);
}
__device__ __forceinline__ unsigned hashCellMagic(int cx, int cy, int cz, unsigned tableSize, unsigned long long magic) {    // = hashCell
    return fastMod((73856093u * (unsigned)cx) ^ (19349663u * (unsigned)cy) ^ (83492791u * (unsigned)cz), tableSize, magic);
}
#endif

struct HashGridDevice {
    int maxPhotons = 0;
    unsigned tableSize = 0;
    uint32_t* keys = nullptr;        // sorted hashes after build
    uint32_t* index = nullptr;       // sorted -> original photon index
    uint32_t* cellStart = nullptr;   // 0xFFFFFFFF = empty
    uint32_t* cellEnd = nullptr;
    uint32_t* prevKeys = nullptr;    // hashes of the previous build: only those slots need clearing
    int prevN = -1;                  // -1: tables never initialised
    int n = 0;
    GridParams params{};
    RadixSortPlan plan;
    int launches = 0;
};

int hashGridAlloc(HashGridDevice& g, int maxPhotons);
void hashGridFree(HashGridDevice& g);
// hashes + sorts + builds the cell tables for n photons given as SoA positions (device pointers)
int hashGridBuild(HashGridDevice& g, const float* d_px, const float* d_py, const float* d_pz, int n, const float sceneMin[3],
                  float mergeRadius, cudaStream_t stream);
int hashGridPreload();      // forces the module load of the grid build's kernels (and the radix sort's)
int nextPrime(int n);
float mergeRadiusAt(float initialRadius, float alpha, int currSample);

}  // namespace nv

namespace nv {
// positions interleaved in records: x at rec[0], y at rec[1], z at rec[2], `strideFloats` floats per record
int hashGridBuildAoS(HashGridDevice& g, const float* d_records, int strideFloats, int n, const float sceneMin[3], float mergeRadius,
                     cudaStream_t stream);
}
