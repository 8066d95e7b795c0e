// wavefront.cuh — state of the wavefront path-tracing engine that replaces the reference's one-thread-per-
// pixel megakernels (Li_unidirectional deviceCode.cu:285-542, Li_naive_unidirectional :158-205).
//
// A fixed pool of path slots lives in HBM as SoA. Each iteration runs
//   advance -> extend (closest hit) -> shade (material-class queues, NEE, BSDF sampling, regeneration) -> shadow
// on persistent warps that pull 32-entry packets from device-side queues; nothing goes back to the host
// except a counter read every few iterations. Terminated paths are regenerated in place from a global
// path counter, so the pool stays full until the last samples drain.
#pragma once
#include "scene_dev.cuh"
#include "trav_stream.cuh"
#include "camera.cuh"

namespace nv {

enum : int { CLS_REGEN = 0, CLS_DIFFUSE = 1, CLS_METAL = 2, CLS_DIELECTRIC = 3, CLS_LEAF = 4, CLS_OTHER = 5, NUM_CLS = 6 };

// device counters (ints), one cache line apart where contended
enum : int {
    C_EXTEND_COUNT = 0,      // entries in the current extend queue
    C_NEXT_COUNT = 1,        // entries appended to the next extend queue
    C_SHADOW_COUNT = 2,
    C_CLASS_COUNT = 4,       // .. +NUM_CLS
    C_FETCH_EXTEND = 16, C_FETCH_SHADE = 32, C_FETCH_SHADOW = 48,
    C_PATH_COUNTER_LO = 64,  // next path id to generate (64-bit, two ints)
    C_TOTAL = 96
};

// Ray binning of the extend queue (NOVUM_RAY_BINS): paths that continue are counted into bins keyed by
// (origin cell of a 2^b x 2^b x 2^b grid over the scene bounds, direction octant) while the shade kernel appends them; after
// the advance kernel's scan of the bin counts a scatter writes the slots bin by bin, so the 32 rays a warp of the extend
// kernel pulls together start in the same part of the scene and point into the same octant. Measured: no gain — a
// ray of these scenes visits ~4.6 BVH8 nodes and 3.8 triangles, there is little divergence to remove. Compiled out.
#ifndef NOVUM_RAY_BINS
#define NOVUM_RAY_BINS 0
#endif
constexpr bool kRayBins = NOVUM_RAY_BINS != 0;
constexpr int kRayCellBits = 3;
constexpr int kNumRayBins = 8 << (3 * kRayCellBits);      // 4096

struct ShadowRec { float4 o_maxT; float4 d_pix; float4 contrib; };   // 48 B

// One 128 B line per path slot (queues hand out slots in material-class order, i.e. at random: with one array per field a
// 16 B access used half of every 32 B sector it touched; a line per slot makes every sector fully used):
//   sector 0  W_RAYO  xyz origin, w = pdf of the last BSDF sample      W_RAYD  xyz direction, w = bits: depth (low 16) | flags (high 16)
//   sector 1  W_HIT   bits tri, t, u, v                                 W_AUX   x = triangle flags of the hit (material id in the low 16 bits), z,w = RNG state
//   sector 2  W_BETA  xyz throughput, w = bits: pixel index             W_PREVP xyz last true-hit point (neePDF needs it), w = bits: medium stack top
//   sector 3  W_MEDIUM 16 x u8 medium stack (deviceCode.cu:306 keeps 16 ints)       (spare)
// Extend reads sector 0 and writes W_HIT + the flags word of sector 1; shade reads the line and writes sectors 0-3.
enum : int { W_RAYO = 0, W_RAYD = 1, W_HIT = 2, W_AUX = 3, W_BETA = 4, W_PREVP = 5, W_MEDIUM = 6, kSlotWords = 8 };

struct WfState {
    float4* slots;       // capacity x kSlotWords
    // queues
    int* extendQ[2];
    int2* binTmp;        // capacity x {slot, bin}: continuing paths in append order (kRayBins)
    int* binPos;         // capacity: rank of the entry inside its bin
    int* binCount;       // kNumRayBins counters of the queue being appended
    int* binStart;       // kNumRayBins: exclusive scan of the counts of the queue being traced
    int* classQ;         // NUM_CLS x capacity
    ShadowRec* shadowQ;
    int* counters;       // C_TOTAL ints
    unsigned long long* rayCounters;   // [0] closest, [1] shadow, [2] paths finished, [3] paths started
    int capacity;
};

enum : unsigned { PF_NONSPEC_SEEN = 1u << 16 };   // hitFirstnonSpecular (deviceCode.cu:314)

struct WfParams {
    NovumCamera cam;
    int w, h;
    int maxDepth;
    int useMIS;
    int naive;
    unsigned long long seed;
    // sample passes rendered by this call: firstSample + rank + k*world, k in [0, localSamples)
    int firstSample, localSamples, rank, world;
    int tilesX, tilesY;                  // 8x4 pixel tiles
    unsigned long long totalIds;         // localSamples * tilesX*tilesY*32
    float4* colors;                      // w*h accumulator (sum over samples)
};

}  // namespace nv
