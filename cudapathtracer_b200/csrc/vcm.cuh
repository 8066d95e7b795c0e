// vcm.cuh — data layout of the bidirectional engine (VCM, SPPM and BDPT share it).
//
// Replaces the reference's per-pixel megakernels doLightPass / doEyePass (deviceCode.cu:2240-2384, 2686-3065) and
// lightPathTracing / Li_bidirectional (:1290-1432, :1768-1815) with a depth-synchronous wavefront:
//   light pass : start -> [extend -> light step -> splat resolve] x lightDepth
//   exchange   : photon slabs of all ranks are all-gathered by the caller (one slab per rank, fixed capacity)
//   grid build : hash -> radix sort -> cell table -> gather of the 32 B photon records into cell order
//   eye pass   : start -> [extend -> eye step (implicit hit, NEE, connections, merging) -> connection resolve] x eyeDepth
// Every shadow ray (t=1 splats, NEE, vertex connections) is deferred into a queue and traced by an any-hit kernel.
//
// Light-vertex cache (what the reference keeps in VCMPathVertices, objects.cuh:958-991, 42 B/vertex in 12 arrays):
// three 16 B words per vertex, indexed [depth * nLocal + localPath], same quantisation as the reference —
// octahedral 2 x snorm16 normal / wo (util.cuh:342-391), fp16 beta, fp16 uv, packed info word:
//   A = { pos.x, pos.y, pos.z, oct(normal) }
//   B = { oct(wo), half2(beta.x, beta.y), half2(beta.z, 0), half2(uv) }
//   C = { info (bit0 delta, bit1 backface, bits 22..31 material), d_vc, d_vcm, 0 }
// Photon record (reference Photons, objects.cuh:1162-1179, 34 B in 10 arrays; its packedNormal is never read): 32 B
//   A = { pos.x, pos.y, pos.z, oct(wi) }     B = { half2(beta.x, beta.y), half2(beta.z, 0), d_vm, d_vcm }
// The grid build gathers the records into cell order in a merge-ready structure of arrays (PhotonSoA, bdReorderPhotons): wi decoded
// once to 3 x fp16, d_vcm pre-divided by eta_vcm.
// Photon slab of rank r: [32 B header: int count, int capacity, ...][capacity x 32 B records]; slabs of all ranks are
// contiguous, so one all-gather on the whole buffer exchanges an iteration's photons.
#pragma once
#include <cuda_fp16.h>
#include "scene_dev.cuh"
#include "trav_stream.cuh"
#include "camera.cuh"
#include "hashgrid.cuh"

namespace nv {

enum : int { BD_MODE_VCM = 0, BD_MODE_BDPT = 1 };
constexpr int kMaxBdDepth = 32;          // eye / light depth limit of the wavefront counters
constexpr int kSlabHeaderBytes = 32;
constexpr int kMaxPeers = 15;            // world <= 16

// per (pass, depth) counters; one memset per pass clears them all
// BC_MERGE counts the Lambert merge records (front of mergeQ), BC_MERGE_GEN the records of every other material (back of mergeQ)
enum : int { BC_QUEUE = 0, BC_SHADOW = 1, BC_FETCH_EXTEND = 2, BC_FETCH_STEP = 3, BC_FETCH_RESOLVE = 4, BC_MERGE = 5, BC_FETCH_MERGE = 6,
             BC_MERGE_GEN = 7, BC_FETCH_MERGE_GEN = 8, BC_STRIDE = 16 };

struct __align__(16) BdShadowRec { float4 o_maxT; float4 d_pix; float4 contrib_skip; };   // 48 B
// eye vertex waiting for its photon gather: {p, pixel} {n, material} {toPrev (local frame), d_vcm} {beta, d_vm} {uv}
struct __align__(16) BdMergeRec { float4 p_pix, n_mat, toPrev_dvcm, beta_dvm, uv; };      // 80 B

// where one eye step appends its merge records and what the merge kernels consume: Lambert vertices from the front
// (cnt[BC_MERGE]), every other material from the back (cnt[BC_MERGE_GEN]). Step 0 (primary hits) is merged at once; deeper
// steps share a queue per group of kMergeGroup steps. The Lambert records of a queue are put in photon-cell order (radix sort
// on the Morton code of the cell, bdMergeKeys) before the gather so that neighbouring warps read the same photon cells (L2 instead
// of HBM) and the 4 vertices a warp gathers together share one cell neighbourhood.
struct MergeQueue { BdMergeRec* recs; int cap; int* cnt; const int* order; };
constexpr int kMergeGroup = 4;
// merge-ready photons in cell order, written by bdReorderPhotons, as two dense 16 B streams so that a warp's candidate loads are
// two fully coalesced 512 B requests:  a = { pos.xyz, half2(wi.x, wi.y) }   b = { half2(wi.z, beta.x), half2(beta.y, beta.z), d_vm, d_vcm / eta_vcm }
struct PhotonSoA { const float4* a; const uint4* b; };

struct BdState {
    // per local path (shared by the light pass and the eye pass, which run one after the other)
    float4* rayO;      // xyz origin,    w = pdf (solid angle) of the last sampled direction
    float4* rayD;      // xyz direction, w = |cos| at the previous vertex
    float4* beta;      // xyz throughput, w = reverse pdf (solid angle) at the previous vertex
    float4* prevP;     // xyz previous vertex, w = bits: 1 = previous vertex was delta
    float4* dv;        // d_vcm, d_vc, d_vm of the previous vertex, w = spatial pdf of the light sample (light pass)
    unsigned long long* rng;
    float4* hit;       // bits tri, t, u, v
    int* queue[2];
    int* pathLen;      // surface vertices stored per light path
    uint4* medium;     // FEAT_NESTED_BIDIR only (else null): per path, the 15 x u8 medium stack of the walk + its height in the top byte
    uint4 *lvA, *lvB, *lvC;
    BdShadowRec* shadowQ;
    BdMergeRec* mergeQ;         // (1 + kMergeGroup) * nLocal records: [0, nLocal) step 0, the rest the current group of deeper steps
    int* mergeOrder;            // processing order of a group's Lambert records (kMergeGroup * nLocal)
    uint32_t* mergeKeys;        // sort keys of a queue's Lambert records (kMergeGroup * nLocal)
    int* counters;     // (kMaxBdDepth + 1) * BC_STRIDE
    unsigned long long* stats;   // [0] closest rays [1] shadow rays [2] photons stored [3] light vertices stored [4] merges accepted [5] merge candidates (photons in the 27 cells) [6] merge queries (eye vertices gathered)
    int nLocal;
    int shadowCap;
};

struct BdParams {
    NovumCamera cam;
    int w, h;
    int eyeDepth, lightDepth;
    int mode;                          // BD_MODE_*
    int lightTrace, nee, naive, connection, doMis, paintWeight, doMerge, doSPPM, sampleEnv;
    float mergeRadius, etaVcm;
    unsigned long long seed;
    int iter;
    int rank, world;                   // pixel rows y with y % world == rank belong to this rank
    int nLocal;
    float4* colors;                    // full-frame sum accumulator of this rank
    // photons
    uint4* slab;                       // this rank's slab (header + records), append target of the light pass
    int slabCap;
    // peer-write exchange (tile-sharded multi-GPU): this rank's slab inside every OTHER rank's slab buffer, mapped through
    // CUDA IPC over NVLink. The light pass stores each photon record locally and into all peers, so the all-gather is done
    // when the light pass is (no separate collective; only a stream-ordered barrier before the grid build).
    uint4* peerSlab[kMaxPeers];
    int numPeers;
    unsigned stamp;                    // written behind the count in the slab headers when the light pass is complete (lane engines wait on it)
    PhotonSoA photons;                 // records in cell order (eye pass), merge-ready
    const uint32_t* cellStart; const uint32_t* cellEnd;
    GridParams grid;
    unsigned long long gridMagic;      // floor((2^64 - 1) / grid.tableSize) + 1: the table slot of a cell without a division (fastMod)
};

NV_D int localToPixel(const BdParams& P, int l, int& x, int& y) {
    int row = l / P.w;
    x = l - row * P.w;
    y = P.rank + row * P.world;
    return y * P.w + x;
}

}  // namespace nv
