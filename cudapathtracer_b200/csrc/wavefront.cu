// wavefront.cu — kernels and host loop of the wavefront unidirectional integrators (NEE+MIS and naive).
#include "wavefront.cuh"
#include "scene.h"
#include <cstdio>
#include <vector>

namespace nv {

constexpr int kWfThreads = 128;
#ifndef NOVUM_TRAV_MINB
#define NOVUM_TRAV_MINB 8          // resident CTAs per SM the traversal kernels' register allocation is held to (64 registers)
#endif

__device__ __forceinline__ int laneId() { return threadIdx.x & 31; }

// warp-aggregated append: returns the slot index for lanes with pred set
__device__ __forceinline__ int warpAppend(int* counter, bool pred) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return -1;
    int leader = __ffs(m) - 1;
    int base = 0;
    if (laneId() == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return pred ? base + __popc(m & ((1u << laneId()) - 1u)) : -1;
}

__device__ __forceinline__ int classOfMaterial(const DMat* __restrict__ mats, int matId) {
    int type = __ldg(&mats[matId].type);
    switch (type) {
        case MAT_DIFFUSE: return CLS_DIFFUSE;
        case MAT_METAL: return CLS_METAL;
        case MAT_SMOOTHDIELECTRIC: return CLS_DIELECTRIC;
        case MAT_LEAF: return CLS_LEAF;
        default: return CLS_OTHER;
    }
}

// ---- advance: rotate queues, clear per-iteration counters; with ray bins also scan the bin counts of the queue about to be traced
constexpr int kAdvanceThreads = 512;
__global__ void __launch_bounds__(kAdvanceThreads) wfAdvance(WfState st) {
    int* counters = st.counters;
    if (threadIdx.x == 0) {
        counters[C_EXTEND_COUNT] = counters[C_NEXT_COUNT];
        counters[C_NEXT_COUNT] = 0;
        counters[C_SHADOW_COUNT] = 0;
        for (int c = 0; c < NUM_CLS; ++c) counters[C_CLASS_COUNT + c] = 0;
        counters[C_FETCH_EXTEND] = 0; counters[C_FETCH_SHADE] = 0; counters[C_FETCH_SHADOW] = 0;
    }
    if (kRayBins) {
        constexpr int kPer = kNumRayBins / kAdvanceThreads;
        __shared__ int warpSum[kAdvanceThreads / 32];
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        int c[kPer], mine = 0;
#pragma unroll
        for (int k = 0; k < kPer; ++k) { c[k] = st.binCount[threadIdx.x * kPer + k]; mine += c[k]; }
        int x = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warpSum[warp] = x;
        __syncthreads();
        int before = x - mine;
        for (int w = 0; w < warp; ++w) before += warpSum[w];
#pragma unroll
        for (int k = 0; k < kPer; ++k) { st.binStart[threadIdx.x * kPer + k] = before; before += c[k]; st.binCount[threadIdx.x * kPer + k] = 0; }
    }
}

// the continuing paths of the last shade pass, written bin by bin into the extend queue
__global__ void __launch_bounds__(256) wfBinScatter(WfState st, int* __restrict__ queue) {
    const int n = st.counters[C_EXTEND_COUNT];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int2 e = st.binTmp[i];
        queue[__ldg(st.binStart + e.y) + st.binPos[i]] = e.x;
    }
}

// first iteration: every slot is handed to the regeneration class
__global__ void wfInit(WfState st, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        st.classQ[CLS_REGEN * st.capacity + i] = i;
        // every slot starts in the regeneration class, which is also where a path that left the scene ends: a zero throughput says
        // "nothing escaped from here" when the environment feature looks at the slot (wfShade)
        st.slots[(size_t)i * kSlotWords + W_BETA] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    if (i == 0) {
        for (int c = 0; c < C_TOTAL; ++c) st.counters[c] = 0;
        st.counters[C_CLASS_COUNT + CLS_REGEN] = n;
    }
}

// ---- extend: closest hit for every queued path, then bin the path by the material class it hit ------------
// Rays stream through the warp's lanes with per-lane refill (trav_stream.cuh). Finished paths are staged in a per-warp
// shared buffer and binned 32 at a time with one ballot-aggregated atomic per class present, as before.
template <bool COUNT>
__global__ void __launch_bounds__(kWfThreads, NOVUM_TRAV_MINB) wfExtend(SceneView sc, WfState st, const int* __restrict__ queue, TravCounters cnt) {
    __shared__ uint2 stackMem[kSmemStack * kWfThreads];
    __shared__ int2 stage[kWfThreads / 32][64];
    __shared__ int stageCnt[kWfThreads / 32];
    TravStack ts; ts.init(stackMem + threadIdx.x, kWfThreads);
    const int lane = laneId(), warp = threadIdx.x >> 5;
    if (lane == 0) stageCnt[warp] = 0;
    __syncwarp();
    const int n = st.counters[C_EXTEND_COUNT];
    auto binStaged = [&](int first, int count) {           // whole warp converged
        int2 e = lane < count ? stage[warp][first + lane] : make_int2(0, -1);
#pragma unroll
        for (int c = 0; c < NUM_CLS; ++c) {
            int pos = warpAppend(&st.counters[C_CLASS_COUNT + c], e.y == c);
            if (e.y == c) st.classQ[c * st.capacity + pos] = e.x;
        }
    };
    unsigned traced = streamClosest<COUNT>(sc.bvh, ts, n, &st.counters[C_FETCH_EXTEND], cnt,
        [&](int i, V3& o, V3& d) {
            int slot = __ldg(queue + i);
            float4 ro = st.slots[(size_t)slot * kSlotWords + W_RAYO], rd = st.slots[(size_t)slot * kSlotWords + W_RAYD];
            o = V3(ro.x, ro.y, ro.z); d = V3(rd.x, rd.y, rd.z);
        },
        [&](int i, const HitRec& h) {
            int slot = __ldg(queue + i);
            st.slots[(size_t)slot * kSlotWords + W_HIT] = make_float4(__int_as_float(h.tri), h.t, h.u, h.v);
            st.slots[(size_t)slot * kSlotWords + W_AUX].x = __int_as_float(h.flags);
            int cls = h.tri < 0 ? CLS_REGEN : classOfMaterial(sc.mats, h.flags & 0xffff);
            stage[warp][atomicAdd(&stageCnt[warp], 1)] = make_int2(slot, cls);
        },
        [&](bool final) {
            __syncwarp();
            int c = stageCnt[warp];
            __syncwarp();
            if (c >= 32) {
                binStaged(0, 32);
                int2 carry = lane < c - 32 ? stage[warp][32 + lane] : make_int2(0, -1);
                __syncwarp();
                if (lane < c - 32) stage[warp][lane] = carry;
                c -= 32;
                if (lane == 0) stageCnt[warp] = c;
                __syncwarp();
            }
            if (final && c > 0) { binStaged(0, c); if (lane == 0) stageCnt[warp] = 0; __syncwarp(); }
        });
    // one atomic per warp for the ray counter
    unsigned long long tr64 = traced;
    for (int o = 16; o > 0; o >>= 1) tr64 += __shfl_xor_sync(0xffffffffu, tr64, o);
    if (lane == 0 && tr64) atomicAdd(&st.rayCounters[0], tr64);
}

// ---- shadow: any-hit for every queued NEE connection; unoccluded ones add their contribution --------------
__global__ void __launch_bounds__(kWfThreads, NOVUM_TRAV_MINB) wfShadow(SceneView sc, WfState st, float4* __restrict__ colors) {
    __shared__ uint2 stackMem[kSmemStack * kWfThreads];
    TravStack ts; ts.init(stackMem + threadIdx.x, kWfThreads);
    const int lane = laneId();
    const int n = st.counters[C_SHADOW_COUNT];
    TravCounters nocnt{nullptr, nullptr};
    unsigned traced = streamShadow<false>(sc, ts, n, &st.counters[C_FETCH_SHADOW], nocnt,
        [&](int i, V3& o, V3& d, float& maxT, int& skip) {
            const float4* rp = reinterpret_cast<const float4*>(st.shadowQ + i);
            float4 a = __ldg(rp), b = __ldg(rp + 1);
            o = V3(a.x, a.y, a.z); d = V3(b.x, b.y, b.z); maxT = a.w; skip = -1;
        },
        [&](int i, V3 thr) {
            if (lengthSq(thr) > 0.0f) {     // deviceCode.cu:131
                const float4* rp = reinterpret_cast<const float4*>(st.shadowQ + i);
                float4 b = __ldg(rp + 1), c = __ldg(rp + 2);
                atomicAdd(&colors[__float_as_int(b.w)], make_float4(c.x * thr.x, c.y * thr.y, c.z * thr.z, 0.f));
            }
        });
    unsigned long long tr64 = traced;
    for (int o = 16; o > 0; o >>= 1) tr64 += __shfl_xor_sync(0xffffffffu, tr64, o);
    if (lane == 0 && tr64) atomicAdd(&st.rayCounters[1], tr64);
}

}  // namespace nv

#include "wavefront_shade.inl"
#include "wavefront_host.inl"
