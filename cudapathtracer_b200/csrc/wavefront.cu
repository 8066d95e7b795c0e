// wavefront.cu — kernels and host loop of the wavefront unidirectional integrators (NEE+MIS and naive).
#include "wavefront.cuh"
#include "scene.h"
#include <cstdio>
#include <vector>

namespace nv {

constexpr int kWfThreads = 128;

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

// ---- advance: rotate queues, clear per-iteration counters (1 thread) --------------------------------------
__global__ void wfAdvance(int* counters) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        counters[C_EXTEND_COUNT] = counters[C_NEXT_COUNT];
        counters[C_NEXT_COUNT] = 0;
        counters[C_SHADOW_COUNT] = 0;
        for (int c = 0; c < NUM_CLS; ++c) counters[C_CLASS_COUNT + c] = 0;
        counters[C_FETCH_EXTEND] = 0; counters[C_FETCH_SHADE] = 0; counters[C_FETCH_SHADOW] = 0;
    }
}

// first iteration: every slot is handed to the regeneration class
__global__ void wfInit(WfState st, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) st.classQ[CLS_REGEN * st.capacity + i] = i;
    if (i == 0) {
        for (int c = 0; c < C_TOTAL; ++c) st.counters[c] = 0;
        st.counters[C_CLASS_COUNT + CLS_REGEN] = n;
    }
}

// ---- extend: closest hit for every queued path, then bin the path by the material class it hit ------------
template <bool COUNT>
__global__ void __launch_bounds__(kWfThreads) wfExtend(SceneView sc, WfState st, const int* __restrict__ queue, TravCounters nocnt) {
    __shared__ uint2 stackMem[kSmemStack * kWfThreads];
    TravStack ts; ts.init(stackMem + threadIdx.x, kWfThreads);
    const int lane = laneId();
    const int n = st.counters[C_EXTEND_COUNT];
    unsigned long long traced = 0;
    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&st.counters[C_FETCH_EXTEND], 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        int i = base + lane;
        bool valid = i < n;
        int slot = valid ? __ldg(queue + i) : 0;
        int cls = -1;
        if (valid) {
            float4 o = st.rayO[slot], d = st.rayD[slot];
            HitRec h = traceClosest<COUNT>(sc.bvh, V3(o.x, o.y, o.z), V3(d.x, d.y, d.z), 999999.0f, -1, ts, nocnt);
            ++traced;
            st.hit[slot] = make_float4(__int_as_float(h.tri), h.t, h.u, h.v);
            st.hitFlags[slot] = h.flags;
            cls = h.tri < 0 ? CLS_REGEN : classOfMaterial(sc.mats, h.flags & 0xffff);
        }
        // bin by class: one aggregated atomic per class present in the warp
#pragma unroll
        for (int c = 0; c < NUM_CLS; ++c) {
            int pos = warpAppend(&st.counters[C_CLASS_COUNT + c], cls == c);
            if (cls == c) st.classQ[c * st.capacity + pos] = slot;
        }
    }
    // one atomic per warp for the ray counter
    for (int o = 16; o > 0; o >>= 1) traced += __shfl_xor_sync(0xffffffffu, traced, o);
    if (lane == 0 && traced) atomicAdd(&st.rayCounters[0], traced);
}

// ---- shadow: any-hit for every queued NEE connection; unoccluded ones add their contribution --------------
__global__ void __launch_bounds__(kWfThreads) wfShadow(SceneView sc, WfState st, float4* __restrict__ colors) {
    __shared__ uint2 stackMem[kSmemStack * kWfThreads];
    TravStack ts; ts.init(stackMem + threadIdx.x, kWfThreads);
    const int lane = laneId();
    const int n = st.counters[C_SHADOW_COUNT];
    TravCounters nocnt{nullptr, nullptr};
    unsigned long long traced = 0;
    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&st.counters[C_FETCH_SHADOW], 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        int i = base + lane;
        if (i < n) {
            const float4* rp = reinterpret_cast<const float4*>(st.shadowQ + i);
            float4 a = rp[0], b = rp[1], c = rp[2];
            V3 thr = traceShadow<false>(sc, V3(a.x, a.y, a.z), V3(b.x, b.y, b.z), a.w, -1, ts, nocnt);
            ++traced;
            if (lengthSq(thr) > 0.0f) {     // deviceCode.cu:131
                int pix = __float_as_int(b.w);
                atomicAdd(&colors[pix], make_float4(c.x * thr.x, c.y * thr.y, c.z * thr.z, 0.f));
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) traced += __shfl_xor_sync(0xffffffffu, traced, o);
    if (lane == 0 && traced) atomicAdd(&st.rayCounters[1], traced);
}

}  // namespace nv

#include "wavefront_shade.inl"
#include "wavefront_host.inl"
