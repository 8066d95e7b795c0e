// bvh8_build.cu — GPU builder: triangle boxes -> 30-bit Morton codes -> radix sort -> Karras radix tree
// -> bottom-up refit -> greedy surface-area collapse into 8-wide nodes with octant-ordered child slots and
// outward-rounded 8-bit child boxes -> 48 B triangle records in node order.
#include "bvh8_build.h"
#include "radix_sort.h"
#include "bvh8.cuh"
#include "vecmath.cuh"
#include <float.h>
#include <vector>

namespace nv {
namespace {

// ---- ordered-uint encoding of floats for atomic min/max ---------------------------------------
__device__ __forceinline__ uint32_t encF(float f) {
    uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ __forceinline__ float decF(uint32_t e) {
    uint32_t b = (e & 0x80000000u) ? (e & 0x7fffffffu) : ~e;
#ifdef __CUDA_ARCH__
    return __uint_as_float(b);
#else
    float f; memcpy(&f, &b, 4); return f;
#endif
}

// bounds[0..2] = min xyz, [3..5] = max xyz of padded triangle boxes; [6..8],[9..11] centroid min/max;
// [12..14],[15..17] raw vertex min/max (the reference's scene box is this +-1e-6, main.cu:29-41, 532-535)
__global__ void initBounds(uint32_t* bounds) {
    int i = threadIdx.x;
    if (i < 18) bounds[i] = ((i / 3) & 1) ? 0u : 0xffffffffu;
}

__global__ void triBoundsKernel(const float4* __restrict__ pos, const NovumTriangle* __restrict__ tris, int n,
                                float4* __restrict__ triLo, float4* __restrict__ triHi, uint32_t* __restrict__ bounds) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float clo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, chi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float rlo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rhi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < n) {
        int4 idx = __ldg(reinterpret_cast<const int4*>(tris + i));
        float4 a = __ldg(pos + idx.x), b = __ldg(pos + idx.y), c = __ldg(pos + idx.z);
        float mn[3] = {fminf(fminf(a.x, b.x), c.x), fminf(fminf(a.y, b.y), c.y), fminf(fminf(a.z, b.z), c.z)};
        float mx[3] = {fmaxf(fmaxf(a.x, b.x), c.x), fmaxf(fmaxf(a.y, b.y), c.y), fmaxf(fmaxf(a.z, b.z), c.z)};
        for (int k = 0; k < 3; ++k) {
            // pad: the reference pads by 1e-6 (main.cu:34, 41); add a few ulps of the coordinate magnitude so the
            // quantised slab test stays conservative against float rounding in the traversal
            float mag = fmaxf(fabsf(mn[k]), fabsf(mx[k]));
            float pad = 1e-6f + mag * 4.8e-7f;
            lo[k] = mn[k] - pad; hi[k] = mx[k] + pad;
            float ce = (mn[k] + mx[k]) * 0.5f;
            clo[k] = ce; chi[k] = ce;
            rlo[k] = mn[k]; rhi[k] = mx[k];
        }
        triLo[i] = make_float4(lo[0], lo[1], lo[2], 0.f);
        triHi[i] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
    // warp reduce then one atomic per warp
    for (int k = 0; k < 3; ++k) {
        for (int o = 16; o > 0; o >>= 1) {
            lo[k] = fminf(lo[k], __shfl_xor_sync(0xffffffffu, lo[k], o));
            hi[k] = fmaxf(hi[k], __shfl_xor_sync(0xffffffffu, hi[k], o));
            clo[k] = fminf(clo[k], __shfl_xor_sync(0xffffffffu, clo[k], o));
            chi[k] = fmaxf(chi[k], __shfl_xor_sync(0xffffffffu, chi[k], o));
            rlo[k] = fminf(rlo[k], __shfl_xor_sync(0xffffffffu, rlo[k], o));
            rhi[k] = fmaxf(rhi[k], __shfl_xor_sync(0xffffffffu, rhi[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
        for (int k = 0; k < 3; ++k) {
            if (lo[k] != FLT_MAX) {
                atomicMin(&bounds[k], encF(lo[k]));
                atomicMax(&bounds[3 + k], encF(hi[k]));
                atomicMin(&bounds[6 + k], encF(clo[k]));
                atomicMax(&bounds[9 + k], encF(chi[k]));
                atomicMin(&bounds[12 + k], encF(rlo[k]));
                atomicMax(&bounds[15 + k], encF(rhi[k]));
            }
        }
    }
}

__device__ __forceinline__ uint32_t expandBits10(uint32_t v) {
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void mortonKernel(const float4* __restrict__ triLo, const float4* __restrict__ triHi, int n,
                             const uint32_t* __restrict__ bounds, uint32_t* __restrict__ keys, uint32_t* __restrict__ vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float4 lo = triLo[i], hi = triHi[i];
    float c[3] = {(lo.x + hi.x) * 0.5f, (lo.y + hi.y) * 0.5f, (lo.z + hi.z) * 0.5f};
    uint32_t q[3];
    for (int k = 0; k < 3; ++k) {
        float mn = decF(bounds[6 + k]), mx = decF(bounds[9 + k]);
        float ext = mx - mn;
        float t = ext > 0.f ? (c[k] - mn) / ext : 0.f;
        t = fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f);
        q[k] = (uint32_t)t;
    }
    keys[i] = (expandBits10(q[0]) << 2) | (expandBits10(q[1]) << 1) | expandBits10(q[2]);
    vals[i] = (uint32_t)i;
}

// ---- Karras 2012 radix tree -----------------------------------------------------------------------
// internal nodes 0..n-2, leaf j is node (n-1)+j.
__device__ __forceinline__ int deltaFn(const uint32_t* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    uint32_t a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clz(a ^ b);
}

__global__ void radixTreeKernel(const uint32_t* __restrict__ keys, int n, int* __restrict__ left, int* __restrict__ right,
                                int* __restrict__ parent, int* __restrict__ rangeFirst, int* __restrict__ rangeLast) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (deltaFn(keys, n, i, i + 1) - deltaFn(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = deltaFn(keys, n, i, i - d);
    int lmax = 2;
    while (deltaFn(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (deltaFn(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = deltaFn(keys, n, i, j);
    int s = 0;
    int t = l;
    do {
        t = (t + 1) >> 1;
        if (deltaFn(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int first = min(i, j), last = max(i, j);
    int lc = (first == gamma) ? (n - 1 + gamma) : gamma;
    int rc = (last == gamma + 1) ? (n - 1 + gamma + 1) : (gamma + 1);
    left[i] = lc; right[i] = rc;
    parent[lc] = i; parent[rc] = i;
    rangeFirst[i] = first; rangeLast[i] = last;
    if (i == 0) parent[0] = -1;
}

// bottom-up refit: node boxes for all 2n-1 BVH2 nodes
__global__ void refitKernel(const float4* __restrict__ triLo, const float4* __restrict__ triHi, const uint32_t* __restrict__ sortedIdx,
                            int n, const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                            float4* __restrict__ nodeLo, float4* __restrict__ nodeHi, int* __restrict__ flags) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    uint32_t t = sortedIdx[j];
    int node = n - 1 + j;
    nodeLo[node] = triLo[t]; nodeHi[node] = triHi[t];
    __threadfence();
    int p = parent[node];
    while (p >= 0) {
        int old = atomicAdd(&flags[p], 1);
        if (old == 0) return;               // first child to arrive: the sibling will finish this node
        __threadfence();
        int l = left[p], r = right[p];
        float4 a = __ldcg(nodeLo + l), b = __ldcg(nodeLo + r), c = __ldcg(nodeHi + l), d = __ldcg(nodeHi + r);
        nodeLo[p] = make_float4(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), 0.f);
        nodeHi[p] = make_float4(fmaxf(c.x, d.x), fmaxf(c.y, d.y), fmaxf(c.z, d.z), 0.f);
        __threadfence();
        p = parent[p];
    }
}

}  // namespace
}  // namespace nv

#include "bvh8_collapse.inl"
