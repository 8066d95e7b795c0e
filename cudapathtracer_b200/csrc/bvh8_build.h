// bvh8_build.h — GPU construction of the compressed 8-wide BVH (replaces the reference's recursive
// CPU binned-SAH BVH2 builder, main.cu:20-233, which the new core ignores at the boundary).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "bsdf.cuh"
#include "../../include/novum_b200.h"

namespace nv {

struct Bvh8Device {
    float4* nodes = nullptr;     // 5 float4 per node
    float4* tris = nullptr;      // 3 float4 per triangle (BVH order)
    size_t bytes = 0;            // size of the one allocation holding nodes, then triangles
    int numNodes = 0;
    int numTris = 0;             // == scene triangle count
    float boundsMin[3] = {0, 0, 0};   // root box of the padded triangle boxes (main.cu:29-41 pads by 1e-6 as well)
    float boundsMax[3] = {0, 0, 0};
    float buildMs = 0.f;
    int kernelLaunches = 0;
    int levels = 0;
};

// positions / tris / mats are DEVICE pointers. mats may be null (no MAT_LEAF flagging).
cudaError_t buildBvh8(const float4* d_positions, const NovumTriangle* d_tris, int numTris, const DMat* d_mats,
                      cudaStream_t stream, Bvh8Device& out);
void freeBvh8(Bvh8Device& b);

}  // namespace nv
