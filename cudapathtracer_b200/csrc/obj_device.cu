// obj_device.cu — the OBJ front end on the device (the step directly before the path; main.cu:936-1068 readObjSimple).
//
// The file's text goes to HBM as it is read (pinned double buffer, copy overlapped with the next pread) and is turned into the
// reference's arrays - positions, normals, uvs, 80 B Triangle records - by nine small kernels, one thread per text LINE:
//   objCountNewlines / objScanInts / objLineStarts   newline positions -> lineStart[] (16 B of text per thread, byte compares with
//                                                    __vcmpeq4, tile counts scanned by one block)
//   objClassify / objScanInts / objPositions         line types; `v` records parsed into positions[startIndex + running count]
//   objFaceCount / objScanSums / objEmit             per line: how many normals / uvs / triangles it contributes (a face's fan is
//                                                    checked against the positions read so far, exactly as the sequential loader
//                                                    does), block sums scanned, then every line writes its records at its offsets
// Per-line rules (number spelling, index lists, dropped triangles, generated normals, the shared zero uv) live in obj_parse.cuh and are
// the host loader's (frontend.cpp loadObj), bit for bit: tests/test_gpu_obj_device.py compares the arrays, tests/test_host.py holds
// the rules themselves to loadObj on the CPU. Spellings outside the rules' fast path (17-digit mantissas, inf/nan, polygons with more
// than 16 corners, files of 4 GiB) return NOVUM_E_UNSUPPORTED and the caller uses novum_obj_parse.
//
// Bound: bytes. One 33 MB dragon = 33 MB H2D + ~3 passes over the text + 120 MB of records written; none of it is arithmetic.
#include "obj_parse.cuh"
#include "scene.h"
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>
#include <chrono>
#include <atomic>
#include <mutex>
#include <thread>
#include <vector>

namespace nv {

constexpr int kObjTileThreads = 128;                      // 16 B of text per thread
constexpr int kObjTileBytes = kObjTileThreads * 16;       // 2 KiB per block
constexpr int kObjLineThreads = 256;                      // one line per thread
constexpr unsigned kObjNoLine = 0xFFFFFFFFu;

struct ObjCounters {
    unsigned numLines;       // lines of the file
    int nPos;                // `v` records
    int nNrm, nUv, nTri;     // records the file adds (nUv without the shared zero uv)
    unsigned firstZeroLine;  // first `f` line that keeps a triangle without uv references (kObjNoLine: none)
    int zeroUvIndex;         // uv records before that line (local to the file)
    int unsupported;         // a spelling outside obj_parse.cuh's rules was seen
};

// ---- small block-wide helpers ---------------------------------------------------------------------------------------------------------
template <int THREADS> __device__ __forceinline__ int blockExclusiveScan(int v, int* total) {
    __shared__ int warpSums[THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    __syncthreads();                        // a previous call's readers are done with warpSums
    if (lane == 31) warpSums[warp] = x;
    __syncthreads();
    int before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < THREADS / 32; ++w) { const int s = warpSums[w]; if (w < warp) before += s; all += s; }
    if (total) *total = all;
    return before + x - v;
}

// newline mask of 16 bytes of text (bit i = byte i is '\n'); a newline in the file's last byte starts no line
__device__ __forceinline__ unsigned newlineMask16(const uint4 w, uint32_t offset, uint32_t bytes) {
    const unsigned words[4] = {w.x, w.y, w.z, w.w};
    unsigned mask = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const unsigned eq = __vcmpeq4(words[k], 0x0a0a0a0au) & 0x01010101u;             // 1 in the low bit of every matching byte
        mask |= (((eq * 0x01020408u) >> 24) & 0xfu) << (4 * k);
    }
    if (offset + 16u >= bytes) {            // the 16 bytes reach the end of the file: drop padding and the last byte
        const unsigned valid = bytes > offset + 1u ? bytes - 1u - offset : 0u;          // bytes of this word that may start a line
        mask &= valid >= 16u ? 0xffffu : ((1u << valid) - 1u);
    }
    return mask;
}

__global__ void __launch_bounds__(kObjTileThreads) objCountNewlines(const uint4* __restrict__ text, uint32_t bytes, int* __restrict__ tileCount) {
    const uint32_t offset = (blockIdx.x * kObjTileThreads + threadIdx.x) * 16u;
    int c = offset < bytes ? __popc(newlineMask16(__ldg(text + offset / 16), offset, bytes)) : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    __shared__ int ws[kObjTileThreads / 32];
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = c;
    __syncthreads();
    if (threadIdx.x == 0) { int s = 0; for (int w = 0; w < kObjTileThreads / 32; ++w) s += ws[w]; tileCount[blockIdx.x] = s; }
}

// exclusive scan of n ints in place by ONE block (n is a few thousand: one entry per tile / per block of lines); *total = their sum + addend
__global__ void __launch_bounds__(1024) objScanInts(int* __restrict__ data, int n, int* __restrict__ total, int addend) {
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n ? data[i] : 0;
        int all;
        const int ex = blockExclusiveScan<1024>(v, &all);
        const int c = carry;
        if (i < n) data[i] = c + ex;
        __syncthreads();
        if (threadIdx.x == 0) carry = c + all;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry + addend;
}

// lineStart[0] = 0, lineStart[1 + k] = position behind the k-th newline, lineStart[numLines] = the end sentinel (see objLineEnd)
__global__ void __launch_bounds__(kObjTileThreads) objLineStarts(const uint4* __restrict__ text, uint32_t bytes, const int* __restrict__ tileBase,
                                                                 uint32_t* __restrict__ lineStart, ObjCounters* __restrict__ counters) {
    const uint32_t offset = (blockIdx.x * kObjTileThreads + threadIdx.x) * 16u;
    unsigned mask = offset < bytes ? newlineMask16(__ldg(text + offset / 16), offset, bytes) : 0u;
    const int ex = blockExclusiveScan<kObjTileThreads>(__popc(mask), nullptr);
    uint32_t slot = 1u + (uint32_t)tileBase[blockIdx.x] + (uint32_t)ex;
    while (mask) {
        const int b = __ffs(mask) - 1;
        mask &= mask - 1u;
        lineStart[slot++] = offset + (uint32_t)b + 1u;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        lineStart[0] = 0u;
        // the last line ends at the file's last byte if that is a newline, else at the end of the file: end(i) = lineStart[i + 1] - 1
        const bool endsWithNewline = reinterpret_cast<const char*>(text)[bytes - 1u] == '\n';
        lineStart[counters->numLines] = endsWithNewline ? bytes : bytes + 1u;
    }
}

// ---- one thread per line ----------------------------------------------------------------------------------------------------------------
struct ObjLines { const char* text; const uint32_t* lineStart; const ObjCounters* counters; };
__device__ __forceinline__ bool objLine(const ObjLines& L, uint32_t* begin, uint32_t* end) {
    const uint32_t i = blockIdx.x * kObjLineThreads + threadIdx.x;
    if (i >= L.counters->numLines) { *begin = *end = 0u; return false; }
    *begin = __ldg(L.lineStart + i);
    *end = __ldg(L.lineStart + i + 1) - 1u;
    return true;
}

// line types + how many `v` lines each block of lines holds
__global__ void __launch_bounds__(kObjLineThreads) objClassify(ObjLines L, unsigned char* __restrict__ types, int* __restrict__ blockPos) {
    uint32_t begin, end;
    const bool live = objLine(L, &begin, &end);
    const int type = live ? objLineType(L.text, begin, end) : OBJ_SKIP;
    if (live) types[blockIdx.x * kObjLineThreads + threadIdx.x] = (unsigned char)type;
    const int n = __syncthreads_count(type == OBJ_V);
    if (threadIdx.x == 0) blockPos[blockIdx.x] = n;
}

// `v` lines -> positions[startIndex + number of `v` lines before this one] (objSweepA)
__global__ void __launch_bounds__(kObjLineThreads) objPositions(ObjLines L, const unsigned char* __restrict__ types, const int* __restrict__ blockPos,
                                                               int startIndex, float yOffset, ObjVec4* __restrict__ positions, ObjCounters* __restrict__ counters) {
    uint32_t begin, end;
    const bool live = objLine(L, &begin, &end);
    const bool isV = live && types[blockIdx.x * kObjLineThreads + threadIdx.x] == OBJ_V;
    const int idx = blockPos[blockIdx.x] + blockExclusiveScan<kObjLineThreads>(isV ? 1 : 0, nullptr);
    if (!isV) return;
    double d[3] = {0.0, 0.0, 0.0};
    int unsupported = 0;
    objScanDoubles(L.text, begin + 2, end, d, 3, &unsupported);
    ObjVec4 q = {(float)d[0], (float)d[1], (float)d[2], 0.f};
    q.y = OBJ_ADD(q.y, yOffset);
    positions[startIndex + idx] = q;
    if (unsupported) counters->unsupported = 1;
}

// what every line adds: packed nrm | uv << 8 | tri << 16 per line, sums per block of lines; the first face line that needs the shared zero uv
__global__ void __launch_bounds__(kObjLineThreads) objFaceCount(ObjLines L, const unsigned char* __restrict__ types, const int* __restrict__ blockPos,
                                                               int startIndex, const ObjVec4* __restrict__ positions,
                                                               uint32_t* __restrict__ lineCounts, int* __restrict__ blockNrm, int* __restrict__ blockUv,
                                                               int* __restrict__ blockTri, ObjCounters* __restrict__ counters) {
    uint32_t begin, end;
    const bool live = objLine(L, &begin, &end);
    const uint32_t line = blockIdx.x * kObjLineThreads + threadIdx.x;
    const int type = live ? types[line] : OBJ_SKIP;
    // positions read before this line (M.pos.size() of the sequential loader)
    const int posSeen = startIndex + blockPos[blockIdx.x] + blockExclusiveScan<kObjLineThreads>(type == OBJ_V ? 1 : 0, nullptr);
    int nNrm = 0, nUv = 0, nTri = 0, unsupported = 0;
    if (type == OBJ_VT) nUv = 1;
    else if (type == OBJ_VN) nNrm = 1;
    else if (type == OBJ_F) {
        ObjFace f;
        objParseFace(L.text, begin + 2, end, f, &unsupported);
        const bool hasUV = f.nt == f.nv, hasN = f.nn == f.nv;
        for (int k = 1; k < f.nv - 1; ++k) {
            int idx[3]; float cp[3], areaSq;
            if (objFanTriangle(f, k, startIndex, posSeen, positions, idx, cp, &areaSq)) ++nTri;
        }
        if (!hasN) nNrm = nTri;
        if (!hasUV && nTri > 0) atomicMin(&counters->firstZeroLine, line);
    }
    if (live) lineCounts[line] = (uint32_t)nNrm | ((uint32_t)nUv << 8) | ((uint32_t)nTri << 16);
    if (unsupported) counters->unsupported = 1;
    int tN, tU, tT;
    blockExclusiveScan<kObjLineThreads>(nNrm, &tN);
    blockExclusiveScan<kObjLineThreads>(nUv, &tU);
    blockExclusiveScan<kObjLineThreads>(nTri, &tT);
    if (threadIdx.x == 0) { blockNrm[blockIdx.x] = tN; blockUv[blockIdx.x] = tU; blockTri[blockIdx.x] = tT; }
}

// after the three block-sum scans: where the shared zero uv goes = uv records before the first line that needs it
__global__ void __launch_bounds__(kObjLineThreads) objZeroUvIndex(const uint32_t* __restrict__ lineCounts, const int* __restrict__ blockUv,
                                                                 ObjCounters* __restrict__ counters) {
    const unsigned first = counters->firstZeroLine;
    if (first == kObjNoLine) return;                       // uniform
    const unsigned block = first / kObjLineThreads;
    const unsigned line = block * kObjLineThreads + threadIdx.x;
    const int nUv = line < counters->numLines ? (int)((lineCounts[line] >> 8) & 0xffu) : 0;
    const int ex = blockExclusiveScan<kObjLineThreads>(nUv, nullptr);
    if (line == first) counters->zeroUvIndex = blockUv[block] + ex;
}

struct ObjEmitArgs {
    int startIndex, nStart, uvStart, triStart;     // records of the meshes appended before this file
    int lightBase, isLight, materialID;
    float emission[3];
    ObjVec4* normals; NovumFloat2* uvs; NovumTriangle* tris;
};

// every line writes its records (objSweepB + the fix-up of loadObj): vt -> uv, vn -> normal, f -> triangles (+ a face normal per
// triangle when the face names no normals; + the file's ONE zero uv in front of the first face that names no uvs)
__global__ void __launch_bounds__(kObjLineThreads) objEmit(ObjLines L, const unsigned char* __restrict__ types, const int* __restrict__ blockPos,
                                                          const uint32_t* __restrict__ lineCounts, const int* __restrict__ blockNrm,
                                                          const int* __restrict__ blockUv, const int* __restrict__ blockTri,
                                                          const ObjVec4* __restrict__ positions, ObjEmitArgs A, ObjCounters* __restrict__ counters) {
    uint32_t begin, end;
    const bool live = objLine(L, &begin, &end);
    const uint32_t line = blockIdx.x * kObjLineThreads + threadIdx.x;
    const int type = live ? types[line] : OBJ_SKIP;
    const uint32_t packed = live ? lineCounts[line] : 0u;
    const int posSeen = A.startIndex + blockPos[blockIdx.x] + blockExclusiveScan<kObjLineThreads>(type == OBJ_V ? 1 : 0, nullptr);
    int nrmAt = blockNrm[blockIdx.x] + blockExclusiveScan<kObjLineThreads>((int)(packed & 0xffu), nullptr);
    int uvAt = blockUv[blockIdx.x] + blockExclusiveScan<kObjLineThreads>((int)((packed >> 8) & 0xffu), nullptr);
    int triAt = blockTri[blockIdx.x] + blockExclusiveScan<kObjLineThreads>((int)(packed >> 16), nullptr);
    const unsigned firstZero = counters->firstZeroLine;
    if (line > firstZero) ++uvAt;                                   // the shared zero uv sits in front of line firstZero's records
    const int sharedZeroUv = A.uvStart + counters->zeroUvIndex;
    int unsupported = 0;
    if (type == OBJ_VT) {
        double d[2] = {0.0, 0.0};
        objScanDoubles(L.text, begin + 3, end, d, 2, &unsupported);
        A.uvs[A.uvStart + uvAt] = {(float)d[0], OBJ_SUB(1.0f, (float)d[1])};
    } else if (type == OBJ_VN) {
        A.normals[A.nStart + nrmAt] = objNormalRecord(L.text, begin, end, &unsupported);
    } else if (type == OBJ_F) {
        if (line == firstZero) A.uvs[sharedZeroUv] = {0.f, 0.f};
        ObjFace f;
        objParseFace(L.text, begin + 2, end, f, &unsupported);
        const bool hasUV = f.nt == f.nv, hasN = f.nn == f.nv;
        for (int k = 1; k < f.nv - 1; ++k) {
            int idx[3]; float cp[3], areaSq;
            if (!objFanTriangle(f, k, A.startIndex, posSeen, positions, idx, cp, &areaSq)) continue;
            NovumTriangle t;
            t.aInd = idx[0]; t.bInd = idx[1]; t.cInd = idx[2];
            if (hasN) { t.naInd = f.ni[0] + A.nStart; t.nbInd = f.ni[k] + A.nStart; t.ncInd = f.ni[k + 1] + A.nStart; }
            else {
                const float il = 1.0f / sqrtf(areaSq);            // IEEE division and square root (nvcc defaults), as on the host
                A.normals[A.nStart + nrmAt] = {OBJ_MUL(cp[0], il), OBJ_MUL(cp[1], il), OBJ_MUL(cp[2], il), 0.f};
                t.naInd = t.nbInd = t.ncInd = A.nStart + nrmAt;
                ++nrmAt;
            }
            if (hasUV) { t.uvaInd = f.ti[0] + A.uvStart; t.uvbInd = f.ti[k] + A.uvStart; t.uvcInd = f.ti[k + 1] + A.uvStart; }
            else t.uvaInd = t.uvbInd = t.uvcInd = sharedZeroUv;
            t.materialID = A.materialID;
            t._pad0[0] = t._pad0[1] = 0;
            t.emission = {A.emission[0], A.emission[1], A.emission[2], 0.f};
            t.lightInd = A.isLight ? A.lightBase + triAt : -51;
            t.triInd = A.triStart + triAt;
            t._pad1[0] = t._pad1[1] = 0;
            // 80 B record as five 16 B stores
            const uint4* src = reinterpret_cast<const uint4*>(&t);
            uint4* dst = reinterpret_cast<uint4*>(A.tris + A.triStart + triAt);
#pragma unroll
            for (int w = 0; w < 5; ++w) dst[w] = src[w];
            ++triAt;
        }
    }
    if (unsupported) counters->unsupported = 1;
}

// novum_scene_create's index checks (scene.cu) for arrays that never visit the host: *bad = 1 when a triangle names a record that does not exist
__global__ void objValidateTris(const NovumTriangle* __restrict__ tris, int n, int nPos, int nNorm, int nUV, int nMat, int* __restrict__ bad) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const NovumTriangle& t = tris[i];
        const bool ok = t.aInd >= 0 && t.bInd >= 0 && t.cInd >= 0 && t.aInd < nPos && t.bInd < nPos && t.cInd < nPos &&
                        t.naInd >= 0 && t.nbInd >= 0 && t.ncInd >= 0 && t.naInd < nNorm && t.nbInd < nNorm && t.ncInd < nNorm &&
                        t.uvaInd >= 0 && t.uvbInd >= 0 && t.uvcInd >= 0 && t.uvaInd < nUV && t.uvbInd < nUV && t.uvcInd < nUV &&
                        t.materialID >= 0 && t.materialID < nMat;
        if (!ok) *bad = 1;
    }
}

// ---- host side --------------------------------------------------------------------------------------------------------------------------
struct ObjDeviceMesh {
    ObjVec4* pos = nullptr; ObjVec4* nrm = nullptr; NovumFloat2* uv = nullptr; NovumTriangle* tris = nullptr;
    size_t capPos = 0, capNrm = 0, capUv = 0, capTri = 0;
    int nPos = 0, nNrm = 0, nUv = 0, nTri = 0, numLights = 0;
    std::vector<std::pair<int, int>> lightRanges;          // {first triangle, count} of every emissive file, in order
    float uploadMs = 0.f, parseMs = 0.f;                   // wall time of reading + copying the text; CUDA-event time of the kernels
    int launches = 0;
    size_t textBytes = 0;
    char* scratch[3] = {nullptr, nullptr, nullptr}; size_t scratchCap[3] = {0, 0, 0};   // tile counts + counters | per-line arrays | text: kept across files
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
};

template <class T> static int objGrow(T*& p, size_t& cap, size_t have, size_t need, cudaStream_t s) {
    if (need <= cap) return 0;
    size_t want = cap ? std::max(need, cap + cap / 2) : need;
    T* q = nullptr;
    cudaError_t e = cudaMalloc((void**)&q, sizeof(T) * (want ? want : 1));
    if (e != cudaSuccess) return (int)e;
    if (have) {
        cudaMemcpyAsync(q, p, sizeof(T) * have, cudaMemcpyDeviceToDevice, s);
        if ((e = cudaStreamSynchronize(s)) != cudaSuccess) { cudaFree(q); return (int)e; }
    }
    cudaFree(p);
    p = q; cap = want;
    return 0;
}

// the kernels over text that is already in device memory (padded with NUL to a multiple of kObjTileBytes)
static int objParseText(ObjDeviceMesh& M, const char* d_text, uint32_t bytes, const float emission[3], int materialID, float yOffset) {
    cudaStream_t s = M.stream;
    cudaError_t e;
    const int nTiles = (int)((bytes + kObjTileBytes - 1) / kObjTileBytes);
    auto scratch = [&](int k, size_t bytes, void** out) -> int {        // the mesh keeps its two scratch buffers (grown by half when too small)
        if (bytes > M.scratchCap[k]) {
            cudaFree(M.scratch[k]); M.scratch[k] = nullptr; M.scratchCap[k] = 0;
            const size_t want = bytes + bytes / 2;
            cudaError_t a = cudaMalloc((void**)&M.scratch[k], want);
            if (a != cudaSuccess) return (int)a;
            M.scratchCap[k] = want;
        }
        *out = M.scratch[k];
        return 0;
    };
    int* d_tile = nullptr;             // [nTiles] ints, then the counters
    int rc = scratch(0, sizeof(int) * (size_t)nTiles + sizeof(ObjCounters) + 16, (void**)&d_tile);
    if (rc) return rc;
    ObjCounters* d_cnt = reinterpret_cast<ObjCounters*>(d_tile + ((nTiles + 3) & ~3));
    char* d_lines = nullptr;           // lineStart | lineCounts | block arrays | types
    auto fail = [&](int code) { cudaStreamSynchronize(s); return code; };
    ObjCounters h{};
    h.firstZeroLine = kObjNoLine;
    cudaMemcpyAsync(d_cnt, &h, sizeof(h), cudaMemcpyHostToDevice, s);
    cudaEventRecord(M.ev0, s);
    const uint4* text16 = reinterpret_cast<const uint4*>(d_text);
    objCountNewlines<<<nTiles, kObjTileThreads, 0, s>>>(text16, bytes, d_tile);
    objScanInts<<<1, 1024, 0, s>>>(d_tile, nTiles, reinterpret_cast<int*>(&d_cnt->numLines), 1);
    cudaMemcpyAsync(&h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail((int)e);
    const size_t numLines = h.numLines;
    const int nBlocks = (int)((numLines + kObjLineThreads - 1) / kObjLineThreads);
    const size_t offCounts = sizeof(uint32_t) * ((numLines + 1 + 3) & ~(size_t)3), offBlocks = offCounts + sizeof(uint32_t) * ((numLines + 3) & ~(size_t)3),
                 offTypes = offBlocks + sizeof(int) * 4 * (size_t)((nBlocks + 3) & ~3);
    if ((rc = scratch(1, offTypes + numLines + 16, (void**)&d_lines))) return fail(rc);
    uint32_t* d_lineStart = reinterpret_cast<uint32_t*>(d_lines);
    uint32_t* d_lineCounts = reinterpret_cast<uint32_t*>(d_lines + offCounts);
    int* d_blockPos = reinterpret_cast<int*>(d_lines + offBlocks);
    int* d_blockNrm = d_blockPos + ((nBlocks + 3) & ~3), *d_blockUv = d_blockNrm + ((nBlocks + 3) & ~3), *d_blockTri = d_blockUv + ((nBlocks + 3) & ~3);
    unsigned char* d_types = reinterpret_cast<unsigned char*>(d_lines + offTypes);
    objLineStarts<<<nTiles, kObjTileThreads, 0, s>>>(text16, bytes, d_tile, d_lineStart, d_cnt);
    ObjLines L{d_text, d_lineStart, d_cnt};
    objClassify<<<nBlocks, kObjLineThreads, 0, s>>>(L, d_types, d_blockPos);
    objScanInts<<<1, 1024, 0, s>>>(d_blockPos, nBlocks, &d_cnt->nPos, 0);
    cudaMemcpyAsync(&h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail((int)e);
    if ((long long)M.nPos + h.nPos > 0x7fffffffLL) return fail(NOVUM_E_UNSUPPORTED);
    if ((rc = objGrow(M.pos, M.capPos, (size_t)M.nPos, (size_t)M.nPos + (size_t)h.nPos, s))) return fail(rc);
    objPositions<<<nBlocks, kObjLineThreads, 0, s>>>(L, d_types, d_blockPos, M.nPos, yOffset, M.pos, d_cnt);
    objFaceCount<<<nBlocks, kObjLineThreads, 0, s>>>(L, d_types, d_blockPos, M.nPos, M.pos, d_lineCounts, d_blockNrm, d_blockUv, d_blockTri, d_cnt);
    objScanInts<<<1, 1024, 0, s>>>(d_blockNrm, nBlocks, &d_cnt->nNrm, 0);
    objScanInts<<<1, 1024, 0, s>>>(d_blockUv, nBlocks, &d_cnt->nUv, 0);
    objScanInts<<<1, 1024, 0, s>>>(d_blockTri, nBlocks, &d_cnt->nTri, 0);
    objZeroUvIndex<<<1, kObjLineThreads, 0, s>>>(d_lineCounts, d_blockUv, d_cnt);
    cudaMemcpyAsync(&h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail((int)e);
    if (h.unsupported) return fail(NOVUM_E_UNSUPPORTED);
    const int addUv = h.nUv + (h.firstZeroLine != kObjNoLine ? 1 : 0);
    if ((long long)M.nNrm + h.nNrm > 0x7fffffffLL || (long long)M.nUv + addUv > 0x7fffffffLL || (long long)M.nTri + h.nTri > 0x7fffffffLL)
        return fail(NOVUM_E_UNSUPPORTED);
    if ((rc = objGrow(M.nrm, M.capNrm, (size_t)M.nNrm, (size_t)M.nNrm + (size_t)h.nNrm, s))) return fail(rc);
    if ((rc = objGrow(M.uv, M.capUv, (size_t)M.nUv, (size_t)M.nUv + (size_t)addUv, s))) return fail(rc);
    if ((rc = objGrow(M.tris, M.capTri, (size_t)M.nTri, (size_t)M.nTri + (size_t)h.nTri, s))) return fail(rc);
    const bool isLight = emission[0] * emission[0] + emission[1] * emission[1] + emission[2] * emission[2] > 0.f;
    ObjEmitArgs A;
    A.startIndex = M.nPos; A.nStart = M.nNrm; A.uvStart = M.nUv; A.triStart = M.nTri;
    A.lightBase = M.numLights; A.isLight = isLight ? 1 : 0; A.materialID = materialID;
    for (int k = 0; k < 3; ++k) A.emission[k] = emission[k];
    A.normals = M.nrm; A.uvs = M.uv; A.tris = M.tris;
    objEmit<<<nBlocks, kObjLineThreads, 0, s>>>(L, d_types, d_blockPos, d_lineCounts, d_blockNrm, d_blockUv, d_blockTri, M.pos, A, d_cnt);
    cudaEventRecord(M.ev1, s);
    cudaMemcpyAsync(&h, d_cnt, sizeof(h), cudaMemcpyDeviceToHost, s);
    if ((e = cudaStreamSynchronize(s)) != cudaSuccess) return fail((int)e);
    if ((e = cudaGetLastError()) != cudaSuccess) return fail((int)e);
    if (h.unsupported) return fail(NOVUM_E_UNSUPPORTED);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, M.ev0, M.ev1);
    M.parseMs += ms; M.launches += 12;
    if (isLight && h.nTri > 0) { M.lightRanges.push_back({M.nTri, h.nTri}); M.numLights += h.nTri; }
    M.nPos += h.nPos; M.nNrm += h.nNrm; M.nUv += addUv; M.nTri += h.nTri;
    return 0;
}

// device buffer for `bytes` of text, padded with NUL to whole tiles (the kernels read 16 B words and whole tiles); the mesh keeps it
static int objAllocText(ObjDeviceMesh& M, size_t bytes, char** d_text) {
    const size_t padded = (bytes + kObjTileBytes - 1) / kObjTileBytes * kObjTileBytes + 16;
    if (padded > M.scratchCap[2]) {
        cudaFree(M.scratch[2]); M.scratch[2] = nullptr; M.scratchCap[2] = 0;
        cudaError_t e = cudaMalloc((void**)&M.scratch[2], padded + padded / 8);
        if (e != cudaSuccess) return (int)e;
        M.scratchCap[2] = padded + padded / 8;
    }
    *d_text = M.scratch[2];
    cudaMemsetAsync(*d_text + bytes, 0, padded - bytes, M.stream);
    return 0;
}

// Pinned staging for the file's text: a few reader threads, each with two 2 MiB pinned buffers, pread the file piece by piece and
// queue the copy of a piece to the device as soon as it is read; a buffer is reused once its copy has finished (event). Allocated at
// the first call and kept for the life of the process (portable pinned memory: any device's copies may use it).
constexpr size_t kObjPiece = (size_t)2 << 20;
constexpr int kObjReaders = 4;
struct ObjStaging { char* buf[kObjReaders][2] = {}; bool ready = false; std::mutex mu; };
static ObjStaging g_objStaging;

static int objUploadFile(ObjDeviceMesh& M, const char* path, char** d_text, size_t* bytesOut) {
    const auto t0 = std::chrono::steady_clock::now();
    int fd = open(path, O_RDONLY);
    if (fd < 0) return NOVUM_E_IO;
    struct stat sb;
    if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { close(fd); return NOVUM_E_IO; }
    const size_t sz = (size_t)sb.st_size;
    *bytesOut = sz;
    if (sz == 0) { close(fd); *d_text = nullptr; return 0; }
    if (sz >= 0x7fffff00ull) { close(fd); return NOVUM_E_UNSUPPORTED; }      // 32-bit text offsets
    std::lock_guard<std::mutex> lock(g_objStaging.mu);                       // one upload at a time through the staging buffers
    if (!g_objStaging.ready) {
        for (int r = 0; r < kObjReaders; ++r)
            for (int k = 0; k < 2; ++k)
                if (!g_objStaging.buf[r][k]) {
                    // not NOVUM_E_IO: the caller takes that for "no such file" and tries the next path (a positive code sends it to the host loader)
                    const cudaError_t pe = cudaHostAlloc((void**)&g_objStaging.buf[r][k], kObjPiece, cudaHostAllocPortable);
                    if (pe != cudaSuccess) { g_objStaging.buf[r][k] = nullptr; close(fd); return (int)pe; }
                }
        g_objStaging.ready = true;
    }
    int rc = objAllocText(M, sz, d_text);
    if (rc) { close(fd); return rc; }
    const int nPieces = (int)((sz + kObjPiece - 1) / kObjPiece);
    const int readers = std::max(1, std::min(kObjReaders, std::min(nPieces, (int)std::thread::hardware_concurrency())));
    int device = 0;
    cudaGetDevice(&device);
    std::atomic<int> next(0);
    std::atomic<int> bad(0);
    char* const dst = *d_text;
    cudaStream_t stream = M.stream;
    auto work = [&](int r) {
        cudaSetDevice(device);
        cudaEvent_t done[2];
        cudaEventCreateWithFlags(&done[0], cudaEventDisableTiming); cudaEventCreateWithFlags(&done[1], cudaEventDisableTiming);
        bool used[2] = {false, false};
        int k = 0;
        for (int i; (i = next.fetch_add(1)) < nPieces && !bad.load();) {
            if (used[k]) cudaEventSynchronize(done[k]);
            size_t off = (size_t)i * kObjPiece;
            const size_t end = std::min(sz, off + kObjPiece), first = off;
            char* b = g_objStaging.buf[r][k];
            while (off < end) { const ssize_t got = pread(fd, b + (off - first), end - off, (off_t)off); if (got <= 0) { bad = 1; break; } off += (size_t)got; }
            if (bad.load()) break;
            if (cudaMemcpyAsync(dst + first, b, end - first, cudaMemcpyHostToDevice, stream) != cudaSuccess) { bad = 2; break; }
            cudaEventRecord(done[k], stream);
            used[k] = true;
            k ^= 1;
        }
        for (int q = 0; q < 2; ++q) { if (used[q]) cudaEventSynchronize(done[q]); cudaEventDestroy(done[q]); }
    };
    std::vector<std::thread> pool;
    for (int r = 1; r < readers; ++r) pool.emplace_back(work, r);
    work(0);
    for (std::thread& t : pool) t.join();
    close(fd);
    if (bad.load()) { cudaStreamSynchronize(M.stream); *d_text = nullptr; return bad.load() == 1 ? NOVUM_E_IO : (int)cudaGetLastError(); }
    M.uploadMs += std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    M.textBytes += sz;
    return 0;
}

static int objMeshInit(ObjDeviceMesh& M) {
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return NOVUM_E_NOGPU;
    cudaError_t e = cudaStreamCreateWithFlags(&M.stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) return (int)e;
    cudaEventCreate(&M.ev0); cudaEventCreate(&M.ev1);
    return 0;
}
static void objMeshRelease(ObjDeviceMesh& M) {
    cudaFree(M.pos); cudaFree(M.nrm); cudaFree(M.uv); cudaFree(M.tris); cudaFree(M.scratch[0]); cudaFree(M.scratch[1]); cudaFree(M.scratch[2]);
    for (int k = 0; k < 3; ++k) { M.scratch[k] = nullptr; M.scratchCap[k] = 0; }
    M.pos = M.nrm = nullptr; M.uv = nullptr; M.tris = nullptr;
    if (M.ev0) { cudaEventDestroy(M.ev0); cudaEventDestroy(M.ev1); M.ev0 = M.ev1 = nullptr; }
    if (M.stream) { cudaStreamDestroy(M.stream); M.stream = nullptr; }
}

int objAppendFile(ObjDeviceMesh& M, const char* path, const float emission[3], int materialID, float yOffset) {
    char* d_text = nullptr; size_t bytes = 0;
    int rc = objUploadFile(M, path, &d_text, &bytes);
    if (rc || bytes == 0) return rc;
    return objParseText(M, d_text, (uint32_t)bytes, emission, materialID, yOffset);
}

int objAppendText(ObjDeviceMesh& M, const char* text, size_t bytes, const float emission[3], int materialID, float yOffset) {
    if (bytes == 0) return 0;
    if (bytes >= 0x7fffff00ull) return NOVUM_E_UNSUPPORTED;
    char* d_text = nullptr;
    int rc = objAllocText(M, bytes, &d_text);
    if (rc) return rc;
    cudaError_t e = cudaMemcpyAsync(d_text, text, bytes, cudaMemcpyHostToDevice, M.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(M.stream);
    if (e != cudaSuccess) return (int)e;
    M.textBytes += bytes;
    return objParseText(M, d_text, (uint32_t)bytes, emission, materialID, yOffset);
}

// text that already lives in device memory: copied once more into a buffer padded to whole tiles (the caller's buffer may end anywhere)
int objAppendDeviceText(ObjDeviceMesh& M, const char* d_src, size_t bytes, const float emission[3], int materialID, float yOffset) {
    if (bytes == 0) return 0;
    if (bytes >= 0x7fffff00ull) return NOVUM_E_UNSUPPORTED;
    char* d_text = nullptr;
    int rc = objAllocText(M, bytes, &d_text);
    if (rc) return rc;
    cudaError_t e = cudaMemcpyAsync(d_text, d_src, bytes, cudaMemcpyDeviceToDevice, M.stream);
    if (e != cudaSuccess) { cudaStreamSynchronize(M.stream); return (int)e; }
    M.textBytes += bytes;
    return objParseText(M, d_text, (uint32_t)bytes, emission, materialID, yOffset);
}

}  // namespace nv

// ---- C ABI ------------------------------------------------------------------------------------------------------------------------------
struct NovumObjDeviceMesh { nv::ObjDeviceMesh M; };

extern "C" {

int novum_obj_device_create(NovumObjDeviceMesh** out) {
    if (!out) return NOVUM_E_ARG;
    NovumObjDeviceMesh* m = new NovumObjDeviceMesh();
    int rc = nv::objMeshInit(m->M);
    if (rc) { delete m; return rc; }
    *out = m;
    return 0;
}

void novum_obj_device_destroy(NovumObjDeviceMesh* m) {
    if (!m) return;
    nv::objMeshRelease(m->M);
    delete m;
}

int novum_obj_device_append_file(NovumObjDeviceMesh* m, const char* path, const float emission[3], int materialID, float yOffset) {
    if (!m || !path) return NOVUM_E_ARG;
    const float zero[3] = {0.f, 0.f, 0.f};
    return nv::objAppendFile(m->M, path, emission ? emission : zero, materialID, yOffset);
}

int novum_obj_device_append_text(NovumObjDeviceMesh* m, const char* text, size_t bytes, const float emission[3], int materialID, float yOffset) {
    if (!m || (!text && bytes)) return NOVUM_E_ARG;
    const float zero[3] = {0.f, 0.f, 0.f};
    return nv::objAppendText(m->M, text, bytes, emission ? emission : zero, materialID, yOffset);
}

int novum_obj_device_append_device_text(NovumObjDeviceMesh* m, const char* d_text, size_t bytes, const float emission[3], int materialID, float yOffset) {
    if (!m || (!d_text && bytes)) return NOVUM_E_ARG;
    const float zero[3] = {0.f, 0.f, 0.f};
    return nv::objAppendDeviceText(m->M, d_text, bytes, emission ? emission : zero, materialID, yOffset);
}

int novum_obj_device_reset(NovumObjDeviceMesh* m) {
    if (!m) return NOVUM_E_ARG;
    nv::ObjDeviceMesh& M = m->M;
    M.nPos = M.nNrm = M.nUv = M.nTri = M.numLights = 0;
    M.lightRanges.clear();
    M.uploadMs = M.parseMs = 0.f; M.launches = 0; M.textBytes = 0;
    return 0;
}

int novum_obj_device_counts(const NovumObjDeviceMesh* m, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights) {
    if (!m) return NOVUM_E_ARG;
    if (nPos) *nPos = m->M.nPos;
    if (nNorm) *nNorm = m->M.nNrm;
    if (nUV) *nUV = m->M.nUv;
    if (nTri) *nTri = m->M.nTri;
    if (nLights) *nLights = m->M.numLights;
    return 0;
}

int novum_obj_device_arrays(const NovumObjDeviceMesh* m, const NovumFloat4** d_positions, const NovumFloat4** d_normals,
                            const NovumFloat2** d_uvs, const NovumTriangle** d_tris) {
    if (!m) return NOVUM_E_ARG;
    if (d_positions) *d_positions = reinterpret_cast<const NovumFloat4*>(m->M.pos);
    if (d_normals) *d_normals = reinterpret_cast<const NovumFloat4*>(m->M.nrm);
    if (d_uvs) *d_uvs = m->M.uv;
    if (d_tris) *d_tris = m->M.tris;
    return 0;
}

int novum_obj_device_download(const NovumObjDeviceMesh* m, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs, NovumTriangle* tris) {
    if (!m) return NOVUM_E_ARG;
    const nv::ObjDeviceMesh& M = m->M;
    cudaError_t e = cudaSuccess;
    auto down = [&](void* dst, const void* src, size_t bytes) { if (e == cudaSuccess && dst && bytes) e = cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToHost); };
    down(positions, M.pos, sizeof(NovumFloat4) * (size_t)M.nPos);
    down(normals, M.nrm, sizeof(NovumFloat4) * (size_t)M.nNrm);
    down(uvs, M.uv, sizeof(NovumFloat2) * (size_t)M.nUv);
    down(tris, M.tris, sizeof(NovumTriangle) * (size_t)M.nTri);
    return (int)e;
}

int novum_obj_device_timings(const NovumObjDeviceMesh* m, float* uploadMs, float* kernelMs, int* launches, long long* textBytes) {
    if (!m) return NOVUM_E_ARG;
    if (uploadMs) *uploadMs = m->M.uploadMs;
    if (kernelMs) *kernelMs = m->M.parseMs;
    if (launches) *launches = m->M.launches;
    if (textBytes) *textBytes = (long long)m->M.textBytes;
    return 0;
}

int novum_scene_from_obj_device(NovumObjDeviceMesh* m, const NovumMaterial* mats, int nMat, const NovumFloat4* textures, int nTexels,
                                NovumScene** out) {
    if (!m || !mats || !out || nMat <= 0 || nTexels < 0) return NOVUM_E_ARG;
    nv::ObjDeviceMesh& M = m->M;
    if (M.nPos <= 0 || M.nTri <= 0) return NOVUM_E_ARG;
    cudaError_t e;
    {   // the index checks of novum_scene_create, on the device
        int* d_bad = nullptr; int bad = 0;
        if ((e = cudaMalloc((void**)&d_bad, sizeof(int))) != cudaSuccess) return (int)e;
        cudaMemsetAsync(d_bad, 0, sizeof(int), M.stream);
        nv::objValidateTris<<<nv::smCount() * 4, 256, 0, M.stream>>>(M.tris, M.nTri, M.nPos, M.nNrm, M.nUv, nMat, d_bad);
        cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, M.stream);
        e = cudaStreamSynchronize(M.stream);
        cudaFree(d_bad);
        if (e != cudaSuccess) return (int)e;
        if (bad) return NOVUM_E_ARG;
    }
    // the parser's text and scratch buffers go back to the driver before the BVH8 build asks for its own scratch
    for (int k = 0; k < 3; ++k) { cudaFree(M.scratch[k]); M.scratch[k] = nullptr; M.scratchCap[k] = 0; }
    NovumScene* s = new NovumScene();
    s->nPos = M.nPos; s->nNorm = M.nNrm; s->nUV = M.nUv; s->nTri = M.nTri; s->nMat = nMat; s->nTexels = nTexels; s->nLights = M.numLights;
    // the scene takes the mesh's arrays over (the mesh is empty afterwards); arrays a file never filled get the 16 B the upload path gives them
    s->d_positions = reinterpret_cast<float4*>(M.pos); s->d_normals = reinterpret_cast<float4*>(M.nrm);
    s->d_uvs = reinterpret_cast<float2*>(M.uv); s->d_tris = M.tris;
    M.pos = M.nrm = nullptr; M.uv = nullptr; M.tris = nullptr;
    M.capPos = M.capNrm = M.capUv = M.capTri = 0;
    M.nPos = M.nNrm = M.nUv = M.nTri = 0;
    e = cudaSuccess;
    if (!s->d_normals) e = cudaMalloc((void**)&s->d_normals, 16);
    if (e == cudaSuccess && !s->d_uvs) e = cudaMalloc((void**)&s->d_uvs, 16);
    if (e == cudaSuccess) e = cudaMalloc((void**)&s->d_lights, sizeof(NovumTriangle) * (size_t)(s->nLights > 0 ? s->nLights : 1));
    size_t at = 0;
    for (const std::pair<int, int>& r : M.lightRanges) {            // main.cu:1018, 1060-1063: the emissive triangles, in order
        if (e != cudaSuccess) break;
        e = cudaMemcpyAsync(s->d_lights + at, s->d_tris + r.first, sizeof(NovumTriangle) * (size_t)r.second, cudaMemcpyDeviceToDevice, M.stream);
        at += (size_t)r.second;
    }
    M.lightRanges.clear(); M.numLights = 0;
    auto up = [&](void** dst, const void* src, size_t bytes) {
        if (e != cudaSuccess) return;
        e = cudaMalloc(dst, bytes ? bytes : 16);
        if (e == cudaSuccess && bytes) e = cudaMemcpy(*dst, src, bytes, cudaMemcpyHostToDevice);
    };
    up((void**)&s->d_matsRef, mats, sizeof(NovumMaterial) * (size_t)nMat);
    up((void**)&s->d_textures, textures, sizeof(float4) * (size_t)nTexels);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { novum_scene_destroy(s); return (int)e; }
    int rc = nv::sceneFinalize(s);
    if (rc) { novum_scene_destroy(s); return rc; }
    *out = s;
    return 0;
}

}  // extern "C"
