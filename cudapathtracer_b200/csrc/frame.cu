// frame.cu — the step directly after the path: accumulator -> displayable frame, on the device.
//
// Replaces, as ONE kernel + one small read-back:
//   * cleanAndFormatImage (integratorUtilities.cuh:337-380): NaN -> pink, Inf -> green, negative -> blue, normalise by
//     1/(currentSampleCount+1), overlay override — the reference runs it, then copies the float4 frame (16 B/pixel) to the
//     host every 5 s (deviceCode.cu:1905-1927, 3449-3463);
//   * the final normalisation of initRender (main.cu:855-884): divide by sampleCount, NaN -> pink, Inf -> green, overlay;
//   * Image::postProcessImage = toneMap (ACES fit, imageUtil.cu:202-211) + gammaCorrect (powf(c, 1/2.2), :213-222), which
//     the reference runs per pixel on ONE host core (push_back loop, :224-232);
//   * the pixel loop of Image::saveImageBMP (imageUtil.cu:82-96): clamp, *255 + 0.5, truncate, BGR order, rows padded to 4 B.
// Only the packed rows (3 B/pixel) go back to the host instead of two float4 frames (32 B/pixel).
//
// Bit-exactness with the reference's HOST arithmetic: the tonemap is evaluated with explicitly rounded, uncontracted
// mul/add/div in the reference's operation order (its host build has no FMA); the gamma curve + quantisation
// byte(t) = (uchar)(clamp(powf(t, 1/2.2f), 0, 1) * 255 + 0.5) is a monotone step function of the tonemapped float t, so the
// device looks the byte up in a table of the 255 step positions (smallest float t that reaches byte k, found once on the
// host by bisection over float bit patterns with the host's own powf) starting from a fast ex2/lg2 estimate.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <algorithm>
#include <utility>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "../../include/novum_b200.h"

namespace nv {

constexpr int FRAME_THREADS = 256;
constexpr int FRAME_PX_PER_THREAD = 2;
constexpr int FRAME_SEG = FRAME_THREADS * FRAME_PX_PER_THREAD;   // pixels of one row per pipeline stage: 1536 output bytes
constexpr int FRAME_CTAS_PER_SM = 5;                             // 5 x 40 KB of staging per SM, 40 resident warps

// ---- the byte of one channel -------------------------------------------------------------------------------------------------
// Reference arithmetic (host, no FMA): c = x / n (or x * 1/(n+1)); q = clamp((c(2.51c+.03)) / (c(2.43c+.59)+.14), 0, 1);
// byte = (uchar)(clamp(powf(q, 1/2.2f), 0, 1) * 255 + 0.5). byte is a step function of q with 255 steps (host-made table).
// Evaluating that literally costs two IEEE divisions and a table walk per channel (324 warp instructions per 32 pixels: the
// kernel was instruction-bound at 47 % of HBM speed, profiles/r1_frameFinish_v1.txt). Instead: an APPROXIMATE q (reciprocal
// multiplies, <= ~8 ulp from the exact q) and an estimated k from the SFU's lg2/ex2; the estimate is accepted only if the
// approximate q lies inside [steps[k], steps[k+1]) shrunk by 2^-17 (32 ulp) on both sides — then the exact q is in the same
// interval and k is the exact byte whatever the estimate's accuracy. Otherwise (p ~ 3e-4 per value) the exact path runs.
struct FrameTables {
    float steps[257];       // steps[k] = smallest q whose byte is >= k; steps[0] = -inf, steps[256] = +inf
    float2 margin[256];     // {steps[k] * (1 + 2^-17), steps[k+1] * (1 - 2^-17)}
};

__device__ __forceinline__ float rcpApproxF(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float lg2ApproxF(float x) { float r; asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float ex2ApproxF(float x) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

__device__ __forceinline__ float toneNum(float c) { return __fmul_rn(c, __fadd_rn(__fmul_rn(2.51f, c), 0.03f)); }                 // imageUtil.cu:210,
__device__ __forceinline__ float toneDen(float c) { return __fadd_rn(__fmul_rn(c, __fadd_rn(__fmul_rn(2.43f, c), 0.59f)), 0.14f); } // host operation order

__device__ __noinline__ unsigned exactGammaByte(float x, float norm, int divide, int k, const FrameTables* __restrict__ T) {
    const float c = divide ? __fdiv_rn(x, norm) : x;
    float q = __fdiv_rn(toneNum(c), toneDen(c));
    if (!(q == q)) return 0u;           // NaN (inf/inf from a huge or non-finite overlay value): the host cast is undefined; 0
    q = q < 0.0f ? 0.0f : (q > 1.0f ? 1.0f : q);     // util.cuh:142-146
    while (k > 0 && q < T->steps[k]) --k;            // from the estimate to the largest k with steps[k] <= q
    while (k < 255 && q >= T->steps[k + 1]) ++k;
    return (unsigned)k;
}

// fast path of one channel: approximate q (contracted arithmetic, reciprocal multiply), estimated k, and whether the estimate
// is proven exact by the margin test. cFast is the channel value, exact unless `divide` (then x * (1/n) for the exact x / n).
struct ByteGuess { int k; bool sure; };
__device__ __forceinline__ ByteGuess guessByte(float cFast, const FrameTables* __restrict__ T) {
    const float num = cFast * fmaf(2.51f, cFast, 0.03f), den = fmaf(cFast, fmaf(2.43f, cFast, 0.59f), 0.14f);
    const float qa = __saturatef(num * rcpApproxF(den));                                   // NaN -> 0
    // k = round(qa^(1/2.2) * 255) read from the low mantissa bits of (value + 2^23); qa in [0,1] -> k in [0,255]
    const int k = __float_as_int(fmaf(ex2ApproxF(lg2ApproxF(qa) * (1.0f / 2.2f)), 255.0f, 8388608.0f)) & 0xff;
    const float2 m = T->margin[k];
    // the margin proof assumes no cancellation in (2.51 c + 0.03): negative channels (FINAL mode keeps them) go to the exact path
    return {k, cFast >= 0.0f && qa > m.x && qa < m.y};
}

__device__ __forceinline__ unsigned plainByte(float c) {     // imageUtil.cu:88-90
    if (!(c == c)) return 0u;                                  // NaN: the host cast is undefined; we write 0
    float t = c < 0.0f ? 0.0f : (c > 1.0f ? 1.0f : c);
    return (unsigned)(__fadd_rn(__fmul_rn(t, 255.0f), 0.5f));
}

// ---- TMA bulk copy + mbarrier (sm_90+ PTX): the accumulator / overlay segments are staged in shared memory by the copy
// engine, two segments ahead of the arithmetic, so the loads cost no registers and overlap the previous segment's math.
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbarInit(uint64_t* bar, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smemAddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smemAddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulkLoad(void* dstSmem, const void* srcGlobal, unsigned bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smemAddr(dstSmem)), "l"(srcGlobal),
                 "r"(bytes), "r"(smemAddr(bar))
                 : "memory");
}
__device__ __forceinline__ void mbarWait(uint64_t* bar, unsigned parity) {
    asm volatile(
        "{\n .reg .pred p;\n WAIT_%=:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra DONE_%=;\n bra WAIT_%=;\n DONE_%=:\n}" ::"r"(smemAddr(bar)),
        "r"(parity)
        : "memory");
}

constexpr int FRAME_STAGES = 2;
struct __align__(128) FrameSmem {
    float4 acc[FRAME_STAGES][FRAME_SEG];       // 2 x 8 KB
    float4 ov[FRAME_STAGES][FRAME_SEG];        // 2 x 8 KB
    uint32_t out[2][FRAME_SEG * 3 / 4 + 4];    // packed B G R bytes of a segment, double-buffered
    FrameTables T;
    uint64_t bar[FRAME_STAGES];
};

// Persistent CTAs (FRAME_CTAS_PER_SM per SM, limited by their 40 KB of staging) walk the frame in segments of FRAME_SEG consecutive pixels of one
// row. mode 0: cleanAndFormatImage (checks on the raw sum, then * 1/(n+1), exact); mode 1: initRender's final pass (/ n, then
// checks). When the cleaned float frame is wanted (`fin`) the final-mode division is IEEE up front; otherwise the byte path
// works on x * (1/n) as described above.
// The five launch options are template parameters (FLAGS bit 0: progressive mode, 1: overlay, 2: cleaned float frame out,
// 3: byte rows out, 4: postProcess): tested at run time they were a quarter of the per-pixel instructions.
template <int FLAGS>
__global__ void __launch_bounds__(FRAME_THREADS, FRAME_CTAS_PER_SM)
frameFinishKernel(const float4* __restrict__ colors, const float4* __restrict__ overlayIn, int w, int h, float norm, float rnorm,
                  const FrameTables* __restrict__ gT, uint8_t* __restrict__ bgrIn, int rowSize, float4* __restrict__ finIn) {
    constexpr int mode = (FLAGS & 1) ? NOVUM_FRAME_PROGRESSIVE : NOVUM_FRAME_FINAL;
    constexpr bool postProcess = (FLAGS & 16) != 0;
    const float4* const overlay = (FLAGS & 2) ? overlayIn : nullptr;
    float4* const fin = (FLAGS & 4) ? finIn : nullptr;
    uint8_t* const bgr = (FLAGS & 8) ? bgrIn : nullptr;
    extern __shared__ __align__(128) unsigned char smemRaw[];
    FrameSmem& S = *reinterpret_cast<FrameSmem*>(smemRaw);
    const int tid = threadIdx.x;
    const int segsPerRow = (w + FRAME_SEG - 1) / FRAME_SEG;
    auto issue = [&](int y, int sx, int stage) { // one thread: arm the stage's barrier and start the copies of segment (row y, sx)
        const int x0 = sx * FRAME_SEG;
        const unsigned bytes = (unsigned)min(FRAME_SEG, w - x0) * 16u;
        const size_t base = (size_t)y * w + x0;
        mbarExpectTx(&S.bar[stage], overlay ? 2u * bytes : bytes);
        bulkLoad(S.acc[stage], colors + base, bytes, &S.bar[stage]);
        if (overlay) bulkLoad(S.ov[stage], overlay + base, bytes, &S.bar[stage]);
    };
    // segments are walked with a stride of gridDim.x; (row, segment-in-row) advance incrementally (one division per kernel, not
    // per segment: the per-segment div/mod was a fifth of all instructions, profiles/r1_frameFinish_v3.txt)
    const int stepY = gridDim.x / segsPerRow, stepX = gridDim.x - stepY * segsPerRow;
    auto advance = [&](int& y, int& sx) { y += stepY; sx += stepX; if (sx >= segsPerRow) { sx -= segsPerRow; ++y; } };
    if (tid == 0) {
        for (int s = 0; s < FRAME_STAGES; ++s) mbarInit(&S.bar[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    int y = blockIdx.x / segsPerRow, sx = blockIdx.x - y * segsPerRow;     // this CTA's current segment
    int py = y, psx = sx;                                                  // the next segment to prefetch (thread 0)
    if (tid == 0)
        for (int s = 0; s < FRAME_STAGES; ++s) {
            if (py < h) issue(py, psx, s);
            advance(py, psx);
        }
    if (bgr && postProcess) {
        const uint32_t* src = reinterpret_cast<const uint32_t*>(gT);
        uint32_t* dst = reinterpret_cast<uint32_t*>(&S.T);
        for (int i = tid; i < (int)(sizeof(FrameTables) / 4); i += FRAME_THREADS) dst[i] = __ldg(src + i);
    }
    __syncthreads();
    const bool exactDiv = fin != nullptr;
    int it = 0;
    for (; y < h; advance(y, sx), ++it) {
        const int stage = it % FRAME_STAGES;
        const int x0 = sx * FRAME_SEG;
        const int npx = min(FRAME_SEG, w - x0);
        const size_t rowBase = (size_t)y * w + x0;
        uint8_t* sBytes = reinterpret_cast<uint8_t*>(S.out[it & 1]);
        mbarWait(&S.bar[stage], (unsigned)(it / FRAME_STAGES) & 1u);
#pragma unroll
        for (int k = 0; k < FRAME_PX_PER_THREAD; ++k) {
            const int x = k * FRAME_THREADS + tid;
            if (x >= npx) continue;
            const float4 a = S.acc[stage][x];
            float4 c;            // the cleaned pixel; `divide`: c holds x * (1/n), the exact value is a / norm
            int divide = 0;
            // 0 * x is NaN exactly when x is NaN or +-Inf: one test for the rare marked pixels
            if (isnan(fmaf(a.x, 0.0f, fmaf(a.y, 0.0f, a.z * 0.0f)))) {
                if (isnan(a.x) || isnan(a.y) || isnan(a.z)) c = make_float4(1.f, 0.f, 1.f, 0.f);      // pink; x / n is NaN / Inf exactly when x is (n >= 1)
                else c = make_float4(0.f, 1.f, 0.f, 0.f);                                             // green
            } else if (mode == NOVUM_FRAME_PROGRESSIVE) {      // integratorUtilities.cuh:356-371
                if (a.x < 0 || a.y < 0 || a.z < 0) c = make_float4(0.f, 0.f, 1.f, 0.f);
                else c = make_float4(__fmul_rn(a.x, norm), __fmul_rn(a.y, norm), __fmul_rn(a.z, norm), 1.0f);
            } else if (exactDiv) {                           // main.cu:862; util.cuh:96-99 leaves w
                c = make_float4(__fdiv_rn(a.x, norm), __fdiv_rn(a.y, norm), __fdiv_rn(a.z, norm), a.w);
            } else { c = make_float4(a.x * rnorm, a.y * rnorm, a.z * rnorm, a.w); divide = 1; }
            if (overlay) {
                const float4 o = S.ov[stage][x];
                if (o.x != 0.0f || o.y != 0.0f || o.z != 0.0f) { c = o; divide = 0; }
            }
            if (fin) fin[rowBase + x] = c;
            if (bgr) {
                unsigned r, g, b;
                if (postProcess) {
                    const ByteGuess gr = guessByte(c.x, &S.T), gg = guessByte(c.y, &S.T), gb = guessByte(c.z, &S.T);
                    r = gr.k; g = gg.k; b = gb.k;
                    if (!(gr.sure && gg.sure && gb.sure)) {        // p ~ 1e-3 per pixel: redo the unsure channels in the reference's arithmetic
                        if (!gr.sure) r = exactGammaByte(divide ? a.x : c.x, norm, divide, gr.k, &S.T);
                        if (!gg.sure) g = exactGammaByte(divide ? a.y : c.y, norm, divide, gg.k, &S.T);
                        if (!gb.sure) b = exactGammaByte(divide ? a.z : c.z, norm, divide, gb.k, &S.T);
                    }
                } else {
                    if (divide) c = make_float4(__fdiv_rn(a.x, norm), __fdiv_rn(a.y, norm), __fdiv_rn(a.z, norm), 0.f);
                    r = plainByte(c.x); g = plainByte(c.y); b = plainByte(c.z);
                }
                sBytes[x * 3 + 0] = (uint8_t)b; sBytes[x * 3 + 1] = (uint8_t)g; sBytes[x * 3 + 2] = (uint8_t)r;
            }
        }
        // the segment's byte count is a multiple of 4 either way: npx * 3 = 1536 for an inner segment, and for the last one of a
        // row (rowSize - x0 * 3) with rowSize % 4 == 0 and x0 * 3 = 1536 * sx; the <= 3 bytes past the pixels are the row padding
        const int nwords = (x0 + npx == w ? rowSize - x0 * 3 : npx * 3) >> 2;
        if (bgr && tid < nwords * 4 - npx * 3) sBytes[npx * 3 + tid] = 0;                      // imageUtil.cu:80
        __syncthreads();        // every thread is done with this stage's staging buffers, and the segment's bytes are complete
        if (tid == 0) {
            if (py < h) issue(py, psx, stage);
            advance(py, psx);
        }
        if (bgr) {
            uint32_t* dst = reinterpret_cast<uint32_t*>(bgr + (size_t)y * rowSize + (size_t)x0 * 3);
            const uint32_t* src = S.out[it & 1];
            if (tid < nwords) dst[tid] = src[tid];
            if (tid + FRAME_THREADS < nwords) dst[tid + FRAME_THREADS] = src[tid + FRAME_THREADS];
        }
    }
}

// ---- host side ----------------------------------------------------------------------------------------------------------
static unsigned char hostGammaByte(float t) {   // imageUtil.cu:213-222 then :88-90, with the host's powf as the reference uses it
    float g = powf(t, 1.0f / 2.2f);
    g = g < 0.0f ? 0.0f : (g > 1.0f ? 1.0f : g);
    return (unsigned char)(g * 255.0f + 0.5f);
}

struct FrameCtx {
    std::mutex mu;
    FrameTables* dTables[16] = {};
    float steps[257];
    bool haveSteps = false;
    uint8_t* dBgr = nullptr; size_t dBgrCap = 0; int dBgrDev = -1;
    uint8_t* hBgr = nullptr; size_t hBgrCap = 0;
    float4* dIn[2] = {nullptr, nullptr}; size_t dInCap = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    float lastMs = 0.f;
};
static FrameCtx g_frame;

static void computeSteps(float* steps) {
    steps[0] = -INFINITY; steps[256] = INFINITY;
    for (int k = 1; k <= 255; ++k) {            // smallest float in [0,1] whose byte is >= k (byte(1.0f) = 255)
        uint32_t lo = 0u, hi = 0x3f800000u;     // non-negative floats order like their bit patterns
        while (lo < hi) {
            uint32_t mid = lo + (hi - lo) / 2; float t; memcpy(&t, &mid, 4);
            if (hostGammaByte(t) >= k) hi = mid; else lo = mid + 1;
        }
        memcpy(&steps[k], &lo, 4);
    }
}

static int tablesOnDevice(const FrameTables** out) {
    int dev = 0; cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 16) return NOVUM_E_ARG;
    if (!g_frame.haveSteps) { computeSteps(g_frame.steps); g_frame.haveSteps = true; }
    if (!g_frame.dTables[dev]) {
        FrameTables t;
        memcpy(t.steps, g_frame.steps, sizeof(t.steps));
        const float d = 1.0f / 131072.0f;       // 2^-17
        for (int k = 0; k < 256; ++k) t.margin[k] = make_float2(t.steps[k] * (1.0f + d), t.steps[k + 1] * (1.0f - d));
        if ((e = cudaMalloc(&g_frame.dTables[dev], sizeof(FrameTables))) != cudaSuccess) return (int)e;
        if ((e = cudaMemcpy(g_frame.dTables[dev], &t, sizeof(FrameTables), cudaMemcpyHostToDevice)) != cudaSuccess) return (int)e;
        if ((e = cudaDeviceSynchronize()) != cudaSuccess) return (int)e;      // pageable upload complete before a kernel on a non-blocking stream reads it
    }
    *out = g_frame.dTables[dev];
    return 0;
}

static int haveGpu() {
    int ndev = 0;
    return cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0;
}

using FrameKernel = void (*)(const float4*, const float4*, int, int, float, float, const FrameTables*, uint8_t*, int, float4*);
template <int... I> static FrameKernel frameKernelAt(int flags, std::integer_sequence<int, I...>) {
    static const FrameKernel table[] = {frameFinishKernel<I>...};
    return table[flags];
}
static FrameKernel frameKernelFor(int flags) { return frameKernelAt(flags & 31, std::make_integer_sequence<int, 32>{}); }
template <typename F> static void forEachFrameKernel(F f) { for (int i = 0; i < 32; ++i) f(frameKernelFor(i)); }

static int launchFinish(const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount, int mode, int postProcess,
                        uint8_t* d_bgr, NovumFloat4* d_final, cudaStream_t st) {
    if (!d_colors || w <= 0 || h <= 0 || (!d_bgr && !d_final)) return NOVUM_E_ARG;
    if (((uintptr_t)d_colors | (uintptr_t)d_overlay | (uintptr_t)d_final) & 15u) return NOVUM_E_ARG;      // float4 arrays: 16 B aligned (bulk copies)
    if (mode != NOVUM_FRAME_PROGRESSIVE && mode != NOVUM_FRAME_FINAL) return NOVUM_E_ARG;
    if ((uintptr_t)d_bgr & 3u) return NOVUM_E_ARG;
    if (mode == NOVUM_FRAME_FINAL ? sampleCount == 0 : sampleCount < 0) return NOVUM_E_ARG;
    const FrameTables* dT = nullptr;
    int rc = tablesOnDevice(&dT);
    if (rc) return rc;
    float norm = mode == NOVUM_FRAME_PROGRESSIVE ? 1.0f / (float)(sampleCount + 1) : (float)sampleCount;
    static bool attrSet[16] = {};
    static int smCount[16] = {};
    int dev = 0; cudaGetDevice(&dev);
    const int flags = (mode == NOVUM_FRAME_PROGRESSIVE ? 1 : 0) | (d_overlay ? 2 : 0) | (d_final ? 4 : 0) | (d_bgr ? 8 : 0) | (postProcess ? 16 : 0);
    if (!attrSet[dev]) {
        cudaError_t e = cudaSuccess;
        forEachFrameKernel([&](FrameKernel k) { if (e == cudaSuccess) e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FrameSmem)); });
        if (e != cudaSuccess) return (int)e;
        cudaDeviceGetAttribute(&smCount[dev], cudaDevAttrMultiProcessorCount, dev);
        attrSet[dev] = true;
    }
    const long nSeg = (long)((w + FRAME_SEG - 1) / FRAME_SEG) * h;
    const int grid = (int)std::min<long>(nSeg, (long)smCount[dev] * FRAME_CTAS_PER_SM);       // persistent CTAs
    frameKernelFor(flags)<<<grid, FRAME_THREADS, sizeof(FrameSmem), st>>>((const float4*)d_colors, (const float4*)d_overlay, w, h, norm, 1.0f / norm, dT,
                                                                           d_bgr, novum_bmp_row_bytes(w), (float4*)d_final);
    return (int)cudaGetLastError();
}

static int writeBmpFile(const char* path, const uint8_t* rows, int w, int h) {   // imageUtil.cu:69-81, 234-262
#pragma pack(push, 1)
    struct FileHeader { uint16_t type; uint32_t size; uint16_t r1, r2; uint32_t offBits; };
    struct InfoHeader { uint32_t size; int32_t w, h; uint16_t planes, bitCount; uint32_t compression, sizeImage; int32_t xppm, yppm; uint32_t clrUsed, clrImportant; };
#pragma pack(pop)
    FILE* f = fopen(path, "wb");
    if (!f) return NOVUM_E_IO;
    uint32_t rowSize = (uint32_t)novum_bmp_row_bytes(w);
    FileHeader fh = {0x4D42, (uint32_t)(sizeof(FileHeader) + sizeof(InfoHeader) + rowSize * h), 0, 0, (uint32_t)(sizeof(FileHeader) + sizeof(InfoHeader))};
    InfoHeader ih = {sizeof(InfoHeader), w, h, 1, 24, 0, rowSize * (uint32_t)h, 0, 0, 0, 0};
    bool ok = fwrite(&fh, sizeof(fh), 1, f) == 1 && fwrite(&ih, sizeof(ih), 1, f) == 1 && fwrite(rows, 1, (size_t)rowSize * h, f) == (size_t)rowSize * h;
    fclose(f);
    return ok ? 0 : NOVUM_E_IO;
}

// device colours (+overlay) -> packed rows in page-locked host memory (cached buffers); returns the host pointer
static int finishToHost(const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount, int mode, int postProcess,
                        const uint8_t** rowsOut) {
    size_t bytes = (size_t)novum_bmp_row_bytes(w) * h;
    int dev = 0; cudaGetDevice(&dev);
    cudaError_t e;
    if (g_frame.dBgrCap < bytes || g_frame.dBgrDev != dev) {
        if (g_frame.dBgr) cudaFree(g_frame.dBgr);
        g_frame.dBgr = nullptr; g_frame.dBgrCap = 0;
        if ((e = cudaMalloc(&g_frame.dBgr, bytes)) != cudaSuccess) return (int)e;
        g_frame.dBgrCap = bytes; g_frame.dBgrDev = dev;
    }
    if (g_frame.hBgrCap < bytes) {
        if (g_frame.hBgr) cudaFreeHost(g_frame.hBgr);
        g_frame.hBgr = nullptr; g_frame.hBgrCap = 0;
        if ((e = cudaMallocHost(&g_frame.hBgr, bytes)) != cudaSuccess) return (int)e;
        g_frame.hBgrCap = bytes;
    }
    if (!g_frame.e0) { cudaEventCreate(&g_frame.e0); cudaEventCreate(&g_frame.e1); }
    cudaEventRecord(g_frame.e0, 0);
    int rc = launchFinish(d_colors, d_overlay, w, h, sampleCount, mode, postProcess, g_frame.dBgr, nullptr, 0);
    if (rc) return rc;
    cudaEventRecord(g_frame.e1, 0);
    if ((e = cudaMemcpyAsync(g_frame.hBgr, g_frame.dBgr, bytes, cudaMemcpyDeviceToHost, 0)) != cudaSuccess) return (int)e;
    if ((e = cudaStreamSynchronize(0)) != cudaSuccess) return (int)e;
    cudaEventElapsedTime(&g_frame.lastMs, g_frame.e0, g_frame.e1);
    *rowsOut = g_frame.hBgr;
    return 0;
}

// ---- asynchronous save: the read-back and the file write overlap the next sample pass ------------------------------------------
// The reference stops its sample loop every 5 s for cleanAndFormatImage + two float4 frame copies + the host post-process + the file
// write (deviceCode.cu:1905-1927; 0.37 s of host work for a 4K frame). Here the finish kernel runs on a stream of its own once the
// render stream has reached the call; the render stream then waits for that kernel only (it has read the accumulators - ~55 us at 4K),
// while the 3 B/pixel copy and the write happen behind it: copy on a second stream, write on a worker thread. Two slots, so a save
// may still be on its way to disk while the next one is being produced; a third call waits for the oldest.
struct SaveSlot {
    uint8_t* dBgr = nullptr; uint8_t* hBgr = nullptr; size_t cap = 0;
    cudaEvent_t copied = nullptr;
    bool busy = false;
    std::string path; int w = 0, h = 0;
};
struct FrameSaver {
    std::mutex mu;
    std::condition_variable cv;
    cudaStream_t stream = nullptr, copyStream = nullptr;      // kernels / read-backs: save k+1's kernel need not queue behind save k's copy
    cudaEvent_t reached = nullptr, kernelDone = nullptr;
    SaveSlot slot[2];
    int next = 0, pending = 0, firstError = 0, dev = -1;
    std::deque<int> jobs;
    bool started = false;
};
static FrameSaver* g_saver = nullptr;       // created on first use, never destroyed (its thread is detached)

static void saverLoop(FrameSaver* S) {
    cudaSetDevice(S->dev);
    while (true) {
        int k;
        {
            std::unique_lock<std::mutex> lk(S->mu);
            S->cv.wait(lk, [&] { return !S->jobs.empty(); });
            k = S->jobs.front(); S->jobs.pop_front();
        }
        SaveSlot& sl = S->slot[k];
        int rc = (int)cudaEventSynchronize(sl.copied);
        if (rc == 0) rc = writeBmpFile(sl.path.c_str(), sl.hBgr, sl.w, sl.h);
        {
            std::lock_guard<std::mutex> lk(S->mu);
            if (rc && !S->firstError) S->firstError = rc;
            sl.busy = false; --S->pending;
        }
        S->cv.notify_all();
    }
}

static int saveAsync(const char* path, const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount, int mode,
                     int postProcess, cudaStream_t renderStream) {
    int dev = 0; cudaGetDevice(&dev);
    if (!g_saver) g_saver = new FrameSaver();
    FrameSaver* S = g_saver;
    cudaError_t e;
    std::unique_lock<std::mutex> lk(S->mu);
    if (S->dev >= 0 && S->dev != dev) return NOVUM_E_ARG;           // one saver per process, bound to the device of its first call
    if (!S->stream) {
        if ((e = cudaStreamCreateWithFlags(&S->stream, cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
        if ((e = cudaStreamCreateWithFlags(&S->copyStream, cudaStreamNonBlocking)) != cudaSuccess) return (int)e;
        cudaEventCreateWithFlags(&S->reached, cudaEventDisableTiming);
        cudaEventCreateWithFlags(&S->kernelDone, cudaEventDisableTiming);
        for (SaveSlot& sl : S->slot) cudaEventCreateWithFlags(&sl.copied, cudaEventDisableTiming);
        S->dev = dev;
    }
    const int k = S->next;
    SaveSlot& sl = S->slot[k];
    S->cv.wait(lk, [&] { return !sl.busy; });                        // its previous file is on disk
    const size_t bytes = (size_t)novum_bmp_row_bytes(w) * h;
    if (sl.cap < bytes) {
        if (sl.dBgr) cudaFree(sl.dBgr);
        if (sl.hBgr) cudaFreeHost(sl.hBgr);
        sl.dBgr = sl.hBgr = nullptr; sl.cap = 0;
        if ((e = cudaMalloc(&sl.dBgr, bytes)) != cudaSuccess) return (int)e;
        if ((e = cudaMallocHost(&sl.hBgr, bytes)) != cudaSuccess) return (int)e;
        sl.cap = bytes;
    }
    if ((e = cudaEventRecord(S->reached, renderStream)) != cudaSuccess) return (int)e;
    cudaStreamWaitEvent(S->stream, S->reached, 0);
    int rc;
    {
        std::lock_guard<std::mutex> fl(g_frame.mu);
        rc = launchFinish(d_colors, d_overlay, w, h, sampleCount, mode, postProcess, sl.dBgr, nullptr, S->stream);
    }
    if (rc) return rc;
    cudaEventRecord(S->kernelDone, S->stream);
    cudaStreamWaitEvent(renderStream, S->kernelDone, 0);             // the next pass may touch the accumulators from here on
    cudaStreamWaitEvent(S->copyStream, S->kernelDone, 0);
    if ((e = cudaMemcpyAsync(sl.hBgr, sl.dBgr, bytes, cudaMemcpyDeviceToHost, S->copyStream)) != cudaSuccess) return (int)e;
    cudaEventRecord(sl.copied, S->copyStream);
    sl.path = path; sl.w = w; sl.h = h; sl.busy = true;
    ++S->pending; S->next = k ^ 1;
    S->jobs.push_back(k);
    if (!S->started) { S->started = true; std::thread(saverLoop, S).detach(); }
    lk.unlock();
    S->cv.notify_all();
    return 0;
}

}  // namespace nv

using namespace nv;

extern "C" {

int novum_bmp_row_bytes(int w) { return (3 * w + 3) & (~3); }

int novum_frame_gamma_steps(float* steps257) {
    if (!steps257) return NOVUM_E_ARG;
    std::lock_guard<std::mutex> lk(g_frame.mu);
    if (!g_frame.haveSteps) { computeSteps(g_frame.steps); g_frame.haveSteps = true; }
    memcpy(steps257, g_frame.steps, sizeof(float) * 257);
    return 0;
}

int novum_frame_finish_device(const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount, int mode,
                              int postProcess, uint8_t* d_bgrRows, NovumFloat4* d_final, void* stream) {
    if (!haveGpu()) return NOVUM_E_NOGPU;
    std::lock_guard<std::mutex> lk(g_frame.mu);
    return launchFinish(d_colors, d_overlay, w, h, sampleCount, mode, postProcess, d_bgrRows, d_final, (cudaStream_t)stream);
}

int novum_frame_finish(const NovumFloat4* colors, const NovumFloat4* overlay, int w, int h, int sampleCount, int mode, int postProcess,
                       uint8_t* bgrRows, NovumFloat4* finalOut) {
    if (!colors || w <= 0 || h <= 0 || (!bgrRows && !finalOut)) return NOVUM_E_ARG;
    if (!haveGpu()) return NOVUM_E_NOGPU;
    std::lock_guard<std::mutex> lk(g_frame.mu);
    size_t n = (size_t)w * h; cudaError_t e;
    if (g_frame.dInCap < n) {
        for (int i = 0; i < 2; ++i) { if (g_frame.dIn[i]) cudaFree(g_frame.dIn[i]); g_frame.dIn[i] = nullptr; }
        g_frame.dInCap = 0;
        for (int i = 0; i < 2; ++i) if ((e = cudaMalloc(&g_frame.dIn[i], n * sizeof(float4))) != cudaSuccess) return (int)e;
        g_frame.dInCap = n;
    }
    if ((e = cudaMemcpyAsync(g_frame.dIn[0], colors, n * sizeof(float4), cudaMemcpyHostToDevice, 0)) != cudaSuccess) return (int)e;
    if (overlay && (e = cudaMemcpyAsync(g_frame.dIn[1], overlay, n * sizeof(float4), cudaMemcpyHostToDevice, 0)) != cudaSuccess) return (int)e;
    const NovumFloat4* dC = (const NovumFloat4*)g_frame.dIn[0];
    const NovumFloat4* dO = overlay ? (const NovumFloat4*)g_frame.dIn[1] : nullptr;
    if (bgrRows) {
        const uint8_t* rows = nullptr;
        int rc = finishToHost(dC, dO, w, h, sampleCount, mode, postProcess, &rows);
        if (rc) return rc;
        memcpy(bgrRows, rows, (size_t)novum_bmp_row_bytes(w) * h);
    }
    if (finalOut) {  // the cleaned linear frame (what the reference hands to Image::setColor)
        float4* dF = nullptr;
        if ((e = cudaMalloc(&dF, n * sizeof(float4))) != cudaSuccess) return (int)e;
        int rc = launchFinish(dC, dO, w, h, sampleCount, mode, postProcess, nullptr, (NovumFloat4*)dF, 0);
        if (rc == 0) rc = (int)cudaMemcpy(finalOut, dF, n * sizeof(float4), cudaMemcpyDeviceToHost);
        cudaFree(dF);
        if (rc) return rc;
    }
    return 0;
}

int novum_frame_save_bmp(const char* path, const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount,
                         int mode, int postProcess) {
    if (!path) return NOVUM_E_ARG;
    if (!haveGpu()) return NOVUM_E_NOGPU;
    std::lock_guard<std::mutex> lk(g_frame.mu);
    const uint8_t* rows = nullptr;
    int rc = finishToHost(d_colors, d_overlay, w, h, sampleCount, mode, postProcess, &rows);
    if (rc) return rc;
    return writeBmpFile(path, rows, w, h);
}

int novum_frame_save_bmp_async(const char* path, const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount,
                               int mode, int postProcess, void* renderStream) {
    if (!path) return NOVUM_E_ARG;
    if (!haveGpu()) return NOVUM_E_NOGPU;
    return saveAsync(path, d_colors, d_overlay, w, h, sampleCount, mode, postProcess, (cudaStream_t)renderStream);
}

int novum_frame_save_wait(void) {
    FrameSaver* S = g_saver;
    if (!S) return 0;
    std::unique_lock<std::mutex> lk(S->mu);
    S->cv.wait(lk, [&] { return S->pending == 0; });
    int rc = S->firstError; S->firstError = 0;
    return rc;
}

int novum_save_bmp(const char* path, const NovumFloat4* colors, const NovumFloat4* overlay, int w, int h, int sampleCount, int mode,
                   int postProcess) {
    if (!path || !colors || w <= 0 || h <= 0) return NOVUM_E_ARG;
    std::vector<uint8_t> rows((size_t)novum_bmp_row_bytes(w) * h);
    int rc = novum_frame_finish(colors, overlay, w, h, sampleCount, mode, postProcess, rows.data(), nullptr);
    if (rc) return rc;
    return writeBmpFile(path, rows.data(), w, h);
}

int novum_write_bmp(const char* path, const uint8_t* bgrRows, int w, int h) {
    if (!path || !bgrRows || w <= 0 || h <= 0) return NOVUM_E_ARG;
    return writeBmpFile(path, bgrRows, w, h);
}

int novum_frame_last_ms(float* ms) {
    if (!ms) return NOVUM_E_ARG;
    *ms = g_frame.lastMs;
    return 0;
}

}  // extern "C"
