// novum_oracle_image.cpp — CPU restatement of the reference's frame finish (the step after the path).
// TEST INFRASTRUCTURE ONLY (checker for tests/, smoke() and bench.py's cpu_baseline leg); the product never links it.
// Pinned: oracle_bmp_rows against the reference's own Image::saveImageBMP run on this container's CPU
// (`oracle/_ref/novum_ref bmp`, fixture tests/golden/frame_bmp.npz); oracle_clean_and_format against the reference's own
// cleanAndFormatImage kernel run on a B200 (`novum_ref clean`, fixture tests/golden/frame_clean.npz);
// oracle_final_normalise restates inline code of initRender (main.cu:855-884) that cannot be called on its own.
// Built with -ffp-contract=off: the reference's host build (g++ behind nvcc, baseline x86-64) has no FMA.
#include <cmath>
#include <cstdint>
#include <cstring>

namespace {

struct F4 { float x, y, z, w; };

inline float clampf(float x, float lo, float hi) {   // util.cuh:142-146
    if (x < lo) return lo;
    if (x > hi) return hi;
    return x;
}

inline float toneMap1(float c) {                     // imageUtil.cu:202-211, one channel of the float4 expression
    const float A = 2.51f, B = 0.03f, C = 2.43f, D = 0.59f, E = 0.14f;
    float num = c * (A * c + B);
    float den = c * (C * c + D) + E;
    return clampf(num / den, 0.0f, 1.0f);
}

inline float gamma1(float c) {                       // imageUtil.cu:213-222
    float invGamma = 1.0f / 2.2f;
    return powf(c, invGamma);
}

inline unsigned char toByte(float c) {               // imageUtil.cu:88-90 (a NaN reaches an undefined cast there; 0 here)
    float v = clampf(c, 0.0f, 1.0f) * 255.0f + 0.5f;
    if (!(v == v)) return 0;
    return static_cast<unsigned char>(v);
}

}  // namespace

extern "C" {

// integratorUtilities.cuh:337-380
void oracle_clean_and_format(const float* accIn, const float* ovIn, float* outF, int w, int h, int currentSampleCount) {
    const F4* acc = reinterpret_cast<const F4*>(accIn); const F4* ov = reinterpret_cast<const F4*>(ovIn); F4* out = reinterpret_cast<F4*>(outF);
    for (long i = 0; i < (long)w * h; ++i) {
        F4 a = acc[i], f;
        if (std::isnan(a.x) || std::isnan(a.y) || std::isnan(a.z)) f = {1.0f, 0.0f, 1.0f, 0.0f};
        else if (std::isinf(a.x) || std::isinf(a.y) || std::isinf(a.z)) f = {0.0f, 1.0f, 0.0f, 0.0f};
        else if (a.x < 0 || a.y < 0 || a.z < 0) f = {0.0f, 0.0f, 1.0f, 0.0f};
        else { float scale = 1.0f / (float)(currentSampleCount + 1); f = {a.x * scale, a.y * scale, a.z * scale, 1.0f}; }
        if (ov) { F4 o = ov[i]; if (o.x != 0.0f || o.y != 0.0f || o.z != 0.0f) f = o; }
        out[i] = f;
    }
}

// main.cu:855-884: /= sampleCount (util.cuh:96-99 divides x,y,z), NaN -> pink, Inf -> green, overlay override
void oracle_final_normalise(const float* accIn, const float* ovIn, float* outF, int w, int h, int sampleCount) {
    const F4* acc = reinterpret_cast<const F4*>(accIn); const F4* ov = reinterpret_cast<const F4*>(ovIn); F4* out = reinterpret_cast<F4*>(outF);
    float d = (float)sampleCount;
    for (long i = 0; i < (long)w * h; ++i) {
        F4 c = acc[i];
        c.x /= d; c.y /= d; c.z /= d;
        if (std::isnan(c.x) || std::isnan(c.y) || std::isnan(c.z)) c = {1.0f, 0.0f, 1.0f, 0.0f};
        if (std::isinf(c.x) || std::isinf(c.y) || std::isinf(c.z)) c = {0.0f, 1.0f, 0.0f, 0.0f};
        if (ov) { F4 o = ov[i]; if (!(o.x == 0 && o.y == 0 && o.z == 0)) c = o; }
        out[i] = c;
    }
}

// Image::saveImageBMP's pixel loop (imageUtil.cu:69-100) with postProcessImage (:224-232): rows of (3w+3)&~3 bytes, B G R
void oracle_bmp_rows(const float* pxIn, int w, int h, int postProcess, uint8_t* rows) {
    const F4* px = reinterpret_cast<const F4*>(pxIn);
    int rowSize = (3 * w + 3) & (~3);
    memset(rows, 0, (size_t)rowSize * h);
#pragma omp parallel for schedule(static)
    for (int y = 0; y < h; ++y) {
        uint8_t* row = rows + (size_t)y * rowSize;
        for (int x = 0; x < w; ++x) {
            F4 c = px[(size_t)y * w + x];
            if (postProcess) { c.x = gamma1(toneMap1(c.x)); c.y = gamma1(toneMap1(c.y)); c.z = gamma1(toneMap1(c.z)); }
            row[x * 3 + 0] = toByte(c.z); row[x * 3 + 1] = toByte(c.y); row[x * 3 + 2] = toByte(c.x);
        }
    }
}

// Exhaustive property behind the product's gamma table: over every float t in [0, 1], byte(t) = toByte(gamma1(t)) never
// decreases as t grows, and equals the table look-up (largest k with steps[k] <= t). Returns the number of violations.
long oracle_gamma_table_check(const float* steps257, uint32_t firstBits, uint32_t lastBits) {
    long bad = 0;
#pragma omp parallel for schedule(static) reduction(+ : bad)
    for (int64_t b = firstBits; b <= (int64_t)lastBits; ++b) {
        uint32_t u = (uint32_t)b; float t; memcpy(&t, &u, 4);
        int want = toByte(gamma1(t));
        int lo = 0, hi = 255;                       // largest k with steps[k] <= t (steps[0] = -inf)
        while (lo < hi) { int mid = (lo + hi + 1) / 2; if (steps257[mid] <= t) lo = mid; else hi = mid - 1; }
        bad += (lo != want);
    }
    return bad;
}

}  // extern "C"
