// vecmath.cuh — 3-vector math for the rendering core (host + device).
// Semantics that results depend on follow the reference: normalize() multiplies by rsqrtf
// (util.cuh:130-133), the orthonormal basis is the branchy |n.x| > |n.z| construction
// (util.cuh:163-185), PI is the truncated 3.141592f (util.cuh:29).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#define NV_HD __host__ __device__ __forceinline__
#define NV_D __device__ __forceinline__

namespace nv {

constexpr float kEps = 0.00001f;        // util.cuh:27 EPSILON
constexpr float kRayEps = 0.001f;       // util.cuh:28 RAY_EPSILON
constexpr float kPi = 3.141592f;        // util.cuh:29 (truncated on purpose)
constexpr float kSkyRadius = 100.0f;    // util.cuh:30
constexpr float kMaxFireflyLum = 5.0f;  // util.cuh:31

struct V3 {
    float x, y, z;
    NV_HD V3() {}
    NV_HD V3(float a) : x(a), y(a), z(a) {}
    NV_HD V3(float a, float b, float c) : x(a), y(b), z(c) {}
    NV_HD explicit V3(const float4& f) : x(f.x), y(f.y), z(f.z) {}
};

NV_HD V3 operator+(V3 a, V3 b) { return V3(a.x + b.x, a.y + b.y, a.z + b.z); }
NV_HD V3 operator-(V3 a, V3 b) { return V3(a.x - b.x, a.y - b.y, a.z - b.z); }
NV_HD V3 operator-(V3 a) { return V3(-a.x, -a.y, -a.z); }
NV_HD V3 operator*(V3 a, V3 b) { return V3(a.x * b.x, a.y * b.y, a.z * b.z); }
NV_HD V3 operator*(V3 a, float s) { return V3(a.x * s, a.y * s, a.z * s); }
NV_HD V3 operator*(float s, V3 a) { return V3(a.x * s, a.y * s, a.z * s); }
NV_HD V3 operator/(V3 a, float s) { return V3(a.x / s, a.y / s, a.z / s); }
NV_HD V3& operator+=(V3& a, V3 b) { a.x += b.x; a.y += b.y; a.z += b.z; return a; }
NV_HD V3& operator*=(V3& a, V3 b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; return a; }
NV_HD V3& operator*=(V3& a, float s) { a.x *= s; a.y *= s; a.z *= s; return a; }
NV_HD V3& operator/=(V3& a, float s) { a.x /= s; a.y /= s; a.z /= s; return a; }

NV_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
NV_HD float lengthSq(V3 a) { return dot(a, a); }
NV_HD float length(V3 a) { return sqrtf(dot(a, a)); }
NV_HD V3 cross(V3 a, V3 b) { return V3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
NV_HD float maxComp(V3 a) { return fmaxf(a.x, fmaxf(a.y, a.z)); }
NV_HD V3 vmin(V3 a, V3 b) { return V3(fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z)); }
NV_HD V3 vmax(V3 a, V3 b) { return V3(fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z)); }
NV_HD float luminance(V3 c) { return 0.2126f * c.x + 0.7152f * c.y + 0.0722f * c.z; }  // util.cuh:283-286

NV_HD float rsqrt_(float x) {
#ifdef __CUDA_ARCH__
    return rsqrtf(x);
#else
    return 1.0f / sqrtf(x);
#endif
}
NV_HD V3 normalize(V3 v) { float il = rsqrt_(dot(v, v)); return V3(v.x * il, v.y * il, v.z * il); }
NV_HD float clampf(float x, float lo, float hi) { return x < lo ? lo : (x > hi ? hi : x); }
NV_HD float4 toF4(V3 v, float w = 0.f) { return make_float4(v.x, v.y, v.z, w); }

// Orthonormal frame around n, identical in construction to util.cuh:163-185.
struct Frame {
    V3 t, b, n;
    NV_HD explicit Frame(V3 nn) : n(nn) {
        if (fabsf(nn.x) > fabsf(nn.z)) t = normalize(V3(-nn.y, nn.x, 0.0f));
        else t = normalize(V3(0.0f, -nn.z, nn.y));
        b = cross(nn, t);
    }
    NV_HD V3 toLocal(V3 w) const { return V3(dot(w, t), dot(w, b), dot(w, n)); }
    NV_HD V3 toWorld(V3 l) const { return l.x * t + l.y * b + l.z * n; }
};

// ---- octahedral 2x snorm16 unit-vector codec (util.cuh:342-391) --------------------------
NV_HD float signNotZero(float k) { return (k >= 0.0f) ? 1.0f : -1.0f; }
NV_HD int f2i_rn(float x) {
#ifdef __CUDA_ARCH__
    return __float2int_rn(x);
#else
    return (int)nearbyintf(x);
#endif
}
NV_HD uint32_t packOct(V3 v) {
    float l1 = fabsf(v.x) + fabsf(v.y) + fabsf(v.z);
    float inv = (l1 > 0.0f) ? (1.0f / l1) : 0.0f;
    float rx = v.x * inv, ry = v.y * inv;
    if (v.z < 0.0f) {
        float tx = rx, ty = ry;
        rx = (1.0f - fabsf(ty)) * signNotZero(tx);
        ry = (1.0f - fabsf(tx)) * signNotZero(ty);
    }
    int16_t qx = (int16_t)f2i_rn(fminf(fmaxf(rx, -1.0f), 1.0f) * 32767.0f);
    int16_t qy = (int16_t)f2i_rn(fminf(fmaxf(ry, -1.0f), 1.0f) * 32767.0f);
    return ((uint32_t)(uint16_t)qx) | (((uint32_t)(uint16_t)qy) << 16);
}
NV_HD V3 unpackOct(uint32_t p) {
    int16_t px = (int16_t)(p & 0xFFFF), py = (int16_t)(p >> 16);
#ifdef __CUDA_ARCH__
    // the reference divides (util.cuh:376-377); on the device the reciprocal multiply (<= 1 ulp apart, far below the 3e-5
    // quantisation step) replaces two IEEE divisions per decode — they were 12 % of the eye step's stall samples
    float x = fmaxf(-1.0f, (float)px * (1.0f / 32767.0f)), y = fmaxf(-1.0f, (float)py * (1.0f / 32767.0f));
#else
    float x = fmaxf(-1.0f, px / 32767.0f), y = fmaxf(-1.0f, py / 32767.0f);
#endif
    float z = 1.0f - (fabsf(x) + fabsf(y));
    float t = fmaxf(-z, 0.0f);
    x += (x >= 0.0f) ? -t : t;
    y += (y >= 0.0f) ? -t : t;
    float il = rsqrt_(x * x + y * y + z * z);
    return V3(x * il, y * il, z * il);
}

// ---- counter-seeded PCG32 stream per path (replaces the reference's 48 B XORWOW state,
//      deviceCode.cu:53-61; statistical parity only) -----------------------------------------
struct Rng {
    uint64_t state;
    NV_HD static uint64_t mix(uint64_t z) {  // splitmix64 finaliser
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        return z ^ (z >> 31);
    }
    NV_HD void seed(uint64_t seed, uint64_t path, uint64_t sample, uint64_t stream) {
        state = mix(seed + 0x9E3779B97F4A7C15ULL * (path + 1)) ^ mix((sample << 8) + stream + 0x632BE59BD9B4E019ULL);
        state = mix(state);
        nextU32();
    }
    NV_HD uint32_t nextU32() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + 1442695040888963407ULL;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((32u - rot) & 31u));
    }
    // uniform in (0, 1], like curand_uniform
    NV_HD float uniform() { return ((nextU32() >> 8) + 1u) * (1.0f / 16777216.0f); }
};

}  // namespace nv
