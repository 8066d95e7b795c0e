// novum_oracle_bsdf.h — CPU restatement of reflectors.cuh (BSDF eval / sample / pdf), written against the
// reference file directly (not against the CUDA core's bsdf.cuh). TEST INFRASTRUCTURE ONLY.
#pragma once
#include "novum_oracle.h"
#include <math.h>

namespace orc {

const float EPS = 0.00001f, RAY_EPS = 0.001f, PI_ = 3.141592f;    // util.cuh:27-29

struct V3 { float x, y, z; };
inline V3 v3(float x, float y, float z) { V3 v = {x, y, z}; return v; }
inline V3 v3(float a) { return v3(a, a, a); }
inline V3 v3(const OF4& f) { return v3(f.x, f.y, f.z); }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return a * s; }
inline V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
inline float lenSq(V3 a) { return dot(a, a); }
inline float len(V3 a) { return sqrtf(dot(a, a)); }
inline V3 normalize(V3 a) { float il = 1.0f / sqrtf(dot(a, a)); return a * il; }     // util.cuh:130 (rsqrtf there)

// util.cuh:163-185
inline void basis(V3 n, V3& t, V3& b) {
    if (fabsf(n.x) > fabsf(n.z)) t = normalize(v3(-n.y, n.x, 0.0f));
    else t = normalize(v3(0.0f, -n.z, n.y));
    b = cross(n, t);
}
inline V3 toLocal(V3 w, V3 n) { V3 t, b; basis(n, t, b); return v3(dot(w, t), dot(w, b), dot(w, n)); }
inline V3 toWorld(V3 l, V3 n) { V3 t, b; basis(n, t, b); return l.x * t + l.y * b + l.z * n; }

// PCG32 stream keyed like the CUDA core (cudapathtracer_b200/csrc/vecmath.cuh Rng): the reference's XORWOW
// (deviceCode.cu:53-61) is replaced on both sides, so CPU and GPU paths draw the same uniforms.
struct Rng {
    uint64_t state;
    static uint64_t mix(uint64_t z) {
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
        return z ^ (z >> 31);
    }
    void seed(uint64_t seed, uint64_t path, uint64_t sample, uint64_t stream) {
        state = mix(seed + 0x9E3779B97F4A7C15ULL * (path + 1)) ^ mix((sample << 8) + stream + 0x632BE59BD9B4E019ULL);
        state = mix(state);
        nextU32();
    }
    uint32_t nextU32() {
        uint64_t old = state;
        state = old * 6364136223846793005ULL + 1442695040888963407ULL;
        uint32_t xs = (uint32_t)(((old >> 18u) ^ old) >> 27u);
        uint32_t rot = (uint32_t)(old >> 59u);
        return (xs >> rot) | (xs << ((32u - rot) & 31u));
    }
    float uniform() { return ((nextU32() >> 8) + 1u) * (1.0f / 16777216.0f); }   // (0,1] like curand_uniform
};

// uniform source: either the PCG stream or a caller-provided array (for unit KATs)
struct Uniforms {
    Rng* rng; const float* arr; int used;
    float next() { ++used; return rng ? rng->uniform() : arr[used - 1]; }
};

inline int wrapi(int v, int d) { int r = v % d; return r < 0 ? r + d : r; }
// reflectors.cuh:371-417
inline V3 sampleTexture(const OMaterial& m, const OF4* tex, float u, float v) {
    int W = m.width, H = m.height;
    float fx = u * W - 0.5f, fy = v * H - 0.5f;
    int xi = (int)floorf(fx), yi = (int)floorf(fy);
    float sx = fx - floorf(fx), sy = fy - floorf(fy);
    int x0 = wrapi(xi, W), y0 = wrapi(yi, H), x1 = wrapi(xi + 1, W), y1 = wrapi(yi + 1, H);
    V3 c00 = v3(tex[m.startInd + y0 * W + x0]), c10 = v3(tex[m.startInd + y0 * W + x1]);
    V3 c01 = v3(tex[m.startInd + y1 * W + x0]), c11 = v3(tex[m.startInd + y1 * W + x1]);
    V3 bot = c00 * (1.0f - sx) + c10 * sx, top = c01 * (1.0f - sx) + c11 * sx;
    return bot * (1.0f - sy) + top * sy;
}

inline float cosinePdf(V3 wo) { return fmaxf(wo.z, EPS) / PI_; }                  // :15-18
inline V3 cosineSample(Uniforms& U) {                                             // :21-39
    float u1 = fminf(U.next(), 1.0f - EPS);
    float u2 = U.next();
    float r = sqrtf(u1), phi = 2.0f * PI_ * u2;
    return v3(r * cosf(phi), r * sinf(phi), sqrtf(1.0f - u1));
}
inline float D_GGX(V3 h, float alpha) {                                           // :80-86
    float a2 = alpha * alpha, d = h.z * h.z * (a2 - 1.0f) + 1.0f;
    return a2 / (PI_ * d * d);
}
inline float G1_GGX(V3 v, float alpha) {                                          // :94-103
    float c = v.z, tanT = sqrtf(1.0f - c * c) / c, a = 1.0f / (alpha * tanT);
    if (a < 1.6f) return (3.535f * a + 2.181f * a * a) / (1.0f + 2.276f * a + 2.577f * a * a);
    return 1.0f;
}
inline float fresnelRs(float c, float eta, float k) {                             // :110-127 (returns Rs only)
    float c2 = c * c, s2 = 1.0f - c2, e2 = eta * eta, k2 = k * k;
    float t0 = e2 - k2 - s2, a2b2 = sqrtf(t0 * t0 + 4.0f * e2 * k2);
    float t1 = a2b2 + c2, a = sqrtf(0.5f * (a2b2 + t0)), t2 = 2.0f * c * a;
    return (t1 - t2) / (t1 + t2);
}
inline float schlick(float c, float ei, float et) {                               // :180-185
    float r0 = (ei - et) / (ei + et); r0 *= r0;
    return r0 + (1.0f - r0) * powf(1.0f - fabsf(c), 5.0f);
}
inline V3 metalF(const OMaterial& m, V3 wi, V3 wo) {                              // :129-151
    if (wi.z <= 0.0f || wo.z <= 0.0f) return v3(0.0f);
    V3 h = normalize(wi + wo);
    if (h.z <= 0.0f) h = -h;
    float alpha = m.roughness * m.roughness;
    float D = D_GGX(h, alpha), G = G1_GGX(wi, alpha) * G1_GGX(wo, alpha), c = dot(wi, h);
    V3 F = v3(fresnelRs(c, m.eta.x, m.k.x), fresnelRs(c, m.eta.y, m.k.y), fresnelRs(c, m.eta.z, m.k.z));
    return (D * G * F) / fmaxf(4.0f * wi.z * wo.z, EPS);
}
inline float microfacetPdf(float rough, V3 wi, V3 wo) {                           // :153-159
    V3 h = normalize(wi + wo);
    return (D_GGX(h, rough * rough) * h.z) / (4.0f * dot(wo, h));
}
inline V3 ggxH(Uniforms& U, float rough) {                                        // :164-175
    float u1 = U.next();
    float alpha = rough * rough;
    float phi = 2.0f * PI_ * U.next();
    float ct = sqrtf((1.0f - u1) / (1.0f + (alpha * alpha - 1.0f) * u1));
    float st = sqrtf(fmaxf(1.0f - ct * ct, 0.0f));
    return v3(st * cosf(phi), st * sinf(phi), ct);
}
inline V3 leafF(V3 albedo, float ior, float cur, float rough, float trans, V3 wi, V3 wo) {    // :420-466
    bool refl = wo.z * wi.z > 0.0f;
    float F = schlick(wi.z, cur, ior);
    if (refl) {
        V3 h = normalize(wi + wo);
        float mF = schlick(dot(wi, h), cur, ior);
        if (h.z <= 0.0f) h = -h;
        float alpha = rough * rough;
        float cut = D_GGX(h, alpha) * G1_GGX(wi, alpha) * G1_GGX(wo, alpha) * mF / fmaxf(4.0f * wi.z * wo.z, EPS);
        return (1.0f - mF) * (1.0f - trans) * (albedo / PI_) + v3(cut);
    }
    return (albedo / PI_) * (trans * (1.0f - F));
}
inline float leafPdf(float ior, float cur, float rough, float trans, V3 wi, V3 wo) {          // :468-510
    bool refl = wo.z * wi.z > 0.0f;
    float F = fminf(schlick(fabsf(wi.z), cur, ior), 1.0f - 0.1f * rough);
    if (refl) {
        V3 h = normalize(wi + wo);
        if (h.z < 0.0f) h = -h;
        float pc = (D_GGX(h, rough * rough) * h.z) / (4.0f * dot(wo, h));
        return F * pc + (1.0f - F) * (1.0f - trans) * cosinePdf(wo);
    }
    return cosinePdf(-wo) * (1.0f - F) * trans;
}

struct Resolved { V3 albedoTex; float trans; };
inline Resolved resolve(const OMaterial& m, const OF4* tex, float u, float v) {    // :553-562 / :597-606
    Resolved r; r.albedoTex = v3(m.albedo); r.trans = m.transmission;
    if (m.hasTexture) r.albedoTex = sampleTexture(m, tex, u, v);
    if (m.hasTransMap) r.trans = sampleTexture(m, tex, u, v).x;
    return r;
}

// f_eval :547-586
inline V3 bsdfEval(const OMaterial& m, const OF4* tex, V3 wiToward, V3 wo, float etaI, float u, float v) {
    Resolved r = resolve(m, tex, u, v);
    switch (m.type) {
        case 0: return v3(m.albedo) / PI_;                       // diffuse: untextured albedo (:566)
        case 1: return metalF(m, -wiToward, wo);
        case 4: return leafF(r.albedoTex, m.ior, etaI, m.roughness, r.trans, -wiToward, wo);
        case 6: return v3(1.0f / fmaxf(wo.z, EPS));
        default: return v3(0.0f);
    }
}
// pdf_eval :631-666
inline float bsdfPdf(const OMaterial& m, const OF4* tex, V3 wiToward, V3 wo, float etaI, float u, float v) {
    Resolved r = resolve(m, tex, u, v);
    switch (m.type) {
        case 0: return cosinePdf(wo);
        case 1: return microfacetPdf(m.roughness, -wiToward, wo);
        case 2: return 0.0f;
        case 4: return leafPdf(m.ior, etaI, m.roughness, r.trans, -wiToward, wo);
        case 6: return 1.0f;
        default: return 0.0f;
    }
}
// sample_f_eval :588-629
inline void bsdfSample(Uniforms& U, const OMaterial& m, const OF4* tex, V3 wiToward, float etaI, bool backface, int mode,
                       float u, float v, V3& wo, V3& f, float& pdf) {
    Resolved r = resolve(m, tex, u, v);
    V3 wi = -wiToward;
    switch (m.type) {
        case 0: wo = cosineSample(U); f = r.albedoTex / PI_; pdf = cosinePdf(wo); break;
        case 1: {
            V3 h = ggxH(U, m.roughness);
            wo = 2.0f * dot(wi, h) * h - wi;
            if (wo.z <= 0.0f) wo.z = -wo.z;
            f = metalF(m, wi, wo); pdf = microfacetPdf(m.roughness, wi, wo);
        } break;
        case 2: {   // dumb_smooth_dielectric_sample_f :304-369
            float eI = backface ? m.ior : 1.0f, eT = backface ? 1.0f : m.ior;
            float cosI = fminf(fmaxf(wi.z, EPS), 1.0f);
            float eta = eI / eT;
            float cosT2 = 1.0f - eta * eta * (1.0f - cosI * cosI);
            float F = schlick(cosI, eI, eT);
            if (cosT2 < 0.0f || F >= 0.99999f) { wo = v3(-wi.x, -wi.y, wi.z); f = v3(1.0f / fmaxf(wo.z, EPS)); pdf = 1.0f; }
            else if (U.next() < F) { wo = v3(-wi.x, -wi.y, wi.z); pdf = F; f = v3(F / fmaxf(wo.z, EPS)); }
            else {
                wo = v3(-eta * wi.x, -eta * wi.y, -sqrtf(cosT2));
                float fv = (1.0f - F) / fmaxf(fabsf(wo.z), EPS);
                if (mode == 1) fv *= eta * eta;
                f = v3(fv); pdf = 1.0f - F;
            }
        } break;
        case 4: {   // :512-543
            float F = schlick(wi.z, etaI, m.ior);
            if (U.next() < F) { V3 h = ggxH(U, m.roughness); wo = 2.0f * dot(wi, h) * h - wi; }
            else { bool through = U.next() < r.trans; wo = cosineSample(U); if (through) wo.z = -wo.z; }
            f = leafF(r.albedoTex, m.ior, etaI, m.roughness, r.trans, wi, wo);
            pdf = leafPdf(m.ior, etaI, m.roughness, r.trans, wi, wo);
        } break;
        case 6: wo = v3(-wi.x, -wi.y, wi.z); f = v3(1.0f / fmaxf(wo.z, EPS)); pdf = 1.0f; break;
        default: wo = v3(0, 0, 1); f = v3(0.0f); pdf = 0.0f; break;
    }
}

}  // namespace orc
