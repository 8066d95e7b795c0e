// bsdf.cuh — BSDF evaluation / sampling / pdf in the local shading frame.
// Restates the behaviour of the reference's reflectors.cuh (dispatch :547-666) including the
// quirks results depend on:
//   * direction convention: `wi` arrives pointing TOWARD the surface and is negated inside (:545-546)
//   * f_eval of a diffuse surface ignores the texture, sample_f uses it (:566 vs :612)
//   * conductor Fresnel returns the s-polarised term only (:126)
//   * the live dielectric sampler assumes air outside (:304-317), Schlick Fresnel, reflection when
//     F >= 0.99999 or TIR, eta^2 radiance scaling only in RADIANCE mode (:360-367)
//   * hasTransMap looks up the ALBEDO texture window and takes .x (:558-562)
//   * pdf of a smooth dielectric is 0, of the delta mirror 1 (:652-662)
#pragma once
#include "vecmath.cuh"

namespace nv {

enum : int { MAT_DIFFUSE = 0, MAT_METAL = 1, MAT_SMOOTHDIELECTRIC = 2, MAT_MICROFACETDIELECTRIC = 3,
             MAT_LEAF = 4, MAT_FLOWER = 5, MAT_DELTAMIRROR = 6 };
enum : int { MODE_IMPORTANCE = 0, MODE_RADIANCE = 1 };          // objects.cuh:578-581

enum : uint32_t { MF_TEXTURE = 1u, MF_TRANSMAP = 2u, MF_SPECULAR = 4u, MF_BOUNDARY = 8u, MF_THINWALLED = 16u };

// compact device material (96 B, six 16 B vector loads) built from the 176 B NovumMaterial
struct __align__(16) DMat {
    float albedo[3]; int type;
    float eta[3];    float roughness;
    float k[3];      float ior;
    float absorption[3]; float transmission;
    int texStart, texW, texH; uint32_t flags;
    int priority; int pad[3];
};

struct MatView {            // registers-only view of the fields one call needs
    V3 albedo, eta, k, absorption;
    float roughness, ior, transmission;
    int type, texStart, texW, texH, priority;
    uint32_t flags;
};

NV_D MatView loadMat(const DMat* __restrict__ mats, int id) {
    const float4* p = reinterpret_cast<const float4*>(mats + id);
    float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2), d = __ldg(p + 3);
    int4 e = __ldg(reinterpret_cast<const int4*>(p + 4));
    int4 f = __ldg(reinterpret_cast<const int4*>(p + 5));
    MatView m;
    m.albedo = V3(a.x, a.y, a.z); m.type = __float_as_int(a.w);
    m.eta = V3(b.x, b.y, b.z); m.roughness = b.w;
    m.k = V3(c.x, c.y, c.z); m.ior = c.w;
    m.absorption = V3(d.x, d.y, d.z); m.transmission = d.w;
    m.texStart = e.x; m.texW = e.y; m.texH = e.z; m.flags = (uint32_t)e.w;
    m.priority = f.x;
    return m;
}

// ---- texture fetch: bilinear with wrap (reflectors.cuh:371-417) ---------------------------
NV_D int wrapi(int v, int dim) { int r = v % dim; return r < 0 ? r + dim : r; }
NV_D V3 sampleTexture(const MatView& m, const float4* __restrict__ tex, float2 uv) {
    int W = m.texW, H = m.texH;
    float fx = uv.x * W - 0.5f, fy = uv.y * H - 0.5f;
    float flx = floorf(fx), fly = floorf(fy);
    int xi = (int)flx, yi = (int)fly;
    float sx = fx - flx, sy = fy - fly;
    int x0 = wrapi(xi, W), y0 = wrapi(yi, H), x1 = wrapi(xi + 1, W), y1 = wrapi(yi + 1, H);
    V3 c00(__ldg(tex + m.texStart + y0 * W + x0)), c10(__ldg(tex + m.texStart + y0 * W + x1));
    V3 c01(__ldg(tex + m.texStart + y1 * W + x0)), c11(__ldg(tex + m.texStart + y1 * W + x1));
    V3 bot = c00 * (1.0f - sx) + c10 * sx, top = c01 * (1.0f - sx) + c11 * sx;
    return bot * (1.0f - sy) + top * sy;
}

// ---- lobes ---------------------------------------------------------------------------------
NV_D float cosinePdf(V3 wo) { return fmaxf(wo.z, kEps) / kPi; }                       // :15-18
NV_D V3 cosineSample(Rng& rng) {                                                       // :21-39
    float u1 = fminf(rng.uniform(), 1.0f - kEps);
    float u2 = rng.uniform();
    float r = sqrtf(u1), phi = 2.0f * kPi * u2;
    float s, c; sincosf(phi, &s, &c);
    return V3(r * c, r * s, sqrtf(1.0f - u1));
}
NV_D float D_GGX(V3 h, float alpha) {                                                  // :80-86
    float a2 = alpha * alpha;
    float d = h.z * h.z * (a2 - 1.0f) + 1.0f;
    return a2 / (kPi * d * d);
}
NV_D float G1_GGX(V3 v, float alpha) {                                                 // :94-103
    float c = v.z;
    float tanT = sqrtf(1.0f - c * c) / c;
    float a = 1.0f / (alpha * tanT);
    if (a < 1.6f) return (3.535f * a + 2.181f * a * a) / (1.0f + 2.276f * a + 2.577f * a * a);
    return 1.0f;
}
NV_D float fresnelRs1(float cosT, float eta, float k) {                                // :110-127, one channel
    float c2 = cosT * cosT, s2 = 1.0f - c2;
    float e2 = eta * eta, k2 = k * k;
    float t0 = e2 - k2 - s2;
    float a2b2 = sqrtf(t0 * t0 + 4.0f * e2 * k2);
    float t1 = a2b2 + c2;
    float a = sqrtf(0.5f * (a2b2 + t0));
    float t2 = 2.0f * cosT * a;
    return (t1 - t2) / (t1 + t2);
}
NV_D float schlick(float cosT, float etaI, float etaT) {                               // :180-185
    float r0 = (etaI - etaT) / (etaI + etaT);
    r0 *= r0;
    return r0 + (1.0f - r0) * powf(1.0f - fabsf(cosT), 5.0f);
}
NV_D V3 metalF(const MatView& m, V3 wi, V3 wo) {                                       // :129-151
    if (wi.z <= 0.0f || wo.z <= 0.0f) return V3(0.0f);
    V3 h = normalize(wi + wo);
    if (h.z <= 0.0f) h = -h;
    float alpha = m.roughness * m.roughness;
    float D = D_GGX(h, alpha);
    float G = G1_GGX(wi, alpha) * G1_GGX(wo, alpha);
    float c = dot(wi, h);
    V3 F(fresnelRs1(c, m.eta.x, m.k.x), fresnelRs1(c, m.eta.y, m.k.y), fresnelRs1(c, m.eta.z, m.k.z));
    return (D * G * F) / fmaxf(4.0f * wi.z * wo.z, kEps);
}
NV_D float microfacetPdf(float roughness, V3 wi, V3 wo) {                              // :153-159
    V3 h = normalize(wi + wo);
    float D = D_GGX(h, roughness * roughness);
    return (D * h.z) / (4.0f * dot(wo, h));
}
NV_D V3 ggxHalfVector(Rng& rng, float roughness) {                                     // :164-175
    float u1 = rng.uniform();
    float alpha = roughness * roughness;
    float phi = 2.0f * kPi * rng.uniform();
    float cosT = sqrtf((1.0f - u1) / (1.0f + (alpha * alpha - 1.0f) * u1));
    float sinT = sqrtf(fmaxf(1.0f - cosT * cosT, 0.0f));
    float s, c; sincosf(phi, &s, &c);
    return V3(sinT * c, sinT * s, cosT);
}
NV_D V3 leafF(V3 albedo, float ior, float currIOR, float roughness, float trans, V3 wi, V3 wo) {   // :420-466
    bool refl = wo.z * wi.z > 0.0f;
    float F = schlick(wi.z, currIOR, ior);
    if (refl) {
        V3 h = normalize(wi + wo);
        float mF = schlick(dot(wi, h), currIOR, ior);
        if (h.z <= 0.0f) h = -h;
        float alpha = roughness * roughness;
        float D = D_GGX(h, alpha);
        float G = G1_GGX(wi, alpha) * G1_GGX(wo, alpha);
        float cut = D * G * mF / fmaxf(4.0f * wi.z * wo.z, kEps);
        return (1.0f - mF) * (1.0f - trans) * (albedo / kPi) + V3(cut);
    }
    return (albedo / kPi) * (trans * (1.0f - F));
}
NV_D float leafPdf(float ior, float currIOR, float roughness, float trans, V3 wi, V3 wo) {         // :468-510
    bool refl = wo.z * wi.z > 0.0f;
    float F = schlick(fabsf(wi.z), currIOR, ior);
    F = fminf(F, 1.0f - 0.1f * roughness);
    if (refl) {
        V3 h = normalize(wi + wo);
        if (h.z < 0.0f) h = -h;
        float D = D_GGX(h, roughness * roughness);
        float pc = (D * h.z) / (4.0f * dot(wo, h));
        return F * pc + (1.0f - F) * (1.0f - trans) * cosinePdf(wo);
    }
    return cosinePdf(-wo) * (1.0f - F) * trans;
}

// resolved per-hit material inputs (texture lookups done once; the reference repeats them per call)
struct Shading {
    MatView m;
    V3 texAlbedo;      // albedo used by sample_f / leaf (texture if present)
    float trans;       // transmission (texture .x if hasTransMap)
};
NV_D Shading resolveShading(const DMat* __restrict__ mats, int id, const float4* __restrict__ tex, float2 uv) {
    Shading s;
    s.m = loadMat(mats, id);
    s.texAlbedo = s.m.albedo;
    s.trans = s.m.transmission;
    if (s.m.flags & (MF_TEXTURE | MF_TRANSMAP)) {
        V3 t = sampleTexture(s.m, tex, uv);
        if (s.m.flags & MF_TEXTURE) s.texAlbedo = t;
        if (s.m.flags & MF_TRANSMAP) s.trans = t.x;
    }
    return s;
}

// f_eval (reflectors.cuh:547-586). wiToward points toward the surface.
NV_D V3 bsdfEval(const Shading& s, V3 wiToward, V3 wo, float etaI) {
    switch (s.m.type) {
        case MAT_DIFFUSE: return s.m.albedo / kPi;
        case MAT_METAL: return metalF(s.m, -wiToward, wo);
        case MAT_LEAF: return leafF(s.texAlbedo, s.m.ior, etaI, s.m.roughness, s.trans, -wiToward, wo);
        case MAT_DELTAMIRROR: return V3(1.0f / fmaxf(wo.z, kEps));
        default: return V3(0.0f);   // dielectric: left unassigned in the reference, never used (:575-578)
    }
}

// pdf_eval (reflectors.cuh:631-666)
NV_D float bsdfPdf(const Shading& s, V3 wiToward, V3 wo, float etaI) {
    switch (s.m.type) {
        case MAT_DIFFUSE: return cosinePdf(wo);
        case MAT_METAL: return microfacetPdf(s.m.roughness, -wiToward, wo);
        case MAT_SMOOTHDIELECTRIC: return 0.0f;
        case MAT_LEAF: return leafPdf(s.m.ior, etaI, s.m.roughness, s.trans, -wiToward, wo);
        case MAT_DELTAMIRROR: return 1.0f;
        default: return 0.0f;
    }
}

// sample_f_eval (reflectors.cuh:588-629). Outputs wo (local), f, pdf.
NV_D void bsdfSample(Rng& rng, const Shading& s, V3 wiToward, float etaI, bool backface, int mode,
                     V3& wo, V3& f, float& pdf) {
    V3 wi = -wiToward;
    switch (s.m.type) {
        case MAT_DIFFUSE: {
            wo = cosineSample(rng);
            f = s.texAlbedo / kPi;
            pdf = cosinePdf(wo);
        } break;
        case MAT_METAL: {                                                                // :161-178
            V3 h = ggxHalfVector(rng, s.m.roughness);
            wo = 2.0f * dot(wi, h) * h - wi;
            if (wo.z <= 0.0f) wo.z = -wo.z;
            f = metalF(s.m, wi, wo);
            pdf = microfacetPdf(s.m.roughness, wi, wo);
        } break;
        case MAT_SMOOTHDIELECTRIC: {                                                     // :304-369
            float eI = backface ? s.m.ior : 1.0f, eT = backface ? 1.0f : s.m.ior;
            float cosI = fminf(fmaxf(wi.z, kEps), 1.0f);
            float eta = eI / eT;
            float cosT2 = 1.0f - eta * eta * (1.0f - cosI * cosI);
            float F = schlick(cosI, eI, eT);
            if (cosT2 < 0.0f || F >= 0.99999f) {
                wo = V3(-wi.x, -wi.y, wi.z);
                f = V3(1.0f / fmaxf(wo.z, kEps));
                pdf = 1.0f;
            } else if (rng.uniform() < F) {
                wo = V3(-wi.x, -wi.y, wi.z);
                pdf = F;
                f = V3(F / fmaxf(wo.z, kEps));
            } else {
                wo = V3(-eta * wi.x, -eta * wi.y, -sqrtf(cosT2));
                float fv = (1.0f - F) / fmaxf(fabsf(wo.z), kEps);
                if (mode == MODE_RADIANCE) fv *= eta * eta;
                f = V3(fv);
                pdf = 1.0f - F;
            }
        } break;
        case MAT_LEAF: {                                                                 // :512-543
            float F = schlick(wi.z, etaI, s.m.ior);
            if (rng.uniform() < F) {
                V3 h = ggxHalfVector(rng, s.m.roughness);
                wo = 2.0f * dot(wi, h) * h - wi;
            } else {
                bool through = rng.uniform() < s.trans;
                wo = cosineSample(rng);
                if (through) wo.z = -wo.z;
            }
            f = leafF(s.texAlbedo, s.m.ior, etaI, s.m.roughness, s.trans, wi, wo);
            pdf = leafPdf(s.m.ior, etaI, s.m.roughness, s.trans, wi, wo);
        } break;
        case MAT_DELTAMIRROR: {                                                          // :69-75
            wo = V3(-wi.x, -wi.y, wi.z);
            f = V3(1.0f / fmaxf(wo.z, kEps));
            pdf = 1.0f;
        } break;
        default: {   // no dispatch exists for MICROFACETDIELECTRIC / FLOWER in the reference: terminate the path
            wo = V3(0.f, 0.f, 1.f); f = V3(0.0f); pdf = 0.0f;
        } break;
    }
}

// ---- 16 x u8 medium stack held in four registers -----------------------------------------------------------
struct MediumStack {
    unsigned w[4]; int top;
    NV_D int get(int i) const {
        unsigned word = (i >> 2) == 0 ? w[0] : ((i >> 2) == 1 ? w[1] : ((i >> 2) == 2 ? w[2] : w[3]));
        return (int)((word >> ((i & 3) * 8)) & 0xffu);
    }
    NV_D void set(int i, int v) {
        unsigned sh = (i & 3) * 8, m = ~(0xffu << sh), b = ((unsigned)v & 0xffu) << sh;
        switch (i >> 2) {
            case 0: w[0] = (w[0] & m) | b; break;
            case 1: w[1] = (w[1] & m) | b; break;
            case 2: w[2] = (w[2] & m) | b; break;
            default: w[3] = (w[3] & m) | b; break;
        }
    }
    NV_D void push(int v) { if (top < 16) { set(top, v); ++top; } }
    NV_D void remove(int mat) {           // integratorUtilities.cuh:414-434: searches down to index 1, never slot 0
        int found = -1;
        for (int i = top - 1; i > 0; --i) if (get(i) == mat) { found = i; break; }
        if (found != -1) { for (int i = found; i < top - 1; ++i) set(i, get(i + 1)); --top; }
    }
};


}  // namespace nv
