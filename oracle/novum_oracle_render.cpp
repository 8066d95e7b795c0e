// novum_oracle_render.cpp — CPU restatement of the reference's unidirectional integrators
// (Li_unidirectional deviceCode.cu:285-542, Li_naive_unidirectional :158-205, nextEventEstimation :87-156,
// neePDF :63-85, Camera::generateCameraRay objects.cuh:268-307). TEST INFRASTRUCTURE ONLY.
#include "novum_oracle.h"
#include "novum_oracle_bsdf.h"
#include <vector>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

struct OScene {   // must match novum_oracle.cpp
    std::vector<OF4> pos, nrm, tex;
    std::vector<OF2> uv;
    std::vector<OTriangle> tris, lights;
    std::vector<OMaterial> mats;
    std::vector<OBVHnode> nodes;
    std::vector<int> indices;
    std::vector<OF4> cent, bmin, bmax;
    int largestLeaf, backup;
};
OHit oracle_closest_first(const OScene& s, const ORay& r, float max_t, int skip);
void oracle_shadow_one(const OScene& s, const ORay& r, float max_t, int skip, float thr[3]);

using namespace orc;

namespace {

struct Isect {   // objects.cuh:550-568
    V3 point, normal, emission; float u, v; int tri, mat; bool valid, backface; float dist;
};

inline ORay mkRay(V3 o, V3 d) { ORay r; r.origin = {o.x, o.y, o.z, 0}; r.direction = {d.x, d.y, d.z, 0}; return r; }

// integratorUtilities.cuh:84-142
Isect sceneIntersect(const OScene& s, V3 o, V3 d, uint64_t* rays) {
    Isect it; it.valid = false; it.tri = -1; it.u = it.v = -1.0f; it.backface = false;
    if (rays) rays[0]++;
    OHit h = oracle_closest_first(s, mkRay(o, d), 999999.0f, -1);
    if (h.tri < 0) return it;
    const OTriangle& t = s.tris[h.tri];
    float w = 1.0f - h.u - h.v;
    it.point = o + h.t * d;
    V3 n = normalize(v3(s.nrm[t.naInd]) * w + v3(s.nrm[t.nbInd]) * h.u + v3(s.nrm[t.ncInd]) * h.v);
    it.u = s.uv[t.uvaInd].x * w + s.uv[t.uvbInd].x * h.u + s.uv[t.uvcInd].x * h.v;
    it.v = s.uv[t.uvaInd].y * w + s.uv[t.uvbInd].y * h.u + s.uv[t.uvcInd].y * h.v;
    if (dot(n, d) > 0.0f) { n = -n; it.backface = true; }
    it.normal = n; it.mat = t.materialID; it.emission = v3(t.emission); it.valid = true; it.tri = h.tri; it.dist = h.t;
    return it;
}

// objects.cuh:268-307
void cameraRay(const OCamera& c, Uniforms& U, int x, int y, V3& o, V3& d) {
    float aspect = (float)c.w / (float)c.h;
    float jx = (U.next() - 0.5f) * c.antiAliasJitterDist;
    float jy = (U.next() - 0.5f) * c.antiAliasJitterDist;
    float u = (2.0f * ((x + jx) / (float)c.w) - 1.0f) * aspect * c.fovScale;
    float v = (2.0f * ((y + jy) / (float)c.h) - 1.0f) * c.fovScale;
    V3 org = v3(c.cameraOrigin), right = v3(c.right), up = v3(c.up), fwd = v3(c.forward);
    V3 focal = org + (right * (u * c.focalDist)) + (up * (v * c.focalDist)) + (fwd * c.focalDist);
    V3 lens = v3(0.0f);
    if (c.aperture > 0.0f) {
        float rr = U.next();
        float theta = 2.0f * 3.141592f * U.next();
        float radius = c.aperture * sqrtf(rr);
        lens = (right * (radius * cosf(theta))) + (up * (radius * sinf(theta)));
    }
    o = org + lens;
    d = normalize(focal - o);
}

// deviceCode.cu:63-85
float neePdf(const OScene& s, int lightNum, int lightTri, const Isect& prev, const Isect& cur) {
    const OTriangle& l = s.tris[lightTri];
    V3 a = v3(s.pos[l.aInd]), b = v3(s.pos[l.bInd]), c = v3(s.pos[l.cInd]);
    V3 stl = cur.point - prev.point;
    V3 wi = normalize(stl);
    float d2 = lenSq(stl);
    V3 ln = v3(s.nrm[l.naInd]);
    float cosL = dot(ln, -wi);
    float area = 0.5f * len(cross(b - a, c - a));
    return d2 / (cosL * lightNum * area);
}

// deviceCode.cu:87-156. Returns light_pdf; contribution and local direction by reference.
float nee(const OScene& s, Uniforms& U, const Isect& it, V3 woTowardLocal, float etaI, V3& contribution, V3& toLightLocal,
          uint64_t* rays) {
    contribution = v3(0.0f);
    int lightNum = (int)s.lights.size();
    if (lightNum == 0) return -1.0f;
    float lightPdf = EPS;
    int index = (int)(U.next() * lightNum); if (index > lightNum - 1) index = lightNum - 1;
    const OTriangle& l = s.lights[index];
    V3 a = v3(s.pos[l.aInd]), b = v3(s.pos[l.bInd]), c = v3(s.pos[l.cInd]);
    float u = sqrtf(U.next());
    float v = U.next();
    V3 p = (1.0f - u) * a + (u * (1.0f - v)) * b + (u * v) * c;
    V3 stl = p - it.point;
    V3 wi = normalize(stl);
    ORay r = mkRay(it.point + wi * EPS, wi);
    float tu, tv, t = 0.f;
    oracle_triangle_intersect(s.pos.data(), &l, &r, &tu, &tv, &t);   // leaves t uninitialised in the reference when it misses
    float thr[3];
    if (rays) rays[1]++;
    oracle_shadow_one(s, r, t * (1.0f - EPS), -1, thr);
    if (thr[0] * thr[0] + thr[1] * thr[1] + thr[2] * thr[2] > 0.0f) {
        float d2 = lenSq(stl);
        V3 ln = v3(s.nrm[l.naInd]);
        float cosL = dot(ln, -wi);
        float cosS = fabsf(dot(it.normal, wi));
        float area = 0.5f * len(cross(b - a, c - a));
        lightPdf = d2 / (cosL * lightNum * area);
        toLightLocal = toLocal(wi, it.normal);
        V3 f = bsdfEval(s.mats[it.mat], s.tex.data(), woTowardLocal, toLightLocal, etaI, it.u, it.v);
        contribution = f * v3(l.emission) * cosS / lightPdf;
        contribution = contribution * v3(thr[0], thr[1], thr[2]);
    }
    return lightPdf;
}

// integratorUtilities.cuh:414-434 (note: never removes slot 0)
void removeFromStack(int* stack, int* top, int mat) {
    int found = -1;
    for (int i = (*top) - 1; i > 0; i--) if (stack[i] == mat) { found = i; break; }
    if (found != -1) { for (int i = found; i < (*top) - 1; i++) stack[i] = stack[i + 1]; (*top)--; }
}

// deviceCode.cu:285-542
V3 liUnidirectional(const OScene& s, const OCamera& cam, Uniforms& U, int x, int y, int maxDepth, bool useMIS, uint64_t* rays) {
    const OMaterial* M = s.mats.data();
    V3 beta = v3(1.0f), Li = v3(0.0f), wiLocal = v3(0.0f), woLocal = v3(0.0f);
    Isect prevReal; memset(&prevReal, 0, sizeof(prevReal));
    int stack[16]; int top = 0; stack[top++] = 0;
    V3 ro, rd;
    cameraRay(cam, U, x, y, ro, rd);
    float pdf = EPS, etaI = EPS, etaT = EPS;
    bool hitFirstNonSpecular = false;
    int lightNum = (int)s.lights.size();
    for (int depth = 0; depth < 100; depth++) {
        Isect it = sceneIntersect(s, ro, rd, rays);
        if (!it.valid) break;                                   // sampleSky returns black (integratorUtilities.cuh:438)
        int mat = it.mat;
        wiLocal = toLocal(rd, it.normal);
        bool isSpecular = M[mat].isSpecular != 0;
        bool trueHit = true;
        int minPrior = M[stack[0]].priority, minPriorID = stack[0];
        for (int i = 1; i < top; i++) if (M[stack[i]].priority < minPrior) { minPrior = M[stack[i]].priority; minPriorID = stack[i]; }
        V3 ab = v3(M[minPriorID].absorption);
        if (it.dist > EPS) beta = beta * v3(expf(-ab.x * it.dist), expf(-ab.y * it.dist), expf(-ab.z * it.dist));
        if (M[mat].boundary) {
            if (M[mat].priority <= minPrior) {
                if (M[mat].type == 2) {
                    etaI = M[minPriorID].ior;
                    if (!it.backface) etaT = M[mat].ior;
                    else if (top == 1) etaT = 1.0f;
                    else {
                        minPrior = 99; int second = stack[0];
                        for (int i = 0; i < top; i++) {
                            if (M[stack[i]].priority) {
                                if (minPrior > M[stack[i]].priority && stack[i] != mat) { second = stack[i]; minPrior = M[stack[i]].priority; }
                            }
                        }
                        etaT = M[second].ior;
                    }
                }
            } else {
                trueHit = false;
                if (!it.backface) { if (top < 16) stack[top++] = mat; }
                else removeFromStack(stack, &top, mat);
            }
        } else etaI = M[minPriorID].ior;
        if (trueHit) {
            if (lenSq(it.emission) > EPS) {
                if (depth == 0 || !hitFirstNonSpecular) Li = Li + beta * it.emission;
                else if (useMIS && !isSpecular) {
                    float lp = neePdf(s, lightNum, it.tri, prevReal, it);
                    if (lp > EPS) { float w = pdf * pdf / (lp * lp + pdf * pdf); Li = Li + beta * it.emission * w; }
                }
            }
            if (useMIS && lenSq(it.emission) < EPS && !isSpecular) {
                V3 contrib;
                float lp = nee(s, U, it, wiLocal, etaI, contrib, woLocal, rays);
                if (lp > EPS) {
                    pdf = bsdfPdf(M[mat], s.tex.data(), wiLocal, woLocal, etaI, it.u, it.v);
                    float w = lp * lp / (pdf * pdf + lp * lp);
                    Li = Li + beta * contrib * w;
                }
            }
            V3 f;
            bsdfSample(U, M[mat], s.tex.data(), wiLocal, etaI, it.backface, 1, it.u, it.v, woLocal, f, pdf);
            V3 woWorld = toWorld(woLocal, it.normal);
            pdf = fmaxf(pdf, 0.01f);
            if (woLocal.z < 0.0f) {
                if (!it.backface) { if (top < 16) stack[top++] = mat; }
                else removeFromStack(stack, &top, mat);
            }
            beta = beta * (f * fabsf(woLocal.z) / pdf);
            ro = (woLocal.z > 0) ? it.point + it.normal * EPS : it.point - it.normal * EPS;
            rd = normalize(woWorld);
            prevReal = it;
        } else {
            woLocal = toLocal(rd, it.normal);
            ro = it.point + rd * RAY_EPS;
            depth--;
        }
        if (depth > maxDepth) {
            float lum = 0.2126f * beta.x + 0.7152f * beta.y + 0.0722f * beta.z;
            float p = lum < 0.05f ? 0.05f : (lum > 0.99f ? 0.99f : lum);
            if (U.next() > p) break;
            beta = beta / p;
        }
        if (!isSpecular) hitFirstNonSpecular = true;
    }
    return Li;
}

// deviceCode.cu:158-205
V3 liNaive(const OScene& s, const OCamera& cam, Uniforms& U, int x, int y, int maxDepth, uint64_t* rays) {
    V3 ro, rd;
    cameraRay(cam, U, x, y, ro, rd);
    V3 Li = v3(0.0f), beta = v3(1.0f);
    for (int depth = 0; depth < maxDepth; depth++) {
        Isect it = sceneIntersect(s, ro, rd, rays);
        if (!it.valid) break;
        V3 toSurf = toLocal(rd, it.normal), toNext, f; float pdf;
        bsdfSample(U, s.mats[it.mat], s.tex.data(), toSurf, 1.0f, it.backface, 1, it.u, it.v, toNext, f, pdf);
        if (pdf <= 0.0f || lenSq(f) < EPS) break;
        Li = Li + it.emission * beta;
        beta = beta * (f * fabsf(toNext.z) / pdf);
        ro = it.point + ((toNext.z > 0.0f) ? (it.normal * RAY_EPS) : (-it.normal * RAY_EPS));
        rd = toWorld(toNext, it.normal);
    }
    return Li;
}

}  // namespace

extern "C" {

void oracle_render_unidirectional(const OScene* s, const OCamera* cam, int maxDepth, int useMIS, int naive,
                                  int firstSample, int numSample, uint64_t seed, OF4* colors, int threads, uint64_t* rayCounts) {
#ifdef _OPENMP
    omp_set_num_threads(threads > 0 ? threads : omp_get_num_procs());
#endif
    int w = cam->w, h = cam->h;
    uint64_t c0 = 0, c1 = 0;
#pragma omp parallel for schedule(dynamic, 1) reduction(+ : c0, c1)
    for (int y = 0; y < h; y++) {
        uint64_t rc[2] = {0, 0};
        for (int x = 0; x < w; x++) {
            int pix = y * w + x;
            V3 acc = v3(0.0f);
            for (int sidx = firstSample; sidx < firstSample + numSample; sidx++) {
                Rng rng; rng.seed(seed, (uint64_t)pix, (uint64_t)sidx, 0);
                Uniforms U = {&rng, nullptr, 0};
                V3 L = naive ? liNaive(*s, *cam, U, x, y, maxDepth, rc) : liUnidirectional(*s, *cam, U, x, y, maxDepth, useMIS != 0, rc);
                acc = acc + L;
            }
            colors[pix].x += acc.x; colors[pix].y += acc.y; colors[pix].z += acc.z;
        }
        c0 += rc[0]; c1 += rc[1];
    }
    if (rayCounts) { rayCounts[0] += c0; rayCounts[1] += c1; }
}

void oracle_bsdf_eval(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], const float wo[3], float etaI,
                      float u, float v, float f[3]) {
    V3 r = bsdfEval(mats[id], tex, v3(wiToward[0], wiToward[1], wiToward[2]), v3(wo[0], wo[1], wo[2]), etaI, u, v);
    f[0] = r.x; f[1] = r.y; f[2] = r.z;
}
float oracle_bsdf_pdf(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], const float wo[3], float etaI,
                      float u, float v) {
    return bsdfPdf(mats[id], tex, v3(wiToward[0], wiToward[1], wiToward[2]), v3(wo[0], wo[1], wo[2]), etaI, u, v);
}
int oracle_bsdf_sample(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], float etaI, int backface,
                       int mode, float u, float v, const float* rng, float wo[3], float f[3], float* pdf) {
    Uniforms U = {nullptr, rng, 0};
    V3 o, ff; float p;
    bsdfSample(U, mats[id], tex, v3(wiToward[0], wiToward[1], wiToward[2]), etaI, backface != 0, mode, u, v, o, ff, p);
    wo[0] = o.x; wo[1] = o.y; wo[2] = o.z; f[0] = ff.x; f[1] = ff.y; f[2] = ff.z; *pdf = p;
    return U.used;
}

}  // extern "C"
