// novum_oracle.cpp — CPU restatement of the reference's intersection path and BVH2 builder.
// TEST INFRASTRUCTURE ONLY (see novum_oracle.h). Built with -ffp-contract=off -mfma: every fused
// multiply-add below is explicit and mirrors the FMA pattern nvcc emits for the reference on sm_100a
// (read off the SASS of triangleIntersect: dot = fma(z,z,fma(x,x,y*y)), cross component = fma(a,b,-(c*d)),
// 1.0/a = correctly rounded float reciprocal), so t/u/v are bit-identical to the GPU reference.
#include "novum_oracle.h"
#include <math.h>
#include <float.h>
#include <string.h>
#include <stdlib.h>
#include <vector>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace {

struct V { float x, y, z; };
inline V mk(float x, float y, float z) { V v = {x, y, z}; return v; }
inline V sub(V a, V b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float dotG(V a, V b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }          // util.cuh:116 as compiled
inline V crossG(V a, V b) {                                                                 // util.cuh:133 as compiled
    return mk(fmaf(a.y, b.z, -(a.z * b.y)), fmaf(a.z, b.x, -(a.x * b.z)), fmaf(a.x, b.y, -(a.y * b.x)));
}
inline V fromF4(const OF4& f) { return mk(f.x, f.y, f.z); }

// integratorUtilities.cuh:8-42
inline bool triIsect(const OF4* pos, const OTriangle& tri, V o, V d, float& u, float& v, float& t) {
    V a = fromF4(pos[tri.aInd]), b = fromF4(pos[tri.bInd]), c = fromF4(pos[tri.cInd]);
    V e1 = sub(b, a), e2 = sub(c, a);
    V h = crossG(d, e2);
    float det = dotG(h, e1);
    if (fabsf(det) < 1e-12f) return false;
    float f = (float)(1.0 / (double)det);
    V s = sub(o, a);
    u = f * dotG(s, h);
    V q = crossG(s, e1);
    v = f * dotG(d, q);
    t = f * dotG(e2, q);
    return (u >= 0) && (v >= 0) && (u + v <= 1) && (t > 0.0f);
}

// integratorUtilities.cuh:44-82
inline bool boxIsect(V o, V d, const OF4& mn, const OF4& mx, float& tmin, float& tmax) {
    tmin = -1e30f; tmax = 1e30f;
    float ix = 1.0f / d.x, iy = 1.0f / d.y, iz = 1.0f / d.z;
    float t1 = (mn.x - o.x) * ix, t2 = (mx.x - o.x) * ix;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (mn.y - o.y) * iy; t2 = (mx.y - o.y) * iy;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    t1 = (mn.z - o.z) * iz; t2 = (mx.z - o.z) * iz;
    tmin = fmaxf(tmin, fminf(t1, t2)); tmax = fminf(tmax, fmaxf(t1, t2));
    return (tmax >= tmin) && (tmax > 0.0f);
}

inline float comp(const OF4& v, int a) { return a == 0 ? v.x : (a == 1 ? v.y : v.z); }
inline OF4 min4(const OF4& a, const OF4& b) { OF4 r = {fminf(a.x, b.x), fminf(a.y, b.y), fminf(a.z, b.z), 0}; return r; }
inline OF4 max4(const OF4& a, const OF4& b) { OF4 r = {fmaxf(a.x, b.x), fmaxf(a.y, b.y), fmaxf(a.z, b.z), 0}; return r; }
inline float sarea(const OF4& mn, const OF4& mx) {
    float dx = mx.x - mn.x, dy = mx.y - mn.y, dz = mx.z - mn.z;
    return 2.0f * (dx * dy + dx * dz + dy * dz);
}

}  // namespace

struct OScene {
    std::vector<OF4> pos, nrm, tex;
    std::vector<OF2> uv;
    std::vector<OTriangle> tris, lights;
    std::vector<OMaterial> mats;
    std::vector<OBVHnode> nodes;
    std::vector<int> indices;
    std::vector<OF4> cent, bmin, bmax;
    int largestLeaf = 0, backup = 0;
};

namespace {

// main.cu:44-58
int partitionPrims(OScene& s, int start, int end, int axis, float splitPos) {
    int mid = start;
    for (int i = start; i < end; i++) {
        if (comp(s.cent[s.indices[i]], axis) < splitPos) { std::swap(s.indices[i], s.indices[mid]); mid++; }
    }
    return mid;
}

// main.cu:60-131 (12-bucket SAH along the chosen axis; nth_element median when no bucket split is valid)
void sahSplit(OScene& s, int start, int end, int axis, const OF4& minB, const OF4& maxB, float& splitPos) {
    const int NB = 12;
    OF4 bmn[NB], bmx[NB]; int cnt[NB];
    for (int i = 0; i < NB; i++) { bmn[i] = {FLT_MAX, FLT_MAX, FLT_MAX, 0}; bmx[i] = {-FLT_MAX, -FLT_MAX, -FLT_MAX, 0}; cnt[i] = 0; }
    float lo = comp(minB, axis), hi = comp(maxB, axis);
    for (int i = start; i < end; i++) {
        int idx = s.indices[i];
        float c = comp(s.cent[idx], axis);
        int b = int(NB * (c - lo) / (hi - lo));
        b = b < 0 ? 0 : (b > NB - 1 ? NB - 1 : b);
        cnt[b]++; bmn[b] = min4(bmn[b], s.bmin[idx]); bmx[b] = max4(bmx[b], s.bmax[idx]);
    }
    float minCost = FLT_MAX; int best = -1;
    for (int i = 1; i < NB; i++) {
        OF4 lmn = bmn[0], lmx = bmx[0]; int lc = cnt[0];
        for (int j = 1; j < i; j++) { lmn = min4(lmn, bmn[j]); lmx = max4(lmx, bmx[j]); lc += cnt[j]; }
        OF4 rmn = bmn[i], rmx = bmx[i]; int rc = cnt[i];
        for (int j = i; j < NB; j++) { rmn = min4(rmn, bmn[j]); rmx = max4(rmx, bmx[j]); rc += cnt[j]; }   // sic: bucket i counted twice (main.cu:104-108)
        float cost = 1.0f + (lc * sarea(lmn, lmx) + rc * sarea(rmn, rmx)) / sarea(minB, maxB);
        if (cost < minCost && (lc > 0 && rc > 0)) { minCost = cost; best = i; }
    }
    if (best == -1) {
        int mid = (start + end) / 2;
        std::nth_element(s.indices.begin() + start, s.indices.begin() + mid, s.indices.begin() + end,
                         [&](int a, int b) { return comp(s.cent[a], axis) < comp(s.cent[b], axis); });
        splitPos = comp(s.cent[s.indices[mid]], axis);
    } else {
        splitPos = lo + (hi - lo) * (float(best) / float(NB));
    }
}

// main.cu:133-233
int buildNode(OScene& s, int start, int end, int maxLeaf) {
    int ni = (int)s.nodes.size();
    s.nodes.push_back(OBVHnode());
    OF4 mn = s.bmin[s.indices[start]], mx = s.bmax[s.indices[start]];
    for (int i = start; i < end; i++) { mn = min4(mn, s.bmin[s.indices[i]]); mx = max4(mx, s.bmax[s.indices[i]]); }
    s.nodes[ni].aabbMIN = mn; s.nodes[ni].aabbMAX = mx;
    int pc = end - start;
    auto makeLeaf = [&]() {
        s.nodes[ni].first = start; s.nodes[ni].primCount = pc; s.nodes[ni].left = s.nodes[ni].right = -1;
        s.largestLeaf = std::max(pc, s.largestLeaf);
        return ni;
    };
    if (pc <= maxLeaf) return makeLeaf();
    float xd = mx.x - mn.x, yd = mx.y - mn.y, zd = mx.z - mn.z;
    int axis = 0;
    if (yd > xd && yd > zd) axis = 1; else if (zd > xd && zd > yd) axis = 2;
    float splitPos;
    sahSplit(s, start, end, axis, mn, mx, splitPos);
    int mid = start, numLeft = 0;
    for (int i = start; i < end; i++) if (comp(s.cent[s.indices[i]], axis) < splitPos) numLeft++;
    if (numLeft > 0 && numLeft < (pc - 1)) mid = partitionPrims(s, start, end, axis, splitPos);
    else {
        float sum = 0.0f; s.backup++;
        for (int i = start; i < end; i++) sum += comp(s.cent[s.indices[i]], axis);
        splitPos = sum / pc;
    }
    numLeft = 0;
    for (int i = start; i < end; i++) if (comp(s.cent[s.indices[i]], axis) < splitPos) numLeft++;
    if (numLeft > 0 && numLeft < (pc - 1)) mid = partitionPrims(s, start, end, axis, splitPos);
    else return makeLeaf();
    int l = buildNode(s, start, mid, maxLeaf);
    s.nodes[ni].left = l;
    int r = buildNode(s, mid, end, maxLeaf);
    s.nodes[ni].right = r;
    s.nodes[ni].primCount = 0; s.nodes[ni].first = -1;
    return ni;
}

enum TieMode { FIRST_FOUND = 0, LOWEST_ID = 1 };

// integratorUtilities.cuh:84-186
template <int TIE>
OHit closestHit(const OScene& s, const ORay& ray, float max_t, int skipTri) {
    V o = fromF4(ray.origin), d = fromF4(ray.direction);
    OHit best = {-1, 0.f, 0.f, 0.f};
    float min_t = 3.402823466e+38f;
    int stack[128]; int top = 0;
    stack[top++] = 0;
    const OBVHnode* B = s.nodes.data();
    while (top > 0) {
        const OBVHnode& node = B[stack[--top]];
        if (node.primCount > 0) {
            for (int i = node.first; i < node.primCount + node.first; i++) {
                int idx = s.indices[i];
                if (idx == skipTri) continue;
                float u, v, t;
                if (triIsect(s.pos.data(), s.tris[idx], o, d, u, v, t) && t < max_t) {
                    bool take = (t < min_t) || (TIE == LOWEST_ID && t == min_t && idx < best.tri);
                    if (take) { min_t = t; best.tri = idx; best.t = t; best.u = u; best.v = v; }
                }
            }
        } else if (node.left >= 0 || node.right >= 0) {
            float tl0 = 0, tl1 = 0, tr0 = 0, tr1 = 0; bool hl = false, hr = false;
            if (node.left >= 0) hl = boxIsect(o, d, B[node.left].aabbMIN, B[node.left].aabbMAX, tl0, tl1);
            if (node.right >= 0) hr = boxIsect(o, d, B[node.right].aabbMIN, B[node.right].aabbMAX, tr0, tr1);
            if (hl && hr) {
                if (tl0 < tr0) { stack[top++] = node.right; stack[top++] = node.left; }
                else { stack[top++] = node.left; stack[top++] = node.right; }
            } else if (hl) stack[top++] = node.left;
            else if (hr) stack[top++] = node.right;
        }
    }
    return best;
}

inline float schlickO(float c, float ei, float et) {
    float r0 = (ei - et) / (ei + et); r0 *= r0;
    return r0 + (1.0f - r0) * powf(1.0f - fabsf(c), 5.0f);
}

// integratorUtilities.cuh:188-288
void shadowRay(const OScene& s, const ORay& ray, float max_t, int skip, float thr[3]) {
    V o = fromF4(ray.origin), d = fromF4(ray.direction);
    thr[0] = thr[1] = thr[2] = 1.0f;
    int stack[128]; int top = 0; stack[top++] = 0;
    const OBVHnode* B = s.nodes.data();
    while (top > 0) {
        const OBVHnode& node = B[stack[--top]];
        if (node.primCount > 0) {
            for (int i = node.first; i < node.primCount + node.first; i++) {
                int idx = s.indices[i];
                float u, v, t;
                bool hit = triIsect(s.pos.data(), s.tris[idx], o, d, u, v, t);
                if (idx == skip) continue;
                if (hit && t < max_t) {
                    const OTriangle& tri = s.tris[idx];
                    const OMaterial& m = s.mats[tri.materialID];
                    if (m.type == 4 /* MAT_LEAF */) {
                        float w = 1.0f - u - v;
                        const OF4 &na = s.nrm[tri.naInd], &nb = s.nrm[tri.nbInd], &nc = s.nrm[tri.ncInd];
                        float nx = na.x * w + nb.x * u + nc.x * v, ny = na.y * w + nb.y * u + nc.y * v, nz = na.z * w + nb.z * u + nc.z * v;
                        float il = 1.0f / sqrtf(nx * nx + ny * ny + nz * nz);
                        float c = fabsf((d.x * nx + d.y * ny + d.z * nz) * il);
                        float F = schlickO(c, 1.0f, m.ior);
                        float k = m.transmission * (1.0f - F);
                        thr[0] *= m.albedo.x * k; thr[1] *= m.albedo.y * k; thr[2] *= m.albedo.z * k;
                        if (fmaxf(thr[0], fmaxf(thr[1], thr[2])) < 0.01f) { thr[0] = thr[1] = thr[2] = 0.f; return; }
                    } else { thr[0] = thr[1] = thr[2] = 0.f; return; }
                }
            }
        } else if (node.left >= 0 || node.right >= 0) {
            float tl0 = 0, tl1 = 0, tr0 = 0, tr1 = 0; bool hl = false, hr = false;
            if (node.left >= 0) hl = boxIsect(o, d, B[node.left].aabbMIN, B[node.left].aabbMAX, tl0, tl1);
            if (node.right >= 0) hr = boxIsect(o, d, B[node.right].aabbMIN, B[node.right].aabbMAX, tr0, tr1);
            if (hl && hr) {
                if (tl0 < tr0) { stack[top++] = node.right; stack[top++] = node.left; }
                else { stack[top++] = node.left; stack[top++] = node.right; }
            } else if (hl) stack[top++] = node.left;
            else if (hr) stack[top++] = node.right;
        }
    }
}

inline void setThreads(int threads) {
#ifdef _OPENMP
    if (threads > 0) omp_set_num_threads(threads);
    else omp_set_num_threads(omp_get_num_procs());
#else
    (void)threads;
#endif
}

}  // namespace

// shared with novum_oracle_render.cpp
OHit oracle_closest_first(const OScene& s, const ORay& r, float max_t, int skip) { return closestHit<FIRST_FOUND>(s, r, max_t, skip); }
void oracle_shadow_one(const OScene& s, const ORay& r, float max_t, int skip, float thr[3]) { shadowRay(s, r, max_t, skip, thr); }

extern "C" {

int oracle_triangle_intersect(const OF4* positions, const OTriangle* tri, const ORay* r, float* u, float* v, float* t) {
    float uu = 0, vv = 0, tt = 0;
    bool h = triIsect(positions, *tri, fromF4(r->origin), fromF4(r->direction), uu, vv, tt);
    if (h) { *u = uu; *v = vv; *t = tt; } else { *u = 0; *v = 0; *t = 0; }
    return h ? 1 : 0;
}

int oracle_aabb_intersect(const ORay* r, OF4 mn, OF4 mx, float* tmin, float* tmax) {
    return boxIsect(fromF4(r->origin), fromF4(r->direction), mn, mx, *tmin, *tmax) ? 1 : 0;
}

OScene* oracle_scene_create(const OF4* positions, int nPos, const OF4* normals, int nNorm, const OF2* uvs, int nUV,
                            const OTriangle* tris, int nTri, const OMaterial* mats, int nMat, const OF4* textures,
                            int nTexels, int maxLeafSize) {
    OScene* s = new OScene();
    s->pos.assign(positions, positions + nPos);
    if (normals) s->nrm.assign(normals, normals + nNorm);
    if (uvs) s->uv.assign(uvs, uvs + nUV);
    s->tris.assign(tris, tris + nTri);
    if (mats) s->mats.assign(mats, mats + nMat);
    if (textures && nTexels > 0) s->tex.assign(textures, textures + nTexels);
    for (int i = 0; i < nTri; i++) {
        const OF4& e = tris[i].emission;
        if (e.x * e.x + e.y * e.y + e.z * e.z > 0) s->lights.push_back(tris[i]);
    }
    // main.cu:17-42 computeInfoForBVH
    s->cent.resize(nTri); s->bmin.resize(nTri); s->bmax.resize(nTri); s->indices.resize(nTri);
    for (int i = 0; i < nTri; i++) {
        const OF4 &a = s->pos[tris[i].aInd], &b = s->pos[tris[i].bInd], &c = s->pos[tris[i].cInd];
        s->cent[i] = {(a.x + b.x + c.x) / 3.0f, (a.y + b.y + c.y) / 3.0f, (a.z + b.z + c.z) / 3.0f, 0};
        s->bmin[i] = {fminf(fminf(a.x, b.x), c.x) - 0.000001f, fminf(fminf(a.y, b.y), c.y) - 0.000001f, fminf(fminf(a.z, b.z), c.z) - 0.000001f, 0};
        s->bmax[i] = {fmaxf(fmaxf(a.x, b.x), c.x) + 0.000001f, fmaxf(fmaxf(a.y, b.y), c.y) + 0.000001f, fmaxf(fmaxf(a.z, b.z), c.z) + 0.000001f, 0};
        s->indices[i] = i;
    }
    s->nodes.reserve(2 * (size_t)nTri);
    if (nTri > 0) buildNode(*s, 0, nTri, maxLeafSize < 1 ? 1 : maxLeafSize);
    return s;
}

void oracle_scene_destroy(OScene* s) { delete s; }

int oracle_scene_bvh_size(const OScene* s, int* numNodes, int* largestLeaf) {
    if (numNodes) *numNodes = (int)s->nodes.size();
    if (largestLeaf) *largestLeaf = s->largestLeaf;
    return 0;
}
int oracle_scene_bvh_copy(const OScene* s, OBVHnode* nodes, int32_t* indices) {
    if (nodes) memcpy(nodes, s->nodes.data(), s->nodes.size() * sizeof(OBVHnode));
    if (indices) memcpy(indices, s->indices.data(), s->indices.size() * sizeof(int));
    return 0;
}
void oracle_scene_bounds(const OScene* s, float mn[3], float mx[3], float* radius) {   // main.cu:531-535
    const OBVHnode& r = s->nodes[0];
    mn[0] = r.aabbMIN.x; mn[1] = r.aabbMIN.y; mn[2] = r.aabbMIN.z;
    mx[0] = r.aabbMAX.x; mx[1] = r.aabbMAX.y; mx[2] = r.aabbMAX.z;
    float c[3], d2 = 0;
    for (int k = 0; k < 3; k++) { c[k] = (mx[k] + mn[k]) * 0.5f; float d = mx[k] - c[k]; d2 += d * d; }
    if (radius) *radius = sqrtf(d2) + 0.01f;
}

void oracle_trace_closest(const OScene* s, const ORay* rays, int n, OHit* hits, int threads) {
    setThreads(threads);
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; i++) hits[i] = closestHit<FIRST_FOUND>(*s, rays[i], 999999.0f, -1);
}
void oracle_trace_closest_lowest_id(const OScene* s, const ORay* rays, int n, OHit* hits, int threads) {
    setThreads(threads);
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; i++) hits[i] = closestHit<LOWEST_ID>(*s, rays[i], 999999.0f, -1);
}
void oracle_trace_brute(const OScene* s, const ORay* rays, int n, OHit* hits, int threads) {
    setThreads(threads);
    int nt = (int)s->tris.size();
#pragma omp parallel for schedule(dynamic, 64)
    for (int i = 0; i < n; i++) {
        V o = fromF4(rays[i].origin), d = fromF4(rays[i].direction);
        OHit best = {-1, 0, 0, 0}; float mt = 3.402823466e+38f;
        for (int j = 0; j < nt; j++) {
            float u, v, t;
            if (triIsect(s->pos.data(), s->tris[j], o, d, u, v, t) && t < 999999.0f && t < mt) { mt = t; best = {j, t, u, v}; }
        }
        hits[i] = best;
    }
}
void oracle_trace_shadow(const OScene* s, const ORay* rays, const float* maxT, const int32_t* skip, int n, OF4* thr, int threads) {
    setThreads(threads);
#pragma omp parallel for schedule(dynamic, 256)
    for (int i = 0; i < n; i++) {
        float t[3];
        shadowRay(*s, rays[i], maxT[i], skip ? skip[i] : -1, t);
        thr[i] = {t[0], t[1], t[2], 0.f};
    }
}

}  // extern "C"
