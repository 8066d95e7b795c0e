// scene_dev.cuh — device-side view of a scene and the two scene queries every integrator uses:
// surface interaction reconstruction after a closest hit, and the attenuating shadow ray.
#pragma once
#include "bvh8.cuh"
#include "bsdf.cuh"
#include "../../include/novum_b200.h"

namespace nv {

struct SceneView {
    Bvh8View bvh;
    const float4* __restrict__ positions;
    const float4* __restrict__ normals;
    const float2* __restrict__ uvs;
    const NovumTriangle* __restrict__ tris;     // reference layout, 80 B
    const float4* __restrict__ shadeTris;       // 6 x 16 B per triangle: vertex normals + uvs + material inlined (ShadeTri)
    const NovumTriangle* __restrict__ lights;   // light list (copies of the emissive triangles)
    const DMat* __restrict__ mats;
    const float4* __restrict__ textures;
    int numTris, numLights, numMats;
    float sceneMin[3];
    float sceneRadius;
    float sceneCenter[3];
    // opt-in features beyond what the reference implements (novum_scene_set_features; all off by default)
    int features;
    float env[3];                               // constant environment radiance (FEAT_ENVIRONMENT)
};
enum : int { FEAT_NESTED_BIDIR = 1, FEAT_ENVIRONMENT = 2 };

// what BVHSceneIntersect fills into `Intersection` (objects.cuh:550-568), minus the unused colour
struct Surface {
    V3 p, n;            // hit point r.at(t); shading normal flipped to face the ray
    V3 emission;
    float2 uv;
    float dist;
    int tri, mat, lightInd;
    bool backface;
};

struct TriRecord { int a, b, c, na, nb, nc, uva, uvb, uvc, mat, lightInd; V3 emission; };
NV_D TriRecord loadTri(const NovumTriangle* __restrict__ tris, int id) {
    const int4* p = reinterpret_cast<const int4*>(tris + id);
    int4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
    float4 em = __ldg(reinterpret_cast<const float4*>(p + 3));
    int4 q4 = __ldg(p + 4);
    TriRecord t;
    t.a = q0.x; t.b = q0.y; t.c = q0.z; t.na = q0.w;
    t.nb = q1.x; t.nc = q1.y; t.uva = q1.z; t.uvb = q1.w;
    t.uvc = q2.x; t.mat = q2.y;
    t.emission = V3(em.x, em.y, em.z);
    t.lightInd = q4.x;
    return t;
}

// Shading record of a triangle: what reconstructing a surface interaction needs, gathered once at scene build so that a hit
// costs ONE dependent fetch of 96 contiguous bytes instead of the 80 B reference triangle followed by six scattered
// normal / uv gathers (two dependent levels, 9 sectors):
//   {na.xyz, nb.x} {nb.yz, nc.xy} {nc.z, uva.xy, uvb.x} {uvb.y, uvc.xy, material} {emission.xyz, lightInd} {spare}
constexpr int kShadeTriWords = 6;
NV_D void packShadeTri(float4* out, V3 na, V3 nb, V3 nc, float2 ua, float2 ub, float2 uc, int mat, V3 em, int lightInd) {
    out[0] = make_float4(na.x, na.y, na.z, nb.x);
    out[1] = make_float4(nb.y, nb.z, nc.x, nc.y);
    out[2] = make_float4(nc.z, ua.x, ua.y, ub.x);
    out[3] = make_float4(ub.y, uc.x, uc.y, __int_as_float(mat));
    out[4] = make_float4(em.x, em.y, em.z, __int_as_float(lightInd));
    out[5] = make_float4(0.f, 0.f, 0.f, 0.f);
}

// integratorUtilities.cuh:111-142: attributes interpolated with (w=1-u-v on a, u on b, v on c),
// normal normalised then flipped against the ray, backface flag.
NV_D Surface makeSurface(const SceneView& sc, const HitRec& h, V3 o, V3 d) {
    Surface s;
    const float4* p = sc.shadeTris + (size_t)h.tri * kShadeTriWords;
    const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2), q3 = __ldg(p + 3), q4 = __ldg(p + 4);
    float w = 1.0f - h.u - h.v;
    s.p = o + h.t * d;
    V3 n = V3(q0.x, q0.y, q0.z) * w + V3(q0.w, q1.x, q1.y) * h.u + V3(q1.z, q1.w, q2.x) * h.v;
    n = normalize(n);
    s.uv = make_float2(q2.y * w + q2.w * h.u + q3.y * h.v, q2.z * w + q3.x * h.u + q3.z * h.v);
    s.backface = dot(n, d) > 0.0f;
    s.n = s.backface ? -n : n;
    s.mat = __float_as_int(q3.w); s.emission = V3(q4.x, q4.y, q4.z); s.tri = h.tri; s.dist = h.t; s.lightInd = __float_as_int(q4.w);
    return s;
}

// Shadow ray with the reference's semantics (integratorUtilities.cuh:188-288): any hit with
// t < maxT on a triangle other than skipTri blocks the ray, except MAT_LEAF surfaces which
// attenuate by albedo*transmission*(1-Schlick) and let it continue; below 1 % it is black.
// Traversal order is irrelevant to the result (every factor is <= 1), so the nearest-first ordering
// the reference pays for is dropped and the first blocking hit ends the lane immediately.
// Resumable like ClosestTrav so persistent warps can refill finished lanes (trav_stream.cuh).
template <bool COUNT>
struct ShadowTrav {
    RaySetup r;
    V3 thr;
    float maxT;
    int skipTri;
    uint2 ngroup, tgroup;
    unsigned nv, tt;

    NV_D void begin(V3 o, V3 d, float maxT_, int skip, TravStack& st) {
        r = setupRay(o, d);
        thr = V3(1.0f); maxT = maxT_; skipTri = skip;
        st.top = 0;
        ngroup = make_uint2(0u, 0x80000000u);
        tgroup = make_uint2(0u, 0u);
        nv = 0; tt = 0;
    }
    NV_D bool step(const SceneView& sc, TravStack& st) {
        const Bvh8View& bvh = sc.bvh;
        if (ngroup.y > 0x00ffffffu) {
            unsigned hits = ngroup.y;
            unsigned bit = 31u - __clz(hits);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00ffffffu) st.push(ngroup);
            unsigned slot = (bit - 24u) ^ (r.octinv4 & 0xffu);
            unsigned rel = __popc(hits & ~(0xffffffffu << slot) & 0xffu);
            const float4* np = bvh.nodes + (size_t)(ngroup.x + rel) * 5;
            float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            if (COUNT) ++nv;
            unsigned hm = intersectNode(r, maxT, n0, n1, n2, n3, n4);
            ngroup.x = __float_as_uint(n1.x);
            tgroup.x = __float_as_uint(n1.y);
            ngroup.y = (hm & 0xff000000u) | (__float_as_uint(n0.w) >> 24);
            tgroup.y = hm & 0x00ffffffu;
        } else {
            tgroup = ngroup;
            ngroup = make_uint2(0u, 0u);
        }
        while (tgroup.y != 0u) {
            if (kTriPostpone > 0 && ngroup.y > 0x00ffffffu && __popc(__activemask()) < kTriPostpone) { st.push(tgroup); tgroup.y = 0u; break; }
            unsigned ti = __ffs(tgroup.y) - 1u;
            tgroup.y &= tgroup.y - 1u;
            const float4* tp = bvh.tris + (size_t)(tgroup.x + ti) * 3;
            float4 v0 = __ldg(tp), e1 = __ldg(tp + 1), e2 = __ldg(tp + 2);
            if (COUNT) ++tt;
            int id = __float_as_int(v0.w);
            float t, u, v;
            if (id != skipTri && triTest(v0, e1, e2, r.o, r.d, t, u, v) && t < maxT) {
                int flags = __float_as_int(e1.w);
                if (flags & TF_LEAFMAT) {
                    TriRecord tr = loadTri(sc.tris, id);
                    MatView m = loadMat(sc.mats, tr.mat);
                    float w = 1.0f - u - v;
                    V3 n = V3(__ldg(sc.normals + tr.na)) * w + V3(__ldg(sc.normals + tr.nb)) * u + V3(__ldg(sc.normals + tr.nc)) * v;
                    float cosT = fabsf(dot(r.d, normalize(n)));
                    float F = schlick(cosT, 1.0f, m.ior);
                    thr *= (m.albedo * m.transmission) * (1.0f - F);     // the reference's association (integratorUtilities.cuh:232)
                    if (maxComp(thr) < 0.01f) { thr = V3(0.0f); return true; }
                } else {
                    thr = V3(0.0f); return true;
                }
            }
        }
        if (ngroup.y <= 0x00ffffffu) {
            if (st.top > 0) ngroup = st.pop();
            else return true;
        }
        return false;
    }
    NV_D void flushCounters(TravCounters cnt) const {
        if (COUNT) { atomicAdd(cnt.nodeVisits, (unsigned long long)nv); atomicAdd(cnt.triTests, (unsigned long long)tt); }
    }
};

template <bool COUNT>
NV_D V3 traceShadow(const SceneView& sc, V3 o, V3 d, float maxT, int skipTri, TravStack& st, TravCounters cnt) {
    ShadowTrav<COUNT> tr;
    tr.begin(o, d, maxT, skipTri, st);
    while (!tr.step(sc, st)) {}
    tr.flushCounters(cnt);
    return tr.thr;
}

}  // namespace nv
