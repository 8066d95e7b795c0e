// vcm.cu — kernels of the bidirectional wavefront engine: VCM, SPPM (= VCM with connections off, one merge per path) and
// BDPT (no merging, the reference's older weighting). See vcm.cuh for the data layout and the stage list.
// Bounce-by-bounce semantics follow generateVCMLightPath / doLightPass / connect* / doEyePass (deviceCode.cu:1958-3065) in
// VCM mode and generateEyePath / generateLightPath / lightPathTracing / connectPath (:751-1765) in BDPT mode, including
// the quirks results depend on: G clamps 15 (NEE, :681) and 2 (connections, :1753, :2675), the firefly clamp lum > 5
// (:1637, :2952), the squared NEE ratio in VCM (:2526) vs the plain one in BDPT (:1519), placeholder etaI = etaT = 1
// (:851, :1177, :2078), un-normalised "to previous vertex" vectors in the VCM connection code (:2493, :2629, :3004),
// every one of the 27 neighbour cells looked up even when two of them hash to the same slot (:2974-2992).
#include "vcm.cuh"
#include "scene.h"
#include <cstdio>
#include <cstring>
#include <vector>
#include <thread>

namespace nv {

constexpr int kBdThreads = 128;

__device__ __forceinline__ int bdLane() { return threadIdx.x & 31; }

__device__ __forceinline__ int bdWarpAppend(int* counter, bool pred) {
    unsigned m = __ballot_sync(0xffffffffu, pred);
    if (m == 0) return -1;
    int leader = __ffs(m) - 1;
    int base = 0;
    if (bdLane() == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    return pred ? base + __popc(m & ((1u << bdLane()) - 1u)) : -1;
}

// opportunistic aggregation for appends issued from divergent code: the lanes that happen to be converged here share one atomic
__device__ __forceinline__ int bdAggAppend(int* counter) {
    unsigned m = __activemask();
    int leader = __ffs(m) - 1;
    int base = 0;
    if (bdLane() == leader) base = atomicAdd(counter, __popc(m));
    base = __shfl_sync(m, base, leader);
    return base + __popc(m & ((1u << bdLane()) - 1u));
}

NV_D float rcpApprox(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
NV_D uint32_t packHalf2(float a, float b) { __half2 h = __halves2half2(__float2half(a), __float2half(b)); return *reinterpret_cast<uint32_t*>(&h); }
NV_D float2 unpackHalf2(uint32_t u) { __half2 h = *reinterpret_cast<__half2*>(&u); return __half22float2(h); }

// BSDF value / pdf for the connection code. The Lambert case is resolved inline; every other material goes through ONE
// out-of-line copy of the full dispatch. Inlined at each of the ~15 call sites the dispatch made bdEyeStep 145 KB of SASS,
// and the kernel stalled on instruction fetch (no_instruction 5.9-10 per issue, profiles/r1_bdEyeStep_inlined.txt).
__device__ __noinline__ V3 bsdfEvalCall(const Shading& s, V3 wiToward, V3 wo, float etaI) { return bsdfEval(s, wiToward, wo, etaI); }
__device__ __noinline__ float bsdfPdfCall(const Shading& s, V3 wiToward, V3 wo, float etaI) { return bsdfPdf(s, wiToward, wo, etaI); }
#ifdef NOVUM_BD_INLINE_BSDF          // A/B build switch: the dispatch inlined at every call site
NV_D V3 bdEval(const Shading& s, V3 wiToward, V3 wo, float etaI) { return bsdfEval(s, wiToward, wo, etaI); }
NV_D float bdPdf(const Shading& s, V3 wiToward, V3 wo, float etaI) { return bsdfPdf(s, wiToward, wo, etaI); }
#else
NV_D V3 bdEval(const Shading& s, V3 wiToward, V3 wo, float etaI) {
    return s.m.type == MAT_DIFFUSE ? s.m.albedo * (1.0f / kPi) : bsdfEvalCall(s, wiToward, wo, etaI);
}
NV_D float bdPdf(const Shading& s, V3 wiToward, V3 wo, float etaI) {
    return s.m.type == MAT_DIFFUSE ? cosinePdf(wo) : bsdfPdfCall(s, wiToward, wo, etaI);
}
#endif

struct LightVtx { V3 p, n, wo, beta; float2 uv; bool delta, backface; int mat; float d_vc, d_vcm; };

// quantises v in place to what the cache holds and writes the three words
NV_D void storeLightVtx(const BdState& st, size_t idx, LightVtx& v) {
    uint32_t pn = packOct(v.n), pw = packOct(v.wo);
    uint32_t bxy = packHalf2(v.beta.x, v.beta.y), bz = packHalf2(v.beta.z, 0.0f), puv = packHalf2(v.uv.x, v.uv.y);
    uint32_t info = (v.delta ? 1u : 0u) | (v.backface ? 2u : 0u) | (0xFFFFEu << 2) | (((uint32_t)v.mat & 0x3FFu) << 22);   // objects.cuh:993-1005
    st.lvA[idx] = make_uint4(__float_as_uint(v.p.x), __float_as_uint(v.p.y), __float_as_uint(v.p.z), pn);
    st.lvB[idx] = make_uint4(pw, bxy, bz, puv);
    st.lvC[idx] = make_uint4(info, __float_as_uint(v.d_vc), __float_as_uint(v.d_vcm), 0u);
    v.n = unpackOct(pn); v.wo = unpackOct(pw);
    float2 a = unpackHalf2(bxy), b = unpackHalf2(bz), c = unpackHalf2(puv);
    v.beta = V3(a.x, a.y, b.x); v.uv = c;
}
struct LightVtxRaw { uint4 A, B, C; };
NV_D LightVtxRaw fetchLightVtx(const BdState& st, size_t idx) {
    LightVtxRaw r; r.A = __ldg(st.lvA + idx); r.B = __ldg(st.lvB + idx); r.C = __ldg(st.lvC + idx);
    return r;
}
NV_D LightVtx decodeLightVtx(const LightVtxRaw& r) {
    const uint4 A = r.A, B = r.B, C = r.C;
    LightVtx v;
    v.p = V3(__uint_as_float(A.x), __uint_as_float(A.y), __uint_as_float(A.z));
    v.n = unpackOct(A.w); v.wo = unpackOct(B.x);
    float2 a = unpackHalf2(B.y), b = unpackHalf2(B.z);
    v.beta = V3(a.x, a.y, b.x); v.uv = unpackHalf2(B.w);
    v.delta = C.x & 1u; v.backface = (C.x >> 1) & 1u; v.mat = (int)((C.x >> 22) & 0x3FFu);
    v.d_vc = __uint_as_float(C.y); v.d_vcm = __uint_as_float(C.z);
    return v;
}
NV_D LightVtx loadLightVtx(const BdState& st, size_t idx) { return decodeLightVtx(fetchLightVtx(st, idx)); }

struct BdPath {
    V3 o, d, beta, prevP;
    float pdfPrevSA, prevCos, pdfOneBefore, d_vcm, d_vc, d_vm, aux;
    bool prevDelta;
    int nReal;                 // FEAT_NESTED_BIDIR: surface vertices this walk has made so far (a false hit consumes a step, not a vertex)
    MediumStack ms;            // FEAT_NESTED_BIDIR: dielectric media the walk is inside of (material ids; slot 0 = air)
    Rng rng;
};
// FEAT_NESTED_BIDIR state: 15 stack entries + the height in the top byte of the fourth word
NV_D void loadBdMedium(const BdState& st, int l, BdPath& p) {
    const uint4 m = st.medium[l];
    p.ms.w[0] = m.x; p.ms.w[1] = m.y; p.ms.w[2] = m.z; p.ms.w[3] = m.w & 0x00ffffffu; p.ms.top = (int)(m.w >> 24);
}
NV_D void storeBdMedium(const BdState& st, int l, const BdPath& p) {
    st.medium[l] = make_uint4(p.ms.w[0], p.ms.w[1], p.ms.w[2], (p.ms.w[3] & 0x00ffffffu) | ((unsigned)p.ms.top << 24));
}
NV_D void loadBdPath(const BdState& st, int l, BdPath& p) {
    float4 a = st.rayO[l], b = st.rayD[l], c = st.beta[l], e = st.prevP[l], f = st.dv[l];
    p.o = V3(a.x, a.y, a.z); p.pdfPrevSA = a.w;
    p.d = V3(b.x, b.y, b.z); p.prevCos = b.w;
    p.beta = V3(c.x, c.y, c.z); p.pdfOneBefore = c.w;
    p.prevP = V3(e.x, e.y, e.z); p.prevDelta = (__float_as_uint(e.w) & 1u) != 0; p.nReal = (int)((__float_as_uint(e.w) >> 8) & 0xffu);
    p.d_vcm = f.x; p.d_vc = f.y; p.d_vm = f.z; p.aux = f.w;
    p.rng.state = st.rng[l];
}
NV_D void storeBdPath(const BdState& st, int l, const BdPath& p) {
    st.rayO[l] = make_float4(p.o.x, p.o.y, p.o.z, p.pdfPrevSA);
    st.rayD[l] = make_float4(p.d.x, p.d.y, p.d.z, p.prevCos);
    st.beta[l] = make_float4(p.beta.x, p.beta.y, p.beta.z, p.pdfOneBefore);
    st.prevP[l] = make_float4(p.prevP.x, p.prevP.y, p.prevP.z, __uint_as_float((p.prevDelta ? 1u : 0u) | (((unsigned)p.nReal & 0xffu) << 8)));
    st.dv[l] = make_float4(p.d_vcm, p.d_vc, p.d_vm, p.aux);
    st.rng[l] = p.rng.state;
}

enum : unsigned { SR_CLAMP = 1u << 31, SR_NEE = 1u << 30, SR_PIXMASK = (1u << 30) - 1u };

NV_D void pushShadow(const BdState& st, int* cnt, V3 o, V3 d, float maxT, int skip, unsigned pixFlags, V3 contrib) {
    if (contrib.x == 0.0f && contrib.y == 0.0f && contrib.z == 0.0f) return;
    if (!(isfinite(contrib.x) && isfinite(contrib.y) && isfinite(contrib.z))) return;     // a non-finite sample is dropped, not accumulated
    int pos = bdAggAppend(&cnt[BC_SHADOW]);
    if (pos >= st.shadowCap) return;
    float4* rec = reinterpret_cast<float4*>(st.shadowQ + pos);
    rec[0] = make_float4(o.x, o.y, o.z, maxT);
    rec[1] = make_float4(d.x, d.y, d.z, __uint_as_float(pixFlags));
    rec[2] = make_float4(contrib.x, contrib.y, contrib.z, __int_as_float(skip));
}

NV_D V3 camFwd(const NovumCamera& c) { return V3(c.forward.x, c.forward.y, c.forward.z); }
NV_D V3 camOrg(const NovumCamera& c) { return V3(c.cameraOrigin.x, c.cameraOrigin.y, c.cameraOrigin.z); }
NV_D float imagePlaneArea(const BdParams& P) { return 4.0f * ((float)P.w / (float)P.h) * P.cam.fovScale * P.cam.fovScale; }
NV_D V3 applyWeight(const BdParams& P, V3 unweighted, float w) {
    if (P.paintWeight) return V3(w);
    return P.doMis ? unweighted * w : unweighted;
}
NV_D V3 fireflyClamp(V3 c) {
    float lum = luminance(c);
    return lum > kMaxFireflyLum ? c * (kMaxFireflyLum / lum) : c;
}
NV_D float triArea(const SceneView& sc, const NovumTriangle* t) {
    int4 idx = __ldg(reinterpret_cast<const int4*>(t));
    V3 a(__ldg(sc.positions + idx.x)), b(__ldg(sc.positions + idx.y)), c(__ldg(sc.positions + idx.z));
    return 0.5f * length(cross(b - a, c - a));
}

// t = 1 strategy: connect a light-path vertex to the camera and splat (deviceCode.cu:1310-1430, :2266-2382).
// v holds the quantised vertex; f is evaluated here unless the vertex is the point on the light itself (BDPT s = 1).
// Returns false when there is nothing to splat; otherwise fills the deferred shadow record.
NV_D bool prepareSplat(const SceneView& sc, const BdParams& P, const LightVtx& v, bool onLight, float pdfSampleY0, BdShadowRec& rec) {
    float rx, ry;
    if (!worldToRaster(P.cam, v.p, rx, ry)) return false;
    int px = (int)rx, py = (int)ry;
    if (px >= P.w || py >= P.h) return false;       // ndc == +1 lands one past the edge in the reference (out-of-range write there)
    if (v.delta) return false;
    V3 toCam = camOrg(P.cam) - v.p;
    V3 dirU = normalize(toCam);
    float cosAtLight = dot(v.n, dirU);
    float cosAtCamera = fabsf(dot(camFwd(P.cam), -dirU));
    if (cosAtLight <= kEps) return false;
    float A = imagePlaneArea(P);
    float We = 1.0f / (A * cosAtCamera * cosAtCamera * cosAtCamera * cosAtCamera);
    float dSq = fmaxf(lengthSq(toCam), kRayEps);
    float G = (cosAtLight * cosAtCamera) / dSq;
    float pdfCurrRevArea = cosAtLight / (dSq * A * cosAtCamera * cosAtCamera * cosAtCamera);
    V3 f; float wLight;
    if (onLight) {
        f = V3(1.0f / kPi);
        wLight = pdfCurrRevArea / pdfSampleY0;
    } else {
        Frame fl(v.n);
        V3 currToPrevL = fl.toLocal(v.wo), toCamL = fl.toLocal(dirU);
        Shading sh = resolveShading(sc.mats, v.mat, sc.textures, v.uv);
        f = bdEval(sh, -currToPrevL, toCamL, 1.0f);
        float pdfOne = bdPdf(sh, -toCamL, currToPrevL, 1.0f);
        wLight = pdfCurrRevArea * (P.etaVcm + v.d_vcm + pdfOne * v.d_vc);
    }
    V3 contrib = v.beta * f * (G * We);
    float w = 1.0f / (1.0f + wLight);
    const V3 ro = v.p + v.n * kRayEps, c = applyWeight(P, contrib, w);
    rec.o_maxT = make_float4(ro.x, ro.y, ro.z, length(toCam) - kRayEps);
    rec.d_pix = make_float4(dirU.x, dirU.y, dirU.z, __uint_as_float((unsigned)(py * P.w + px)));
    rec.contrib_skip = make_float4(c.x, c.y, c.z, __int_as_float(-1));
    return true;
}
NV_D void storeShadow(const BdState& st, int pos, const BdShadowRec& rec) {
    if (pos >= st.shadowCap) return;
    float4* out = reinterpret_cast<float4*>(st.shadowQ + pos);
    out[0] = rec.o_maxT; out[1] = rec.d_pix; out[2] = rec.contrib_skip;
}
NV_D void lightTraceSplat(const SceneView& sc, const BdState& st, const BdParams& P, int* cnt, const LightVtx& v, bool onLight, float pdfSampleY0) {
    BdShadowRec rec;
    if (prepareSplat(sc, P, v, onLight, pdfSampleY0, rec)) storeShadow(st, bdAggAppend(&cnt[BC_SHADOW]), rec);
}

// ---- light pass -----------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBdThreads) bdLightStart(SceneView sc, BdState st, BdParams P) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    bool active = false;
    int* cnt = st.counters;
    if (l < P.nLocal) {
        int x, y;
        int pix = localToPixel(P, l, x, y);
        st.pathLen[l] = 0;
        BdPath p;
        p.rng.seed(P.seed, (unsigned long long)pix, (unsigned long long)P.iter, 1);
        int nl = sc.numLights;
        int lightInd = P.sampleEnv ? (min((int)(p.rng.uniform() * (nl + 1)), nl) - 1) : min((int)(p.rng.uniform() * nl), nl - 1);
        float pdfChoose = 1.0f / (float)(P.sampleEnv ? nl + 1 : nl);
        if (lightInd >= 0) {        // -1 = the (black) environment: the reference returns an empty path (:1982)
            const NovumTriangle* lt = sc.lights + lightInd;
            int4 q0 = __ldg(reinterpret_cast<const int4*>(lt)), q1 = __ldg(reinterpret_cast<const int4*>(lt) + 1);
            float4 em = __ldg(reinterpret_cast<const float4*>(lt) + 3);
            V3 a(__ldg(sc.positions + q0.x)), b(__ldg(sc.positions + q0.y)), c(__ldg(sc.positions + q0.z));
            V3 an(__ldg(sc.normals + q0.w)), bn(__ldg(sc.normals + q1.x)), cn(__ldg(sc.normals + q1.y));
            float area = 0.5f * length(cross(b - a, c - a));
            float pdfPos = pdfChoose / area;
            float u = sqrtf(p.rng.uniform()), v = p.rng.uniform();
            float w0 = 1.0f - u, w1 = u * (1.0f - v), w2 = u * v;
            V3 y0 = w0 * a + w1 * b + w2 * c;
            V3 n0 = normalize(w0 * an + w1 * bn + w2 * cn);
            V3 woL = cosineSample(p.rng);
            float pdfSA = cosinePdf(woL);
            V3 dirW = Frame(n0).toWorld(woL);
            p.o = y0 + n0 * kRayEps; p.d = dirW;
            p.beta = V3(em.x, em.y, em.z) * kPi / pdfPos;
            p.prevCos = fabsf(dot(normalize(dirW), n0));
            p.pdfPrevSA = P.mode == BD_MODE_BDPT ? p.prevCos / kPi : pdfSA;       // :1101 vs :2010
            p.pdfOneBefore = -1.0f; p.prevP = y0; p.prevDelta = false;
            p.d_vcm = p.d_vc = p.d_vm = -1.0f; p.aux = pdfPos; p.nReal = 0;
            storeBdPath(st, l, p);
            if (st.medium) st.medium[l] = make_uint4(0u, 0u, 0u, 1u << 24);          // in air (material 0), as Li_unidirectional starts (:308)
            active = true;
            if (P.mode == BD_MODE_BDPT && P.lightTrace) {      // s = 1, t = 1: the sampled light point seen directly by the camera (:1357-1389)
                LightVtx v; v.p = y0; v.n = n0; v.wo = V3(0.0f); v.beta = p.beta; v.uv = make_float2(0.f, 0.f); v.delta = false; v.backface = false;
                v.mat = 0; v.d_vc = 0.f; v.d_vcm = 0.f;
                lightTraceSplat(sc, st, P, cnt, v, true, pdfPos);
            }
        }
    }
    int pos = bdWarpAppend(&cnt[BC_QUEUE], active);
    if (active) st.queue[0][pos] = l;
}

// closest hit for every queued path of this depth (rays streamed through the lanes with per-lane refill, trav_stream.cuh)
template <bool COUNT>
__global__ void __launch_bounds__(kBdThreads) bdExtend(SceneView sc, BdState st, const int* __restrict__ queue, int* cnt, TravCounters tc) {
    __shared__ uint2 stackMem[kSmemStack * kBdThreads];
    TravStack ts; ts.init(stackMem + threadIdx.x, kBdThreads);
    const int lane = bdLane();
    const int n = cnt[BC_QUEUE];
    unsigned traced = streamClosest<COUNT>(sc.bvh, ts, n, &cnt[BC_FETCH_EXTEND], tc,
        [&](int i, V3& o, V3& d) {
            int l = __ldg(queue + i);
            float4 ro = st.rayO[l], rd = st.rayD[l];
            o = V3(ro.x, ro.y, ro.z); d = V3(rd.x, rd.y, rd.z);
        },
        [&](int i, const HitRec& h) { st.hit[__ldg(queue + i)] = make_float4(__int_as_float(h.tri), h.t, h.u, h.v); },
        [](bool) {});
    unsigned long long tr64 = traced;
    for (int o = 16; o > 0; o >>= 1) tr64 += __shfl_xor_sync(0xffffffffu, tr64, o);
    if (lane == 0 && tr64) atomicAdd(&st.stats[0], tr64);
}

// vertex processing shared by both walks: forward area pdf, BSDF sample, reverse pdf, d_vcm / d_vc / d_vm recurrences
struct StepOut { Surface s; Shading sh; V3 woL, wiL, wiW, currBeta; float d_vcm, d_vc, d_vm; bool delta; };

// Returns WALK_END, WALK_VERTEX or - NESTED only - WALK_THROUGH: a false hit on a dielectric interface of lower priority than the medium
// the walk is in (deviceCode.cu:413-425 in Li_unidirectional): the medium stack is updated, the ray goes on from the hit point, no vertex
// is made. NESTED is FEAT_NESTED_BIDIR: the walk carries the medium stack, attenuates by the dominant medium's absorption along the
// segment (:347-358), hands the dominant medium's index of refraction to the BSDFs as the incident one (:359) and pushes / pops media on
// refraction (:484-497) - exactly what shadeMIS does for the unidirectional integrator. Without it the reference's placeholders stand
// (etaI = etaT = 1, deviceCode.cu:851-852, 1177-1178, 2078-2079).
enum : int { WALK_END = 0, WALK_VERTEX = 1, WALK_THROUGH = 2 };
template <bool NESTED>
NV_D int walkVertex(const SceneView& sc, const BdParams& P, BdPath& p, const float4 hf, bool lightWalk, bool first, StepOut& o) {
    HitRec h; h.tri = __float_as_int(hf.x); h.t = hf.y; h.u = hf.z; h.v = hf.w; h.flags = 0;
    o.s = makeSurface(sc, h, p.o, p.d);
    o.sh = resolveShading(sc.mats, o.s.mat, sc.textures, o.s.uv);
    o.delta = (o.sh.m.flags & MF_SPECULAR) != 0;
    float etaI = 1.0f;
    if (NESTED) {
        int minId = p.ms.get(0);
        int minPr = __ldg(&sc.mats[minId].priority);
        for (int i = 1; i < p.ms.top; ++i) {
            const int id = p.ms.get(i), pr = __ldg(&sc.mats[id].priority);
            if (pr < minPr) { minPr = pr; minId = id; }
        }
        const MatView dom = loadMat(sc.mats, minId);
        if (o.s.dist > kEps) p.beta *= V3(expf(-dom.absorption.x * o.s.dist), expf(-dom.absorption.y * o.s.dist), expf(-dom.absorption.z * o.s.dist));
        etaI = dom.ior;
        if ((o.sh.m.flags & MF_BOUNDARY) && o.sh.m.priority > minPr) {
            if (!o.s.backface) { if (p.ms.top < 15) p.ms.push(o.s.mat); } else p.ms.remove(o.s.mat);
            p.o = o.s.p + p.d * kRayEps;
            return WALK_THROUGH;
        }
    }
    Frame fr(o.s.n);
    o.woL = fr.toLocal(p.d);
    float distSq = fmaxf(lengthSq(o.s.p - p.prevP), kRayEps);
    float pdfFwdArea = p.pdfPrevSA * fabsf(o.woL.z) / distSq;
    V3 f; float pdfFwdSA;
    bsdfSample(p.rng, o.sh, o.woL, etaI, o.s.backface, lightWalk ? MODE_IMPORTANCE : MODE_RADIANCE, o.wiL, f, pdfFwdSA);
    o.wiW = fr.toWorld(o.wiL);
    float pdfRevSA = bdPdf(o.sh, -o.wiL, -o.woL, etaI);
    if (NESTED && pdfFwdSA >= kEps && o.wiL.z < 0.0f) {        // refracted (or transmitted): enter / leave the medium
        if (!o.s.backface) { if (p.ms.top < 15) p.ms.push(o.s.mat); } else p.ms.remove(o.s.mat);
    }
    if (pdfFwdSA < kEps) return WALK_END;
    o.currBeta = p.beta;
    p.beta = p.beta * f * (fabsf(o.wiL.z) / pdfFwdSA);
    const float eta = P.etaVcm;                   // 0 in BDPT mode
    const bool vm = P.mode == BD_MODE_VCM;
    float G = p.prevCos / distSq;
    if (first) {
        if (lightWalk) {                          // :2111-2142, :1211-1239 (pdf_connect == pdf_trace == spatial pdf of the light sample)
            float pdfTrace = p.aux;
            o.d_vcm = pdfTrace / (pdfTrace * pdfFwdArea);
            o.d_vc = G / (pdfTrace * pdfFwdArea);
            o.d_vm = vm ? G / (pdfTrace * pdfFwdArea * eta) : 0.0f;
        } else {                                  // :2802-2819, :876-891
            o.d_vcm = 1.0f / pdfFwdArea; o.d_vc = 0.0f; o.d_vm = 0.0f;
        }
    } else if (p.prevDelta) {                     // :2143-2159
        o.d_vcm = 0.0f;
        o.d_vc = (G / pdfFwdArea) * (p.pdfOneBefore * p.d_vc);
        o.d_vm = vm ? (G / pdfFwdArea) * (p.pdfOneBefore * p.d_vm) : 0.0f;
    } else {                                      // :2160-2181
        o.d_vcm = 1.0f / pdfFwdArea;
        o.d_vc = (G / pdfFwdArea) * (eta + p.d_vcm + p.pdfOneBefore * p.d_vc);
        o.d_vm = vm ? (G / pdfFwdArea) * (1.0f + p.d_vcm / eta + p.pdfOneBefore * p.d_vm) : 0.0f;
    }
    p.d_vcm = o.d_vcm; p.d_vc = o.d_vc; p.d_vm = o.d_vm;
    p.pdfOneBefore = pdfRevSA;
    p.pdfPrevSA = pdfFwdSA;
    p.prevCos = fabsf(o.wiL.z);
    return WALK_VERTEX;
}

template <bool NESTED>
__global__ void __launch_bounds__(kBdThreads) bdLightStep(SceneView sc, BdState st, BdParams P, int step, int numSteps,
                                                         const int* __restrict__ queue, int* __restrict__ nextQueue) {
    const int lane = bdLane();
    int* cnt = st.counters + step * BC_STRIDE;
    int* nextCnt = st.counters + (step + 1) * BC_STRIDE;
    const int n = cnt[BC_QUEUE];
    unsigned long long nVerts = 0, nPhot = 0;
    // The three appends of a packet (photon slab, splat shadow rays, next queue) are each ONE ballot-aggregated atomic at a
    // converged point instead of an atomic per photon and per splat inside the divergent walk (profiles/r1_bdLightStep_atomics.txt).
    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&cnt[BC_FETCH_STEP], 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        int i = base + lane;
        bool cont = false, wantPhoton = false, wantSplat = false;
        int l = 0;
        uint4 ph0 = make_uint4(0u, 0u, 0u, 0u), ph1 = ph0;
        BdShadowRec splat;
        if (i < n) {
            l = __ldg(queue + i);
            float4 hf = st.hit[l];
            if (__float_as_int(hf.x) >= 0) {
                BdPath p; loadBdPath(st, l, p);
                if (NESTED) loadBdMedium(st, l, p);
                V3 inDir = p.d;
                StepOut o;
                // vertex number of this path: the step, unless false hits (NESTED) have used up steps without making vertices
                const int vi = NESTED ? st.pathLen[l] : step;
                const int status = walkVertex<NESTED>(sc, P, p, hf, true, vi == 0, o);
                if (NESTED && status == WALK_THROUGH) {
                    cont = step + 1 < numSteps;
                    if (cont) { storeBdPath(st, l, p); storeBdMedium(st, l, p); }
                }
                if (status == WALK_VERTEX) {
                    LightVtx v;
                    v.p = o.s.p; v.n = o.s.n; v.wo = normalize(-inDir); v.beta = o.currBeta; v.uv = o.s.uv;
                    v.delta = o.delta; v.backface = o.s.backface; v.mat = o.s.mat; v.d_vc = o.d_vc; v.d_vcm = o.d_vcm;
                    storeLightVtx(st, (size_t)vi * P.nLocal + l, v);             // v now holds the quantised values
                    st.pathLen[l] = vi + 1;
                    ++nVerts;
                    if (!o.delta && P.mode == BD_MODE_VCM && P.doMerge) {        // photon for merging (:2206-2221)
                        wantPhoton = true;
                        ph0 = make_uint4(__float_as_uint(o.s.p.x), __float_as_uint(o.s.p.y), __float_as_uint(o.s.p.z), packOct(normalize(-inDir)));
                        ph1 = make_uint4(packHalf2(o.currBeta.x, o.currBeta.y), packHalf2(o.currBeta.z, 0.0f), __float_as_uint(o.d_vm), __float_as_uint(o.d_vcm));
                    }
                    if (P.lightTrace) wantSplat = prepareSplat(sc, P, v, false, 0.0f, splat);
                    bool through = P.mode == BD_MODE_BDPT ? (dot(o.wiW, o.s.n) < 0.0f) : (o.wiL.z < kEps);    // :1274 vs :2229
                    p.o = through ? o.s.p - o.s.n * kRayEps : o.s.p + o.s.n * kRayEps;
                    p.d = o.wiW;
                    p.prevDelta = o.delta; p.prevP = o.s.p;
                    cont = step + 1 < numSteps;
                    if (cont) { storeBdPath(st, l, p); if (NESTED) storeBdMedium(st, l, p); }
                }
            }
        }
        const int idx = bdWarpAppend(reinterpret_cast<int*>(P.slab), wantPhoton);
        if (wantPhoton && idx < P.slabCap) {
            uint4* rec = P.slab + (kSlabHeaderBytes / 16) + (size_t)idx * 2;
            rec[0] = ph0; rec[1] = ph1;
            for (int k = 0; k < P.numPeers; ++k) {               // the same record into every peer's copy of this slab (NVLink stores)
                uint4* prec = P.peerSlab[k] + (kSlabHeaderBytes / 16) + (size_t)idx * 2;
                prec[0] = ph0; prec[1] = ph1;
            }
            ++nPhot;
        }
        const int spos = bdWarpAppend(&cnt[BC_SHADOW], wantSplat);
        if (wantSplat) storeShadow(st, spos, splat);
        const int pos = bdWarpAppend(&nextCnt[BC_QUEUE], cont);
        if (cont) nextQueue[pos] = l;
    }
    for (int o = 16; o > 0; o >>= 1) { nVerts += __shfl_xor_sync(0xffffffffu, nVerts, o); nPhot += __shfl_xor_sync(0xffffffffu, nPhot, o); }
    if (lane == 0) { if (nVerts) atomicAdd(&st.stats[3], nVerts); if (nPhot) atomicAdd(&st.stats[2], nPhot); }
}

// any-hit for every deferred connection of this depth; unoccluded ones add their (already MIS-weighted) contribution
__global__ void __launch_bounds__(kBdThreads) bdResolve(SceneView sc, BdState st, float4* __restrict__ colors, int* cnt) {
    __shared__ uint2 stackMem[kSmemStack * kBdThreads];
    TravStack ts; ts.init(stackMem + threadIdx.x, kBdThreads);
    const int lane = bdLane();
    const int n = min(cnt[BC_SHADOW], st.shadowCap);
    TravCounters nocnt{nullptr, nullptr};
    unsigned traced = streamShadow<false>(sc, ts, n, &cnt[BC_FETCH_RESOLVE], nocnt,
        [&](int i, V3& o, V3& d, float& maxT, int& skip) {
            const float4* rp = reinterpret_cast<const float4*>(st.shadowQ + i);
            float4 a = __ldg(rp), b = __ldg(rp + 1), c = __ldg(rp + 2);
            o = V3(a.x, a.y, a.z); d = V3(b.x, b.y, b.z); maxT = a.w; skip = __float_as_int(c.w);
        },
        [&](int i, V3 thr) {
            const float4* rp = reinterpret_cast<const float4*>(st.shadowQ + i);
            float4 b = __ldg(rp + 1), c = __ldg(rp + 2);
            unsigned pf = __float_as_uint(b.w);
            if (lengthSq(thr) > ((pf & SR_NEE) ? 0.0f : kEps)) {       // :668 (NEE: > 0) vs :1336, :2304, :2599 (> EPSILON)
                V3 v = V3(c.x, c.y, c.z) * thr;
                if (pf & SR_CLAMP) v = fireflyClamp(v);
                atomicAdd(&colors[pf & SR_PIXMASK], make_float4(v.x, v.y, v.z, 0.f));
            }
        });
    unsigned long long tr64 = traced;
    for (int o = 16; o > 0; o >>= 1) tr64 += __shfl_xor_sync(0xffffffffu, tr64, o);
    if (lane == 0 && tr64) atomicAdd(&st.stats[1], tr64);
}

// ---- eye pass ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBdThreads) bdEyeStart(BdState st, BdParams P) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    bool active = l < P.nLocal;
    if (active) {
        int x, y;
        int pix = localToPixel(P, l, x, y);
        BdPath p;
        p.rng.seed(P.seed, (unsigned long long)pix, (unsigned long long)P.iter, 2);
        cameraRay(P.cam, p.rng, x, y, p.o, p.d);
        float cosAtCamera = fabsf(dot(camFwd(P.cam), p.d));
        p.pdfPrevSA = 1.0f / (imagePlaneArea(P) * cosAtCamera * cosAtCamera * cosAtCamera);      // :2706, :778
        p.prevCos = cosAtCamera;
        p.beta = V3(1.0f); p.pdfOneBefore = -1.0f;
        p.prevP = camOrg(P.cam);
        p.prevDelta = P.mode == BD_MODE_VCM;        // :2719 (true) vs :792 (false)
        p.d_vcm = p.d_vc = p.d_vm = -1.0f; p.aux = 0.f; p.nReal = 0;
        storeBdPath(st, l, p);
        if (st.medium) st.medium[l] = make_uint4(0u, 0u, 0u, 1u << 24);
    }
    int pos = bdWarpAppend(&st.counters[BC_QUEUE], active);
    if (active) st.queue[0][pos] = l;
}

template <bool NESTED>
__global__ void __launch_bounds__(kBdThreads) bdEyeStep(SceneView sc, BdState st, BdParams P, int step, int numSteps,
                                                       const int* __restrict__ queue, int* __restrict__ nextQueue, MergeQueue mq) {
    const int lane = bdLane();
    int* cnt = st.counters + step * BC_STRIDE;
    int* nextCnt = st.counters + (step + 1) * BC_STRIDE;
    const int n = cnt[BC_QUEUE];
    const bool vcm = P.mode == BD_MODE_VCM;
    const float eta = P.etaVcm;
    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&cnt[BC_FETCH_STEP], 32);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        int i = base + lane;
        bool cont = false;
        int l = 0;
        if (i < n) {
            l = __ldg(queue + i);
            float4 hf = st.hit[l];
            if (__float_as_int(hf.x) < 0) {
                // left the scene. FEAT_ENVIRONMENT: the escape is the s = 0 strategy for the constant environment and, as nothing samples
                // the environment from the light side, the only one (weight 1); it is switched with the other implicit hits (BDPT_NAIVE)
                if ((sc.features & FEAT_ENVIRONMENT) && P.naive) {
                    const float4 b = st.beta[l];
                    const V3 c = V3(b.x, b.y, b.z) * V3(sc.env[0], sc.env[1], sc.env[2]);
                    int x, y;
                    const int pix = localToPixel(P, l, x, y);
                    if ((c.x != 0.0f || c.y != 0.0f || c.z != 0.0f) && isfinite(c.x) && isfinite(c.y) && isfinite(c.z))
                        atomicAdd(&P.colors[pix], make_float4(c.x, c.y, c.z, 0.f));
                }
            } else {
                BdPath p; loadBdPath(st, l, p);
                if (NESTED) loadBdMedium(st, l, p);
                const V3 inDir = p.d, prevP = p.prevP;
                const bool prevDelta = p.prevDelta;
                StepOut o;
                const bool firstVertex = NESTED ? p.nReal == 0 : step == 0;
                const int status = walkVertex<NESTED>(sc, P, p, hf, false, firstVertex, o);
                if (NESTED && status == WALK_THROUGH) {
                    cont = step + 1 < numSteps;
                    if (cont) { storeBdPath(st, l, p); storeBdMedium(st, l, p); }
                }
                if (status == WALK_VERTEX) {
                    if (NESTED) ++p.nReal;
                    int x, y;
                    const int pix = localToPixel(P, l, x, y);
                    const Surface& s = o.s;
                    Frame fr(s.n);
                    bool endPath = false;
                    V3 sum(0.0f);
                    if (!o.delta) {
                        // first light vertex of this pixel: requested here so that the fetch overlaps the s = 0 / s = 1 strategies
                        const int len = P.connection ? st.pathLen[l] : 0;
                        LightVtxRaw raw = fetchLightVtx(st, (size_t)l);
                        const unsigned clampFlag = vcm ? SR_CLAMP : 0u;
                        const float pdfChoose = 1.0f / (float)(P.sampleEnv ? sc.numLights + 1 : sc.numLights);
                        // vector from this vertex back to the previous one as the reference hands it to the BSDFs:
                        // un-normalised position difference in VCM (:2493, :2629, :3004), the stored unit wo in BDPT (:1490, :1708)
                        const V3 toPrevW = vcm ? (prevP - s.p) : normalize(-inDir);
                        const V3 toPrevL = fr.toLocal(toPrevW);
                        // ---- s = 0: the walk hit an emitter from its front side ----
                        if (P.naive && lengthSq(s.emission) > kEps && !s.backface && s.lightInd >= 0) {
                            float area = triArea(sc, sc.tris + s.tri);
                            float w; V3 c = s.emission * o.currBeta;
                            bool clampIt = true;
                            if (vcm) {                                                           // connectImplicitHit :2389-2458
                                float cosLight = dot(s.n, normalize(prevP - s.p));
                                float pdfConnect = prevDelta ? 0.0f : pdfChoose / area;
                                w = 1.0f / (1.0f + pdfConnect * o.d_vcm + pdfConnect * (cosLight / kPi) * o.d_vc);
                            } else if (firstVertex) {                                            // t == 2 :1567-1595
                                V3 lightToPrevU = normalize(-inDir);
                                float cosThetaLight = fabsf(dot(s.n, lightToPrevU));
                                float cosAtCamera = fabsf(dot(camFwd(P.cam), -lightToPrevU));
                                float pdfTraceFromCamera = cosThetaLight / (lengthSq(s.p - prevP) * imagePlaneArea(P) * cosAtCamera * cosAtCamera * cosAtCamera);
                                w = 1.0f / (1.0f + (pdfChoose / area) / pdfTraceFromCamera);
                                clampIt = false;
                            } else {                                                             // :1596-1642
                                float cosThetaLight = fabsf(dot(s.n, normalize(-inDir)));
                                float pdfConnect = prevDelta ? 0.0f : pdfChoose / area;
                                w = 1.0f / (1.0f + pdfConnect * o.d_vcm + pdfConnect * (cosThetaLight / kPi) * o.d_vc);
                                c = fireflyClamp(c);                                             // clamped before weighting in BDPT
                                clampIt = false;
                            }
                            V3 cw = applyWeight(P, c, w);
                            sum += clampIt ? fireflyClamp(cw) : cw;
                        }
                        // ---- s = 1: next-event estimation (BDPTnextEventEstimation :622-705, connectNEE :2460-2546, connectPath :1479-1546) ----
                        if (P.nee && sc.numLights > 0) {
                            int nl = sc.numLights;
                            int lightInd = P.sampleEnv ? (min((int)(p.rng.uniform() * (nl + 1)), nl) - 1) : min((int)(p.rng.uniform() * nl), nl - 1);
                            if (lightInd >= 0) {
                                const NovumTriangle* lt = sc.lights + lightInd;
                                int4 q0 = __ldg(reinterpret_cast<const int4*>(lt)), q1 = __ldg(reinterpret_cast<const int4*>(lt) + 1);
                                float4 em = __ldg(reinterpret_cast<const float4*>(lt) + 3);
                                int lightTri = __ldg(&lt->triInd);
                                V3 a(__ldg(sc.positions + q0.x)), b(__ldg(sc.positions + q0.y)), c(__ldg(sc.positions + q0.z));
                                V3 an(__ldg(sc.normals + q0.w)), bn(__ldg(sc.normals + q1.x)), cn(__ldg(sc.normals + q1.y));
                                float u = sqrtf(p.rng.uniform()), v = p.rng.uniform();
                                float w0 = 1.0f - u, w1 = u * (1.0f - v), w2 = u * v;
                                V3 lp = w0 * a + w1 * b + w2 * c;
                                V3 ln = normalize(w0 * an + w1 * bn + w2 * cn);
                                V3 toL = lp - s.p;
                                V3 toLU = normalize(toL);
                                float dSq = fmaxf(lengthSq(toL), kRayEps);
                                float cosLight = dot(ln, -toLU);
                                if (cosLight >= kEps) {
                                    float cosSurf = fabsf(dot(s.n, toLU));
                                    float G = fminf(cosLight * cosSurf / dSq, 15.0f);
                                    float area = 0.5f * length(cross(b - a, c - a));
                                    float pdfConnect = pdfChoose / area;
                                    float pdfEmit = cosLight / kPi;
                                    V3 prevToCurrL = -toPrevL;                   // VCM: toLocal(eyePos - prevPos); BDPT: toLocal(-wo)
                                    V3 toLL = fr.toLocal(toLU);
                                    V3 f = bdEval(o.sh, prevToCurrL, toLL, 1.0f);
                                    V3 unweighted = f * V3(em.x, em.y, em.z) * (G / pdfConnect) * o.currBeta;
                                    float pdfBsdfArea = bdPdf(o.sh, prevToCurrL, toLL, 1.0f) * fabsf(cosLight) / dSq;
                                    float ratio = pdfBsdfArea / pdfConnect;
                                    float wLight = vcm ? ratio * ratio : ratio;               // :2526 vs :1519
                                    float pdfCurrRevArea = pdfEmit * fabsf(toLL.z) / dSq;
                                    float pdfOne = bdPdf(o.sh, -toLL, -prevToCurrL, 1.0f);
                                    float wEye = pdfCurrRevArea * (eta + o.d_vcm + pdfOne * o.d_vc);
                                    float w = 1.0f / (1.0f + wLight + wEye);
                                    pushShadow(st, cnt, s.p + s.n * kRayEps, toLU, sqrtf(dSq) - kEps, lightTri, (unsigned)pix | clampFlag | SR_NEE,
                                               applyWeight(P, unweighted, w));
                                }
                            }
                        }
                        // ---- s >= 2: connect to every stored vertex of this pixel's light path (connectGeneral :2549-2684, connectPath :1653-1762) ----
                        if (P.connection) {
                            // the next vertex's three words are requested before this one is decoded and connected: the wait
                            // for them was 30 % of this kernel's stall samples
                            for (int k = 0; k < len; ++k) {
                                const LightVtxRaw cur = raw;
                                if (k + 1 < len) raw = fetchLightVtx(st, (size_t)(k + 1) * P.nLocal + l);
                                LightVtx lv = decodeLightVtx(cur);
                                if (lv.delta) continue;
                                V3 e2l = lv.p - s.p;
                                float dSq = fmaxf(lengthSq(e2l), kRayEps);
                                float dist = length(e2l);
                                V3 e2lU = e2l * (1.0f / dist);                       // one division instead of three
                                float cosLight = fabsf(dot(lv.n, -e2lU)), cosEye = fabsf(dot(s.n, e2lU));
                                if (cosLight < kEps || cosEye < kEps) continue;
                                const float invDSq = 1.0f / dSq;
                                Frame fl(lv.n);
                                Shading shL = resolveShading(sc.mats, lv.mat, sc.textures, lv.uv);
                                V3 l2eAtLight = fl.toLocal(-e2lU), toLightFromPrevAtLight = fl.toLocal(-lv.wo);
                                float pdfEyeRevArea = bdPdf(shL, toLightFromPrevAtLight, l2eAtLight, 1.0f) * cosEye * invDSq;
                                V3 l2eAtEye = fr.toLocal(-e2lU);
                                float pdfOneBeforeEye = bdPdf(o.sh, l2eAtEye, toPrevL, 1.0f);
                                float pdfLightRevArea = bdPdf(o.sh, -toPrevL, -l2eAtEye, 1.0f) * cosLight * invDSq;
                                float pdfOneBeforeLight = bdPdf(shL, -l2eAtLight, -toLightFromPrevAtLight, 1.0f);
                                float wEye = pdfEyeRevArea * (eta + o.d_vcm + pdfOneBeforeEye * o.d_vc);
                                float wLight = pdfLightRevArea * (eta + lv.d_vcm + pdfOneBeforeLight * lv.d_vc);
                                V3 fEye = bdEval(o.sh, l2eAtEye, toPrevL, 1.0f);
                                V3 fLight = bdEval(shL, -l2eAtLight, -toLightFromPrevAtLight, 1.0f);
                                float G = fminf(cosEye * cosLight * invDSq, 2.0f);
                                V3 unweighted = o.currBeta * lv.beta * fEye * fLight * G;
                                float w = 1.0f / (1.0f + wEye + wLight);
                                pushShadow(st, cnt, s.p + s.n * kRayEps, e2lU, dist - kRayEps, -1, (unsigned)pix | clampFlag, applyWeight(P, unweighted, w));
                            }
                        }
                        // ---- merging with the photons of the 27 neighbouring cells (:2968-3048) ----
                        if (vcm && P.doMerge) {
                          if (P.mergeRadius > 0.0f) {
                            // deferred: the photon gather runs warp-cooperatively in bdMerge on this record
                            // Lambert vertices fill mergeQ from the front (bdMergeLambert), every other material from the back (bdMerge)
                            const bool lambert = o.sh.m.type == MAT_DIFFUSE && !P.paintWeight;
                            int pos = lambert ? bdAggAppend(&mq.cnt[BC_MERGE]) : mq.cap - 1 - bdAggAppend(&mq.cnt[BC_MERGE_GEN]);
                            float4* rec = reinterpret_cast<float4*>(mq.recs + pos);
                            rec[0] = make_float4(s.p.x, s.p.y, s.p.z, __int_as_float(pix));
                            rec[1] = make_float4(s.n.x, s.n.y, s.n.z, __int_as_float(s.mat));
                            rec[2] = make_float4(toPrevL.x, toPrevL.y, toPrevL.z, o.d_vcm);
                            rec[3] = make_float4(o.currBeta.x, o.currBeta.y, o.currBeta.z, o.d_vm);
                            rec[4] = make_float4(s.uv.x, s.uv.y, 0.f, 0.f);
                          }
                          if (P.doSPPM) endPath = true;              // SPPM gathers density once (:3046)
                        }
                    }
                    if ((sum.x != 0.0f || sum.y != 0.0f || sum.z != 0.0f) && isfinite(sum.x) && isfinite(sum.y) && isfinite(sum.z)) atomicAdd(&P.colors[pix], make_float4(sum.x, sum.y, sum.z, 0.f));
                    bool through = dot(o.wiW, s.n) < 0.0f;                                       // :3053, :936
                    p.o = through ? s.p - s.n * kRayEps : s.p + s.n * kRayEps;
                    p.d = vcm ? o.wiW : normalize(o.wiW);                                        // :3056 vs :944
                    p.prevDelta = o.delta; p.prevP = s.p;
                    cont = !endPath && step + 1 < numSteps;
                    if (cont) { storeBdPath(st, l, p); if (NESTED) storeBdMedium(st, l, p); }
                }
            }
        }
        int pos = bdWarpAppend(&nextCnt[BC_QUEUE], cont);
        if (cont) nextQueue[pos] = l;
    }
}

// ---- merging (deviceCode.cu:2968-3048), one warp per eye vertex ----------------------------------------------------------------
// Photon records are read in the merge-ready form written by bdReorderPhotons; the cells looked at are described at kMergeFine below. Two kernels, because the eye vertex's material decides how much work one accepted photon is:
//   bdMergeLambert  Lambert vertices (front of mergeQ): the reverse pdfs and the BSDF value reduce to one cosine, so the
//                   accept path is ~20 instructions and is run in place, predicated, with no compaction;
//   bdMerge         everything else (back of mergeQ): accepted candidates are compacted into a per-warp buffer so that the
//                   expensive part - two reverse pdfs, the BSDF value, the VM MIS weight - always runs on full warps.
constexpr int kMergeThreads = 128;
constexpr int kMergeWarps = kMergeThreads / 32;
constexpr int kMergeCtasPerSm = 8;        // 64 registers per thread; the 90-register build at 5 CTAs/SM was 14 % slower
constexpr int kMergeGather = 4;           // eye vertices a warp gathers together

// The photon grid of the engine has cells of r / kMergeFine (the reference's cells are r wide, integratorUtilities.cuh:382-388, with a
// 27-cell walk, deviceCode.cu:2968-2995): a photon within r of a vertex then lies in one of the (2 kMergeFine + 1)^3 cells around the
// vertex's own, and of those only the cells whose box comes within r of the vertex are looked at. On a surface the 27 r-cells cover
// 9 r^2 of which the disc of radius r is pi r^2 (2.9 photons scanned per photon merged at best, 3.9 measured with the 4-vertex union);
// the reachable r/2-cells cover ~5.3 r^2. The photon SET a vertex merges is unchanged - the distance test decides - so images differ
// by summation order only.
constexpr int kMergeFine = 2;
constexpr int kMergeReach = kMergeFine;                       // cells on either side of the vertex's own cell
constexpr int kMergeSide = 2 * kMergeReach + 1;               // 5
constexpr int kMergeNbr = kMergeSide * kMergeSide * kMergeSide;   // 125
constexpr int kMergeMaxBox = 256;                             // largest cell box a group of vertices may span and still be gathered together
constexpr int kMergeMaxList = kMergeMaxBox;
constexpr int kMergeSeen = 2 * kMergeMaxBox;                  // slots of the per-warp set of ranges already listed (load factor <= 1/2)

// squared distance from p to the box of cell c (slightly shrunk distances: 1e-4 cell of slack absorbs the rounding of the cell function,
// nothing within r is ever dropped)
NV_D float cellDist2(const GridParams& g, int cx, int cy, int cz, V3 p) {
    const float lx = g.minX + (float)cx * g.cell, ly = g.minY + (float)cy * g.cell, lz = g.minZ + (float)cz * g.cell;
    const float slack = 1e-4f * g.cell;
    const float dx = fmaxf(fmaxf(lx - p.x, p.x - (lx + g.cell)) - slack, 0.0f);
    const float dy = fmaxf(fmaxf(ly - p.y, p.y - (ly + g.cell)) - slack, 0.0f);
    const float dz = fmaxf(fmaxf(lz - p.z, p.z - (lz + g.cell)) - slack, 0.0f);
    return fmaf(dx, dx, fmaf(dy, dy, dz * dz));
}

// first photon and photon count of neighbour cell `i` (< kMergeNbr) of the vertex at eyeP, {0, 0} when the cell is empty or out of reach
struct MergeCells { uint32_t cs, cc; };
NV_D MergeCells mergeCellFine(const BdParams& P, V3 eyeP, int3 c0, int i, float r2) {
    MergeCells m{0u, 0u};
    if (i < kMergeNbr) {
        const int dz = i / (kMergeSide * kMergeSide) - kMergeReach, dy = (i / kMergeSide) % kMergeSide - kMergeReach, dx = i % kMergeSide - kMergeReach;
        if (cellDist2(P.grid, c0.x + dx, c0.y + dy, c0.z + dz, eyeP) <= r2) {
            const unsigned hsh = hashCellMagic(c0.x + dx, c0.y + dy, c0.z + dz, P.grid.tableSize, P.gridMagic);
            const uint32_t s0 = __ldg(P.cellStart + hsh);
            if (s0 != 0xFFFFFFFFu) { m.cs = s0; m.cc = __ldg(P.cellEnd + hsh) - s0; }
        }
    }
    return m;
}

NV_D void mergeSplat(const BdParams& P, int pix, V3 sum, int lane) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        sum.x += __shfl_xor_sync(0xffffffffu, sum.x, o); sum.y += __shfl_xor_sync(0xffffffffu, sum.y, o); sum.z += __shfl_xor_sync(0xffffffffu, sum.z, o);
    }
    if (lane == 0 && (sum.x != 0.0f || sum.y != 0.0f || sum.z != 0.0f) && isfinite(sum.x) && isfinite(sum.y) && isfinite(sum.z))
        atomicAdd(&P.colors[pix], make_float4(sum.x, sum.y, sum.z, 0.f));
}

// Lambert eye vertex: pdf_eval = max(cos, eps)/pi of the outgoing argument only, f_eval = albedo/pi (reflectors.cuh:10-18, :566,
// :644), so per photon   w = 1 / (1 + wEye + d_vcm/eta + max(cos(wi, n), eps)/pi * d_vm)   and the vertex's contribution is
// (currBeta * albedo/pi / (pi r^2 w h)) * sum(beta_photon * w).
//
// One warp gathers G consecutive records TOGETHER (on the primary-hit queue and on the binned deeper queues consecutive records are
// neighbours in space): every lane loads ONE candidate photon (two coalesced 16 B words) and tests it against all G vertices, whose
// constants live in registers, so a load and its address arithmetic serve G (vertex, photon) pairs.
//   list    the cells of the box spanned by the G neighbourhoods are enumerated across the lanes (distinct cells by construction);
//           a cell is kept if its box comes within reach of the group (one sphere around the group: centre = mean vertex, radius =
//           r + largest offset) and its table slot is not empty; kept cells are appended to a shared-memory list with ballot / popc,
//           together with the running sum of their photon counts. Two distinct cells of the box that collide in the hash table name
//           the same photon range: a range is listed ONCE per group (a small open-addressing set of range starts in shared memory,
//           claimed with atomicCAS), so that no photon is merged twice into a vertex. (The reference sweeps a colliding slot once per
//           colliding neighbour, deviceCode.cu:2968-2995, and so counts those photons twice; with 125 r/2-cells per neighbourhood
//           instead of 27 r-cells that would no longer be a 1e-6 effect on small scenes, whose tables are small.)
//   sweep   the listed ranges are walked as ONE sequence: candidate j of the group belongs to lane j % 32, which finds its range by
//           advancing a pointer through the prefix sums - every trip is a full warp whatever the sizes of the cells (per-cell strides
//           left the last trip of each cell partly empty, which matters once cells hold ~40 photons instead of ~170).
// A group whose vertices are so far apart that the box exceeds kMergeMaxBox cells is gathered one vertex at a time.
template <bool MIS, int MINB>
__global__ void __launch_bounds__(kMergeThreads, MINB) bdMergeLambert(SceneView sc, BdState st, BdParams P, MergeQueue mq) {
    int* cnt = mq.cnt;
    const int lane = bdLane();
    const int n = min(cnt[BC_MERGE], mq.cap);
    const float r2 = P.mergeRadius * P.mergeRadius;
    const float norm = 1.0f / (kPi * r2 * (float)P.w * (float)P.h);
    const float invEta = 1.0f / P.etaVcm;
    const PhotonSoA ph = P.photons;
    const unsigned ltMask = (1u << lane) - 1u;
    __shared__ uint2 listAll[kMergeWarps][kMergeMaxList + 1];         // {first photon - candidates listed before this range, candidates listed up to its end}
    __shared__ uint32_t seenAll[kMergeWarps][kMergeSeen];            // first photon of every range listed for the current group (0xFFFFFFFF = free)
    __shared__ float4 vtxAll[kMergeWarps][kMergeGather];             // {n/pi, 1 + d_vcm/eta + eyeRev * d_vm} of the gathered vertices: warp-uniform, read where a photon is accepted
    uint2* list = listAll[threadIdx.x >> 5];
    uint32_t* seen = seenAll[threadIdx.x >> 5];
    float4* vtx = vtxAll[threadIdx.x >> 5];
    // shared-memory addresses kept in registers (the asm keeps the compiler from recomputing them from %tid at every use in the sweep)
    uint32_t listAddr, vtxAddr;
    asm volatile("mov.u32 %0, %1;" : "=r"(listAddr) : "r"((uint32_t)__cvta_generic_to_shared(list)));
    asm volatile("mov.u32 %0, %1;" : "=r"(vtxAddr) : "r"((uint32_t)__cvta_generic_to_shared(vtx)));
    unsigned long long nCand = 0;
    unsigned nAcc = 0;
    while (true) {
        __syncwarp();
        int base = 0;
        if (lane == 0) base = atomicAdd(&cnt[BC_FETCH_MERGE], kMergeGather);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const int real = min(kMergeGather, n - base);
        V3 eyeP[kMergeGather], sum[kMergeGather];
        int3 lo = make_int3(0x7fffffff, 0x7fffffff, 0x7fffffff), hi = make_int3(-0x7fffffff, -0x7fffffff, -0x7fffffff);
#pragma unroll
        for (int g = 0; g < kMergeGather; ++g) {
            // a short last group repeats its last record (its result is dropped below)
            const int r = min(base + g, n - 1);
            const float4* rp = reinterpret_cast<const float4*>(mq.recs + (mq.order ? __ldg(mq.order + r) : r));
            const float4 r0 = __ldg(rp);
            eyeP[g] = V3(r0.x, r0.y, r0.z);
            sum[g] = V3(0.0f);
            const int3 c = gridIndex(eyeP[g].x, eyeP[g].y, eyeP[g].z, P.grid);
            lo.x = min(lo.x, c.x); lo.y = min(lo.y, c.y); lo.z = min(lo.z, c.z);
            hi.x = max(hi.x, c.x); hi.y = max(hi.y, c.y); hi.z = max(hi.z, c.z);
        }
        const bool together = (long long)(hi.x - lo.x + kMergeSide) * (hi.y - lo.y + kMergeSide) * (hi.z - lo.z + kMergeSide) <= kMergeMaxBox;
        const int passes = together ? 1 : real;
#pragma unroll 1
        for (int pass = 0; pass < passes; ++pass) {
            // the vertices gathered in this pass: all of them, or vertex `pass` alone (the others are parked at NaN for the pass, so
            // that every distance test on them fails; their positions wait in the list's tail slots of shared memory)
            V3 ctr(0.0f);
            int bx0, by0, bz0, nx, ny, nz;
            float reach;
            if (together) {
#pragma unroll
                for (int k = 0; k < kMergeGather; ++k) ctr += eyeP[k];
                ctr = ctr * (1.0f / kMergeGather);               // a short group repeats its last vertex: still inside the hull
                float far2 = 0.0f;
#pragma unroll
                for (int k = 0; k < kMergeGather; ++k) far2 = fmaxf(far2, lengthSq(eyeP[k] - ctr));
                reach = P.mergeRadius + sqrtf(far2);
                bx0 = lo.x - kMergeReach; by0 = lo.y - kMergeReach; bz0 = lo.z - kMergeReach;
                nx = hi.x - lo.x + kMergeSide; ny = hi.y - lo.y + kMergeSide; nz = hi.z - lo.z + kMergeSide;
            } else {
                float* park = reinterpret_cast<float*>(list + kMergeMaxList + 1 - 2 * kMergeGather);          // 125 list entries at most in this mode
                __syncwarp();
                if (pass == 0 && lane == 0) {
#pragma unroll
                    for (int k = 0; k < kMergeGather; ++k) { park[3 * k] = eyeP[k].x; park[3 * k + 1] = eyeP[k].y; park[3 * k + 2] = eyeP[k].z; }
                }
                __syncwarp();
#pragma unroll
                for (int k = 0; k < kMergeGather; ++k)
                    eyeP[k] = k == pass ? V3(park[3 * k], park[3 * k + 1], park[3 * k + 2]) : V3(__int_as_float(0x7fc00000));
#pragma unroll
                for (int k = 0; k < kMergeGather; ++k) if (k == pass) ctr = eyeP[k];
                const int3 c = gridIndex(ctr.x, ctr.y, ctr.z, P.grid);
                reach = P.mergeRadius;
                bx0 = c.x - kMergeReach; by0 = c.y - kMergeReach; bz0 = c.z - kMergeReach;
                nx = ny = nz = kMergeSide;
            }
            const int nCells = nx * ny * nz, nxy = nx * ny;
            const float invXY = 1.0f / (float)nxy, invX = 1.0f / (float)nx, reach2 = reach * reach;
            // ---- list ----
            uint32_t total = 0;
            int m = 0;
#pragma unroll
            for (int k = 0; k < kMergeSeen / 128; ++k) reinterpret_cast<uint4*>(seen)[lane + 32 * k] = make_uint4(~0u, ~0u, ~0u, ~0u);
            __syncwarp();
#pragma unroll 1
            for (int i0 = 0; i0 < nCells; i0 += 32) {
                const int i = i0 + lane;
                uint32_t cs = 0u, cc = 0u;
                if (i < nCells) {
                    const int cz = (int)(((float)i + 0.5f) * invXY), rem = i - cz * nxy;          // exact: i < 256
                    const int cy = (int)(((float)rem + 0.5f) * invX), cx = rem - cy * nx;
                    if (cellDist2(P.grid, bx0 + cx, by0 + cy, bz0 + cz, ctr) <= reach2) {
                        const unsigned hsh = hashCellMagic(bx0 + cx, by0 + cy, bz0 + cz, P.grid.tableSize, P.gridMagic);
                        const uint32_t s0 = __ldg(P.cellStart + hsh);
                        if (s0 != 0xFFFFFFFFu) { cs = s0; cc = __ldg(P.cellEnd + hsh) - s0; }
                    }
                }
                if (cc) {      // claim the range for this group: listed once however many cells of the box share its table slot
                    uint32_t at = (cs * 2654435761u) >> (32 - 9);
                    static_assert(kMergeSeen == 512, "9 hash bits");
                    while (true) {
                        const uint32_t old = atomicCAS(seen + at, 0xFFFFFFFFu, cs);
                        if (old == 0xFFFFFFFFu) break;
                        if (old == cs) { cc = 0u; break; }
                        at = (at + 1u) & (kMergeSeen - 1);
                    }
                }
                const unsigned keep = __ballot_sync(0xffffffffu, cc != 0u);
                uint32_t x = cc;                                                                // inclusive scan of the counts
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) { const uint32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
                if (cc) {
                    const int slot = m + __popc(keep & ltMask);
                    list[slot] = make_uint2(cs - (total + x - cc), total + x);
                }
                total += __shfl_sync(0xffffffffu, x, 31);
                m += __popc(keep);
            }
            if (lane == 0) list[m] = make_uint2(0u, 0xFFFFFFFFu);
            __syncwarp();
            if (lane == 0) nCand += (unsigned long long)total * (unsigned)(together ? real : 1);
            // ---- sweep ----
            // the vertices' other constants are fetched only now (the records are in L1): they are not live while the list is built
            if (MIS && lane < kMergeGather) {                  // lane g writes vertex g's
                const int r = min(base + lane, n - 1);
                const float4* rp = reinterpret_cast<const float4*>(mq.recs + (mq.order ? __ldg(mq.order + r) : r));
                const float4 r1 = __ldg(rp + 1), r2q = __ldg(rp + 2), r3 = __ldg(rp + 3);
                // cos/pi = dot(wi, n/pi);   1 + d_vcm/eta + eyeRev * d_vm
                vtx[lane] = make_float4(r1.x * (1.0f / kPi), r1.y * (1.0f / kPi), r1.z * (1.0f / kPi),
                                        1.0f + r2q.w * invEta + cosinePdf(V3(r2q.x, r2q.y, r2q.z)) * r3.w);
            }
            __syncwarp();
            auto test = [&](const float4 Ac, const uint4 Bc) {
                const float2 wxy = unpackHalf2(__float_as_uint(Ac.w)), wzb = unpackHalf2(Bc.x), byz = unpackHalf2(Bc.y);
#pragma unroll
                for (int k = 0; k < kMergeGather; ++k) {
                    const float dx = eyeP[k].x - Ac.x, dy = eyeP[k].y - Ac.y, dz = eyeP[k].z - Ac.z;
                    if (fmaf(dx, dx, fmaf(dy, dy, dz * dz)) <= r2) {
                        float w = 1.0f;
                        if (MIS) {
                            float4 v;
                            asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(vtxAddr + k * 16));
                            const float rev = fmaxf(fmaf(wxy.x, v.x, fmaf(wxy.y, v.y, wzb.x * v.z)), kEps * (1.0f / kPi));
                            w = rcpApprox(fmaf(rev, __uint_as_float(Bc.z), v.w + __uint_as_float(Bc.w)));      // the argument is >= 1
                        }
                        sum[k].x = fmaf(wzb.y, w, sum[k].x); sum[k].y = fmaf(byz.x, w, sum[k].y); sum[k].z = fmaf(byz.y, w, sum[k].z);
                        ++nAcc;
                    }
                }
            };
            // candidate j of the group -> photon index (j only ever grows): the current range's {bias, end} stay in registers and the
            // pointer moves on only when j has left the range
            uint32_t lp = listAddr;
            uint32_t bias, end;
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(bias), "=r"(end) : "r"(lp));
            auto locate = [&](uint32_t j) -> uint32_t {
                while (j >= end) {
                    lp += 8u;
                    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(bias), "=r"(end) : "r"(lp));
                }
                return bias + j;
            };
            uint32_t j = (uint32_t)lane;
            if (j < total) {
                uint32_t q = locate(j);
                float4 A0 = __ldg(ph.a + q), A1; uint4 B0 = __ldg(ph.b + q), B1;
                // the next trip's two words are requested before this trip's tests (two trips per loop turn: no register moves)
                while (true) {
                    const bool more1 = j + 32 < total;
                    if (more1) { q = locate(j + 32); A1 = __ldg(ph.a + q); B1 = __ldg(ph.b + q); }
                    test(A0, B0);
                    if (!more1) break;
                    const bool more2 = j + 64 < total;
                    if (more2) { q = locate(j + 64); A0 = __ldg(ph.a + q); B0 = __ldg(ph.b + q); }
                    test(A1, B1);
                    if (!more2) break;
                    j += 64;
                }
            }
        }
        // the G sums are reduced over the warp; lane g then finishes vertex g (albedo, normalisation, one atomic per vertex)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
            for (int g = 0; g < kMergeGather; ++g) {
                sum[g].x += __shfl_xor_sync(0xffffffffu, sum[g].x, o); sum[g].y += __shfl_xor_sync(0xffffffffu, sum[g].y, o); sum[g].z += __shfl_xor_sync(0xffffffffu, sum[g].z, o);
            }
        }
        if (lane < real) {
            V3 mine = sum[0];
#pragma unroll
            for (int g = 1; g < kMergeGather; ++g) if (lane == g) mine = sum[g];
            const float4* rp = reinterpret_cast<const float4*>(mq.recs + (mq.order ? __ldg(mq.order + base + lane) : base + lane));
            const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1), r3 = __ldg(rp + 3), r4 = __ldg(rp + 4);
            const Shading sh = resolveShading(sc.mats, __float_as_int(r1.w), sc.textures, make_float2(r4.x, r4.y));
            mine = mine * (V3(r3.x, r3.y, r3.z) * sh.m.albedo * (norm * (1.0f / kPi)));
            if ((mine.x != 0.0f || mine.y != 0.0f || mine.z != 0.0f) && isfinite(mine.x) && isfinite(mine.y) && isfinite(mine.z))
                atomicAdd(&P.colors[__float_as_int(r0.w)], make_float4(mine.x, mine.y, mine.z, 0.f));
        }
    }
    unsigned long long acc = nAcc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { acc += __shfl_xor_sync(0xffffffffu, acc, o); nCand += __shfl_xor_sync(0xffffffffu, nCand, o); }
    if (lane == 0) { if (acc) atomicAdd(&st.stats[4], acc); if (nCand) atomicAdd(&st.stats[5], nCand); }
    if (blockIdx.x == 0 && threadIdx.x == 0 && n > 0) atomicAdd(&st.stats[6], (unsigned long long)n);
}

// per-eye-vertex constants of the general merge, hoisted out of the photon loop
struct MergeCtx {
    Shading sh; Frame fr; V3 toPrevL, scale;     // scale = currBeta * 1/(pi r^2 w h)
    float eyeDvcmOverEta, eyeDvm;
    NV_D MergeCtx(const Shading& s, V3 n) : sh(s), fr(n) {}
};

NV_D V3 mergeOne(const BdParams& P, const MergeCtx& c, uint32_t q) {
    const uint4 B = __ldg(P.photons.b + q);
    const float2 wxy = unpackHalf2(__float_as_uint(__ldg(&P.photons.a[q].w))), wzb = unpackHalf2(B.x), byz = unpackHalf2(B.y);
    const float lightDvm = __uint_as_float(B.z), lightDvcmOverEta = __uint_as_float(B.w);
    const V3 wiPhoton(wxy.x, wxy.y, wzb.x);                 // photon -> its previous vertex, world space
    const V3 beta(wzb.y, byz.x, byz.y);
    V3 photonPrevToPhotonL = c.fr.toLocal(-wiPhoton);
    float eyeRev = bdPdf(c.sh, photonPrevToPhotonL, c.toPrevL, 1.0f);
    float lightRev = bdPdf(c.sh, -c.toPrevL, -photonPrevToPhotonL, 1.0f);
    float wEye = c.eyeDvcmOverEta + eyeRev * c.eyeDvm;
    float wLight = lightDvcmOverEta + lightRev * lightDvm;
    float w = __fdividef(1.0f, 1.0f + wEye + wLight);
    V3 f = bdEval(c.sh, photonPrevToPhotonL, c.toPrevL, 1.0f);
    return applyWeight(P, beta * f * c.scale, w);
}

__global__ void __launch_bounds__(kMergeThreads) bdMerge(SceneView sc, BdState st, BdParams P, MergeQueue mq) {
    __shared__ uint32_t accBuf[kMergeWarps][64];
    int* cnt = mq.cnt;
    const int lane = bdLane(), warp = threadIdx.x >> 5;
    const int n = min(cnt[BC_MERGE_GEN], mq.cap);
    const float r2 = P.mergeRadius * P.mergeRadius;
    const float norm = 1.0f / (kPi * r2 * (float)P.w * (float)P.h);
    const unsigned ltMask = (1u << lane) - 1u;
    const PhotonSoA ph = P.photons;
    unsigned long long nMerged = 0, nCand = 0;
    while (true) {
        int base = 0;
        if (lane == 0) base = atomicAdd(&cnt[BC_FETCH_MERGE_GEN], 4);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= n) break;
        const int end = min(base + 4, n);
        for (int r = base; r < end; ++r) {
            const float4* rp = reinterpret_cast<const float4*>(mq.recs + (mq.cap - 1 - r));
            const float4 r0 = __ldg(rp), r1 = __ldg(rp + 1), r2q = __ldg(rp + 2), r3 = __ldg(rp + 3), r4 = __ldg(rp + 4);
            const V3 eyeP(r0.x, r0.y, r0.z);
            const int pix = __float_as_int(r0.w);
            MergeCtx ctx(resolveShading(sc.mats, __float_as_int(r1.w), sc.textures, make_float2(r4.x, r4.y)), V3(r1.x, r1.y, r1.z));
            ctx.toPrevL = V3(r2q.x, r2q.y, r2q.z); ctx.scale = V3(r3.x, r3.y, r3.z) * norm;
            ctx.eyeDvcmOverEta = r2q.w / P.etaVcm; ctx.eyeDvm = r3.w;
            const int3 c0 = gridIndex(eyeP.x, eyeP.y, eyeP.z, P.grid);
            int count = 0;
            V3 sum(0.0f);
#pragma unroll 1
            for (int round = 0; round < (kMergeNbr + 31) / 32; ++round) {
            const MergeCells mc = mergeCellFine(P, eyeP, c0, round * 32 + lane, r2);
            nCand += mc.cc;
            unsigned nonEmpty = __ballot_sync(0xffffffffu, mc.cc != 0u);
            while (nonEmpty) {
                const int c = __ffs(nonEmpty) - 1;
                nonEmpty &= nonEmpty - 1;
                const uint32_t ccount = __shfl_sync(0xffffffffu, mc.cc, c);
                const uint32_t cstart = __shfl_sync(0xffffffffu, mc.cs, c);
                for (uint32_t j0 = 0; j0 < ccount; j0 += 32) {
                    // one sweep step: test a candidate per lane, compact the accepted ones, flush a full warp of them
                    const uint32_t q = cstart + j0 + lane;
                    bool accept = false;
                    if (j0 + lane < ccount) {
                        const float4 A = __ldg(ph.a + q);
                        V3 diff = eyeP - V3(A.x, A.y, A.z);
                        accept = dot(diff, diff) <= r2;
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, accept);
                    if (accept) accBuf[warp][count + __popc(m & ltMask)] = q;
                    count += __popc(m);
                    __syncwarp();
                    if (count >= 32) {
                        sum += mergeOne(P, ctx, accBuf[warp][lane]);
                        uint32_t carry = accBuf[warp][32 + lane];
                        __syncwarp();
                        accBuf[warp][lane] = carry;
                        count -= 32; nMerged += 32;
                        __syncwarp();
                    }
                }
            }
            }
            if (lane < count) sum += mergeOne(P, ctx, accBuf[warp][lane]);
            nMerged += count;
            __syncwarp();
            mergeSplat(P, pix, sum, lane);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) nCand += __shfl_xor_sync(0xffffffffu, nCand, o);
    if (lane == 0) { if (nMerged) atomicAdd(&st.stats[4], nMerged); if (nCand) atomicAdd(&st.stats[5], nCand); }
    if (blockIdx.x == 0 && threadIdx.x == 0 && n > 0) atomicAdd(&st.stats[6], (unsigned long long)n);
}

// Order of a queue's Lambert merge records for the gather: sorted by the Morton code of the photon-grid cell the vertex lies in (10 bits
// per axis; `shift` drops low cell bits when the scene spans more than 1024 cells), so that the 4 consecutive records a warp gathers
// together lie in one cell or in Morton-adjacent ones, and neighbouring warps read neighbouring photon cells (L2). The count is on the
// device (cnt[BC_MERGE]): keys past it are never read by the device-count radix sort.
__global__ void bdMergeKeys(MergeQueue mq, GridParams g, int shift, uint32_t* __restrict__ keys, int* __restrict__ order) {
    const int n = min(mq.cnt[BC_MERGE], mq.cap);
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = __ldg(reinterpret_cast<const float4*>(mq.recs + i));
        const int3 c = gridIndex(p.x, p.y, p.z, g);
        const uint32_t cx = (uint32_t)min(max(c.x >> shift, 0), 1023), cy = (uint32_t)min(max(c.y >> shift, 0), 1023), cz = (uint32_t)min(max(c.z >> shift, 0), 1023);
        auto spread = [](uint32_t v) { v = (v | (v << 16)) & 0x030000FFu; v = (v | (v << 8)) & 0x0300F00Fu; v = (v | (v << 4)) & 0x030C30C3u; v = (v | (v << 2)) & 0x09249249u; return v; };
        keys[i] = spread(cx) | (spread(cy) << 1) | (spread(cz) << 2);
        order[i] = i;
    }
}

// end of a light pass: publish this rank's photon count in every peer's copy of the slab header and then, behind a system-scope
// fence, the stamp that says "the records and the count of light pass `stamp` are all there" - in the peers' copies and in its own.
// Header = two 16 B words: {count, 0, 0, 0} {stamp, 0, 0, 0}.
__global__ void bdPushHeader(BdParams P) {
    __threadfence_system();
    const int t = threadIdx.x;
    if (t < P.numPeers) {
        const uint4 h = *P.slab;
        *P.peerSlab[t] = h;
        __threadfence_system();
        *reinterpret_cast<volatile unsigned*>(P.peerSlab[t] + 1) = P.stamp;
    } else if (t == P.numPeers) {
        __threadfence();
        *reinterpret_cast<volatile unsigned*>(P.slab + 1) = P.stamp;
    }
}

// A lane engine's rendezvous with its peers, on the device: spins until every rank's slab header in this buffer carries `stamp`
// (the peers store it over NVLink, bdPushHeader). Replaces a host-issued collective per iteration. Gives up after ~4 s
// (a peer that died must not hang this GPU) and reports through flag[0].
__global__ void bdWaitSlabs(const uint4* __restrict__ slabs, size_t slabWords, int world, unsigned stamp, int* __restrict__ flag) {
    const int r = threadIdx.x;
    if (r >= world) return;
    const volatile unsigned* p = reinterpret_cast<const volatile unsigned*>(slabs + (size_t)r * slabWords + 1);
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    while (*p != stamp) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        if (t1 - t0 > 4000000000ull) { atomicExch(flag, 1 + r); return; }
        __nanosleep(200);
    }
    __threadfence_system();
}

// Clears inside an iteration are KERNELS, and the photon counts reach the host through mapped page-locked memory, not a copy: with
// several lanes and ranks waiting on each other on the device (bdWaitSlabs), a cudaMemsetAsync / cudaMemcpyAsync of one lane can sit in
// the copy engine's queue behind another lane's copy that waits for a peer, whose own clear sits behind ITS other lane's copy ...
// (a cross-rank, cross-lane cycle that only the wait kernel's time-out broke: seen at 8 x 3 lanes and at 4 x 2 lanes).
__global__ void bdFillWords(uint4* __restrict__ p, size_t n16, unsigned v) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) p[i] = make_uint4(v, v, v, v);
}

__global__ void bdZeroInts(int* __restrict__ p, int n) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) p[i] = 0;
}

// ---- photon slabs -> contiguous array -> cell order -----------------------------------------------------------------------------
// counts[0, 16): photons per slab; hostCounts (mapped page-locked memory) receives them and the wait kernel's time-out flag
__global__ void bdSlabCounts(const uint4* __restrict__ slabs, size_t slabWords, int world, int cap, int* __restrict__ counts,
                             int* __restrict__ hostCounts) {
    int r = threadIdx.x;
    int c = 0;
    if (r < world) { c = min(*reinterpret_cast<const int*>(slabs + (size_t)r * slabWords), cap); counts[r] = c; }
    if (r < 16) hostCounts[r] = c;
    if (r == 16) hostCounts[16] = counts[16];
    __threadfence_system();
}
__global__ void bdCompactSlabs(const uint4* __restrict__ slabs, size_t slabWords, int world, const int* __restrict__ counts,
                               uint4* __restrict__ out) {
    // blockIdx.y = slab; records copied word by word (two 16 B words per photon)
    int r = blockIdx.y;
    int n = counts[r];
    int off = 0;
    for (int k = 0; k < r; ++k) off += counts[k];
    const uint4* src = slabs + (size_t)r * slabWords + kSlabHeaderBytes / 16;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < (size_t)n * 2; i += (size_t)gridDim.x * blockDim.x)
        out[(size_t)off * 2 + i] = src[i];
}
// gathers the records into cell order and rewrites them for the merge kernels: the octahedral direction is decoded here, once
// per photon, instead of once per accepted (photon, eye vertex) pair, and d_vcm is pre-divided by eta_vcm
__global__ void bdReorderPhotons(const uint4* __restrict__ in, const uint32_t* __restrict__ index, int n, float invEta,
                                 float4* __restrict__ oa, uint4* __restrict__ ob) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    size_t s = (size_t)index[i] * 2;
    uint4 a = __ldg(in + s), b = __ldg(in + s + 1);
    V3 wi = unpackOct(a.w);
    float2 bxy = unpackHalf2(b.x), bz = unpackHalf2(b.y);
    oa[i] = make_float4(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(packHalf2(wi.x, wi.y)));
    b.x = packHalf2(wi.z, bxy.x);
    b.y = packHalf2(bxy.y, bz.x);
    b.w = __float_as_uint(__uint_as_float(b.w) * invEta);
    ob[i] = b;
}

}  // namespace nv

#include "vcm_host.inl"
