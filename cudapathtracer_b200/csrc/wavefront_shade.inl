// wavefront_shade.inl — the shading stage: one warp-uniform material class per 32-path packet.
// Follows Li_unidirectional (deviceCode.cu:285-542) bounce by bounce, including its quirks: pdf floor 0.01
// (:481), EPSILON ray offsets (:504-506) but RAY_EPSILON pass-through on false hits (:517), emission counted
// directly until the first non-specular vertex (:438) and MIS-weighted with neePDF afterwards (:442-453), the
// un-normalised first-vertex light normal in the NEE pdf (:76, :134), Russian roulette only past maxDepth (:523).

namespace nv {

struct PathRegs {
    V3 o, d, beta, prevP;
    float pdf;
    int depth; unsigned flags;
    int pix;
    MediumStack ms;
    Rng rng;
};

NV_D void loadPath(const WfState& st, int slot, PathRegs& p) {
    const float4* w = st.slots + (size_t)slot * kSlotWords;
    float4 a = w[W_RAYO], b = w[W_RAYD], c = w[W_BETA], e = w[W_PREVP], x = w[W_AUX];
    uint4 m = *reinterpret_cast<const uint4*>(w + W_MEDIUM);
    p.o = V3(a.x, a.y, a.z); p.pdf = a.w;
    p.d = V3(b.x, b.y, b.z);
    unsigned df = __float_as_uint(b.w); p.depth = (int)(df & 0xffffu); p.flags = df & 0xffff0000u;
    p.beta = V3(c.x, c.y, c.z); p.pix = __float_as_int(c.w);
    p.prevP = V3(e.x, e.y, e.z); p.ms.top = __float_as_int(e.w);
    p.ms.w[0] = m.x; p.ms.w[1] = m.y; p.ms.w[2] = m.z; p.ms.w[3] = m.w;
    p.rng.state = (unsigned long long)__float_as_uint(x.z) | ((unsigned long long)__float_as_uint(x.w) << 32);
}
NV_D void storePath(const WfState& st, int slot, const PathRegs& p) {
    float4* w = st.slots + (size_t)slot * kSlotWords;
    w[W_RAYO] = make_float4(p.o.x, p.o.y, p.o.z, p.pdf);
    w[W_RAYD] = make_float4(p.d.x, p.d.y, p.d.z, __uint_as_float(((unsigned)p.depth & 0xffffu) | p.flags));
    w[W_BETA] = make_float4(p.beta.x, p.beta.y, p.beta.z, __int_as_float(p.pix));
    w[W_PREVP] = make_float4(p.prevP.x, p.prevP.y, p.prevP.z, __int_as_float(p.ms.top));
    *reinterpret_cast<uint4*>(w + W_MEDIUM) = make_uint4(p.ms.w[0], p.ms.w[1], p.ms.w[2], p.ms.w[3]);
    // the RNG state shares a word with the hit flags extend writes: store only its two lanes
    *reinterpret_cast<float2*>(&w[W_AUX].z) = make_float2(__uint_as_float((unsigned)p.rng.state), __uint_as_float((unsigned)(p.rng.state >> 32)));
}

// A non-finite sample (e.g. the NEE weight inf/inf when the shading point lies exactly in the light's plane, deviceCode.cu:468-471)
// poisons the pixel for good in the reference, which paints such pixels magenta at the end (main.cu:863-868); here the one
// sample is dropped instead so the accumulator stays usable.
NV_D void addColor(float4* colors, int pix, V3 c) {
    if (!(isfinite(c.x) && isfinite(c.y) && isfinite(c.z))) return;
    if (c.x != 0.0f || c.y != 0.0f || c.z != 0.0f) atomicAdd(&colors[pix], make_float4(c.x, c.y, c.z, 0.f));
}

// neePDF (deviceCode.cu:63-85): area-to-solid-angle pdf of having picked the hit point on light triangle `tri`
NV_D float neePdfOf(const SceneView& sc, int tri, V3 prevPoint, V3 hitPoint) {
    int4 idx = __ldg(reinterpret_cast<const int4*>(sc.tris + tri));     // a b c na
    V3 a(__ldg(sc.positions + idx.x)), b(__ldg(sc.positions + idx.y)), c(__ldg(sc.positions + idx.z));
    V3 stl = hitPoint - prevPoint;
    V3 wi = normalize(stl);
    float d2 = lengthSq(stl);
    V3 ln(__ldg(sc.normals + idx.w));
    float cosL = dot(ln, -wi);
    float area = 0.5f * length(cross(b - a, c - a));
    return d2 / (cosL * sc.numLights * area);
}

// One bounce of the NEE+MIS integrator. Returns true if the path continues (p holds the next ray).
NV_D bool shadeMIS(const SceneView& sc, const WfState& st, const WfParams& P, PathRegs& p, const HitRec& h, int hitFlags) {
    const DMat* mats = sc.mats;
    Surface s = makeSurface(sc, h, p.o, p.d);
    Frame fr(s.n);
    V3 wiLocal = fr.toLocal(p.d);
    Shading sh = resolveShading(mats, s.mat, sc.textures, s.uv);
    const bool isSpecular = (sh.m.flags & MF_SPECULAR) != 0;
    bool trueHit = true;
    // dominant medium = lowest priority on the stack (:337-346)
    int minPriorID = p.ms.get(0);
    int minPrior = __ldg(&mats[minPriorID].priority);
    for (int i = 1; i < p.ms.top; ++i) {
        int id = p.ms.get(i); int pr = __ldg(&mats[id].priority);
        if (pr < minPrior) { minPrior = pr; minPriorID = id; }
    }
    MatView dom = loadMat(mats, minPriorID);
    if (s.dist > kEps) {     // Beer-Lambert through the dominant medium (:347-358)
        p.beta *= V3(expf(-dom.absorption.x * s.dist), expf(-dom.absorption.y * s.dist), expf(-dom.absorption.z * s.dist));
    }
    float etaI = dom.ior;
    if (sh.m.flags & MF_BOUNDARY) {
        if (sh.m.priority > minPrior) {       // false hit: a lower-priority surface inside the dominant medium (:413-425)
            trueHit = false;
            if (!s.backface) p.ms.push(s.mat);
            else p.ms.remove(s.mat);
        }
        // true hit: the reference derives etaT here (:372-411) but the live dielectric sampler ignores it
    }
    int dEff = p.depth;
    if (trueHit) {
        bool emissive = lengthSq(s.emission) > kEps;
        if (emissive) {
            if (p.depth == 0 || !(p.flags & PF_NONSPEC_SEEN)) addColor(P.colors, p.pix, p.beta * s.emission);
            else if (P.useMIS && !isSpecular) {
                float lp = neePdfOf(sc, s.tri, p.prevP, s.p);
                if (lp > kEps) {
                    float wgt = p.pdf * p.pdf / (lp * lp + p.pdf * p.pdf);
                    addColor(P.colors, p.pix, p.beta * s.emission * wgt);
                }
            }
        }
        if (P.useMIS && lengthSq(s.emission) < kEps && !isSpecular && sc.numLights > 0) {
            // nextEventEstimation (:87-156); the shadow ray itself is deferred to the shadow stage
            int index = min((int)(p.rng.uniform() * sc.numLights), sc.numLights - 1);
            const NovumTriangle* lt = sc.lights + index;
            int4 idx = __ldg(reinterpret_cast<const int4*>(lt));
            float4 em = __ldg(reinterpret_cast<const float4*>(lt) + 3);
            float4 pa4 = __ldg(sc.positions + idx.x), pb4 = __ldg(sc.positions + idx.y), pc4 = __ldg(sc.positions + idx.z);
            V3 a(pa4), b(pb4), c(pc4);
            float u = sqrtf(p.rng.uniform());
            float v = p.rng.uniform();
            V3 lp = (1.0f - u) * a + (u * (1.0f - v)) * b + (u * v) * c;
            V3 stl = lp - s.p;
            V3 wi = normalize(stl);
            V3 ro = s.p + wi * kEps;
            float tl = 0.f, tu, tv;
            float4 e1 = make_float4(pb4.x - pa4.x, pb4.y - pa4.y, pb4.z - pa4.z, 0.f);
            float4 e2 = make_float4(pc4.x - pa4.x, pc4.y - pa4.y, pc4.z - pa4.z, 0.f);
            if (!triTest(pa4, e1, e2, ro, wi, tl, tu, tv)) tl = 0.f;
            float d2 = lengthSq(stl);
            V3 ln(__ldg(sc.normals + idx.w));
            float cosL = dot(ln, -wi);
            float cosS = fabsf(dot(s.n, wi));
            float area = 0.5f * length(cross(b - a, c - a));
            float lightPdf = d2 / (cosL * sc.numLights * area);
            if (lightPdf > kEps) {
                V3 toLight = fr.toLocal(wi);
                V3 f = bsdfEval(sh, wiLocal, toLight, etaI);
                float bp = bsdfPdf(sh, wiLocal, toLight, etaI);
                float wgt = lightPdf * lightPdf / (bp * bp + lightPdf * lightPdf);
                V3 contrib = p.beta * (f * V3(em.x, em.y, em.z) * cosS / lightPdf) * wgt;
                bool push = (contrib.x != 0.0f || contrib.y != 0.0f || contrib.z != 0.0f) && isfinite(contrib.x) && isfinite(contrib.y) && isfinite(contrib.z);
                // (no warp aggregation: lanes reach this point divergently inside a class only via lightPdf)
                if (push) {
                    int pos = atomicAdd(&st.counters[C_SHADOW_COUNT], 1);
                    float4* rec = reinterpret_cast<float4*>(st.shadowQ + pos);
                    rec[0] = make_float4(ro.x, ro.y, ro.z, tl * (1.0f - kEps));
                    rec[1] = make_float4(wi.x, wi.y, wi.z, __int_as_float(p.pix));
                    rec[2] = make_float4(contrib.x, contrib.y, contrib.z, 0.f);
                }
            }
        }
        V3 wo, f; float pdf;
        bsdfSample(p.rng, sh, wiLocal, etaI, s.backface, MODE_RADIANCE, wo, f, pdf);
        V3 woWorld = fr.toWorld(wo);
        pdf = fmaxf(pdf, 0.01f);
        if (wo.z < 0.0f) {       // refracted: enter / leave the medium (:484-497)
            if (!s.backface) p.ms.push(s.mat);
            else p.ms.remove(s.mat);
        }
        p.beta *= (f * fabsf(wo.z) / pdf);
        p.o = (wo.z > 0.0f) ? s.p + s.n * kEps : s.p - s.n * kEps;
        p.d = normalize(woWorld);
        p.pdf = pdf;
        p.prevP = s.p;
    } else {
        p.o = s.p + p.d * kRayEps;       // pass straight through (:517); depth is not consumed (:520)
        dEff = p.depth - 1;
    }
    if (dEff > P.maxDepth) {             // Russian roulette (:523-532)
        float lum = luminance(p.beta);
        float q = clampf(lum, 0.05f, 0.99f);
        if (p.rng.uniform() > q) return false;
        p.beta /= q;
    }
    if (!isSpecular) p.flags |= PF_NONSPEC_SEEN;
    p.depth = dEff + 1;
    if (p.depth >= 100) return false;    // hard bounce cap (:318)
    // a path whose throughput is exactly zero can never contribute again: stop tracing it (the reference keeps going)
    if (p.beta.x == 0.0f && p.beta.y == 0.0f && p.beta.z == 0.0f) return false;
    return true;
}

// One bounce of the naive integrator (deviceCode.cu:158-205)
NV_D bool shadeNaive(const SceneView& sc, const WfParams& P, PathRegs& p, const HitRec& h) {
    Surface s = makeSurface(sc, h, p.o, p.d);
    Frame fr(s.n);
    V3 toSurf = fr.toLocal(p.d);
    Shading sh = resolveShading(sc.mats, s.mat, sc.textures, s.uv);
    V3 wo, f; float pdf;
    bsdfSample(p.rng, sh, toSurf, 1.0f, s.backface, MODE_RADIANCE, wo, f, pdf);
    if (pdf <= 0.0f || lengthSq(f) < kEps) return false;
    addColor(P.colors, p.pix, s.emission * p.beta);
    p.beta *= f * fabsf(wo.z) / pdf;
    p.o = s.p + ((wo.z > 0.0f) ? (s.n * kRayEps) : (-s.n * kRayEps));
    p.d = fr.toWorld(wo);
    p.depth += 1;
    return p.depth < P.maxDepth;
}

// bin of a continuing ray: origin cell of the 2^b-per-axis grid over the scene bounds (clamped: the camera may stand outside), direction octant
NV_D int rayBin(const SceneView& sc, V3 o, V3 d) {
    constexpr int kCells = 1 << kRayCellBits;
    const float sx = (float)kCells * 0.5f / fmaxf(sc.sceneCenter[0] - sc.sceneMin[0], 1e-20f);
    const float sy = (float)kCells * 0.5f / fmaxf(sc.sceneCenter[1] - sc.sceneMin[1], 1e-20f);
    const float sz = (float)kCells * 0.5f / fmaxf(sc.sceneCenter[2] - sc.sceneMin[2], 1e-20f);
    const int cx = min(max((int)((o.x - sc.sceneMin[0]) * sx), 0), kCells - 1);
    const int cy = min(max((int)((o.y - sc.sceneMin[1]) * sy), 0), kCells - 1);
    const int cz = min(max((int)((o.z - sc.sceneMin[2]) * sz), 0), kCells - 1);
    const int oct = (d.x < 0.0f ? 1 : 0) | (d.y < 0.0f ? 2 : 0) | (d.z < 0.0f ? 4 : 0);
    return ((((cz << kRayCellBits) | cy) << kRayCellBits | cx) << 3) | oct;
}

#ifndef NOVUM_SHADE_MINB
#define NOVUM_SHADE_MINB 5
#endif
__global__ void __launch_bounds__(kWfThreads, NOVUM_SHADE_MINB) wfShade(SceneView sc, WfState st, WfParams P, int* __restrict__ nextQueue) {
    const int lane = laneId();
    int cnt[NUM_CLS], chunkStart[NUM_CLS + 1];
    chunkStart[0] = 0;
#pragma unroll
    for (int c = 0; c < NUM_CLS; ++c) { cnt[c] = st.counters[C_CLASS_COUNT + c]; chunkStart[c + 1] = chunkStart[c] + (cnt[c] + 31) / 32; }
    const int totalChunks = chunkStart[NUM_CLS];
    const unsigned long long idsPerSample = (unsigned long long)P.tilesX * P.tilesY * 32ull;
    unsigned long long* pathCounter = reinterpret_cast<unsigned long long*>(st.counters + C_PATH_COUNTER_LO);
    unsigned long long finished = 0, started = 0;
    while (true) {
        int chunk = 0;
        if (lane == 0) chunk = atomicAdd(&st.counters[C_FETCH_SHADE], 1);
        chunk = __shfl_sync(0xffffffffu, chunk, 0);
        if (chunk >= totalChunks) break;
        int cls = 0;
#pragma unroll
        for (int c = 1; c < NUM_CLS; ++c) if (chunk >= chunkStart[c]) cls = c;
        int i = (chunk - chunkStart[cls]) * 32 + lane;
        bool valid = i < cnt[cls];
        int slot = valid ? st.classQ[cls * st.capacity + i] : -1;
        bool alive = false;
        PathRegs p;
        if (valid && cls != CLS_REGEN) {
            loadPath(st, slot, p);
            float4 hf = st.slots[(size_t)slot * kSlotWords + W_HIT];
            HitRec h; h.tri = __float_as_int(hf.x); h.t = hf.y; h.u = hf.z; h.v = hf.w;
            h.flags = __float_as_int(st.slots[(size_t)slot * kSlotWords + W_AUX].x);
            alive = P.naive ? shadeNaive(sc, P, p, h) : shadeMIS(sc, st, P, p, h, h.flags);
            if (!alive) ++finished;
        } else if (valid && cls == CLS_REGEN) {
            // missed the scene: sampleSky is black (integratorUtilities.cuh:438); slots of the first iteration land here too.
            // FEAT_ENVIRONMENT: the escaping path collects the constant environment radiance. No strategy samples the environment from
            // the light side, so this is the only way such a path is found: full weight.
            if (sc.features & FEAT_ENVIRONMENT) {
                const float4 b = st.slots[(size_t)slot * kSlotWords + W_BETA];
                if (b.x != 0.0f || b.y != 0.0f || b.z != 0.0f) addColor(P.colors, __float_as_int(b.w), V3(b.x, b.y, b.z) * V3(sc.env[0], sc.env[1], sc.env[2]));
            }
        }
        // regeneration: dead lanes pull new path ids until they get a pixel inside the image or the work runs out
        bool need = valid && !alive;
        while (true) {
            unsigned m = __ballot_sync(0xffffffffu, need);
            if (m == 0) break;
            int leader = __ffs(m) - 1;
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(pathCounter, (unsigned long long)__popc(m));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (need) {
                unsigned long long id = base + __popc(m & ((1u << lane) - 1u));
                if (id >= P.totalIds) need = false;
                else {
                    unsigned long long k = id / idsPerSample;
                    unsigned l = (unsigned)(id - k * idsPerSample);
                    unsigned tile = l >> 5, in = l & 31u;
                    int x = (int)(tile % P.tilesX) * 8 + (int)(in & 7u), y = (int)(tile / P.tilesX) * 4 + (int)(in >> 3);
                    if (x < P.w && y < P.h) {
                        int sample = P.firstSample + P.rank + (int)k * P.world;
                        p.pix = y * P.w + x;
                        p.rng.seed(P.seed, (unsigned long long)p.pix, (unsigned long long)sample, 0);
                        cameraRay(P.cam, p.rng, x, y, p.o, p.d);
                        p.beta = V3(1.0f); p.prevP = V3(0.0f); p.pdf = kEps; p.depth = 0; p.flags = 0;
                        p.ms.w[0] = p.ms.w[1] = p.ms.w[2] = p.ms.w[3] = 0; p.ms.top = 1;   // slot 0 = material 0 = air (:308)
                        alive = true; need = false; ++started;
                    }
                }
            }
        }
        int pos = warpAppend(&st.counters[C_NEXT_COUNT], alive);
        if (!kRayBins) {
            if (alive) { storePath(st, slot, p); nextQueue[pos] = slot; }
        } else {
            const unsigned am = __ballot_sync(0xffffffffu, alive);
            if (alive) {
                storePath(st, slot, p);
                const int bin = rayBin(sc, p.o, p.d);
                const unsigned peers = __match_any_sync(am, bin);
                const int leader = __ffs(peers) - 1;
                int first = 0;
                if (lane == leader) first = atomicAdd(&st.binCount[bin], __popc(peers));
                first = __shfl_sync(peers, first, leader);
                st.binTmp[pos] = make_int2(slot, bin);
                st.binPos[pos] = first + __popc(peers & ((1u << lane) - 1u));
            }
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        finished += __shfl_xor_sync(0xffffffffu, finished, o);
        started += __shfl_xor_sync(0xffffffffu, started, o);
    }
    if (lane == 0) { if (finished) atomicAdd(&st.rayCounters[2], finished); if (started) atomicAdd(&st.rayCounters[3], started); }
}

}  // namespace nv
