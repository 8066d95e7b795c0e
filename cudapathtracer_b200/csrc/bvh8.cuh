// bvh8.cuh — compressed 8-wide BVH: node/triangle layout and the per-ray traversal routines.
//
// Replaces the reference's BVH2 stack traversal (integratorUtilities.cuh:84-186 closest hit,
// :188-288 shadow ray) and its 48 B AoS nodes (objects.cuh:12-25). Design:
//   * 80 B nodes = five 128-bit loads: origin p (3 f32), per-axis power-of-two scale exponents,
//     internal-child mask, child/triangle base indices, 8 meta bytes, 8x6 quantised child-box bytes
//     (outward rounded, so no hit the reference's float boxes admit is ever culled).
//   * 48 B triangles = three 128-bit loads: v0 | original triangle id, e1 | flags, e2.
//     e1 = b - a and e2 = c - a are the same float subtractions triangleIntersect performs
//     (integratorUtilities.cuh:13-14), so the Moller-Trumbore below is arithmetically identical.
//   * octant-ordered traversal with a bit-mask hit group, short stack in shared memory with a
//     local-memory overflow, pruning against the current closest t (the reference never prunes).
//   * ties on bit-equal t resolve to the LOWEST original triangle id (the reference keeps the
//     first found in its own BVH2 order, which no other structure can reproduce).
#pragma once
#include "vecmath.cuh"

namespace nv {

struct Bvh8View {
    const float4* __restrict__ nodes;   // 5 float4 per node, node 0 = root
    const float4* __restrict__ tris;    // 3 float4 per triangle, BVH order
    int numNodes, numTris;
};

// triangle flag bits (stored in e1.w as int)
enum : int { TF_LEAFMAT = 1 << 30 };    // material type is MAT_LEAF (shadow rays pass through, attenuated)
// low 16 bits of flags = materialID

struct HitRec { int tri; float t, u, v; int flags; };

constexpr int kSmemStack = 12;          // entries per thread kept in shared memory
constexpr int kLocalStack = 40;         // overflow entries in local memory
#ifndef NOVUM_TRI_POSTPONE
#define NOVUM_TRI_POSTPONE 0
#endif
// Triangle postponing (Ylitie et al. 2017, sec. 4.4): when fewer than this many lanes of the warp are in the leaf loop and this
// lane still has inner nodes to visit, its leaf group goes on the stack and the lane returns to node tests with the others.
constexpr int kTriPostpone = NOVUM_TRI_POSTPONE;

struct TravCounters { unsigned long long* nodeVisits; unsigned long long* triTests; };

NV_D unsigned extractByte(unsigned x, unsigned i) { return (x >> (i * 8u)) & 0xffu; }
NV_D unsigned signExtendS8x4(unsigned x) {   // per byte: 0x80 set -> 0xff else 0x00
    unsigned m = (x >> 7) & 0x01010101u;
    return m * 0xffu;
}

// Moller-Trumbore in the reference's exact operation order (integratorUtilities.cuh:8-42).
// f = 1.0/a is a double divide rounded to float there; with 53 >= 2*24+2 bits the double rounding
// is innocuous, so the correctly rounded float reciprocal is the same number.
// The FMA contraction is spelled out with intrinsics so it cannot drift with inlining context: it is the
// pattern nvcc 12.9 -O3 (default -fmad=true) emits for the reference's dot()/cross3() (util.cuh:116-139),
// read off the SASS of triangleIntersect compiled for sm_100a:
//   dot(a,b)   = fma(a.z,b.z, fma(a.x,b.x, a.y*b.y))
//   cross(a,b) = ( fma(a.y,b.z,-(a.z*b.y)), fma(a.z,b.x,-(a.x*b.z)), fma(a.x,b.y,-(a.y*b.x)) )
// and `1.0/a` becomes the correctly rounded float reciprocal (MUFU.RCP + Newton step + slow path).
NV_D float dotRef(V3 a, V3 b) { return __fmaf_rn(a.z, b.z, __fmaf_rn(a.x, b.x, __fmul_rn(a.y, b.y))); }
NV_D V3 crossRef(V3 a, V3 b) {
    return V3(__fmaf_rn(a.y, b.z, -__fmul_rn(a.z, b.y)), __fmaf_rn(a.z, b.x, -__fmul_rn(a.x, b.z)),
              __fmaf_rn(a.x, b.y, -__fmul_rn(a.y, b.x)));
}
NV_D bool triTest(const float4 v0, const float4 e1f, const float4 e2f, V3 o, V3 d, float& t, float& u, float& v) {
    V3 e1(e1f.x, e1f.y, e1f.z), e2(e2f.x, e2f.y, e2f.z);
    V3 h = crossRef(d, e2);
    float a = dotRef(h, e1);
    if (fabsf(a) < 1e-12f) return false;
    float f = __frcp_rn(a);
    V3 s(__fsub_rn(o.x, v0.x), __fsub_rn(o.y, v0.y), __fsub_rn(o.z, v0.z));
    u = __fmul_rn(f, dotRef(s, h));
    V3 q = crossRef(s, e1);
    v = __fmul_rn(f, dotRef(d, q));
    t = __fmul_rn(f, dotRef(e2, q));
    return (u >= 0.0f) && (v >= 0.0f) && (__fadd_rn(u, v) <= 1.0f) && (t > 0.0f);
}

struct RaySetup {
    V3 o, d, idir;
    unsigned octinv4;
};
NV_D RaySetup setupRay(V3 o, V3 d) {
    RaySetup r; r.o = o; r.d = d;
    const float tiny = 1e-30f;
    // the sign of the substitute must agree with the `d < 0` tests below and in intersectNode: -0.0f counts as positive
    r.idir.x = 1.0f / (fabsf(d.x) > tiny ? d.x : (d.x < 0.0f ? -tiny : tiny));
    r.idir.y = 1.0f / (fabsf(d.y) > tiny ? d.y : (d.y < 0.0f ? -tiny : tiny));
    r.idir.z = 1.0f / (fabsf(d.z) > tiny ? d.z : (d.z < 0.0f ? -tiny : tiny));
    unsigned oct = (d.x < 0.0f ? 4u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 1u : 0u);
    r.octinv4 = (7u - oct) * 0x01010101u;
    return r;
}

// Intersects the 8 quantised child boxes of one node; returns the hit mask
// (bits 24..31: internal children in octant order, bits 0..23: leaf triangles).
NV_D unsigned intersectNode(const RaySetup& r, float tmax, const float4 n0, const float4 n1, const float4 n2,
                            const float4 n3, const float4 n4) {
    unsigned eimask = __float_as_uint(n0.w);
    float ax = __uint_as_float((eimask & 0xffu) << 23) * r.idir.x;
    float ay = __uint_as_float(((eimask >> 8) & 0xffu) << 23) * r.idir.y;
    float az = __uint_as_float(((eimask >> 16) & 0xffu) << 23) * r.idir.z;
    float ox = (n0.x - r.o.x) * r.idir.x, oy = (n0.y - r.o.y) * r.idir.y, oz = (n0.z - r.o.z) * r.idir.z;
    unsigned hitmask = 0;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
        unsigned meta4 = __float_as_uint(half == 0 ? n1.z : n1.w);
        unsigned isInner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        unsigned innerMask4 = signExtendS8x4(isInner4 << 3);
        unsigned bitIndex4 = (meta4 ^ (r.octinv4 & innerMask4)) & 0x1f1f1f1fu;
        unsigned childBits4 = (meta4 >> 5) & 0x07070707u;
        unsigned qlox = __float_as_uint(half == 0 ? n2.x : n2.y), qloy = __float_as_uint(half == 0 ? n2.z : n2.w);
        unsigned qloz = __float_as_uint(half == 0 ? n3.x : n3.y), qhix = __float_as_uint(half == 0 ? n3.z : n3.w);
        unsigned qhiy = __float_as_uint(half == 0 ? n4.x : n4.y), qhiz = __float_as_uint(half == 0 ? n4.z : n4.w);
        unsigned nx = r.d.x < 0.0f ? qhix : qlox, fx = r.d.x < 0.0f ? qlox : qhix;
        unsigned ny = r.d.y < 0.0f ? qhiy : qloy, fy = r.d.y < 0.0f ? qloy : qhiy;
        unsigned nz = r.d.z < 0.0f ? qhiz : qloz, fz = r.d.z < 0.0f ? qloz : qhiz;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float t0x = fmaf((float)extractByte(nx, j), ax, ox), t1x = fmaf((float)extractByte(fx, j), ax, ox);
            float t0y = fmaf((float)extractByte(ny, j), ay, oy), t1y = fmaf((float)extractByte(fy, j), ay, oy);
            float t0z = fmaf((float)extractByte(nz, j), az, oz), t1z = fmaf((float)extractByte(fz, j), az, oz);
            float cmin = fmaxf(fmaxf(t0x, t0y), fmaxf(t0z, 0.0f));
            float cmax = fminf(fminf(t1x, t1y), fminf(t1z, tmax));
            // small outward slack: the slab arithmetic above rounds, the boxes must stay conservative
            if (cmin <= cmax * 1.0000024f + 1e-30f) {
                unsigned bits = extractByte(childBits4, j);
                unsigned idx = extractByte(bitIndex4, j);
                hitmask |= bits << idx;
            }
        }
    }
    return hitmask;
}

// Per-thread traversal stack: first kSmemStack entries in shared memory (strided by blockDim.x so that
// a warp's accesses are conflict-free), overflow in local memory.
struct TravStack {
    uint2* smem;          // base for this thread: entry i at smem[i * stride]
    int stride;
    uint2 local[kLocalStack];
    int top;
    NV_D void init(uint2* s, int st) { smem = s; stride = st; top = 0; }
    NV_D void push(uint2 v) {
        if (top < kSmemStack) smem[top * stride] = v;
        else if (top - kSmemStack < kLocalStack) local[top - kSmemStack] = v;
        ++top;
    }
    NV_D uint2 pop() {
        --top;
        if (top < kSmemStack) return smem[top * stride];
        if (top - kSmemStack < kLocalStack) return local[top - kSmemStack];
        return make_uint2(0, 0);
    }
};

// Closest hit as a resumable state machine, so that a persistent warp can refill finished lanes with new rays between
// steps (trav_stream.cuh). tmaxIn mirrors the reference's max_t argument (999999 by default), skipTri its skipTri.
template <bool COUNT>
struct ClosestTrav {
    RaySetup r;
    HitRec best;
    float tlimit;                          // strict: t < max_t and t < min_t
    int skipTri;
    uint2 ngroup, tgroup;
    unsigned nv, tt;                       // node visits / triangle tests of this ray (COUNT builds)

    NV_D void begin(V3 o, V3 d, float tmaxIn, int skip, TravStack& st) {
        r = setupRay(o, d);
        best.tri = -1; best.t = 3.402823466e+38f; best.u = 0.f; best.v = 0.f; best.flags = 0;
        tlimit = tmaxIn; skipTri = skip;
        st.top = 0;
        ngroup = make_uint2(0u, 0x80000000u);
        tgroup = make_uint2(0u, 0u);
        nv = 0; tt = 0;
    }
    // One step: at most one node fetch + 8 box tests, then every triangle of the current leaf group. Returns true when done.
    NV_D bool step(const Bvh8View& bvh, TravStack& st) {
        if (ngroup.y > 0x00ffffffu) {
            unsigned hits = ngroup.y;
            unsigned bit = 31u - __clz(hits);
            ngroup.y &= ~(1u << bit);
            if (ngroup.y > 0x00ffffffu) st.push(ngroup);
            unsigned slot = (bit - 24u) ^ (r.octinv4 & 0xffu);
            unsigned rel = __popc(hits & ~(0xffffffffu << slot) & 0xffu);
            unsigned nodeIdx = ngroup.x + rel;
            const float4* np = bvh.nodes + (size_t)nodeIdx * 5;
            float4 n0 = __ldg(np), n1 = __ldg(np + 1), n2 = __ldg(np + 2), n3 = __ldg(np + 3), n4 = __ldg(np + 4);
            if (COUNT) ++nv;
            unsigned hm = intersectNode(r, fminf(tlimit, best.t), n0, n1, n2, n3, n4);
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
            if (id != skipTri && triTest(v0, e1, e2, r.o, r.d, t, u, v)) {
                if (t < tlimit && (t < best.t || (t == best.t && id < best.tri))) {
                    best.t = t; best.u = u; best.v = v; best.tri = id; best.flags = __float_as_int(e1.w);
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
NV_D HitRec traceClosest(const Bvh8View& bvh, V3 o, V3 d, float tmaxIn, int skipTri, TravStack& st, TravCounters cnt) {
    ClosestTrav<COUNT> tr;
    tr.begin(o, d, tmaxIn, skipTri, st);
    while (!tr.step(bvh, st)) {}
    tr.flushCounters(cnt);
    return tr.best;
}

}  // namespace nv
