// obj_parse.cuh — the per-line rules of the OBJ front end (main.cu:936-1068 as restated by frontend.cpp's loadObj), written once for
// the device parser (obj_device.cu, one thread per line). Everything here is a pure function of the text of ONE line, so the same
// source compiles as plain C++ for the CPU tests that hold these rules to loadObj (tests/obj_rules_host.cpp).
//
// A line is text[begin, end): `end` is the position of its '\n' (or the end of the file). Characters at or beyond `end` read as NUL,
// which is also what the host loader sees (it terminates every line in place); a NUL inside the line ends it early on both sides.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define OBJ_HD __host__ __device__ __forceinline__
#else
#define OBJ_HD inline
#endif

namespace nv {

// float arithmetic exactly as the host loader's: one rounding per operation, never contracted into an fma
#if defined(__CUDA_ARCH__)
#define OBJ_MUL(a, b) __fmul_rn((a), (b))
#define OBJ_ADD(a, b) __fadd_rn((a), (b))
#define OBJ_SUB(a, b) __fsub_rn((a), (b))
#else
#define OBJ_MUL(a, b) ((a) * (b))
#define OBJ_ADD(a, b) ((a) + (b))
#define OBJ_SUB(a, b) ((a) - (b))
#endif

enum ObjLineType { OBJ_SKIP = 0, OBJ_V = 1, OBJ_VT = 2, OBJ_VN = 3, OBJ_F = 4 };
constexpr int kObjMaxCorners = 16;        // a face with more corners is left to the host loader (NOVUM_E_UNSUPPORTED)

// 10^k, k = 0..22: exactly representable doubles (the device copy sits in constant memory instead of every thread's stack)
#define OBJ_POW10_LIST {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11, 1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22}
#if defined(__CUDACC__)
static __constant__ double kObjPow10Device[23] = OBJ_POW10_LIST;
#endif
static const double kObjPow10Host[23] = OBJ_POW10_LIST;
OBJ_HD double objPow10(int k) {
#if defined(__CUDA_ARCH__)
    return kObjPow10Device[k];
#else
    return kObjPow10Host[k];
#endif
}

OBJ_HD int objChr(const char* t, uint32_t p, uint32_t end) { return p < end ? (int)(unsigned char)t[p] : 0; }
OBJ_HD bool objSpace(int c) { return c == ' ' || (c >= '\t' && c <= '\r'); }          // isspace in the C locale
OBJ_HD bool objDigit(int c) { return c >= '0' && c <= '9'; }

// forEachLine + the dispatch of objSweepA / objSweepB: empty lines, comments and 's' lines are skipped; "vt" and "vn" are recognised by
// their two letters alone, "v" and "f" need a blank or a tab behind them
OBJ_HD int objLineType(const char* t, uint32_t begin, uint32_t end) {
    const int c0 = objChr(t, begin, end), c1 = objChr(t, begin + 1, end);
    if (begin >= end || c0 == '#' || c0 == 's') return OBJ_SKIP;
    if (c0 == 'v') {
        if (c1 == ' ' || c1 == '\t') return OBJ_V;
        if (c1 == 't') return OBJ_VT;
        if (c1 == 'n') return OBJ_VN;
        return OBJ_SKIP;
    }
    if (c0 == 'f' && (c1 == ' ' || c1 == '\t')) return OBJ_F;
    return OBJ_SKIP;
}

// One number as strtod reads it, for the spellings of frontend.cpp's fastStrtod fast path: [ws][sign]digits[.digits][e[sign]digits]
// with at most 15 significant digits and a decimal exponent within +-22 - an integer below 2^53 times or divided by an exact power of
// ten, ONE correctly rounded double operation (IEEE on both sides). Returns the position behind the number, or p when there is none.
// Spellings strtod would read some other way (inf, nan, hex, longer mantissas, larger exponents) set *unsupported: the caller gives
// the file to the host loader rather than guess.
OBJ_HD uint32_t objStrtod(const char* t, uint32_t p, uint32_t end, double* v, int* unsupported) {
    uint32_t s = p;
    while (objSpace(objChr(t, s, end))) ++s;
    bool neg = false;
    int c = objChr(t, s, end);
    if (c == '+' || c == '-') { neg = c == '-'; ++s; c = objChr(t, s, end); }
    const int c1 = objChr(t, s + 1, end);
    if (!(objDigit(c) || (c == '.' && objDigit(c1)))) {
        if (c == 'i' || c == 'I' || c == 'n' || c == 'N') *unsupported = 1;       // strtod reads "inf" / "nan(...)"
        return p;
    }
    if (c == '0' && (c1 == 'x' || c1 == 'X')) { *unsupported = 1; return p; }
    unsigned long long mant = 0; int sig = 0, exp10 = 0;
    for (; objDigit(c = objChr(t, s, end)); ++s) {
        if (sig > 0 || c != '0') { if (sig < 19) mant = mant * 10 + (unsigned)(c - '0'); else ++exp10; ++sig; }
    }
    if (c == '.') {
        ++s;
        for (; objDigit(c = objChr(t, s, end)); ++s) {
            if (sig > 0 || c != '0') { if (sig < 19) { mant = mant * 10 + (unsigned)(c - '0'); --exp10; } ++sig; }
            else --exp10;
        }
    }
    if (c == 'e' || c == 'E') {
        uint32_t u = s + 1; bool eneg = false;
        int d = objChr(t, u, end);
        if (d == '+' || d == '-') { eneg = d == '-'; ++u; d = objChr(t, u, end); }
        if (objDigit(d)) {
            int e = 0;
            for (; objDigit(d = objChr(t, u, end)); ++u) if (e < 100000) e = e * 10 + (d - '0');
            exp10 += eneg ? -e : e;
            s = u;
        }
    }
    if (sig > 15 || exp10 < -22 || exp10 > 22) { *unsupported = 1; return p; }
    double x = (double)mant;
    x = exp10 < 0 ? x / objPow10(-exp10) : x * objPow10(exp10);
    *v = neg ? -x : x;
    return s;
}

// up to n numbers; stops at the first field that does not parse (sscanf("%lf %lf %lf") semantics); the rest of out[] is left alone
OBJ_HD int objScanDoubles(const char* t, uint32_t p, uint32_t end, double* out, int n, int* unsupported) {
    int got = 0;
    while (got < n) {
        double v = 0.0;
        const uint32_t e = objStrtod(t, p, end, &v, unsupported);
        if (e == p) break;
        out[got++] = v; p = e;
    }
    return got;
}

// atoi: optional sign, digits, stops at the first other character. More than 9 digits (atoi's overflow is undefined) -> unsupported.
OBJ_HD int objAtoi(const char* t, uint32_t p, uint32_t end, int* unsupported) {
    int c = objChr(t, p, end);
    while (objSpace(c)) c = objChr(t, ++p, end);
    bool neg = false;
    if (c == '+' || c == '-') { neg = c == '-'; c = objChr(t, ++p, end); }
    int v = 0, digits = 0;
    for (; objDigit(c); c = objChr(t, ++p, end)) {          // digits = significant digits (leading zeros do not count)
        if (digits > 0 || c != '0') ++digits;
        if (digits <= 9) v = v * 10 + (c - '0');
    }
    if (digits > 9) *unsupported = 1;
    return neg ? -v : v;
}

// The corners of an `f` line as objSweepB collects them: three independent lists (position / uv / normal references, 0-based), a
// reference of 0 or a missing one adds nothing to its list.
struct ObjFace { int nv, nt, nn; int vi[kObjMaxCorners], ti[kObjMaxCorners], ni[kObjMaxCorners]; };

OBJ_HD void objParseFace(const char* t, uint32_t p, uint32_t end, ObjFace& f, int* unsupported) {
    f.nv = f.nt = f.nn = 0;
    uint32_t s = p;
    while (true) {
        while (objSpace(objChr(t, s, end))) ++s;
        if (!objChr(t, s, end)) break;
        uint32_t tokEnd = s;
        while (objChr(t, tokEnd, end) && !objSpace(objChr(t, tokEnd, end))) ++tokEnd;
        int a = objAtoi(t, s, tokEnd, unsupported), b = 0, c = 0;
        uint32_t s1 = s;
        while (s1 < tokEnd && t[s1] != '/') ++s1;
        if (s1 < tokEnd) {
            if (s1 + 1 != tokEnd && t[s1 + 1] != '/') b = objAtoi(t, s1 + 1, tokEnd, unsupported);
            uint32_t s2 = s1 + 1;
            while (s2 < tokEnd && t[s2] != '/') ++s2;
            if (s2 < tokEnd && s2 + 1 != tokEnd) c = objAtoi(t, s2 + 1, tokEnd, unsupported);
        }
        if (a != 0) { if (f.nv < kObjMaxCorners) f.vi[f.nv] = a - 1; ++f.nv; }
        if (b != 0) { if (f.nt < kObjMaxCorners) f.ti[f.nt] = b - 1; ++f.nt; }
        if (c != 0) { if (f.nn < kObjMaxCorners) f.ni[f.nn] = c - 1; ++f.nn; }
        s = tokEnd;
    }
    if (f.nv > kObjMaxCorners || f.nt > kObjMaxCorners || f.nn > kObjMaxCorners) { *unsupported = 1; f.nv = f.nt = f.nn = 0; }
}

struct ObjVec4 { float x, y, z, w; };

// triangle k (1 <= k < nv - 1) of the face's fan: false when the loader drops it (a corner that names no position read so far, or an
// area below 1e-9). cp = the unnormalised face normal, areaSq its squared length.
OBJ_HD bool objFanTriangle(const ObjFace& f, int k, int startIndex, int posSeen, const ObjVec4* allPos, int idx[3], float cp[3], float* areaSq) {
    const int i0 = f.vi[0] + startIndex, i1 = f.vi[k] + startIndex, i2 = f.vi[k + 1] + startIndex;
    if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= posSeen || i1 >= posSeen || i2 >= posSeen) return false;
    const ObjVec4 p0 = allPos[i0], p1 = allPos[i1], p2 = allPos[i2];
    const float e1x = OBJ_SUB(p1.x, p0.x), e1y = OBJ_SUB(p1.y, p0.y), e1z = OBJ_SUB(p1.z, p0.z);
    const float e2x = OBJ_SUB(p2.x, p0.x), e2y = OBJ_SUB(p2.y, p0.y), e2z = OBJ_SUB(p2.z, p0.z);
    cp[0] = OBJ_SUB(OBJ_MUL(e1y, e2z), OBJ_MUL(e1z, e2y));
    cp[1] = OBJ_SUB(OBJ_MUL(e1z, e2x), OBJ_MUL(e1x, e2z));
    cp[2] = OBJ_SUB(OBJ_MUL(e1x, e2y), OBJ_MUL(e1y, e2x));
    *areaSq = OBJ_ADD(OBJ_ADD(OBJ_MUL(cp[0], cp[0]), OBJ_MUL(cp[1], cp[1])), OBJ_MUL(cp[2], cp[2]));
    if (*areaSq < 1e-18f) return false;
    idx[0] = i0; idx[1] = i1; idx[2] = i2;
    return true;
}

// a `vn` record: fewer than three numbers or a length below 1e-6 gives (0, 1, 0)
OBJ_HD ObjVec4 objNormalRecord(const char* t, uint32_t begin, uint32_t end, int* unsupported) {
    double d[3] = {0.0, 0.0, 0.0};
    const int got = objScanDoubles(t, begin + 3, end, d, 3, unsupported);
    ObjVec4 n = {(float)d[0], (float)d[1], (float)d[2], 0.f};
    if (got != 3 || OBJ_ADD(OBJ_ADD(OBJ_MUL(n.x, n.x), OBJ_MUL(n.y, n.y)), OBJ_MUL(n.z, n.z)) < 1e-12f) n = {0.f, 1.f, 0.f, 0.f};
    return n;
}

}  // namespace nv
