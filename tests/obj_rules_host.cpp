// Test infrastructure (not shipped, never linked into libnovum_b200.so): the per-line rules of the DEVICE OBJ parser
// (cudapathtracer_b200/csrc/obj_parse.cuh) compiled as plain C++ and driven line by line in file order, with the same three passes and
// the same offset arithmetic as the kernels of obj_device.cu (classify + positions, per-line counts, emit). tests/test_host.py compares
// its arrays with the host loader's (novum_obj_parse) on adversarial files, so that the rules are pinned on the CPU before the kernels
// run them on a GPU. Build: g++ -O2 -shared -fPIC -I../include -I../cudapathtracer_b200/csrc obj_rules_host.cpp -o _build/libobjrules.so
#include "novum_b200.h"
#include "obj_parse.cuh"
#include <cmath>
#include <cstring>
#include <vector>

using namespace nv;

struct RulesMesh {
    std::vector<ObjVec4> pos, nrm;
    std::vector<NovumFloat2> uv;
    std::vector<NovumTriangle> tris;
    int numLights = 0;
    int unsupported = 0;
};

extern "C" {

RulesMesh* obj_rules_parse(const char* text, size_t bytes, const float emission[3], int materialID, float yOffset) {
    RulesMesh* M = new RulesMesh();
    if (bytes == 0) return M;
    // objLineStarts: a newline in the last byte starts no line; end(i) = position of the line's newline, or the end of the file
    std::vector<uint32_t> start{0u};
    for (size_t p = 0; p + 1 < bytes; ++p) if (text[p] == '\n') start.push_back((uint32_t)p + 1u);
    const size_t n = start.size();
    start.push_back(text[bytes - 1] == '\n' ? (uint32_t)bytes : (uint32_t)bytes + 1u);
    auto lineEnd = [&](size_t i) { return start[i + 1] - 1u; };
    const int startIndex = 0, nStart = 0, uvStart = 0, triStart = 0, lightBase = 0;
    const bool isLight = emission[0] * emission[0] + emission[1] * emission[1] + emission[2] * emission[2] > 0.f;
    // objClassify + objPositions
    std::vector<unsigned char> types(n);
    std::vector<int> posSeen(n);
    for (size_t i = 0; i < n; ++i) {
        types[i] = (unsigned char)objLineType(text, start[i], lineEnd(i));
        posSeen[i] = startIndex + (int)M->pos.size();
        if (types[i] == OBJ_V) {
            double d[3] = {0.0, 0.0, 0.0};
            objScanDoubles(text, start[i] + 2, lineEnd(i), d, 3, &M->unsupported);
            ObjVec4 q = {(float)d[0], (float)d[1], (float)d[2], 0.f};
            q.y = q.y + yOffset;
            M->pos.push_back(q);
        }
    }
    // objFaceCount
    std::vector<int> cN(n, 0), cU(n, 0), cT(n, 0);
    size_t firstZero = (size_t)-1;
    for (size_t i = 0; i < n; ++i) {
        if (types[i] == OBJ_VT) cU[i] = 1;
        else if (types[i] == OBJ_VN) cN[i] = 1;
        else if (types[i] == OBJ_F) {
            ObjFace f;
            objParseFace(text, start[i] + 2, lineEnd(i), f, &M->unsupported);
            const bool hasUV = f.nt == f.nv, hasN = f.nn == f.nv;
            for (int k = 1; k < f.nv - 1; ++k) {
                int idx[3]; float cp[3], areaSq;
                if (objFanTriangle(f, k, startIndex, posSeen[i], M->pos.data(), idx, cp, &areaSq)) ++cT[i];
            }
            if (!hasN) cN[i] = cT[i];
            if (!hasUV && cT[i] > 0 && firstZero == (size_t)-1) firstZero = i;
        }
    }
    // scans
    std::vector<int> aN(n), aU(n), aT(n);
    int sN = 0, sU = 0, sT = 0;
    for (size_t i = 0; i < n; ++i) { aN[i] = sN; aU[i] = sU; aT[i] = sT; sN += cN[i]; sU += cU[i]; sT += cT[i]; }
    const int zeroUvIndex = firstZero != (size_t)-1 ? aU[firstZero] : 0;
    M->nrm.resize((size_t)sN); M->uv.resize((size_t)sU + (firstZero != (size_t)-1 ? 1 : 0)); M->tris.resize((size_t)sT);
    const int sharedZeroUv = uvStart + zeroUvIndex;
    // objEmit
    for (size_t i = 0; i < n; ++i) {
        const uint32_t begin = start[i], end = lineEnd(i);
        int nrmAt = aN[i], uvAt = aU[i] + ((firstZero != (size_t)-1 && i > firstZero) ? 1 : 0), triAt = aT[i];
        if (types[i] == OBJ_VT) {
            double d[2] = {0.0, 0.0};
            objScanDoubles(text, begin + 3, end, d, 2, &M->unsupported);
            M->uv[(size_t)(uvStart + uvAt)] = {(float)d[0], 1.0f - (float)d[1]};
        } else if (types[i] == OBJ_VN) {
            M->nrm[(size_t)(nStart + nrmAt)] = objNormalRecord(text, begin, end, &M->unsupported);
        } else if (types[i] == OBJ_F) {
            if (i == firstZero) M->uv[(size_t)sharedZeroUv] = {0.f, 0.f};
            ObjFace f;
            objParseFace(text, begin + 2, end, f, &M->unsupported);
            const bool hasUV = f.nt == f.nv, hasN = f.nn == f.nv;
            for (int k = 1; k < f.nv - 1; ++k) {
                int idx[3]; float cp[3], areaSq;
                if (!objFanTriangle(f, k, startIndex, posSeen[i], M->pos.data(), idx, cp, &areaSq)) continue;
                NovumTriangle t; memset(&t, 0, sizeof(t));
                t.aInd = idx[0]; t.bInd = idx[1]; t.cInd = idx[2];
                if (hasN) { t.naInd = f.ni[0] + nStart; t.nbInd = f.ni[k] + nStart; t.ncInd = f.ni[k + 1] + nStart; }
                else {
                    const float il = 1.0f / sqrtf(areaSq);
                    M->nrm[(size_t)(nStart + nrmAt)] = {cp[0] * il, cp[1] * il, cp[2] * il, 0.f};
                    t.naInd = t.nbInd = t.ncInd = nStart + nrmAt;
                    ++nrmAt;
                }
                if (hasUV) { t.uvaInd = f.ti[0] + uvStart; t.uvbInd = f.ti[k] + uvStart; t.uvcInd = f.ti[k + 1] + uvStart; }
                else t.uvaInd = t.uvbInd = t.uvcInd = sharedZeroUv;
                t.materialID = materialID;
                t.emission = {emission[0], emission[1], emission[2], 0.f};
                t.lightInd = isLight ? lightBase + triAt : -51;
                t.triInd = triStart + triAt;
                M->tris[(size_t)(triStart + triAt)] = t;
                ++triAt;
            }
        }
    }
    if (isLight) M->numLights = sT;
    return M;
}

void obj_rules_free(RulesMesh* m) { delete m; }
int obj_rules_unsupported(const RulesMesh* m) { return m->unsupported; }
void obj_rules_counts(const RulesMesh* m, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights) {
    *nPos = (int)m->pos.size(); *nNorm = (int)m->nrm.size(); *nUV = (int)m->uv.size(); *nTri = (int)m->tris.size(); *nLights = m->numLights;
}
void obj_rules_download(const RulesMesh* m, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs, NovumTriangle* tris) {
    if (!m->pos.empty()) memcpy(positions, m->pos.data(), m->pos.size() * sizeof(ObjVec4));
    if (!m->nrm.empty()) memcpy(normals, m->nrm.data(), m->nrm.size() * sizeof(ObjVec4));
    if (!m->uv.empty()) memcpy(uvs, m->uv.data(), m->uv.size() * sizeof(NovumFloat2));
    if (!m->tris.empty()) memcpy(tris, m->tris.data(), m->tris.size() * sizeof(NovumTriangle));
}

}  // extern "C"
