// frontend.cpp — host front end kept from the reference so the core is a drop-in:
//   .rendertron config parser   (objects.cuh:844-943, accepts both spellings of the radius key)
//   OBJ loader                  (main.cu:936-1068: v/vt/vn/f, fan triangulation, area cull, light list)
//   24-entry material table     (main.cu:397-467)
//   camera                      (objects.cuh:221-266, 309-325)
//   BMP texture reader / BMP + PFM writers with the ACES-fit tonemap (imageUtil.cu:69-100, 144-232)
// Differences from the reference, all deliberate: Material fields the reference leaves indeterminate are
// zero; faces without vn/vt get a flat face normal / a zero uv instead of index -1.
#include "scene.h"
#include <algorithm>
#include <cctype>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <atomic>
#include <fcntl.h>
#include <memory>
#include <sys/stat.h>
#include <unistd.h>
#include <string>
#include <thread>
#include <vector>

namespace nv {

static std::string trim(const std::string& s) {
    size_t a = s.find_first_not_of(" \t\r\n");
    if (a == std::string::npos) return s;
    size_t b = s.find_last_not_of(" \t\r\n");
    return s.substr(a, b - a + 1);
}
static bool parseBool(std::string v) {
    std::transform(v.begin(), v.end(), v.begin(), ::tolower);
    return v == "true";
}
static void parseVec3(const std::string& v, float out[3]) {
    std::stringstream ss(v);
    out[0] = out[1] = out[2] = 0.f;
    ss >> out[0] >> out[1] >> out[2];
}

struct MeshCfg { std::string path; float mult = 0.f; float color[3] = {0, 0, 0}; int materialID = 0; };

int matchIntegrator(const std::string& name) {   // objects.cuh:583-593
    if (name == "UNIDIRECTIONAL") return NOVUM_UNIDIRECTIONAL;
    if (name == "BIDIRECTIONAL" || name == "BDPT") return NOVUM_BIDIRECTIONAL;
    if (name == "NAIVE_UNIDIRECTIONAL") return NOVUM_NAIVE_UNIDIRECTIONAL;
    if (name == "VCM") return NOVUM_VCM;
    if (name == "SPPM") return NOVUM_SPPM;
    return -1;
}

int loadConfigFull(const char* path, NovumRenderConfig* cfg, std::vector<MeshCfg>* meshes) {
    std::ifstream file(path);
    if (!file.is_open()) return NOVUM_E_IO;
    memset(cfg, 0, sizeof(*cfg));
    cfg->integrator = -1;
    std::string line;
    bool inMeshes = false;
    try {
    while (std::getline(file, line)) {
        line = trim(line);
        if (line.empty()) continue;
        if (line.rfind("Meshes", 0) == 0) { inMeshes = true; continue; }
        if (inMeshes) {   // path; multiplier * (r, g, b); materialID
            MeshCfg m;
            std::stringstream ss(line);
            std::string seg;
            if (std::getline(ss, seg, ';')) m.path = trim(seg);
            if (std::getline(ss, seg, ';')) {
                std::string e = trim(seg);
                size_t star = e.find('*'), op = e.find('('), cp = e.find(')');
                if (star != std::string::npos && op != std::string::npos) {
                    m.mult = std::stof(e.substr(0, star));
                    std::string vec = e.substr(op + 1, cp - op - 1);
                    std::replace(vec.begin(), vec.end(), ',', ' ');
                    parseVec3(vec, m.color);
                }
            }
            if (std::getline(ss, seg, ';')) m.materialID = std::stoi(trim(seg));
            if (meshes) meshes->push_back(m);
            cfg->numMeshes++;
            continue;
        }
        size_t d = line.find(':');
        if (d == std::string::npos) continue;
        std::string key = trim(line.substr(0, d)), val = trim(line.substr(d + 1));
        if (val.empty()) continue;
        if (key == "width") cfg->width = std::stoi(val);
        else if (key == "height") cfg->height = std::stoi(val);
        else if (key == "Integrator") cfg->integrator = matchIntegrator(val);
        else if (key == "Name") { strncpy(cfg->name, val.c_str(), sizeof(cfg->name) - 1); }
        else if (key == "Sample Count") cfg->sampleCount = std::stoi(val);
        else if (key == "Unidirectional Max Depth") cfg->maxDepth = std::stoi(val);
        else if (key == "BVH recommended leaf size") cfg->bvhLeafSize = std::stoi(val);
        else if (key == "Bidirectional Eye Depth") cfg->bdptEyeDepth = std::stoi(val);
        else if (key == "Bidirectional Light Depth") cfg->bdptLightDepth = std::stoi(val);
        else if (key == "BDPT_LIGHTTRACE") cfg->bdptLightTrace = parseBool(val);
        else if (key == "BDPT_NEE") cfg->bdptNee = parseBool(val);
        else if (key == "BDPT_NAIVE") cfg->bdptNaive = parseBool(val);
        else if (key == "BDPT_CONNECTION") cfg->bdptConnection = parseBool(val);
        else if (key == "BDPT_DRAWPATH") cfg->bdptDrawPath = parseBool(val);
        else if (key == "BDPT_DOMIS") cfg->bdptDoMis = parseBool(val);
        else if (key == "BDPT_PAINTWEIGHT") cfg->bdptPaintWeight = parseBool(val);
        else if (key == "Pinhole Camera") cfg->pinholeCamera = parseBool(val);
        else if (key == "SAMPLE_ENVIRONMENT") cfg->sampleEnvironment = parseBool(val);
        else if (key == "Post Process") cfg->postProcess = parseBool(val);
        else if (key == "VCM_DOMERGE") cfg->vcmDoMerge = parseBool(val);
        else if (key == "Camera Position") parseVec3(val, cfg->camPos);
        else if (key == "Camera Rotation") parseVec3(val, cfg->camRot);
        else if (key == "Camera FOV") cfg->camFov = std::stof(val);
        else if (key == "Camera Apeture") cfg->camApeture = std::stof(val);
        else if (key == "Camera FocalDist") cfg->camFocalDist = std::stof(val);
        else if (key == "VCM Merge Radius Power Factor") cfg->vcmMergeConst = std::stof(val);
        else if (key == "VCM Initial Merge Radius Multiplier" || key == "VCM Initial Merge Radius Multipler")
            cfg->vcmInitialMergeRadiusMultiplier = std::stof(val);   // the shipped config misspells the key (config.rendertron:27)
    }
    } catch (...) { return NOVUM_E_IO; }
    if (cfg->integrator == NOVUM_SPPM) {   // main.cu:314-333
        cfg->bdptConnection = cfg->bdptNaive = cfg->bdptNee = cfg->bdptLightTrace = cfg->bdptDoMis = 0;
        cfg->vcmDoMerge = 1; cfg->doSPPM = 1;
    }
    return 0;
}

// ---- material table (main.cu:397-467); every field the factories leave unset is zero ---------------
static NovumMaterial matBase(int type) {
    NovumMaterial m; memset(&m, 0, sizeof(m));
    m.type = type;
    m.albedo = {0.8f, 0.8f, 0.8f, 0.f}; m.roughness = 0.5f; m.ior = 1.5f; m.metallic = 0.f; m.specular = 1.f;   // objects.cuh:640-643
    return m;
}
static NovumMaterial matDiffuse(float r, float g, float b) {   // objects.cuh:645-657
    NovumMaterial m = matBase(NOVUM_MAT_DIFFUSE);
    m.albedo = {r, g, b, 0.f}; m.roughness = 1.0f;
    return m;
}
static NovumMaterial matDiffuseTex(int start, int w, int h) {   // objects.cuh:659-675
    NovumMaterial m = matBase(NOVUM_MAT_DIFFUSE);
    m.hasTexture = (w > 0 && h > 0) ? 1 : 0; m.startInd = start; m.width = w; m.height = h; m.roughness = 1.0f;
    return m;
}
static NovumMaterial matMetal(const float eta[3], const float k[3], float rough) {   // objects.cuh:677-693
    NovumMaterial m = matBase(NOVUM_MAT_METAL);
    m.eta = {eta[0], eta[1], eta[2], 0.f}; m.k = {k[0], k[1], k[2], 0.f}; m.roughness = rough;
    m.albedo = {1.f, 1.f, 1.f, 0.f}; m.metallic = 1.0f;
    return m;
}
static NovumMaterial matDielectric(float ior, float ax, float ay, float az, int pri) {   // objects.cuh:695-710
    NovumMaterial m = matBase(NOVUM_MAT_SMOOTHDIELECTRIC);
    m.ior = ior; m.albedo = {1.f, 1.f, 1.f, 0.f}; m.priority = pri; m.isSpecular = 1; m.boundary = 1;
    m.absorption = {ax, ay, az, 0.f};
    return m;
}
static NovumMaterial matLeaf(int start, int w, int h, float ior, float rough, float trans) {   // objects.cuh:726-749
    NovumMaterial m = matBase(NOVUM_MAT_LEAF);
    m.hasTexture = (w > 0 && h > 0) ? 1 : 0; m.ior = ior; m.roughness = rough; m.albedo = {0.22f, 0.75f, 0.28f, 0.f};
    m.transmission = trans; m.startInd = start; m.width = w; m.height = h; m.thinWalled = 1;
    return m;
}

void buildMaterialTable(NovumMaterial* out, const int texStart[4], const int texW[4], const int texH[4]) {
    const float etaSteel[3] = {0.14f, 0.16f, 0.13f}, etaGold[3] = {0.17f, 0.35f, 1.5f};
    out[0] = matDielectric(1.0f, 0, 0, 0, 99);                       // air
    out[1] = matDiffuse(0.4f, 0.4f, 0.8f);
    out[2] = matDiffuse(0.9f, 0.9f, 0.9f);
    out[3] = matDiffuse(0.2f, 0.6f, 0.6f);
    out[4] = matMetal(etaGold, etaGold, 0.05f);                      // k = eta, as in main.cu:419
    out[5] = matDielectric(1.5f, 0, 0, 0, 1);                        // glass
    out[6] = matDiffuse(0.90f, 0.1f, 0.1f);
    out[7] = matMetal(etaSteel, etaSteel, 0.15f);
    out[8] = matDielectric(1.333f, 2.5f * 0.180f, 2.5f * 1.5f, 2.5f * 2.996f, 2);   // tea
    out[9] = matDielectric(1.31f, 0.2f, 0.2f, 0.2f, 0);             // ice
    out[10] = matDielectric(1.333f, 0, 0, 0, 2);                     // water
    out[11] = matDiffuseTex(texStart[0], texW[0], texH[0]);
    out[12] = matDiffuseTex(texStart[1], texW[1], texH[1]);
    out[13] = matLeaf(texStart[2], texW[2], texH[2], 1.5f, 0.10f, 0.15f);
    out[14] = matDiffuse(0.90f, 0.9f, 0.83f);
    out[15] = matDiffuse(0.4f, 0.4f, 1.00f);
    out[16] = matLeaf(texStart[3], texW[3], texH[3], 1.5f, 0.8f, 0.6f);
    out[17] = matDiffuse(0.8f, 0.8f, 0.8f);
    out[18] = matDielectric(2.42f, 0, 0, 0, 1);                      // diamond
    out[19] = matBase(NOVUM_MAT_DELTAMIRROR); out[19].isSpecular = 1;
    out[20] = matDiffuse(0.0f, 0.0f, 0.0f);
    out[21] = matDiffuse(0.95f, 0.95f, 0.95f);
    out[22] = matDiffuse(0.5f, 0.5f, 0.5f);
    out[23] = matDiffuse(0.1f, 0.9f, 0.1f);
}

// ---- camera (objects.cuh:221-266, 309-325) ------------------------------------------------------------
static void rotX(float v[3], float a) { float c = cosf(a), s = sinf(a); float y = v[1] * c - v[2] * s, z = v[1] * s + v[2] * c; v[1] = y; v[2] = z; }
static void rotY(float v[3], float a) { float c = cosf(a), s = sinf(a); float x = v[0] * c + v[2] * s, z = -v[0] * s + v[2] * c; v[0] = x; v[2] = z; }
static void rotZ(float v[3], float a) { float c = cosf(a), s = sinf(a); float x = v[0] * c - v[1] * s, y = v[0] * s + v[1] * c; v[0] = x; v[1] = y; }
static NovumFloat4 rotAll(float x, float y, float z, const NovumCamera& c) {
    float v[3] = {x, y, z};
    rotX(v, c.xRot); rotY(v, c.yRot); rotZ(v, c.zRot);
    float il = 1.0f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
    return {v[0] * il, v[1] * il, v[2] * il, 0.f};
}

void makeCamera(const NovumRenderConfig* cfg, NovumCamera* c) {
    memset(c, 0, sizeof(*c));
    c->w = cfg->width; c->h = cfg->height;
    c->cameraOrigin = {cfg->camPos[0], cfg->camPos[1], cfg->camPos[2], 0.f};
    c->fovScale = tanf((cfg->camFov * 0.5f) * (3.141592f / 180.0f));
    c->xRot = cfg->camRot[0] * (3.14159265f / 180.0f);
    c->yRot = cfg->camRot[1] * (3.14159265f / 180.0f);
    c->zRot = cfg->camRot[2] * (3.14159265f / 180.0f);
    if (cfg->pinholeCamera) { c->aperture = 0.000001f; c->focalDist = 1.0f / cfg->camFov; }   // objects.cuh:235-236
    else { c->aperture = cfg->camApeture; c->focalDist = cfg->camFocalDist; }
    c->antiAliasJitterDist = 2.0f;
    c->forward = rotAll(0.f, 0.f, -1.f, *c);
    c->right = rotAll(1.f, 0.f, 0.f, *c);
    c->up = rotAll(0.f, 1.f, 0.f, *c);
}

// ---- OBJ (main.cu:936-1068) ---------------------------------------------------------------------------
// resize() of these arrays leaves new elements unwritten (default-initialised PODs): the loader fills them from several threads, a
// zero-fill by the calling thread first would touch - page-fault - every new page serially (45 ms for the 871k-triangle dragon)
template <class T> struct DefaultInitAlloc : std::allocator<T> {
    template <class U> struct rebind { using other = DefaultInitAlloc<U>; };
    using std::allocator<T>::allocator;
    template <class U> void construct(U* p) noexcept { ::new ((void*)p) U; }
    template <class U, class... A> void construct(U* p, A&&... a) { ::new ((void*)p) U(std::forward<A>(a)...); }
};
template <class T> using RawVec = std::vector<T, DefaultInitAlloc<T>>;
struct HostMesh {
    RawVec<NovumFloat4> pos, nrm;
    RawVec<NovumFloat2> uv;
    RawVec<NovumTriangle> tris;
    int numLights = 0;
};

// up to n doubles exactly as sscanf("%lf ...") reads them (strtod is what %lf calls; stops at the first field that does not parse)
// strtod for the common OBJ spellings: [ws][sign]digits[.digits][e[sign]digits] with at most 15 significant digits and a decimal
// exponent within +-22 is an integer below 2^53 times or divided by an exactly representable power of ten, i.e. ONE correctly
// rounded double operation — the same value strtod returns (Clinger's fast path). Anything else goes to strtod itself.
static double fastStrtod(const char* p, char** end) {
    static const double kPow10[23] = {1e0,  1e1,  1e2,  1e3,  1e4,  1e5,  1e6,  1e7,  1e8,  1e9,  1e10, 1e11,
                                      1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};
    const char* s = p;
    while (isspace((unsigned char)*s)) ++s;
    bool neg = false;
    if (*s == '+' || *s == '-') { neg = *s == '-'; ++s; }
    if (!((*s >= '0' && *s <= '9') || (*s == '.' && s[1] >= '0' && s[1] <= '9'))) return strtod(p, end);   // inf, nan, hex, junk
    if (*s == '0' && (s[1] == 'x' || s[1] == 'X')) return strtod(p, end);
    unsigned long long mant = 0; int sig = 0, exp10 = 0;
    for (; *s >= '0' && *s <= '9'; ++s) {
        if (sig > 0 || *s != '0') { if (sig < 19) mant = mant * 10 + (unsigned)(*s - '0'); else ++exp10; ++sig; }
    }
    if (*s == '.') {
        ++s;
        for (; *s >= '0' && *s <= '9'; ++s) {
            if (sig > 0 || *s != '0') { if (sig < 19) { mant = mant * 10 + (unsigned)(*s - '0'); --exp10; } ++sig; }
            else --exp10;
        }
    }
    if (*s == 'e' || *s == 'E') {
        const char* t = s + 1; bool eneg = false;
        if (*t == '+' || *t == '-') { eneg = *t == '-'; ++t; }
        if (*t >= '0' && *t <= '9') {
            int e = 0;
            for (; *t >= '0' && *t <= '9'; ++t) if (e < 100000) e = e * 10 + (*t - '0');
            exp10 += eneg ? -e : e;
            s = t;
        }
    }
    if (sig > 15 || exp10 < -22 || exp10 > 22) return strtod(p, end);
    double v = (double)mant;
    v = exp10 < 0 ? v / kPow10[-exp10] : v * kPow10[exp10];
    *end = const_cast<char*>(s);
    return neg ? -v : v;
}

static int scanDoubles(const char* p, double* out, int n) {
    int got = 0;
    while (got < n) {
        char* e = nullptr;
        double v = fastStrtod(p, &e);
        if (e == p) break;
        out[got++] = v; p = e;
    }
    return got;
}

// Same records as the reference's readObjSimple (main.cu:936-1068) in the same order. The file is read in one piece, cut into chunks at
// line ends and parsed by several host threads in two sweeps (the getline + sscanf + istringstream form of the first version took
// 0.85 s for the 871k-triangle dragon, the single-threaded buffer walk 0.30 s):
//   sweep A  every chunk parses its `v` lines; positions are concatenated, so that sweep B may look up any vertex of the file;
//   sweep B  every chunk parses its vt / vn / f lines into chunk-local arrays in file order. What the sequential loader decides from
//            "how far the file has been read" is kept: a face may only use positions declared before it (chunk base + running count),
//            face normals generated for faces without `vn` references sit between the `vn` records in file order, and the single
//            all-zero uv shared by faces without `vt` references is inserted where the FIRST such face stands;
//   fix-up   prefix sums over the chunks turn chunk-local normal / triangle / light numbers into global ones.
// One chunk (small files, threads = 1) is the sequential loader; tests/test_host.py holds chunked parses of adversarial files equal to it.
struct ObjChunk {
    char* begin = nullptr; char* end = nullptr;
    std::vector<NovumFloat4> pos, nrm;
    std::vector<NovumFloat2> uv;
    std::vector<NovumTriangle> tris;
    std::vector<unsigned char> flags;       // per triangle: 1 = normal index is chunk-local (generated), 2 = uses the shared zero uv
    int posBase = 0, zeroUvAt = -1;
};

template <class F> static void forEachLine(char* cur, char* const endAll, F f) {
    while (cur && cur < endAll) {
        char* p = cur;
        char* nl = (char*)memchr(cur, '\n', (size_t)(endAll - cur));
        char* lineEnd = nl ? nl : endAll;
        cur = nl ? nl + 1 : endAll;
        const char saved = *lineEnd;            // terminate the line in place (std::string keeps a writable NUL at the end)
        *lineEnd = 0;
        if (!(p == lineEnd || p[0] == '#' || p[0] == 's')) f(p);
        *lineEnd = saved;
    }
}
static inline bool isPosLine(const char* p) { return p[0] == 'v' && (p[1] == ' ' || p[1] == '\t'); }

static void objSweepA(ObjChunk& c, float yOffset) {
    // one allocation per array, sized from the chunk's bytes (a record line has at least ~8 characters per number): growing by
    // doubling would map, touch and unmap every array several times, and with many threads the page faults serialise in the kernel
    if ((size_t)(c.end - c.begin) > 65536) c.pos.reserve((size_t)(c.end - c.begin) / 16);
    forEachLine(c.begin, c.end, [&](char* p) {
        if (!isPosLine(p)) return;
        double d[3] = {0, 0, 0};
        scanDoubles(p + 2, d, 3);
        NovumFloat4 q = {(float)d[0], (float)d[1], (float)d[2], 0.f};
        q.y = q.y + yOffset;
        c.pos.push_back(q);
    });
}

static void objSweepB(ObjChunk& c, const NovumFloat4* allPos, int startIndex, int nStart, int uvStart,
                      const float emission[3], int materialID) {
    std::vector<int> vi, ti, ni;
    int posSeen = startIndex + c.posBase;        // M.pos.size() of the sequential loader at the current line
    if ((size_t)(c.end - c.begin) > 65536) {
        const size_t lines = (size_t)(c.end - c.begin) / 8;      // upper bound on triangles / normals of the chunk ("f 1 2 3\n" = 8 bytes)
        c.tris.reserve(lines); c.flags.reserve(lines); c.nrm.reserve(lines); c.uv.reserve(lines / 2);
    }
    forEachLine(c.begin, c.end, [&](char* p) {
        if (isPosLine(p)) { ++posSeen; return; }
        if (p[0] == 'v' && p[1] == 't') {
            double d[2] = {0, 0};
            scanDoubles(p + 3, d, 2);
            c.uv.push_back({(float)d[0], 1.0f - (float)d[1]});
        } else if (p[0] == 'v' && p[1] == 'n') {
            double d[3] = {0, 0, 0};
            int got = scanDoubles(p + 3, d, 3);
            NovumFloat4 n = {(float)d[0], (float)d[1], (float)d[2], 0.f};
            if (got != 3 || std::isnan(d[0]) || std::isnan(d[1]) || std::isnan(d[2]) || n.x * n.x + n.y * n.y + n.z * n.z < 1e-12f)
                n = {0.f, 1.f, 0.f, 0.f};
            c.nrm.push_back(n);
        } else if (p[0] == 'f' && (p[1] == ' ' || p[1] == '\t')) {
            vi.clear(); ti.clear(); ni.clear();
            const char* s = p + 2;
            while (true) {
                while (*s && isspace((unsigned char)*s)) ++s;          // operator>> on a string: whitespace-separated tokens
                if (!*s) break;
                const char* tokEnd = s;
                while (*tokEnd && !isspace((unsigned char)*tokEnd)) ++tokEnd;
                int a = 0, b = 0, cc = 0;
                a = atoi(s);
                const char* s1 = (const char*)memchr(s, '/', (size_t)(tokEnd - s));
                if (s1) {
                    if (s1 + 1 != tokEnd && s1[1] != '/') b = atoi(s1 + 1);
                    const char* s2 = (const char*)memchr(s1 + 1, '/', (size_t)(tokEnd - (s1 + 1)));
                    if (s2 && s2 + 1 != tokEnd) cc = atoi(s2 + 1);
                }
                if (a != 0) vi.push_back(a - 1);
                if (b != 0) ti.push_back(b - 1);
                if (cc != 0) ni.push_back(cc - 1);
                s = tokEnd;
            }
            bool hasUV = ti.size() == vi.size(), hasN = ni.size() == vi.size();
            int n = (int)vi.size();
            for (int i = 1; i < n - 1; ++i) {
                int i0 = vi[0] + startIndex, i1 = vi[i] + startIndex, i2 = vi[i + 1] + startIndex;
                if (i0 < 0 || i1 < 0 || i2 < 0 || i0 >= posSeen || i1 >= posSeen || i2 >= posSeen) continue;
                const NovumFloat4 &p0 = allPos[i0], &p1 = allPos[i1], &p2 = allPos[i2];
                float e1[3] = {p1.x - p0.x, p1.y - p0.y, p1.z - p0.z}, e2[3] = {p2.x - p0.x, p2.y - p0.y, p2.z - p0.z};
                float cp[3] = {e1[1] * e2[2] - e1[2] * e2[1], e1[2] * e2[0] - e1[0] * e2[2], e1[0] * e2[1] - e1[1] * e2[0]};
                float areaSq = cp[0] * cp[0] + cp[1] * cp[1] + cp[2] * cp[2];
                if (areaSq < 1e-18f) continue;
                NovumTriangle t; memset(&t, 0, sizeof(t));
                unsigned char fl = 0;
                t.aInd = i0; t.bInd = i1; t.cInd = i2;
                if (hasN) { t.naInd = ni[0] + nStart; t.nbInd = ni[i] + nStart; t.ncInd = ni[i + 1] + nStart; }
                else {   // the reference would index normals[-1]; use the face normal instead
                    float il = 1.0f / sqrtf(areaSq);
                    c.nrm.push_back({cp[0] * il, cp[1] * il, cp[2] * il, 0.f});
                    t.naInd = t.nbInd = t.ncInd = (int)c.nrm.size() - 1;
                    fl |= 1;
                }
                if (hasUV) { t.uvaInd = ti[0] + uvStart; t.uvbInd = ti[i] + uvStart; t.uvcInd = ti[i + 1] + uvStart; }
                else {
                    if (c.zeroUvAt < 0) { c.uv.push_back({0.f, 0.f}); c.zeroUvAt = (int)c.uv.size() - 1; }
                    fl |= 2;
                }
                t.materialID = materialID;
                t.emission = {emission[0], emission[1], emission[2], 0.f};
                c.tris.push_back(t);
                c.flags.push_back(fl);
            }
        }
    });
}

template <class F> static void runChunks(int nChunks, int threads, F f) {
    if (threads <= 1 || nChunks <= 1) { for (int i = 0; i < nChunks; ++i) f(i); return; }
    std::atomic<int> next(0);
    auto work = [&]() { for (int i; (i = next.fetch_add(1)) < nChunks;) f(i); };
    std::vector<std::thread> pool;
    for (int t = 1; t < std::min(threads, nChunks); ++t) pool.emplace_back(work);
    work();
    for (std::thread& t : pool) t.join();
}

int loadObj(const std::string& file, HostMesh& M, const float emission[3], int materialID, float yOffset, int threads = 0, size_t chunkBytes = 0) {
    // The text is read by several threads (pread, 4 MiB pieces) into a buffer that is not zero-filled first; one byte more than the
    // file so that the last line can be NUL-terminated in place like the others. (A private copy-on-write mapping of the file instead
    // was 4x SLOWER with threads: every page is written once - the line ends - and the copy-on-write faults serialise in the kernel.)
    struct Text { char* base = nullptr; size_t size = 0; RawVec<char> owned; } text;
    {
        int fd = open(file.c_str(), O_RDONLY);
        if (fd < 0) return NOVUM_E_IO;
        struct stat sb;
        if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode)) { close(fd); return NOVUM_E_IO; }
        const size_t sz = (size_t)sb.st_size, piece = (size_t)4 << 20;
        text.owned.resize(sz + 1);
        std::atomic<bool> bad(false);
        int readers = threads > 0 ? threads : (int)std::thread::hardware_concurrency();
        if (readers > 8) readers = 8;
        runChunks((int)((sz + piece - 1) / piece), readers, [&](int i) {
            size_t off = (size_t)i * piece, end = std::min(sz, off + piece);
            while (off < end) { ssize_t r = pread(fd, &text.owned[off], end - off, (off_t)off); if (r <= 0) { bad = true; break; } off += (size_t)r; }
        });
        close(fd);
        if (bad) return NOVUM_E_IO;
        text.owned[sz] = 0;
        text.base = sz ? text.owned.data() : nullptr; text.size = sz;
    }
    const bool timing = getenv("NOVUM_OBJ_TIMING") != nullptr;
    auto tick = std::chrono::steady_clock::now();
    auto lap = [&](const char* what) {
        if (!timing) return;
        auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "loadObj %-10s %.1f ms\n", what, std::chrono::duration<double, std::milli>(now - tick).count());
        tick = now;
    };
    lap("read");
    if (threads <= 0) { threads = (int)std::thread::hardware_concurrency(); if (threads < 1) threads = 1; if (threads > 16) threads = 16; }
    // chunks of 256 KiB .. 1 MiB handed out from a queue, about eight per thread: positions sit at the start of a file and faces at its
    // end, so the two sweeps load different chunks; files below 256 KiB are one chunk, i.e. the sequential loader
    if (chunkBytes == 0) chunkBytes = std::min<size_t>(std::max<size_t>(text.size / ((size_t)threads * 8) + 1, (size_t)256 << 10), (size_t)1 << 20);
    const int startIndex = (int)M.pos.size(), nStart = (int)M.nrm.size(), uvStart = (int)M.uv.size();
    const bool isLight = emission[0] * emission[0] + emission[1] * emission[1] + emission[2] * emission[2] > 0.f;
    std::vector<ObjChunk> chunks;
    if (text.size) {
        char* const base = text.base;
        char* const endAll = base + text.size;
        char* cur = base;
        while (cur < endAll) {
            char* want = (size_t)(endAll - cur) > chunkBytes ? cur + chunkBytes : endAll;
            if (want < endAll) { char* nl = (char*)memchr(want, '\n', (size_t)(endAll - want)); want = nl ? nl + 1 : endAll; }
            ObjChunk c; c.begin = cur; c.end = want;
            chunks.push_back(std::move(c));
            cur = want;
        }
    }
    const int nc = (int)chunks.size();
    lap("split");
    runChunks(nc, threads, [&](int i) { objSweepA(chunks[i], yOffset); });
    lap("sweepA");
    size_t addPos = 0;
    for (ObjChunk& c : chunks) { c.posBase = (int)addPos; addPos += c.pos.size(); }
    M.pos.resize((size_t)startIndex + addPos);
    runChunks(nc, threads, [&](int i) { if (!chunks[i].pos.empty()) memcpy(&M.pos[(size_t)startIndex + chunks[i].posBase], chunks[i].pos.data(), chunks[i].pos.size() * sizeof(NovumFloat4)); });
    lap("gatherPos");
    runChunks(nc, threads, [&](int i) { objSweepB(chunks[i], M.pos.data(), startIndex, nStart, uvStart, emission, materialID); });
    lap("sweepB");
    // fix-up: only the first chunk (in file order) that needed the shared zero uv keeps its copy
    int firstZero = -1;
    for (int i = 0; i < nc && firstZero < 0; ++i) if (chunks[i].zeroUvAt >= 0) firstZero = i;
    std::vector<size_t> nrmBase(nc), uvBase(nc), triBase(nc);
    size_t nN = M.nrm.size(), nU = M.uv.size(), nT = M.tris.size();
    int sharedZeroUV = -1;
    for (int i = 0; i < nc; ++i) {
        nrmBase[i] = nN; uvBase[i] = nU; triBase[i] = nT;
        if (i == firstZero) sharedZeroUV = (int)(nU + (size_t)chunks[i].zeroUvAt);
        nN += chunks[i].nrm.size();
        nU += chunks[i].uv.size() - ((chunks[i].zeroUvAt >= 0 && i != firstZero) ? 1 : 0);
        nT += chunks[i].tris.size();
    }
    const int lightBase = M.numLights;
    const size_t triStart = M.tris.size();
    M.nrm.resize(nN); M.uv.resize(nU); M.tris.resize(nT);
    lap("resize");
    runChunks(nc, threads, [&](int i) {
        ObjChunk& c = chunks[i];
        if (!c.nrm.empty()) memcpy(&M.nrm[nrmBase[i]], c.nrm.data(), c.nrm.size() * sizeof(NovumFloat4));
        size_t o = uvBase[i];
        for (size_t k = 0; k < c.uv.size(); ++k) { if ((int)k == c.zeroUvAt && i != firstZero) continue; M.uv[o++] = c.uv[k]; }
        for (size_t k = 0; k < c.tris.size(); ++k) {
            NovumTriangle t = c.tris[k];
            if (c.flags[k] & 1) { t.naInd += (int)nrmBase[i]; t.nbInd = t.ncInd = t.naInd; }
            if (c.flags[k] & 2) t.uvaInd = t.uvbInd = t.uvcInd = sharedZeroUV;
            const size_t g = triBase[i] + k;
            t.triInd = (int)g;
            t.lightInd = isLight ? lightBase + (int)(g - triStart) : -51;
            M.tris[g] = t;
        }
    });
    lap("fixup");
    if (isLight) M.numLights += (int)(nT - triStart);
    return 0;
}

// ---- images -------------------------------------------------------------------------------------------
#pragma pack(push, 1)
struct BmpFileHeader { uint16_t type; uint32_t size; uint16_t r1, r2; uint32_t offBits; };
struct BmpInfoHeader { uint32_t size; int32_t w, h; uint16_t planes, bitCount; uint32_t compression, sizeImage; int32_t xppm, yppm; uint32_t clrUsed, clrImportant; };
#pragma pack(pop)

bool loadBmpTexture(const std::string& file, std::vector<NovumFloat4>& px, int& w, int& h) {   // imageUtil.cu:144-190
    w = h = 0;
    std::ifstream in(file, std::ios::binary);
    if (!in.is_open()) return false;
    BmpFileHeader fh; BmpInfoHeader ih;
    in.read((char*)&fh, sizeof(fh)); in.read((char*)&ih, sizeof(ih));
    if (fh.type != 0x4D42 || ih.bitCount != 24) return false;
    w = ih.w; h = ih.h;
    size_t base = px.size();
    px.resize(base + (size_t)w * h);
    int rowSize = (3 * w + 3) & (~3);
    std::vector<unsigned char> row(rowSize);
    for (int y = 0; y < h; ++y) {
        in.read((char*)row.data(), rowSize);
        for (int x = 0; x < w; ++x) {
            float b = powf(row[x * 3 + 0] / 255.0f, 2.2f), g = powf(row[x * 3 + 1] / 255.0f, 2.2f), r = powf(row[x * 3 + 2] / 255.0f, 2.2f);
            px[base + (size_t)(h - 1 - y) * w + x] = {r, g, b, 1.0f};
        }
    }
    return true;
}

}  // namespace nv

using namespace nv;

extern "C" {

int novum_abi_version(void) { return NOVUM_ABI_VERSION; }

const char* novum_error_string(int code) {
    switch (code) {
        case NOVUM_OK: return "ok";
        case NOVUM_E_ARG: return "invalid argument";
        case NOVUM_E_IO: return "file could not be read or written";
        case NOVUM_E_NOSCENE: return "null scene handle";
        case NOVUM_E_NOGPU: return "no CUDA device";
        case NOVUM_E_UNSUPPORTED: return "unsupported";
        case NOVUM_E_TIMEOUT: return "timed out waiting for a peer rank";
        case NOVUM_E_TOODEEP: return "BVH deeper than the traversal stack";
        case NOVUM_E_TOOBIG: return "size beyond the entry point's built-in limit";
        default: return code > 0 ? cudaGetErrorString((cudaError_t)code) : "unknown error";
    }
}

int novum_device_count(void) { int n = 0; if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; } return n; }
int novum_set_device(int device) { return (int)cudaSetDevice(device); }
int novum_sync(void) { return (int)cudaDeviceSynchronize(); }

int novum_config_load(const char* path, NovumRenderConfig* out) {
    if (!path || !out) return NOVUM_E_ARG;
    return loadConfigFull(path, out, nullptr);
}

int novum_material_table(NovumMaterial* out24) {
    if (!out24) return NOVUM_E_ARG;
    int z[4] = {0, 0, 0, 0};
    buildMaterialTable(out24, z, z, z);
    return 0;
}

int novum_camera_make(const NovumRenderConfig* cfg, NovumCamera* out) {
    if (!cfg || !out) return NOVUM_E_ARG;
    makeCamera(cfg, out);
    return 0;
}

int novum_scene_load_config(const char* configPath, int renderNumber, NovumScene** out, NovumRenderConfig* cfgOut) {
    if (!configPath || !out) return NOVUM_E_ARG;
    NovumRenderConfig cfg; std::vector<MeshCfg> meshes;
    int rc = loadConfigFull(configPath, &cfg, &meshes);
    if (rc) return rc;
    if (cfgOut) *cfgOut = cfg;
    // textures (main.cu:371-391): missing files give empty images, like the reference
    std::vector<NovumFloat4> texels; int ts[4], tw[4], th[4];
    const char* texNames[4] = {"textures/enkidutexture.bmp", "textures/enkiduchibitexture.bmp", "textures/leaftex2.bmp", "textures/leafautumn.bmp"};
    // paths in the config are relative to the working directory in the reference; also try the config's directory
    std::string cfgDir = configPath;
    size_t slash = cfgDir.find_last_of('/');
    cfgDir = slash == std::string::npos ? std::string(".") : cfgDir.substr(0, slash);
    for (int i = 0; i < 4; ++i) {
        ts[i] = (int)texels.size();
        if (!loadBmpTexture(texNames[i], texels, tw[i], th[i]) && !loadBmpTexture(cfgDir + "/" + texNames[i], texels, tw[i], th[i]))
            loadBmpTexture(cfgDir + "/../" + texNames[i], texels, tw[i], th[i]);
    }
    NovumMaterial mats[24];
    buildMaterialTable(mats, ts, tw, th);
    const auto parseT0 = std::chrono::steady_clock::now();
    // The OBJ files are parsed on the device (obj_device.cu) unless NOVUM_OBJ_DEVICE=0; a file with spellings its rules do not read
    // themselves (NOVUM_E_UNSUPPORTED) sends the whole config through the host loader below, which gives the same arrays.
    const char* objDev = getenv("NOVUM_OBJ_DEVICE");
    NovumObjDeviceMesh* dm = nullptr;
    if (!(objDev && objDev[0] == '0') && novum_obj_device_create(&dm) == 0) {
        bool ok = true;
        for (const MeshCfg& m : meshes) {
            float e[3] = {m.mult * m.color[0], m.mult * m.color[1], m.mult * m.color[2]};
            bool emissive = m.color[0] * m.color[0] + m.color[1] * m.color[1] + m.color[2] * m.color[2] > 0.f;
            float yoff = emissive ? -0.01f * renderNumber : 0.f;   // main.cu:476-478
            if (m.materialID < 0 || m.materialID >= 24) { novum_obj_device_destroy(dm); return NOVUM_E_ARG; }
            int r = novum_obj_device_append_file(dm, m.path.c_str(), e, m.materialID, yoff);
            if (r == NOVUM_E_IO) r = novum_obj_device_append_file(dm, (cfgDir + "/" + m.path).c_str(), e, m.materialID, yoff);
            if (r == NOVUM_E_IO) r = novum_obj_device_append_file(dm, (cfgDir + "/../" + m.path).c_str(), e, m.materialID, yoff);
            if (r == NOVUM_E_IO) fprintf(stderr, "Error: Could not open OBJ file %s\n", m.path.c_str());   // main.cu:941-944: keeps going
            else if (r) { ok = false; break; }
        }
        int nTri = 0;
        novum_obj_device_counts(dm, nullptr, nullptr, nullptr, &nTri, nullptr);
        if (ok && nTri > 0) {
            const float parseMs = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - parseT0).count();
            if (getenv("NOVUM_OBJ_TIMING")) {
                float up = 0.f, k = 0.f; int l = 0; long long b = 0;
                novum_obj_device_timings(dm, &up, &k, &l, &b);
                fprintf(stderr, "device OBJ front end: %.1f ms for %lld bytes (read + copy %.1f ms, kernels %.2f ms in %d launches; the rest is allocation)\n",
                        parseMs, b, up, k, l);
            }
            rc = novum_scene_from_obj_device(dm, mats, 24, texels.data(), (int)texels.size(), out);
            novum_obj_device_destroy(dm);
            if (rc == 0) for (const MeshCfg& m : meshes) (*out)->meshPaths.push_back(m.path);
            if (rc == 0) (*out)->parseMs = parseMs;
            return rc;
        }
        novum_obj_device_destroy(dm);
    }
    HostMesh M;
    for (const MeshCfg& m : meshes) {
        float e[3] = {m.mult * m.color[0], m.mult * m.color[1], m.mult * m.color[2]};
        bool emissive = m.color[0] * m.color[0] + m.color[1] * m.color[1] + m.color[2] * m.color[2] > 0.f;
        float yoff = emissive ? -0.01f * renderNumber : 0.f;   // main.cu:476-478
        if (m.materialID < 0 || m.materialID >= 24) return NOVUM_E_ARG;
        int r = loadObj(m.path, M, e, m.materialID, yoff);
        if (r) r = loadObj(cfgDir + "/" + m.path, M, e, m.materialID, yoff);
        if (r) r = loadObj(cfgDir + "/../" + m.path, M, e, m.materialID, yoff);
        if (r) fprintf(stderr, "Error: Could not open OBJ file %s\n", m.path.c_str());   // main.cu:941-944: keeps going
    }
    if (M.tris.empty()) return NOVUM_E_IO;   // main.cu:505-508
    const float parseMs = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - parseT0).count();
    rc = novum_scene_create(M.pos.data(), (int)M.pos.size(), M.nrm.data(), (int)M.nrm.size(), M.uv.data(), (int)M.uv.size(),
                            M.tris.data(), (int)M.tris.size(), mats, 24, texels.data(), (int)texels.size(), out);
    if (rc == 0) for (const MeshCfg& m : meshes) (*out)->meshPaths.push_back(m.path);
    if (rc == 0) (*out)->parseMs = parseMs;
    return rc;
}

double novum_parse_double(const char* text, int* consumed) {
    if (!text) { if (consumed) *consumed = 0; return 0.0; }
    char* e = nullptr;
    double v = fastStrtod(text, &e);
    if (consumed) *consumed = (int)(e - text);
    return v;
}

int novum_save_pfm(const char* path, const NovumFloat4* colors, int w, int h, float scale) {
    if (!path || !colors || w <= 0 || h <= 0) return NOVUM_E_ARG;
    FILE* f = fopen(path, "wb");
    if (!f) return NOVUM_E_IO;
    fprintf(f, "PF\n%d %d\n-1.0\n", w, h);
    std::vector<float> row((size_t)w * 3);
    for (int y = h - 1; y >= 0; --y) {
        for (int x = 0; x < w; ++x) {
            const NovumFloat4& c = colors[(size_t)y * w + x];
            row[x * 3 + 0] = c.x * scale; row[x * 3 + 1] = c.y * scale; row[x * 3 + 2] = c.z * scale;
        }
        fwrite(row.data(), sizeof(float), row.size(), f);
    }
    fclose(f);
    return 0;
}

// ---- host-only view of the OBJ front end (inspection / tests: no GPU, no BVH) ------------------------------------------------------------
struct NovumObjMesh { HostMesh M; };

int novum_obj_parse(const char* path, const float emission[3], int materialID, float yOffset, int threads, int chunkBytes, NovumObjMesh** out) {
    if (!path || !out || threads < 0 || chunkBytes < 0) return NOVUM_E_ARG;
    const float zero[3] = {0.f, 0.f, 0.f};
    NovumObjMesh* m = new NovumObjMesh();
    int rc = loadObj(path, m->M, emission ? emission : zero, materialID, yOffset, threads, (size_t)chunkBytes);
    if (rc) { delete m; return rc; }
    *out = m;
    return 0;
}
void novum_obj_free(NovumObjMesh* m) { delete m; }
int novum_obj_counts(const NovumObjMesh* m, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights) {
    if (!m) return NOVUM_E_ARG;
    if (nPos) *nPos = (int)m->M.pos.size();
    if (nNorm) *nNorm = (int)m->M.nrm.size();
    if (nUV) *nUV = (int)m->M.uv.size();
    if (nTri) *nTri = (int)m->M.tris.size();
    if (nLights) *nLights = m->M.numLights;
    return 0;
}
int novum_obj_download(const NovumObjMesh* m, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs, NovumTriangle* tris) {
    if (!m) return NOVUM_E_ARG;
    if (positions && !m->M.pos.empty()) memcpy(positions, m->M.pos.data(), m->M.pos.size() * sizeof(NovumFloat4));
    if (normals && !m->M.nrm.empty()) memcpy(normals, m->M.nrm.data(), m->M.nrm.size() * sizeof(NovumFloat4));
    if (uvs && !m->M.uv.empty()) memcpy(uvs, m->M.uv.data(), m->M.uv.size() * sizeof(NovumFloat2));
    if (tris && !m->M.tris.empty()) memcpy(tris, m->M.tris.data(), m->M.tris.size() * sizeof(NovumTriangle));
    return 0;
}

// ---- resumable accumulators -------------------------------------------------------------------------------------
// The reference keeps its running sums only in device memory: a stopped render starts again from sample 0. The sums of this library
// are a pure function of (scene, config, seed, sample index) - counter-based RNG - so a render resumes by loading the file and calling
// novum_render with firstSample = the stored count. Layout: 64-byte little-endian header, then w*h float4 sums, row-major.
namespace {
#pragma pack(push, 1)
struct AccumHeader {
    char magic[8];               // "NOVUMACC"
    uint32_t version;            // 1
    int32_t w, h, sampleCount, integrator;
    uint32_t reserved0;
    uint64_t seed, payloadBytes, checksum;     // FNV-1a 64 over the payload
    uint64_t reserved1;
};
#pragma pack(pop)
static_assert(sizeof(AccumHeader) == 64, "accumulator header is 64 bytes");
uint64_t fnv1a64(const void* data, size_t n) {
    const uint64_t* p8 = (const uint64_t*)data;                 // word-wise variant: 8 bytes per multiply (the payload is 16 B per pixel)
    uint64_t hsh = 1469598103934665603ull;
    for (size_t i = 0; i < n / 8; ++i) { hsh ^= p8[i]; hsh *= 1099511628211ull; }
    return hsh;
}
int readAccumHeader(FILE* f, AccumHeader& hd) {
    if (fread(&hd, sizeof(hd), 1, f) != 1) return NOVUM_E_IO;
    if (memcmp(hd.magic, "NOVUMACC", 8) != 0 || (hd.version != 1 && hd.version != 2) || hd.w <= 0 || hd.h <= 0 || hd.sampleCount < 0) return NOVUM_E_IO;
    // version 2: the last header word is a checksum of the 56 bytes before it (a flipped sampleCount or seed would resume at the wrong pass)
    if (hd.version == 2 && hd.reserved1 != fnv1a64(&hd, 56)) return NOVUM_E_IO;
    if (hd.payloadBytes != (uint64_t)hd.w * (uint64_t)hd.h * sizeof(NovumFloat4)) return NOVUM_E_IO;
    return 0;
}
}  // namespace

int novum_accum_save(const char* path, const NovumFloat4* colors, int w, int h, int sampleCount, uint64_t seed, int integrator) {
    if (!path || !colors || w <= 0 || h <= 0 || sampleCount < 0) return NOVUM_E_ARG;
    AccumHeader hd;
    memset(&hd, 0, sizeof(hd));
    memcpy(hd.magic, "NOVUMACC", 8);
    hd.version = 2; hd.w = w; hd.h = h; hd.sampleCount = sampleCount; hd.integrator = integrator; hd.seed = seed;
    hd.payloadBytes = (uint64_t)w * (uint64_t)h * sizeof(NovumFloat4);
    hd.checksum = fnv1a64(colors, (size_t)hd.payloadBytes);
    hd.reserved1 = fnv1a64(&hd, 56);
    // written under a temporary name and renamed: a crash during the write never leaves a truncated checkpoint under `path`
    std::string tmp = std::string(path) + ".part";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return NOVUM_E_IO;
    bool ok = fwrite(&hd, sizeof(hd), 1, f) == 1 && fwrite(colors, 1, (size_t)hd.payloadBytes, f) == (size_t)hd.payloadBytes;
    ok = ok && fflush(f) == 0 && fsync(fileno(f)) == 0;          // on disk before the rename makes it the checkpoint
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return NOVUM_E_IO; }
    return 0;
}

int novum_accum_info(const char* path, int* w, int* h, int* sampleCount, uint64_t* seed, int* integrator) {
    if (!path) return NOVUM_E_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return NOVUM_E_IO;
    AccumHeader hd;
    int rc = readAccumHeader(f, hd);
    fclose(f);
    if (rc) return rc;
    if (w) *w = hd.w;
    if (h) *h = hd.h;
    if (sampleCount) *sampleCount = hd.sampleCount;
    if (seed) *seed = hd.seed;
    if (integrator) *integrator = hd.integrator;
    return 0;
}

int novum_accum_load(const char* path, NovumFloat4* colors, int w, int h, int* sampleCount, uint64_t* seed) {
    if (!path || !colors || w <= 0 || h <= 0) return NOVUM_E_ARG;
    FILE* f = fopen(path, "rb");
    if (!f) return NOVUM_E_IO;
    AccumHeader hd;
    int rc = readAccumHeader(f, hd);
    if (rc == 0 && (hd.w != w || hd.h != h)) rc = NOVUM_E_ARG;              // a checkpoint of another frame size
    if (rc == 0 && fread(colors, 1, (size_t)hd.payloadBytes, f) != (size_t)hd.payloadBytes) rc = NOVUM_E_IO;
    fclose(f);
    if (rc == 0 && fnv1a64(colors, (size_t)hd.payloadBytes) != hd.checksum) rc = NOVUM_E_IO;
    if (rc) return rc;
    if (sampleCount) *sampleCount = hd.sampleCount;
    if (seed) *seed = hd.seed;
    return 0;
}

}  // extern "C"
