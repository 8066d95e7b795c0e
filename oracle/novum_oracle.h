/*
 * novum_oracle.h — CPU restatement of the reference's hot path (TEST INFRASTRUCTURE ONLY).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
 * this library. The product (libnovum_b200.so) never links or calls it.
 *
 * Parity status: the reference (DanielQ-51/cudapathtracer) ships no tests, golden vectors or fixtures
 * (SURVEY.md §4), and it is GPU-only, so it cannot run in the build container. The oracle is pinned by
 * fixtures produced by the reference's own device code run on a B200 through oracle/ref/ (see
 * tests/golden/README.md): intersection results of BVHSceneIntersect / BVHShadowRay on fixed ray batches.
 * Until those fixtures are committed the header of tests/golden/README.md says "parity unpinned".
 */
#ifndef NOVUM_ORACLE_H
#define NOVUM_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, z, w; } OF4;
typedef struct { float x, y; } OF2;

/* same 80 B layout as the reference's Triangle (objects.cuh:159-183) */
typedef struct __attribute__((aligned(16))) {
    int32_t aInd, bInd, cInd, naInd, nbInd, ncInd, uvaInd, uvbInd, uvcInd, materialID, pad0[2];
    OF4 emission;
    int32_t lightInd, triInd, pad1[2];
} OTriangle;

/* same 48 B layout as the reference's BVHnode (objects.cuh:12-25) */
typedef struct __attribute__((aligned(16))) { OF4 aabbMIN, aabbMAX; int32_t left, right, first, primCount; } OBVHnode;

/* same 176 B layout as the reference's Material (objects.cuh:605-643) */
typedef struct __attribute__((aligned(16))) {
    uint8_t hasTexture, p0[3]; int32_t startInd, width, height;
    uint8_t hasTransMap, p1[3]; int32_t tstartInd, twidth, theight;
    int32_t type, p2[3];
    OF4 albedo; float roughness, p3[3];
    OF4 eta, k; float ior, metallic, specular, transmission;
    uint8_t isSpecular, boundary, thinWalled, p4; int32_t p5[3];
    OF4 absorption; int32_t priority, p6[3];
} OMaterial;

typedef struct { OF4 origin, direction; } ORay;
typedef struct { int32_t tri; float t, u, v; } OHit;

typedef struct {   /* objects.cuh:199-219 */
    OF4 cameraOrigin; int32_t w, h; float xRot, yRot, zRot, aperture, focalDist, fovScale, antiAliasJitterDist;
    int32_t pad[3]; OF4 forward, right, up;
} OCamera;

typedef struct OScene OScene;

/* integratorUtilities.cuh:8-42 */
int  oracle_triangle_intersect(const OF4* positions, const OTriangle* tri, const ORay* r, float* u, float* v, float* t);
/* integratorUtilities.cuh:44-82 */
int  oracle_aabb_intersect(const ORay* r, OF4 mn, OF4 mx, float* tmin, float* tmax);

/* scene = geometry + the reference's CPU BVH2 (main.cu:20-233) */
OScene* oracle_scene_create(const OF4* positions, int nPos, const OF4* normals, int nNorm, const OF2* uvs, int nUV,
                            const OTriangle* tris, int nTri, const OMaterial* mats, int nMat, const OF4* textures,
                            int nTexels, int maxLeafSize);
void oracle_scene_destroy(OScene* s);
int  oracle_scene_bvh_size(const OScene* s, int* numNodes, int* largestLeaf);
int  oracle_scene_bvh_copy(const OScene* s, OBVHnode* nodes, int32_t* indices);
void oracle_scene_bounds(const OScene* s, float mn[3], float mx[3], float* radius);

/* integratorUtilities.cuh:84-186 (closest hit) on a batch; threads = OpenMP threads (0 = all) */
void oracle_trace_closest(const OScene* s, const ORay* rays, int n, OHit* hits, int threads);
/* same with "lowest triangle id wins among bit-equal t" instead of first-found (our tie policy) */
void oracle_trace_closest_lowest_id(const OScene* s, const ORay* rays, int n, OHit* hits, int threads);
/* brute force over all triangles (integratorUtilities.cuh:290-335), lowest id among equal t */
void oracle_trace_brute(const OScene* s, const ORay* rays, int n, OHit* hits, int threads);
/* integratorUtilities.cuh:188-288 */
void oracle_trace_shadow(const OScene* s, const ORay* rays, const float* maxT, const int32_t* skip, int n, OF4* thr,
                         int threads);

/* ---- hash grid (integratorUtilities.cuh:382-412, deviceCode.cu:3067-3198, util.cuh:393-418) ---- */
int      oracle_next_prime(int n);
float    oracle_merge_radius(float initialRadius, float alpha, int currSample);
uint32_t oracle_grid_hash(float px, float py, float pz, const float sceneMin[3], float mergeRadius, int tableSize);
void     oracle_grid_index(float px, float py, float pz, const float sceneMin[3], float cell, int32_t out[3]);
uint32_t oracle_hash_grid_index(int32_t gx, int32_t gy, int32_t gz, int tableSize);
/* stable sort by hash + cell tables, as buildHashGrid produces them */
void oracle_hashgrid_build(const float* px, const float* py, const float* pz, int n, const float sceneMin[3],
                           float mergeRadius, int tableSize, uint32_t* sortedIndex, uint32_t* sortedHash,
                           uint32_t* cellStart, uint32_t* cellEnd);
/* 27-cell count query exactly as doEyePass walks it (deviceCode.cu:2968-2995), duplicates included */
void oracle_hashgrid_query_count(const float* spx, const float* spy, const float* spz, const uint32_t* cellStart,
                                 const uint32_t* cellEnd, int tableSize, const float sceneMin[3], float mergeRadius,
                                 const float* qx, const float* qy, const float* qz, int nq, uint32_t* counts);

/* ---- codecs (util.cuh:342-391) ---- */
uint32_t oracle_pack_oct(float x, float y, float z);
void     oracle_unpack_oct(uint32_t p, float out[3]);

/* ---- BSDF (reflectors.cuh:547-666) with an explicit stream of uniforms in (0,1] ---- */
void  oracle_bsdf_eval(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], const float wo[3],
                       float etaI, float u, float v, float f[3]);
float oracle_bsdf_pdf(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], const float wo[3],
                      float etaI, float u, float v);
/* draws from rng[]: returns how many uniforms were consumed */
int   oracle_bsdf_sample(const OMaterial* mats, int id, const OF4* tex, const float wiToward[3], float etaI,
                         int backface, int mode, float u, float v, const float* rng, float wo[3], float f[3], float* pdf);

/* ---- integrators on the CPU (deviceCode.cu:285-542 Li_unidirectional, :158-205 naive) ---- */
/* adds numSample passes (indices firstSample..) into colors (w*h OF4 sums); rng keyed like the CUDA core */
void oracle_render_unidirectional(const OScene* s, const OCamera* cam, int maxDepth, int useMIS, int naive,
                                  int firstSample, int numSample, uint64_t seed, OF4* colors, int threads,
                                  uint64_t* rayCounts /* [0] closest, [1] shadow, may be NULL */);

#ifdef __cplusplus
}
#endif
#endif
