/*
 * novum_b200.h — C ABI of the B200-native rendering core that replaces the hot path of
 * DanielQ-51/cudapathtracer ("Novum") behind its own launch boundary.
 *
 * Every entry point is extern "C", takes plain pointers and sizes, and returns an int
 * (0 = ok, >0 = cudaError_t, <0 = NOVUM_E_*).  No C++ / torch types cross this boundary.
 *
 * Reference interfaces replaced (citations into the reference tree):
 *   deviceCode.cuh:6        updateConstants(RenderConfig&)            -> novum_update_constants
 *   deviceCode.cuh:8-9      launch_unidirectional(...)                -> novum_launch_unidirectional
 *   deviceCode.cuh:11-12    launch_naive_unidirectional(...)          -> novum_launch_naive_unidirectional
 *   deviceCode.cuh:14-15    launch_bidirectional(...)                 -> novum_launch_bidirectional
 *   deviceCode.cuh:17-18    launch_VCM(...)                           -> novum_launch_vcm
 *   integratorUtilities.cuh:84   BVHSceneIntersect (per ray)          -> novum_trace_closest (ray batch)
 *   integratorUtilities.cuh:188  BVHShadowRay (per ray)               -> novum_trace_shadow  (ray batch)
 *   deviceCode.cu:3140      buildHashGrid(...)                        -> novum_hashgrid_build
 *   integratorUtilities.cuh:337  cleanAndFormatImage, main.cu:855-884,
 *   imageUtil.cu:69-100, 202-232  Image::saveImageBMP / postProcessImage -> novum_frame_finish[_device] / novum_frame_save_bmp
 *   objects.cuh:844         loadConfig, main.cu:936 readObjSimple,
 *   main.cu:397-467         material table, main.cu:235 initRender    -> novum_scene_load_config / novum_render
 *
 * The POD structs below are layout-compatible (sizeof and every offset, checked by
 * tests/test_abi.py against numbers printed from the reference headers) with the reference's
 * Triangle (80 B), Material (176 B), Camera (112 B), Vertices (32 B), BVHnode (48 B) so that the
 * reference's main.cu can hand its device buffers to these entry points unchanged.
 */
#ifndef NOVUM_B200_H
#define NOVUM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NOVUM_ABI_VERSION 6

/* error codes (negative; positive return values are cudaError_t) */
#define NOVUM_OK             0
#define NOVUM_E_ARG        (-1)
#define NOVUM_E_IO         (-2)
#define NOVUM_E_NOSCENE    (-3)
#define NOVUM_E_NOGPU      (-4)
#define NOVUM_E_UNSUPPORTED (-5)
#define NOVUM_E_TIMEOUT    (-8)   /* a peer rank never published its light pass (tile renderer) */
#define NOVUM_E_TOODEEP    (-7)   /* the BVH8 built from the scene is deeper than the traversal stack (52 levels): degenerate geometry */
#define NOVUM_E_TOOBIG     (-6)   /* size beyond a fixed built-in limit (novum_hashgrid_build_static) */

typedef struct NovumFloat4 { float x, y, z, w; } NovumFloat4;   /* 16 B, 16-aligned in device arrays */
typedef struct NovumFloat2 { float x, y; } NovumFloat2;

/* objects.cuh:159-183 Triangle, 80 B */
typedef struct __attribute__((aligned(16))) NovumTriangle {
    int32_t aInd, bInd, cInd;          /* position indices            off 0  */
    int32_t naInd, nbInd, ncInd;       /* normal indices              off 12 */
    int32_t uvaInd, uvbInd, uvcInd;    /* uv indices                  off 24 */
    int32_t materialID;                /*                             off 36 */
    int32_t _pad0[2];
    NovumFloat4 emission;              /*                             off 48 */
    int32_t lightInd;                  /* index in light list, -51 if not a light   off 64 */
    int32_t triInd;                    /* index in the scene triangle list          off 68 */
    int32_t _pad1[2];
} NovumTriangle;

/* objects.cuh:605-643 Material, 176 B (bools are 1 byte) */
typedef struct __attribute__((aligned(16))) NovumMaterial {
    uint8_t hasTexture; uint8_t _p0[3];
    int32_t startInd, width, height;
    uint8_t hasTransMap; uint8_t _p1[3];
    int32_t tstartInd, twidth, theight;
    int32_t type;                      /* off 32 */
    int32_t _p2[3];
    NovumFloat4 albedo;                /* off 48 */
    float roughness;                   /* off 64 */
    float _p3[3];
    NovumFloat4 eta;                   /* off 80 */
    NovumFloat4 k;                     /* off 96 */
    float ior, metallic, specular, transmission;   /* off 112 */
    uint8_t isSpecular, boundary, thinWalled, _p4; /* off 128 */
    int32_t _p5[3];
    NovumFloat4 absorption;            /* off 144 */
    int32_t priority;                  /* off 160 */
    int32_t _p6[3];
} NovumMaterial;

enum NovumMaterialType {               /* objects.cuh:595-603 */
    NOVUM_MAT_DIFFUSE = 0, NOVUM_MAT_METAL = 1, NOVUM_MAT_SMOOTHDIELECTRIC = 2,
    NOVUM_MAT_MICROFACETDIELECTRIC = 3, NOVUM_MAT_LEAF = 4, NOVUM_MAT_FLOWER = 5,
    NOVUM_MAT_DELTAMIRROR = 6
};

/* objects.cuh:199-219 Camera, 112 B */
typedef struct __attribute__((aligned(16))) NovumCamera {
    NovumFloat4 cameraOrigin;
    int32_t w, h;
    float xRot, yRot, zRot;
    float aperture, focalDist, fovScale;
    float antiAliasJitterDist;
    int32_t _pad[3];
    NovumFloat4 forward, right, up;
} NovumCamera;

/* objects.cuh:151-157 Vertices, 32 B: struct of four device pointers */
typedef struct NovumVertices {
    NovumFloat4* positions;
    NovumFloat4* normals;
    NovumFloat4* colors;
    NovumFloat2* uvs;
} NovumVertices;

/* objects.cuh:12-25 BVHnode, 48 B — accepted at the boundary and ignored (we build our own BVH8) */
typedef struct __attribute__((aligned(16))) NovumBVHnode {
    NovumFloat4 aabbMIN, aabbMAX;
    int32_t left, right, first, primCount;
} NovumBVHnode;

/* objects.cuh:185-197 Ray, 32 B */
typedef struct NovumRay { NovumFloat4 origin, direction; } NovumRay;

/* result of a closest-hit query: the fields of the reference's Intersection (objects.cuh:550-568)
 * that identify the hit; barycentrics follow triangleIntersect's convention
 * (integratorUtilities.cuh:33: u weights vertex b, v weights vertex c). tri = -1 on a miss. */
typedef struct NovumHit { int32_t tri; float t, u, v; } NovumHit;

/* integrator ids, objects.cuh:570-576 */
enum NovumIntegrator {
    NOVUM_UNIDIRECTIONAL = 0, NOVUM_BIDIRECTIONAL = 1, NOVUM_NAIVE_UNIDIRECTIONAL = 2,
    NOVUM_VCM = 3, NOVUM_SPPM = 4
};

/* scalar part of RenderConfig (objects.cuh:801-842); mesh list stays inside the scene object */
typedef struct NovumRenderConfig {
    int32_t width, height;
    int32_t integrator;                /* NovumIntegrator, -1 if the name was not recognised */
    int32_t sampleCount, maxDepth, bvhLeafSize;
    int32_t sampleEnvironment, postProcess;
    int32_t bdptEyeDepth, bdptLightDepth;
    int32_t bdptLightTrace, bdptNee, bdptNaive, bdptConnection, bdptDrawPath, bdptDoMis, bdptPaintWeight;
    int32_t vcmDoMerge, doSPPM;
    float vcmMergeConst, vcmInitialMergeRadiusMultiplier;
    int32_t pinholeCamera;
    float camPos[3], camRot[3];
    float camFov, camApeture, camFocalDist;
    int32_t numMeshes;
    char name[64];
} NovumRenderConfig;

/* the ten strategy toggles updateConstants uploads (deviceCode.cu:9-51) */
typedef struct NovumConstants {
    int32_t sampleEnvironment;
    int32_t bdptLightTrace, bdptNee, bdptNaive, bdptConnection;
    int32_t vcmDoMerge, doSPPM;
    int32_t bdptDrawPath, bdptDoMis, bdptPaintWeight;
} NovumConstants;

/* sharding of one render across ranks (one process per GPU): this rank renders sample passes /
 * light-path slabs / pixel tiles numbered rank, rank+world, ... ; world=1 is the reference's
 * single-GPU behaviour. */
typedef struct NovumShard { int32_t rank, world; } NovumShard;

/* counters returned by a render (device ray counters, summed over the call) */
typedef struct NovumRenderStats {
    uint64_t closestRays, shadowRays, paths, photons, lightVertices;
    uint64_t kernelLaunches;
    float    deviceMs;                 /* CUDA-event time of the sample loop on the launch stream */
    float    _pad;
    uint64_t bvhNodeVisits, triTests;  /* closest-hit traversal only; filled when profile bit 1 is set */
    float    stageMs[4];               /* profile bit 0: CUDA-event time of [0] extend (closest hit) [1] shade / light+eye step kernels
                                          (+ photon grid build) [2] shadow / connection resolves [3] other / photon merge kernels */
    uint64_t stageLaunches[4];
    uint64_t mergeQueries, mergeCandidates, mergeAccepted;   /* VCM / SPPM: eye vertices gathered, photons scanned in their 27 cells,
                                                                photons inside the radius */
} NovumRenderStats;

typedef struct NovumScene NovumScene;  /* opaque: device geometry + BVH8 + materials + lights */

/* ---- library -------------------------------------------------------------------------- */
int  novum_abi_version(void);
const char* novum_error_string(int code);
int  novum_device_count(void);
int  novum_set_device(int device);

/* ---- scene (host front end: config / OBJ / material table / camera) --------------------- */
/* objects.cuh:844 loadConfig — accepts both spellings of "...Merge Radius Multipl(i)er". */
int  novum_config_load(const char* path, NovumRenderConfig* out);
/* main.cu:235-557: config -> meshes -> material table -> device upload -> BVH8 build. */
int  novum_scene_load_config(const char* configPath, int renderNumber, NovumScene** out, NovumRenderConfig* cfgOut);
/* scene from caller arrays (host pointers) in the reference's layouts */
int  novum_scene_create(const NovumFloat4* positions, int nPos, const NovumFloat4* normals, int nNorm,
                        const NovumFloat2* uvs, int nUV, const NovumTriangle* tris, int nTri,
                        const NovumMaterial* mats, int nMat, const NovumFloat4* textures, int nTexels,
                        NovumScene** out);
/* scene from the reference's DEVICE buffers (what launch_* receives); nothing is copied to host
 * except the 32 B Vertices struct. */
int  novum_scene_from_device(const NovumVertices* d_vertices, int vertNum, const NovumTriangle* d_scene, int triNum,
                             const NovumTriangle* d_lights, int lightNum, const NovumMaterial* d_materials, int nMat,
                             const NovumFloat4* d_textures, NovumScene** out);
void novum_scene_destroy(NovumScene* s);
int  novum_scene_info(const NovumScene* s, int* nTri, int* nLights, int* nBvh8Nodes, float sceneMin[3], float sceneMax[3],
                      float* sceneRadius);
/* the step before the path: host time of the config / OBJ front end (main.cu:936-1068 equivalent; 0 unless the scene came
 * from novum_scene_load_config) and CUDA-event time + launch count of the device BVH8 build that replaces main.cu:20-233 */
int  novum_scene_timings(const NovumScene* s, float* parseMs, float* bvhBuildMs, int* bvhLaunches);
/* ---- opt-in features the reference only half-implements. All OFF by default: with them off every integrator
 * reproduces the reference's behaviour (and the goldens); they change results, so a caller switches them on deliberately.
 *   NOVUM_FEATURE_NESTED_BIDIR  the nested-dielectric priority stack (Beer-Lambert absorption through the dominant medium, false hits
 *                               on lower-priority interfaces, incident index of refraction of the dominant medium) inside the BDPT / VCM /
 *                               SPPM walks, as Li_unidirectional has it (deviceCode.cu:296-425, 484-497). The reference leaves
 *                               `etaI = etaT = 1.0f // TEMPORARY` there (deviceCode.cu:851-852, 1177-1178, 2078-2079).
 *   NOVUM_FEATURE_ENVIRONMENT   a constant-radiance environment (novum_scene_set_environment): a path that leaves the scene collects it,
 *                               in every integrator. The reference's sampleSky returns black (integratorUtilities.cuh:436-461) and its
 *                               bidirectional environment strategy is "currently unimplemented" (deviceCode.cu:1539-1542). No strategy
 *                               samples the environment from the light side, so the escape is the only strategy for those paths (MIS weight 1). */
#define NOVUM_FEATURE_NESTED_BIDIR 1
#define NOVUM_FEATURE_ENVIRONMENT  2
int  novum_scene_set_features(NovumScene* s, int features);
int  novum_scene_set_environment(NovumScene* s, const float rgb[3]);
/* depth of the BVH8 (levels of 8-wide nodes, root = 1). The traversal stack holds 52 entries, one per level at most: scene creation
 * fails with NOVUM_E_TOODEEP beyond that instead of losing subtrees silently. Negative = error code. */
int  novum_scene_bvh_depth(const NovumScene* s);
/* the OBJ front end's number reader (what sscanf("%lf") / operator>> return in main.cu:936-1068, i.e. strtod): exported so that its
 * fast path can be held to strtod by the CPU tests. *consumed = characters read (0: no number). */
double novum_parse_double(const char* text, int* consumed);
/* Host-only view of the OBJ front end (main.cu:936-1068 restated; no GPU, no BVH): parses ONE file into the arrays a scene load would
 * append. threads 0 = one per host core (at most 16), chunkBytes 0 = 256 KiB .. 1 MiB, about eight chunks per thread; any (threads, chunkBytes)
 * gives the arrays of the sequential loader (threads = 1, one chunk) - the tests parse adversarial files in 48-byte chunks. */
typedef struct NovumObjMesh NovumObjMesh;
int  novum_obj_parse(const char* path, const float emission[3], int materialID, float yOffset, int threads, int chunkBytes,
                     NovumObjMesh** out);
void novum_obj_free(NovumObjMesh* m);
int  novum_obj_counts(const NovumObjMesh* m, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights);
int  novum_obj_download(const NovumObjMesh* m, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs, NovumTriangle* tris);
/* ---- the OBJ front end on the device (the step directly before the path; replaces main.cu:936-1068 readObjSimple) ----
 * The text of the file is copied to the GPU as it is read and parsed there, one thread per line (csrc/obj_device.cu): the arrays are
 * those of novum_obj_parse bit for bit (positions, generated face normals, the shared zero uv, dropped triangles, light indices), but
 * they are born in device memory and go straight into the BVH8 build. A mesh object accumulates the files of one scene in order, as
 * the reference's initRender appends meshes (main.cu:470-503). Files with spellings the device rules do not read themselves (more than
 * 15 significant digits, inf / nan / hex numbers, faces with more than 16 corners, 2 GiB of text) return NOVUM_E_UNSUPPORTED and leave
 * the mesh unchanged: parse those with novum_obj_parse. novum_scene_load_config uses this path by itself (NOVUM_OBJ_DEVICE=0 turns
 * it off) and falls back to the host loader for such files. */
typedef struct NovumObjDeviceMesh NovumObjDeviceMesh;
int  novum_obj_device_create(NovumObjDeviceMesh** out);
void novum_obj_device_destroy(NovumObjDeviceMesh* m);
int  novum_obj_device_append_file(NovumObjDeviceMesh* m, const char* path, const float emission[3], int materialID, float yOffset);
/* the same for text that is already in (host) memory */
int  novum_obj_device_append_text(NovumObjDeviceMesh* m, const char* text, size_t bytes, const float emission[3], int materialID, float yOffset);
/* the same for text that is already in DEVICE memory (copied once into a padded buffer of the parser's own) */
int  novum_obj_device_append_device_text(NovumObjDeviceMesh* m, const char* d_text, size_t bytes, const float emission[3], int materialID, float yOffset);
/* empties the mesh but keeps its device memory (arrays and scratch) for the next files; timings restart at zero */
int  novum_obj_device_reset(NovumObjDeviceMesh* m);
int  novum_obj_device_counts(const NovumObjDeviceMesh* m, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights);
/* DEVICE pointers to the accumulated arrays (valid until the next append / destroy / novum_scene_from_obj_device) */
int  novum_obj_device_arrays(const NovumObjDeviceMesh* m, const NovumFloat4** d_positions, const NovumFloat4** d_normals,
                             const NovumFloat2** d_uvs, const NovumTriangle** d_tris);
int  novum_obj_device_download(const NovumObjDeviceMesh* m, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs, NovumTriangle* tris);
/* wall time of reading + copying the text, CUDA-event time of the parse kernels, their launch count, bytes of text parsed so far */
int  novum_obj_device_timings(const NovumObjDeviceMesh* m, float* uploadMs, float* kernelMs, int* launches, long long* textBytes);
/* scene from the mesh: the arrays MOVE into the scene (the mesh is empty afterwards and may be reused or destroyed); materials and
 * textures are host arrays as in novum_scene_create, whose index checks run on the device here (NOVUM_E_ARG). */
int  novum_scene_from_obj_device(NovumObjDeviceMesh* m, const NovumMaterial* mats, int nMat, const NovumFloat4* textures, int nTexels,
                                 NovumScene** out);
int  novum_material_table(NovumMaterial* out24);                      /* main.cu:397-467, 24 entries */
int  novum_camera_make(const NovumRenderConfig* cfg, NovumCamera* out); /* objects.cuh:221-266, 309-325 */
/* copies of the scene arrays back to host, for oracles and tests */
int  novum_scene_counts(const NovumScene* s, int* nPos, int* nNorm, int* nUV, int* nTri, int* nLights);
int  novum_scene_download(const NovumScene* s, NovumFloat4* positions, NovumFloat4* normals, NovumFloat2* uvs,
                          NovumTriangle* tris, NovumTriangle* lights);

/* ---- intersection on ray batches (integratorUtilities.cuh:84, :188) --------------------- */
/* device pointers; asynchronous on the library stream unless sync != 0 */
int  novum_trace_closest_device(NovumScene* s, const NovumRay* d_rays, int n, NovumHit* d_hits, int sync);
int  novum_trace_shadow_device(NovumScene* s, const NovumRay* d_rays, const float* d_maxT, const int32_t* d_skipTri,
                               int n, NovumFloat4* d_throughput, int sync);
/* host pointers: H2D of rays, trace, D2H of results, all inside the call */
int  novum_trace_closest(NovumScene* s, const NovumRay* rays, int n, NovumHit* hits);
int  novum_trace_shadow(NovumScene* s, const NovumRay* rays, const float* maxT, const int32_t* skipTri, int n,
                        NovumFloat4* throughput);
/* last trace's device time in ms (CUDA events on the library stream) and counters */
int  novum_trace_last_ms(NovumScene* s, float* ms);
int  novum_trace_counters(NovumScene* s, int enable, uint64_t* nodeVisits, uint64_t* triTests);
/* profiling switches for subsequent renders on this scene: bit 0 = per-stage CUDA-event timing (4 events per
 * wavefront iteration), bit 1 = count BVH8 node visits / triangle tests in the closest-hit stage (slower kernel). */
int  novum_set_profile(NovumScene* s, int flags);

/* ---- integrators ------------------------------------------------------------------------ */
int  novum_update_constants(const NovumConstants* c);                 /* deviceCode.cu:38 */
/* Scene-object form: adds `numSample` sample passes (pass indices firstSample .. firstSample+numSample-1,
 * strided by shard.world) into d_colors (w*h float4 SUM accumulator, like the reference's `colors`). */
int  novum_render_device(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample,
                         int numSample, NovumShard shard, uint64_t seed, NovumFloat4* d_colors, NovumRenderStats* stats);
/* Host form: renders and copies the accumulator to host memory (w*h float4). */
int  novum_render(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, int firstSample, int numSample,
                  NovumShard shard, uint64_t seed, NovumFloat4* colors, NovumRenderStats* stats);

/* The reference's launch boundary (deviceCode.cuh:8-18; device pointers exactly as main.cu:565, 641, 677, 786 pass them) + `numGpus`.
 * Blocking; `colors` receives the SUM over samples (the caller divides, main.cu:862). BVH / BVHindices are accepted and ignored (a BVH8 is
 * built on the device from `scene` + `vertices`). PathVertices / VCMPathVertices / Photons scratch of the reference is accepted as void* and
 * ignored: this core owns its own compact light-vertex cache. The strategy toggles are the ones novum_update_constants stored.
 *
 * Scene cache: the reference calls one launcher per render, 75 times (main.cu:925-933), re-uploading the scene each time. The boundary
 * keeps the last scene: a call with the same buffers and the same content (64-bit checksum of positions / triangles / lights / materials /
 * referenced normals + uvs, computed on the device) reuses BVH8 + engines without allocating; new content rebuilds the scene part and
 * moves the engines over (novum_launch_cache_stats counts both; novum_launch_cache_clear frees everything).
 * sceneCenter / sceneRadius / sceneMin are the caller's root-box values and are used as given (merge radius schedule, grid origin).
 * overlay: read by the progressive save and cleared after it (deviceCode.cu:1931, 3466); the reference writes it only in its debug path
 * drawing (BDPT_DRAWPATH), which this core lacks: with that constant set the bidirectional launchers return NOVUM_E_UNSUPPORTED.
 * Progressive output: every 5 s (novum_launch_set_progressive) "render.bmp" is written in the working directory from colors / overlay
 * (cleanAndFormatImage + saveImageBMP, deviceCode.cu:240-267, 1893-1934, 3447-3468), without stalling the sample loop; no CSV side file.
 * numGpus <= 1: the current device. numGpus > 1: devices 0..numGpus-1 of this process (peer access required), see DESIGN.md 6. */
int  novum_launch_unidirectional(int maxDepth, NovumCamera camera, NovumMaterial* materials, NovumFloat4* textures,
                                 NovumBVHnode* BVH, int* BVHindices, NovumVertices* vertices, int vertNum,
                                 NovumTriangle* scene, int triNum, NovumTriangle* lights, int lightNum, int numSample,
                                 int useMIS, int w, int h, NovumFloat4* colors, int numGpus);
int  novum_launch_naive_unidirectional(int maxDepth, NovumCamera camera, NovumMaterial* materials, NovumFloat4* textures,
                                 NovumBVHnode* BVH, int* BVHindices, NovumVertices* vertices, int vertNum,
                                 NovumTriangle* scene, int triNum, NovumTriangle* lights, int lightNum, int numSample,
                                 int useMIS, int w, int h, NovumFloat4* colors, int numGpus);
int  novum_launch_bidirectional(int eyeDepth, int lightDepth, NovumCamera camera, void* eyePath, void* lightPath,
                                 NovumMaterial* materials, NovumFloat4* textures, NovumBVHnode* BVH, int* BVHindices,
                                 NovumVertices* vertices, int vertNum, NovumTriangle* scene, int triNum,
                                 NovumTriangle* lights, int lightNum, int numSample, int w, int h,
                                 NovumFloat4 sceneCenter, float sceneRadius, NovumFloat4* colors, NovumFloat4* overlay,
                                 int postProcess, int numGpus);
int  novum_launch_vcm(int eyeDepth, int lightDepth, NovumCamera camera, void* lightPath, void* photons,
                                 void* photons_sorted, NovumMaterial* materials, NovumFloat4* textures, NovumBVHnode* BVH,
                                 int* BVHindices, NovumVertices* vertices, int vertNum, NovumTriangle* scene, int triNum,
                                 NovumTriangle* lights, int lightNum, int numSample, int w, int h,
                                 NovumFloat4 sceneCenter, float sceneRadius, NovumFloat4 sceneMin, NovumFloat4* colors,
                                 NovumFloat4* overlay, int postProcess, float mergeRadiusPower,
                                 float initialRadiusMultiplier, int numGpus);
typedef struct NovumLaunchCacheStats {
    uint64_t launches;          /* novum_launch_* calls that ran */
    uint64_t hits;              /* same buffers, same content: nothing built, nothing allocated */
    uint64_t rebinds;           /* same content in new allocations: borrowed pointers refreshed, nothing built */
    uint64_t sceneBuilds;       /* BVH8 + shading data built from the call's buffers */
    uint64_t enginesMoved;      /* scene rebuilds that kept the integrator engines' buffers */
    uint64_t progressiveSaves;  /* render.bmp files written by the launchers */
} NovumLaunchCacheStats;
int  novum_launch_cache_stats(NovumLaunchCacheStats* out);
void novum_launch_cache_clear(void);
/* seconds <= 0 disables the progressive save; path NULL keeps the current one ("render.bmp") */
int  novum_launch_set_progressive(float seconds, const char* path);

/* ---- photon hash grid (deviceCode.cu:3067-3198; integratorUtilities.cuh:382-412) -------- */
typedef struct NovumHashGrid NovumHashGrid;
int  novum_hashgrid_create(int maxPhotons, NovumHashGrid** out);      /* table size = nextPrime(2*maxPhotons) */
void novum_hashgrid_destroy(NovumHashGrid* g);
/* positions: device float x/y/z arrays (SoA, like Photons.pos_*). Produces sorted order + cell tables. */
int  novum_hashgrid_build_device(NovumHashGrid* g, const float* d_px, const float* d_py, const float* d_pz, int n,
                                 const float sceneMin[3], float mergeRadius, int sync);
int  novum_hashgrid_table_size(const NovumHashGrid* g);
/* host-facing: positions in, sorted photon indices + cell_start/cell_end out (tableSize entries each) */
int  novum_hashgrid_build(NovumHashGrid* g, const float* px, const float* py, const float* pz, int n,
                          const float sceneMin[3], float mergeRadius, uint32_t* sortedIndex, uint32_t* sortedHash,
                          uint32_t* cellStart, uint32_t* cellEnd);
/* radius query on the last-built grid: for each query point count photons within mergeRadius the way
 * doEyePass's 27-cell loop does (deviceCode.cu:2968-3048; duplicates from hash collisions included) */
int  novum_hashgrid_query_count(NovumHashGrid* g, const float* qx, const float* qy, const float* qz, int nq,
                                uint32_t* counts);
int  novum_hashgrid_last_ms(NovumHashGrid* g, float* ms);
/* Allocation-free grid build. No counterpart in the reference: added for callers that cannot
 * call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it). Same pipeline and same results as
 * novum_hashgrid_create(maxPhotons) + novum_hashgrid_build_device, but it owns no object and calls no cudaMalloc / cudaFree / allocating
 * library routine: all buffers are the caller's DEVICE buffers (positions in; d_sortedHash / d_sortedIndex: n entries each out;
 * d_cellStart / d_cellEnd: nextPrime(2 * maxPhotons) entries each out, cellStart = 0xFFFFFFFF for empty slots, cellEnd defined only
 * where cellStart is set), and the radix sort's scratch is a file-scope device array of the library (8.25 MiB held by every GPU that
 * loads it, until exit). Limit: n <= maxPhotons <= NOVUM_STATIC_GRID_MAX_PHOTONS, otherwise NOVUM_E_TOOBIG. Asynchronous on `stream`
 * (a cudaStream_t, NULL = default stream). Every call on a device shares the one scratch array: calls that overlap on the GPU race;
 * serialise them (one stream, or events). */
#define NOVUM_STATIC_GRID_MAX_PHOTONS (1 << 20)
int  novum_hashgrid_build_static(const float* d_px, const float* d_py, const float* d_pz, int n, int maxPhotons,
                                 const float sceneMin[3], float mergeRadius, uint32_t* d_sortedHash, uint32_t* d_sortedIndex,
                                 uint32_t* d_cellStart, uint32_t* d_cellEnd, void* stream);

/* ---- multi-GPU phase API (one process per GPU; the exchange itself is done by the caller with
 *      NCCL / torch.distributed on the device pointers returned here) --------------------------
 * Tile sharding: rank r owns pixel rows y % world == r for the light walk AND the eye walk of an iteration, so vertex
 * connections (paired by pixel in the reference, deviceCode.cu:2866-2942) stay local; what every rank needs from every
 * other rank is the photons (merging looks at all of them, :2968-3048). They live in `world` fixed-capacity slabs
 * ([32 B header: int count][32 B records]) laid out back to back: one in-place all-gather per iteration exchanges them.
 * The t=1 splats land on arbitrary pixels: every rank accumulates into its own full-frame buffer and the frames are summed
 * once at the end. (novum_render_device instead stripes whole iterations over ranks and needs no exchange at all.) */
typedef struct NovumVcmState NovumVcmState;
int  novum_vcm_create(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, NovumShard shard,
                      uint64_t seed, NovumVcmState** out);
void novum_vcm_destroy(NovumVcmState* v);
/* light pass for this rank's slab of light paths of iteration `iter`; fills the local slab of the
 * light-vertex cache and splats t=1 contributions into d_splat (w*h float4). */
int  novum_vcm_light_pass(NovumVcmState* v, int iter, NovumFloat4* d_splat);
/* device address + byte size of this rank's fixed-capacity slab inside the world-sized cache, so the caller
 * can all-gather in place (ncclAllGather / dist.all_gather_into_tensor on the whole cache). */
int  novum_vcm_cache_buffer(NovumVcmState* v, void** d_cacheBase, uint64_t* slabBytes, uint64_t* totalBytes);
/* after the gather: hash grid over all ranks' photons, then the eye pass over this rank's pixel tiles. */
int  novum_vcm_build_grid(NovumVcmState* v, int iter);
int  novum_vcm_eye_pass(NovumVcmState* v, int iter, NovumFloat4* d_colors);
int  novum_vcm_stats(NovumVcmState* v, NovumRenderStats* stats);
int  novum_vcm_reset_stats(NovumVcmState* v);     /* zero the counters; picks up novum_set_profile's flags for the stage timers */
/* the CUDA stream (cudaStream_t) the passes are enqueued on, so the caller can order its collective after the light pass */
int  novum_vcm_stream(NovumVcmState* v, void** stream);
/* Peer-write exchange (the fused form of "light pass -> all-gather"): every rank exports its slab buffer as a 64-byte CUDA
 * IPC handle, the caller carries the handles between the processes and attaches them, and with peer_write enabled each
 * light pass stores every photon record into its own slab AND into all peers' copies of it over NVLink as it appends them,
 * then publishes its count in the peers' headers. No collective moves photon data; the caller only orders the ranks with a
 * tiny stream-ordered collective (a) between light pass and grid build and (b) between grid build and the next light pass. */
int  novum_vcm_ipc_export(NovumVcmState* v, void* handle64);
int  novum_vcm_ipc_attach(NovumVcmState* v, int peerRank, const void* handle64);
int  novum_vcm_attach_pointer(NovumVcmState* v, int peerRank, void* d_peerCacheBase);   /* same-process ranks: novum_vcm_cache_buffer's base of the peer */
int  novum_vcm_peer_write(NovumVcmState* v, int enable);
int  novum_sync(void);

/* ---- tile renderer: the multi-GPU decomposition of the bidirectional integrators as ONE call per rank ---------------------------
 * Rank r owns pixel rows y % world == r for both walks of every iteration (as the phase API above). Every rank runs `lanes` iterations
 * concurrently (lane k: iterations first + k, first + k + lanes, ...; own stream + host thread each), which fills the latency-bound tails
 * of a depth-synchronous iteration - they do not shrink with the rank's share of the pixels. The photon exchange is fused into the light
 * pass (peer stores over NVLink into the same lane's slab on every peer) and the per-iteration rendezvous is a device-side wait on
 * stamps in the slab headers: no collective and no host round trip between ranks. Setup: create on every rank, exchange the lanes' slab
 * buffers (CUDA IPC handles across processes, pointers inside one process), then every rank calls novum_tiles_render with the same
 * (firstIter, numIter); the full-frame accumulators are summed by the caller at the end. world = 1 is valid (one GPU, lanes in flight). */
typedef struct NovumTiles NovumTiles;
int  novum_tiles_create(NovumScene* s, const NovumRenderConfig* cfg, const NovumCamera* cam, NovumShard shard, uint64_t seed, int lanes,
                        NovumTiles** out);
void novum_tiles_destroy(NovumTiles* t);
int  novum_tiles_lanes(const NovumTiles* t);
/* camera, strategy toggles, merge parameters and seed of the following renders (resolution, depths, integrator family stay as created) */
int  novum_tiles_set_view(NovumTiles* t, const NovumRenderConfig* cfg, const NovumCamera* cam, uint64_t seed);
int  novum_tiles_ipc_export(NovumTiles* t, void* handles /* lanes x 64 B */);
int  novum_tiles_ipc_attach(NovumTiles* t, int peerRank, const void* handles /* lanes x 64 B */);
int  novum_tiles_bases(NovumTiles* t, void** bases /* lanes */);
int  novum_tiles_attach_pointers(NovumTiles* t, int peerRank, void* const* bases /* lanes */);
int  novum_tiles_render(NovumTiles* t, int firstIter, int numIter, NovumFloat4* d_colors, NovumRenderStats* stats);

/* ---- synthetic scene generator (stand-ins for the reference's LFS-pointer OBJ files) ----- */
/* writes OBJ files + a .rendertron config under outDir; name in {cornell, dragon, water, rainbow}. */
int  novum_make_scene(const char* name, const char* outDir, int detail);

/* ---- frame finish + image output: the step directly after the path ----------------
 * One device kernel replaces cleanAndFormatImage (integratorUtilities.cuh:337-380) or the final normalisation of
 * initRender (main.cu:855-884), Image::postProcessImage = ACES toneMap + gammaCorrect (imageUtil.cu:202-232) and the
 * clamp / *255+0.5 / BGR / 4-byte-padded-row loop of Image::saveImageBMP (imageUtil.cu:82-96). Bytes are bit-exact with the
 * reference's host arithmetic. Only the packed rows (3 B/pixel) are read back, not two float4 frames. */
#define NOVUM_FRAME_PROGRESSIVE 0   /* cleanAndFormatImage: NaN/Inf/negative checks on the raw sum, then * 1/(sampleCount+1);
                                       sampleCount = the reference's currentSampleCount argument */
#define NOVUM_FRAME_FINAL       1   /* main.cu:862-870: / sampleCount, then NaN -> pink, Inf -> green */
int  novum_bmp_row_bytes(int w);                                      /* (3w+3)&~3, imageUtil.cu:80 */
/* device buffers in, device buffers out; d_overlay, d_bgrRows (h rows of novum_bmp_row_bytes(w), 4-byte aligned) and d_final
 * (the cleaned linear float4 frame the reference hands to Image::setColor) may each be NULL, not both outputs. Asynchronous on
 * `stream` (a cudaStream_t, NULL = default stream). */
int  novum_frame_finish_device(const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount, int mode,
                               int postProcess, uint8_t* d_bgrRows, NovumFloat4* d_final, void* stream);
/* host buffers in and out (uploads, runs the kernel, reads back); blocking */
int  novum_frame_finish(const NovumFloat4* colors, const NovumFloat4* overlay, int w, int h, int sampleCount, int mode, int postProcess,
                        uint8_t* bgrRows, NovumFloat4* finalOut);
/* device accumulators -> .bmp file: kernel, 3 B/pixel read-back into page-locked memory, headers + rows (imageUtil.cu:69-100) */
int  novum_frame_save_bmp(const char* path, const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h, int sampleCount,
                          int mode, int postProcess);
/* The same, off the critical path (the reference stalls its sample loop for this every 5 s, deviceCode.cu:1905-1927):
 * the kernel runs on the library's save stream once `renderStream` (a cudaStream_t, NULL = default stream) has reached this call,
 * `renderStream` then waits for the kernel only - it has read the accumulators - and the read-back and the file write overlap whatever
 * is queued next. Returns once the work is queued; at most two saves in flight (a third call waits for the oldest).
 * novum_frame_save_wait blocks until every queued file is on disk and returns the first error since the last wait. One saver per
 * process (one process per GPU): it belongs to the device current at its first call, calls from another device return NOVUM_E_ARG. */
int  novum_frame_save_bmp_async(const char* path, const NovumFloat4* d_colors, const NovumFloat4* d_overlay, int w, int h,
                                int sampleCount, int mode, int postProcess, void* renderStream);
int  novum_frame_save_wait(void);
/* host accumulators -> .bmp file (same kernel; no CPU tonemap exists in the product) */
int  novum_save_bmp(const char* path, const NovumFloat4* colors, const NovumFloat4* overlay, int w, int h, int sampleCount, int mode,
                    int postProcess);
int  novum_write_bmp(const char* path, const uint8_t* bgrRows, int w, int h);   /* headers + already packed rows */
int  novum_frame_last_ms(float* ms);                                  /* CUDA-event time of the last kernel run by the file/host variants */
int  novum_frame_gamma_steps(float* steps257);                        /* the 255 step positions of byte(t) (host powf), for inspection */
int  novum_save_pfm(const char* path, const NovumFloat4* colors, int w, int h, float scale);
/* Resumable accumulators (the reference has no equivalent - its sums live in device memory only, main.cu:818-853).
 * File = 64-byte header {"NOVUMACC", version 1, w, h, sampleCount, integrator, seed, payload bytes, checksum} + w*h float4 sums.
 * Resume: novum_accum_load, then novum_render(..., firstSample = sampleCount, ...) and add - sample k of a pixel depends only on
 * (seed, pixel, k). Host only, no GPU needed. Written via "<path>.part" + rename; load verifies size and checksum (NOVUM_E_IO). */
int  novum_accum_save(const char* path, const NovumFloat4* colors, int w, int h, int sampleCount, uint64_t seed, int integrator);
int  novum_accum_info(const char* path, int* w, int* h, int* sampleCount, uint64_t* seed, int* integrator);
int  novum_accum_load(const char* path, NovumFloat4* colors, int w, int h, int* sampleCount, uint64_t* seed);

#ifdef __cplusplus
}
#endif
#endif /* NOVUM_B200_H */
