// dropin_all.cu — the reference's OWN main.cu + imageUtil.cu (unmodified, included where they lie under /root/reference, exactly as
// ref_all.cu does) built against libnovum_b200.so: deviceCode.cu is replaced by integration/novum_shim.cu, nothing else changes.
// Output: oracle/_ref/novum_dropin (oracle/build_ref.sh). TEST INFRASTRUCTURE: this is what proves the drop-in boundary -
//   novum_dropin render <sceneDir> [--spp N] [--out f.raw] [--repeat R] [--no-progressive]
//        the driver of ref_render.inl (the buffers initRender allocates, main.cu:560-780) calling launch_* -> the shim -> novum_launch_*
//   novum_dropin initrender <sceneDir> [count]
//        the reference's own initRender("configs/config.rendertron", i) for i < count, untouched: config, textures, materials, OBJ
//        files, CPU BVH2, uploads, launch, final host pass, renders/<name><i>.bmp (main.cu:235-923)
// No reference source is copied into the repository; reference device code is not compiled into this binary at all.
#include <cstdint>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cuda_runtime.h>
#include <cuda.h>
#include <curand_kernel.h>
#include <cuda_fp16.h>
#include <algorithm>
#include <sstream>
#include <numeric>
#include <iostream>
#include <fstream>
#include <string>
#include <vector>
#include <random>
#include <ctime>
#include <chrono>
#include <set>
#include <iomanip>
#include <unistd.h>
#include <utility>
#include <sys/stat.h>
#undef __forceinline__
#define __forceinline__ __attribute__((always_inline))
#include "../../integration/novum_shim.cu"
#define main novum_reference_main_unused
#include "main.cu"
#undef main
#include "imageUtil.cu"
#define NOVUM_DROPIN 1
#include "ref_scene.inl"
#include "ref_render.inl"

static void printCacheStats() {
    NovumLaunchCacheStats cs;
    novum_launch_cache_stats(&cs);
    printf("\nNOVUM_DROPIN_CACHE {\"launches\": %llu, \"hits\": %llu, \"rebinds\": %llu, \"scene_builds\": %llu, \"engines_moved\": %llu, "
           "\"progressive_saves\": %llu}\n", (unsigned long long)cs.launches, (unsigned long long)cs.hits, (unsigned long long)cs.rebinds,
           (unsigned long long)cs.sceneBuilds, (unsigned long long)cs.enginesMoved, (unsigned long long)cs.progressiveSaves);
}

static int cmdInitRender(int argc, char** argv) {
    int count = argc > 3 ? atoi(argv[3]) : 1;
    if (chdir(argv[2]) != 0) { fprintf(stderr, "cannot chdir to %s\n", argv[2]); return 1; }
    mkdir("renders", 0755);
    for (int i = 0; i < count; i++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        int rc = initRender("configs/config.rendertron", i);
        double s = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
        printf("\nNOVUM_DROPIN_INITRENDER {\"render\": %d, \"rc\": %d, \"seconds\": %.4f}\n", i, rc, s);
        if (rc) return 2;
    }
    printCacheStats();
    cudaError_t err = cudaDeviceSynchronize();
    return err == cudaSuccess ? 0 : 2;
}

int main(int argc, char** argv) {
    if (argc >= 3 && !strcmp(argv[1], "render")) { int rc = cmdRender(argc, argv); printCacheStats(); return rc; }
    if (argc >= 3 && !strcmp(argv[1], "initrender")) return cmdInitRender(argc, argv);
    fprintf(stderr, "usage: novum_dropin render <sceneDir> [--spp N] [--out f.raw] [--repeat R] [--no-progressive] | initrender <sceneDir> [count]\n");
    return 64;
}
