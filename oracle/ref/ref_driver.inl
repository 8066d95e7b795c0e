// ref_driver.inl — command line around the reference's launch_* functions and intersection routines.
//   novum_ref render <sceneDir> [--spp N] [--out file.raw]      -> raw float4 accumulator (sum over samples) + JSON timing line
//   novum_ref trace  <sceneDir> <rays.bin> <out.bin> [--shadow] -> per-ray {tri,t,u,v} (closest) or throughput float4 (shadow)
// rays.bin: int32 n, then n x 8 floats (origin.xyzw, direction.xyzw); with --shadow followed by n floats maxT, n int32 skipTri.

static int cmdTrace(int argc, char** argv) {
    if (argc < 5) return 1;
    const char* dir = argv[2]; std::string raysPath = argv[3], outPath = argv[4];
    bool shadow = argc > 5 && !strcmp(argv[5], "--shadow");
    char cwd[4096]; if (!getcwd(cwd, sizeof(cwd))) return 1;
    auto abs = [&](const std::string& p) { return p[0] == '/' ? p : std::string(cwd) + "/" + p; };
    raysPath = abs(raysPath); outPath = abs(outPath);
    FILE* f = fopen(raysPath.c_str(), "rb");
    if (!f) { fprintf(stderr, "cannot open %s\n", raysPath.c_str()); return 1; }
    int n = 0;
    if (fread(&n, 4, 1, f) != 1 || n < 0) return 1;
    std::vector<Ray> rays(n); std::vector<float> maxT(n, 999999.0f); std::vector<int> skip(n, -1);
    if (n && fread(rays.data(), sizeof(Ray), n, f) != (size_t)n) return 1;
    if (shadow && n) { if (fread(maxT.data(), 4, n, f) != (size_t)n) return 1; if (fread(skip.data(), 4, n, f) != (size_t)n) return 1; }
    fclose(f);
    if (chdir(dir) != 0) return 1;
    RefScene S;
    if (buildRefScene(S, -1)) return 1;
    Ray* d_rays; float4* d_out; float* d_maxT; int* d_skip;
    cudaMalloc(&d_rays, sizeof(Ray) * (n + 1)); cudaMalloc(&d_out, sizeof(float4) * (n + 1));
    cudaMalloc(&d_maxT, 4 * (n + 1)); cudaMalloc(&d_skip, 4 * (n + 1));
    cudaMemcpy(d_rays, rays.data(), sizeof(Ray) * n, cudaMemcpyHostToDevice);
    cudaMemcpy(d_maxT, maxT.data(), 4 * n, cudaMemcpyHostToDevice);
    cudaMemcpy(d_skip, skip.data(), 4 * n, cudaMemcpyHostToDevice);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 3; rep++) {       // best of 3 (first run warms caches / clocks)
        cudaEventRecord(e0);
        if (n) {
            if (shadow) refShadowKernel<<<(n + 255) / 256, 256>>>(d_rays, d_maxT, d_skip, n, S.BVH_d, S.BVHindices_d, S.verts_d, S.scene_d, S.mats_d, d_out);
            else refClosestKernel<<<(n + 255) / 256, 256>>>(d_rays, n, S.BVH_d, S.BVHindices_d, S.verts_d, S.scene_d, d_out);
        }
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
        best = fminf(best, ms);
    }
    cudaError_t err = cudaDeviceSynchronize();
    std::vector<float4> host(n);
    cudaMemcpy(host.data(), d_out, sizeof(float4) * n, cudaMemcpyDeviceToHost);
    f = fopen(outPath.c_str(), "wb");
    if (f) { fwrite(host.data(), sizeof(float4), n, f); fclose(f); }
    printf("\nNOVUM_REF_JSON {\"mode\": \"%s\", \"rays\": %d, \"kernel_ms\": %.4f, \"mrays_per_s\": %.2f, \"tris\": %d, \"bvh_nodes\": %d, \"cuda_error\": %d}\n",
           shadow ? "shadow" : "closest", n, best, n / (best * 1e3), (int)S.mesh.size(), (int)S.bvh.size(), (int)err);
    return err == cudaSuccess ? 0 : 2;
}

// ---- frame finish: the reference's own post-path code on a raw float4 frame --------------------------------------------
//   novum_ref bmp   <frame.raw> <w> <h> <postProcess> <out.bmp> [--time N]   CPU ONLY: Image::setColor + Image::saveImageBMP
//                   (imageUtil.cu:60-100, 202-232: ACES toneMap, gammaCorrect, clamp/*255+0.5, BGR rows). Runs without a GPU.
//   novum_ref clean <acc.raw> <overlay.raw> <w> <h> <currentSampleCount> <out.raw>   GPU: cleanAndFormatImage
//                   (integratorUtilities.cuh:337-380), launched with the reference's 16x16 blocks (deviceCode.cu:3313, 3449).
static bool readRaw(const char* path, std::vector<float4>& v, size_t n) {
    FILE* f = fopen(path, "rb"); if (!f) return false;
    v.resize(n); size_t got = fread(v.data(), sizeof(float4), n, f); fclose(f);
    return got == n;
}

static int cmdBmp(int argc, char** argv) {
    int w = atoi(argv[3]), h = atoi(argv[4]); bool post = atoi(argv[5]) != 0; int reps = 1;
    if (argc >= 9 && !strcmp(argv[7], "--time")) reps = atoi(argv[8]);
    std::vector<float4> px;
    if (w <= 0 || h <= 0 || !readRaw(argv[2], px, (size_t)w * h)) { fprintf(stderr, "cannot read %s\n", argv[2]); return 1; }
    double best = 1e30;
    for (int r = 0; r < reps; r++) {
        auto t0 = std::chrono::high_resolution_clock::now();
        Image image(w, h);
        image.postProcess = post;
        for (int y = 0; y < h; y++) for (int x = 0; x < w; x++) image.setColor(x, y, px[(size_t)y * w + x]);
        image.saveImageBMP(argv[6]);
        best = std::min(best, std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count());
    }
    printf("\nNOVUM_REF_JSON {\"mode\": \"bmp\", \"w\": %d, \"h\": %d, \"post\": %d, \"host_ms\": %.3f}\n", w, h, (int)post, best * 1e3);
    return 0;
}

static int cmdClean(int argc, char** argv) {
    int w = atoi(argv[4]), h = atoi(argv[5]), n = atoi(argv[6]);
    std::vector<float4> acc, ov;
    if (w <= 0 || h <= 0 || !readRaw(argv[2], acc, (size_t)w * h) || !readRaw(argv[3], ov, (size_t)w * h)) { fprintf(stderr, "cannot read inputs\n"); return 1; }
    size_t bytes = sizeof(float4) * (size_t)w * h;
    float4 *dA, *dO, *dF; cudaMalloc(&dA, bytes); cudaMalloc(&dO, bytes); cudaMalloc(&dF, bytes);
    cudaMemcpy(dA, acc.data(), bytes, cudaMemcpyHostToDevice); cudaMemcpy(dO, ov.data(), bytes, cudaMemcpyHostToDevice);
    dim3 blockSize(16, 16); dim3 gridSize((w + 15) / 16, (h + 15) / 16);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        cleanAndFormatImage<<<gridSize, blockSize>>>(dA, dO, dF, w, h, n);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = fminf(best, ms);
    }
    cudaError_t err = cudaDeviceSynchronize();
    std::vector<float4> out((size_t)w * h);
    cudaMemcpy(out.data(), dF, bytes, cudaMemcpyDeviceToHost);
    FILE* f = fopen(argv[7], "wb"); if (f) { fwrite(out.data(), sizeof(float4), out.size(), f); fclose(f); }
    printf("\nNOVUM_REF_JSON {\"mode\": \"clean\", \"w\": %d, \"h\": %d, \"kernel_ms\": %.4f, \"cuda_error\": %d}\n", w, h, best, (int)err);
    return err == cudaSuccess ? 0 : 2;
}

// ---- photon hash grid: the reference's own buildHashGrid (deviceCode.cu:3140-3198) on a given photon set ------------------
//   novum_ref grid <photons.bin> <out.bin>
// photons.bin: int32 n, int32 maxPhotons, float radius, float sceneMin[3], then x[n], y[n], z[n] (float32).
// out.bin: uint32 sortedHash[n], uint32 sortedIndex[n], uint32 cellStart[sortedHash[i]] (n), uint32 cellEnd[sortedHash[i]] (n).
// Buffers are allocated as launch_VCM does (deviceCode.cu:3276-3296): table = GetNextPrime(2 * maxPhotons), cub scratch sized for maxPhotons.
__global__ void refGatherTable(const unsigned int* hashes, const unsigned int* cs, const unsigned int* ce, int n, unsigned int* o0, unsigned int* o1) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { o0[i] = cs[hashes[i]]; o1[i] = ce[hashes[i]]; }
}

static int cmdGrid(int argc, char** argv) {
    FILE* f = fopen(argv[2], "rb"); if (!f) { fprintf(stderr, "cannot open %s\n", argv[2]); return 1; }
    int n = 0, maxPhotons = 0; float radius = 0, mn[3];
    if (fread(&n, 4, 1, f) != 1 || fread(&maxPhotons, 4, 1, f) != 1 || fread(&radius, 4, 1, f) != 1 || fread(mn, 4, 3, f) != 3 || n <= 0 || n > maxPhotons) return 1;
    std::vector<float> h((size_t)n * 3);
    if (fread(h.data(), 4, h.size(), f) != h.size()) return 1;
    fclose(f);
    auto mkPhotons = [&](size_t cap) {
        Photons p;
        p.pos_x = devAllocZero<float>(cap); p.pos_y = devAllocZero<float>(cap); p.pos_z = devAllocZero<float>(cap);
        p.packedWi = devAllocZero<unsigned int>(cap); p.packedNormal = devAllocZero<unsigned int>(cap);
        p.beta_x = devAllocZero<half>(cap); p.beta_y = devAllocZero<half>(cap); p.beta_z = devAllocZero<half>(cap);
        p.d_vm = devAllocZero<float>(cap); p.d_vcm = devAllocZero<float>(cap);
        return p;
    };
    Photons ph = mkPhotons(n), sorted = mkPhotons(n);
    cudaMemcpy(ph.pos_x, h.data(), 4 * (size_t)n, cudaMemcpyHostToDevice);
    cudaMemcpy(ph.pos_y, h.data() + n, 4 * (size_t)n, cudaMemcpyHostToDevice);
    cudaMemcpy(ph.pos_z, h.data() + 2 * (size_t)n, 4 * (size_t)n, cudaMemcpyHostToDevice);
    unsigned int *kin = devAllocZero<unsigned int>(n), *kout = devAllocZero<unsigned int>(n), *iin = devAllocZero<unsigned int>(n), *iout = devAllocZero<unsigned int>(n);
    int hashTableSize = GetNextPrime(maxPhotons * 2);
    unsigned int *cs, *ce; cudaMalloc(&cs, (size_t)hashTableSize * 4); cudaMalloc(&ce, (size_t)hashTableSize * 4);
    void* tmp = NULL; size_t tmpBytes = 0;
    cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, kin, kout, iin, iout, n);
    cudaMalloc(&tmp, tmpBytes);
    float4 sceneMin = f4(mn[0], mn[1], mn[2]);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f, bestReorder = 1e30f;
    for (int rep = 0; rep < 3; rep++) {
        cudaDeviceSynchronize();
        cudaEventRecord(e0);
        buildHashGrid(ph, sorted, n, kin, kout, iin, iout, tmp, tmpBytes, cs, ce, sceneMin, radius, hashTableSize);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); best = fminf(best, ms);
        cudaEventRecord(e0);
        reorderPhotons<<<(n + 255) / 256, 256>>>(ph, sorted, n, iout);      // its share of the total, timed on its own
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        cudaEventElapsedTime(&ms, e0, e1); bestReorder = fminf(bestReorder, ms);
    }
    unsigned int *o0 = devAllocZero<unsigned int>(n), *o1 = devAllocZero<unsigned int>(n);
    refGatherTable<<<(n + 255) / 256, 256>>>(kout, cs, ce, n, o0, o1);
    cudaError_t err = cudaDeviceSynchronize();
    std::vector<unsigned int> out((size_t)n * 4);
    cudaMemcpy(out.data(), kout, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    cudaMemcpy(out.data() + n, iout, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    cudaMemcpy(out.data() + 2 * (size_t)n, o0, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    cudaMemcpy(out.data() + 3 * (size_t)n, o1, 4 * (size_t)n, cudaMemcpyDeviceToHost);
    f = fopen(argv[3], "wb"); if (f) { fwrite(out.data(), 4, out.size(), f); fclose(f); }
    printf("\nNOVUM_REF_JSON {\"mode\": \"grid\", \"photons\": %d, \"table_size\": %d, \"build_ms\": %.4f, \"reorder_ms\": %.4f, \"cuda_error\": %d}\n",
           n, hashTableSize, best, bestReorder, (int)err);
    return err == cudaSuccess ? 0 : 2;
}

int main(int argc, char** argv) {
    if (argc >= 3 && !strcmp(argv[1], "render")) return cmdRender(argc, argv);
    if (argc >= 5 && !strcmp(argv[1], "trace")) return cmdTrace(argc, argv);
    if (argc >= 4 && !strcmp(argv[1], "grid")) return cmdGrid(argc, argv);
    if (argc >= 7 && !strcmp(argv[1], "bmp")) return cmdBmp(argc, argv);
    if (argc >= 8 && !strcmp(argv[1], "clean")) return cmdClean(argc, argv);
    fprintf(stderr, "usage: novum_ref render <sceneDir> [--spp N] [--out f.raw] | trace <sceneDir> <rays.bin> <out.bin> [--shadow]\n"
                    "       novum_ref bmp <frame.raw> <w> <h> <postProcess> <out.bmp> [--time N] | clean <acc.raw> <overlay.raw> <w> <h> <n> <out.raw>\n"
                    "       novum_ref grid <photons.bin> <out.bin>\n");
    return 64;
}
