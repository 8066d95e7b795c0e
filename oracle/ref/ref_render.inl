// ref_render.inl — `render <sceneDir> [--spp N] [--out file.raw]`: the reference's launch_* (its own deviceCode.cu in novum_ref, the
// shim over libnovum_b200.so in novum_dropin) driven with the buffers initRender allocates (main.cu:560-780); dumps the raw float4
// SUM accumulator and a JSON timing line. TEST / BASELINE INFRASTRUCTURE ONLY.
template <typename T> static T* devAllocZero(size_t count) {
    T* p = nullptr;
    cudaMalloc(&p, sizeof(T) * (count ? count : 1));
    cudaMemset(p, 0, sizeof(T) * (count ? count : 1));
    return p;
}

static int cmdRender(int argc, char** argv) {
    const char* dir = argv[2]; int spp = -1, repeat = 1; std::string out = "ref.raw";
    for (int i = 3; i < argc; i++) {
        if (!strcmp(argv[i], "--spp") && i + 1 < argc) spp = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--out") && i + 1 < argc) out = argv[++i];
        else if (!strcmp(argv[i], "--repeat") && i + 1 < argc) repeat = atoi(argv[++i]);
#ifdef NOVUM_DROPIN
        else if (!strcmp(argv[i], "--no-progressive")) novum_launch_set_progressive(0.0f, nullptr);
#endif
    }
    if (repeat < 1) repeat = 1;
    char cwd[4096]; if (!getcwd(cwd, sizeof(cwd))) return 1;
    std::string outAbs = out[0] == '/' ? out : std::string(cwd) + "/" + out;
    if (chdir(dir) != 0) { fprintf(stderr, "cannot chdir to %s\n", dir); return 1; }
    RefScene S;
    int rc = buildRefScene(S, spp);
    if (rc) { fprintf(stderr, "scene build failed (%d)\n", rc); return 1; }
    RenderConfig& config = S.config;
    int w = config.width, h = config.height, integ = matchIntegrator(config.integratorType);
    float4* out_colors = devAllocZero<float4>((size_t)w * h);
    float4* out_overlay = devAllocZero<float4>((size_t)w * h);
    int nv = (int)S.points.size(), nt = (int)S.mesh.size(), nl = (int)S.lightsvec.size();
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    std::string eachMs = "[";
    auto t0 = std::chrono::high_resolution_clock::now();
    // --repeat R: the same launch R times on the same buffers (what main.cu:925-933 does 75 times); the accumulator is cleared in between
    // and the LAST launch is the one timed and dumped. The reference's scratch (PathVertices / VCMPathVertices / Photons) is allocated once.
    PathVertices *bdEye = nullptr, *bdLight = nullptr;
    VCMPathVertices lp; Photons ph, phs; bool vcmScratch = false;
    for (int rep = 0; rep < repeat; rep++) {
    cudaMemset(out_colors, 0, sizeof(float4) * (size_t)w * h);
    cudaDeviceSynchronize();
    t0 = std::chrono::high_resolution_clock::now();
    cudaEventRecord(e0);
    if (integ == UNIDIRECTIONAL)
        launch_unidirectional(config.maxDepth, S.camera, S.mats_d, S.textures_d, S.BVH_d, S.BVHindices_d, S.verts_d, nv, S.scene_d, nt, S.lights_d, nl,
                              config.sampleCount, true, w, h, out_colors);
    else if (integ == NAIVE_UNIDIRECTIONAL)
        launch_naive_unidirectional(config.maxDepth, S.camera, S.mats_d, S.textures_d, S.BVH_d, S.BVHindices_d, S.verts_d, nv, S.scene_d, nt, S.lights_d, nl,
                                    config.sampleCount, true, w, h, out_colors);
    else if (integ == BIDIRECTIONAL) {
        auto mk = [&](size_t n) {
            PathVertices t;
            t.materialID = devAllocZero<int>(n); t.pt = devAllocZero<float4>(n); t.n = devAllocZero<float4>(n); t.wo = devAllocZero<float4>(n);
            t.beta = devAllocZero<float4>(n); t.d_vcm = devAllocZero<float>(n); t.d_vc = devAllocZero<float>(n); t.pdfFwd = devAllocZero<float>(n);
            t.misWeight = devAllocZero<float>(n); t.isDelta = devAllocZero<bool>(n); t.lightInd = devAllocZero<int>(n); t.uv = devAllocZero<float2>(n);
            t.backface = devAllocZero<bool>(n);
            PathVertices* d; cudaMalloc(&d, sizeof(PathVertices)); cudaMemcpy(d, &t, sizeof(PathVertices), cudaMemcpyHostToDevice);
            return d;
        };
        if (!bdEye) { bdEye = mk((size_t)w * h * config.bdptEyeDepth); bdLight = mk((size_t)w * h * config.bdptLightDepth); }
        launch_bidirectional(config.bdptEyeDepth, config.bdptLightDepth, S.camera, bdEye, bdLight, S.mats_d, S.textures_d, S.BVH_d, S.BVHindices_d, S.verts_d, nv,
                             S.scene_d, nt, S.lights_d, nl, config.sampleCount, w, h, S.sceneCenter, S.sceneRadius, out_colors, out_overlay, config.postProcess);
    } else if (integ == VCM || integ == SPPM) {
        size_t n = (size_t)w * h * config.bdptLightDepth;
        if (!vcmScratch) {
        lp.pos_x = devAllocZero<float>(n); lp.pos_y = devAllocZero<float>(n); lp.pos_z = devAllocZero<float>(n);
        lp.beta_x = devAllocZero<half>(n); lp.beta_y = devAllocZero<half>(n); lp.beta_z = devAllocZero<half>(n);
        lp.packedNormal = devAllocZero<unsigned int>(n); lp.packedWo = devAllocZero<unsigned int>(n); lp.packedInfo = devAllocZero<unsigned int>(n);
        lp.packedUV = devAllocZero<half2>(n); lp.d_vc = devAllocZero<float>(n); lp.d_vcm = devAllocZero<float>(n);
        auto mkPh = [&]() {
            Photons p;
            p.pos_x = devAllocZero<float>(n); p.pos_y = devAllocZero<float>(n); p.pos_z = devAllocZero<float>(n);
            p.packedWi = devAllocZero<unsigned int>(n); p.packedNormal = devAllocZero<unsigned int>(n);
            p.beta_x = devAllocZero<half>(n); p.beta_y = devAllocZero<half>(n); p.beta_z = devAllocZero<half>(n);
            p.d_vm = devAllocZero<float>(n); p.d_vcm = devAllocZero<float>(n);
            return p;
        };
        ph = mkPh(); phs = mkPh(); vcmScratch = true;
        }
        launch_VCM(config.bdptEyeDepth, config.bdptLightDepth, S.camera, &lp, &ph, &phs, S.mats_d, S.textures_d, S.BVH_d, S.BVHindices_d, S.verts_d, nv,
                   S.scene_d, nt, S.lights_d, nl, config.sampleCount, w, h, S.sceneCenter, S.sceneRadius, S.sceneMin, out_colors, out_overlay,
                   config.postProcess, config.vcmMergeConst, config.vcmInitialMergeRadiusMultiplier);
    } else { fprintf(stderr, "unknown integrator\n"); return 1; }
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    { float m = 0; cudaEventElapsedTime(&m, e0, e1); char b[32]; snprintf(b, sizeof(b), "%s%.3f", rep ? ", " : "", m); eachMs += b; }
    }
    eachMs += "]";
    cudaError_t err = cudaDeviceSynchronize();
    double wall = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    std::vector<float4> host((size_t)w * h);
    cudaMemcpy(host.data(), out_colors, host.size() * sizeof(float4), cudaMemcpyDeviceToHost);
    FILE* f = fopen(outAbs.c_str(), "wb");
    if (f) { fwrite(host.data(), sizeof(float4), host.size(), f); fclose(f); }
    unsigned long long rayCnt[2] = {0, 0};
#ifdef NOVUM_REF_COUNT_RAYS
    cudaMemcpyFromSymbol(rayCnt, g_refRayCounters, sizeof(rayCnt));
#endif
    printf("\nNOVUM_REF_JSON {\"integrator\": \"%s\", \"w\": %d, \"h\": %d, \"spp\": %d, \"tris\": %d, \"lights\": %d, \"bvh_nodes\": %d, "
           "\"launch_ms\": %.3f, \"launch_ms_each\": %s, \"launch_wall_s\": %.4f, \"bvh_build_s\": %.4f, \"obj_parse_s\": %.4f, \"closest_rays\": %llu, \"shadow_rays\": %llu, \"cuda_error\": %d}\n",
           config.integratorType.c_str(), w, h, config.sampleCount, nt, nl, (int)S.bvh.size(), ms, eachMs.c_str(), wall, S.buildSeconds, S.parseSeconds, rayCnt[0], rayCnt[1], (int)err);
    return err == cudaSuccess ? 0 : 2;
}

