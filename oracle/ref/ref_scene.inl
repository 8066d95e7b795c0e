// ref_scene.inl — scene set-up of the reference drivers (novum_ref and novum_dropin): restates initRender's flow (main.cu:235-557) with the
// reference's own loadConfig / loadBMPToImage / readObjSimple / computeInfoForBVH / buildBVH / Camera / Material factories.
// TEST / BASELINE INFRASTRUCTURE ONLY. Material fields the reference leaves indeterminate are zeroed.
static void sanitize(Material& m) {
    if (m.type != MAT_LEAF && m.type != MAT_DELTAMIRROR) m.hasTransMap = false;
    if (m.type != MAT_SMOOTHDIELECTRIC) m.priority = 0;
    if (m.type == MAT_DELTAMIRROR) { m.boundary = false; m.thinWalled = false; m.absorption = f4(); }
    if (m.type == MAT_LEAF) m.absorption = f4();
    if (!m.hasTexture) { m.startInd = 0; m.width = 0; m.height = 0; }
    m.tstartInd = 0; m.twidth = 0; m.theight = 0;
}

struct RefScene {
    RenderConfig config; Camera camera;
    vector<float4> points, normals, colors; vector<float2> uvs; vector<Triangle> mesh, lightsvec; vector<BVHnode> bvh; vector<int> ind;
    float4 sceneCenter, sceneMin; float sceneRadius;
    Material* mats_d; float4* textures_d; Vertices* verts_d; Vertices tmp; Triangle *scene_d, *lights_d; BVHnode* BVH_d; int* BVHindices_d;
    double buildSeconds;
    double parseSeconds;      // the reference's readObjSimple over the config's meshes (main.cu:474-482)
};

static int buildRefScene(RefScene& S, int sppOverride) {
    RenderConfig& config = S.config;
    if (!loadConfig("configs/config.rendertron", config)) return 1;
    if (sppOverride > 0) config.sampleCount = sppOverride;
    int w = config.width, h = config.height;
    if (config.pinholeCamera) S.camera = Camera::Pinhole(config.camPos, w, h, config.camRot.x, config.camRot.y, config.camRot.z, config.camFov);
    else S.camera = Camera::NotPinhole(config.camPos, w, h, config.camRot.x, config.camRot.y, config.camRot.z, config.camFov, config.camApeture, config.camFocalDist);
    if (matchIntegrator(config.integratorType) == SPPM) {          // main.cu:314-333
        config.bdptConnection = false; config.bdptNaive = false; config.bdptNee = false; config.bdptLightTrace = false;
        config.bdptDoMis = false; config.vcmDoMerge = true; config.doSPPM = true;
    }
    updateConstants(config);
    // textures, main.cu:364-391 (files that do not exist give zero-size images, as in the reference; such materials are then
    // marked texture-less below, where the reference would take a modulo by zero in sampleTexture)
    vector<float4> pixels; vector<int> widths, heights, startIndices; int currentStartIndex = 0;
    const char* texNames[4] = {"textures/enkidutexture.bmp", "textures/enkiduchibitexture.bmp", "textures/leaftex2.bmp", "textures/leafautumn.bmp"};
    for (int t = 0; t < 4; t++) {
        Image i = loadBMPToImage(texNames[t], false);
        vector<float4> pix = i.data();
        pixels.insert(pixels.end(), pix.begin(), pix.end());
        widths.push_back(i.width); heights.push_back(i.height); startIndices.push_back(currentStartIndex);
        currentStartIndex += i.width * i.height;
    }
    // material table, main.cu:397-467
    vector<Material> mats;
    float4 eta_steel = f4(0.14f, 0.16f, 0.13f, 1.0f), eta_gold = f4(0.17f, 0.35f, 1.5f);
    mats.push_back(Material::SmoothDielectric(1.0f, f4(0.0f), 99));
    mats.push_back(Material::Diffuse(f4(0.4f, 0.4f, 0.8f)));
    mats.push_back(Material::Diffuse(f4(0.9f, 0.9f, 0.9f)));
    mats.push_back(Material::Diffuse(f4(0.2f, 0.6f, 0.6f)));
    mats.push_back(Material::Metal(eta_gold, eta_gold, 0.05f));
    mats.push_back(Material::SmoothDielectric(1.5f, f4(0.0f), 1));
    mats.push_back(Material::Diffuse(f4(0.90f, 0.1f, 0.1f)));
    mats.push_back(Material::Metal(eta_steel, eta_steel, 0.15f));
    mats.push_back(Material::SmoothDielectric(1.333f, 2.5f * f4(0.180f, 1.5f, 2.996f), 2));
    mats.push_back(Material::SmoothDielectric(1.31f, f4(0.2f), 0));
    mats.push_back(Material::SmoothDielectric(1.333f, f4(), 2));
    mats.push_back(Material::DiffuseTextured(startIndices[0], widths[0], heights[0]));
    mats.push_back(Material::DiffuseTextured(startIndices[1], widths[1], heights[1]));
    mats.push_back(Material::Leaf(startIndices[2], widths[2], heights[2], 1.5f, 0.10f, f4(0.22f, 0.75f, 0.28f), 0.15f));
    mats.push_back(Material::Diffuse(f4(0.90f, 0.9f, 0.83f)));
    mats.push_back(Material::Diffuse(f4(0.4f, 0.4f, 1.00f)));
    mats.push_back(Material::Leaf(startIndices[3], widths[3], heights[3], 1.5f, 0.8f, f4(0.22f, 0.75f, 0.28f), 0.6f));
    mats.push_back(Material::Diffuse(f4(0.8f, 0.8f, 0.8f)));
    mats.push_back(Material::SmoothDielectric(2.42f, f4(0.0f), 1));
    mats.push_back(Material::Mirror());
    mats.push_back(Material::Diffuse(f4(0.0f, 0.0f, 0.0f)));
    mats.push_back(Material::Diffuse(f4(0.95f, 0.95f, 0.95f)));
    mats.push_back(Material::Diffuse(f4(0.5f, 0.5f, 0.5f)));
    mats.push_back(Material::Diffuse(f4(0.1f, 0.9f, 0.1f)));
    for (Material& m : mats) sanitize(m);
    for (Material& m : mats) if (m.hasTexture && (m.width <= 0 || m.height <= 0)) m.hasTexture = false;          // no texels to sample
    cudaMalloc(&S.mats_d, mats.size() * sizeof(Material));
    cudaMemcpy(S.mats_d, mats.data(), mats.size() * sizeof(Material), cudaMemcpyHostToDevice);
    cudaMalloc(&S.textures_d, pixels.size() * sizeof(float4) + 16);
    cudaMemcpy(S.textures_d, pixels.data(), pixels.size() * sizeof(float4), cudaMemcpyHostToDevice);
    auto tParse = std::chrono::high_resolution_clock::now();
    for (MeshConfig c : config.meshes)      // main.cu:474-482 (render number 0: no light offset)
        readObjSimple(c.path, S.points, S.normals, S.colors, S.uvs, S.mesh, S.lightsvec, f4(), c.emissionMultiplier * c.emissionColor, c.materialID);
    S.parseSeconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - tParse).count();
    if (S.mesh.empty()) return 2;
    S.colors.assign(S.points.size(), f4());        // the reference leaves `colors` empty and reads it out of bounds (dead code after DCE)
    cudaMalloc(&S.verts_d, sizeof(Vertices));
    cudaMalloc(&S.tmp.positions, sizeof(float4) * S.points.size());
    cudaMalloc(&S.tmp.normals, sizeof(float4) * S.normals.size());
    cudaMalloc(&S.tmp.colors, sizeof(float4) * S.colors.size());
    cudaMalloc(&S.tmp.uvs, sizeof(float2) * S.uvs.size());
    cudaMemcpy(S.tmp.positions, S.points.data(), S.points.size() * sizeof(float4), cudaMemcpyHostToDevice);
    cudaMemcpy(S.tmp.normals, S.normals.data(), S.normals.size() * sizeof(float4), cudaMemcpyHostToDevice);
    cudaMemcpy(S.tmp.colors, S.colors.data(), S.colors.size() * sizeof(float4), cudaMemcpyHostToDevice);
    cudaMemcpy(S.tmp.uvs, S.uvs.data(), S.uvs.size() * sizeof(float2), cudaMemcpyHostToDevice);
    cudaMemcpy(S.verts_d, &S.tmp, sizeof(Vertices), cudaMemcpyHostToDevice);
    auto t0 = std::chrono::high_resolution_clock::now();
    Vertices hostV; hostV.positions = S.points.data(); hostV.normals = S.normals.data(); hostV.colors = S.colors.data(); hostV.uvs = S.uvs.data();
    vector<float4> centroids, minboxes, maxboxes;
    S.ind.resize(S.mesh.size());
    for (size_t i = 0; i < S.mesh.size(); i++) S.ind[i] = (int)i;
    computeInfoForBVH(hostV, S.mesh, centroids, minboxes, maxboxes);
    int largest = 0, backup = 0;
    buildBVH(S.bvh, S.ind, centroids, minboxes, maxboxes, 0, (int)S.mesh.size(), config.bvhLeafSize, largest, backup);
    S.buildSeconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
    BVHnode root = S.bvh[0];
    S.sceneCenter = (root.aabbMAX + root.aabbMIN) * 0.5f;
    S.sceneRadius = length(root.aabbMAX - S.sceneCenter) + 0.01f;
    S.sceneMin = root.aabbMIN;
    cudaMalloc(&S.scene_d, S.mesh.size() * sizeof(Triangle));
    cudaMalloc(&S.lights_d, (S.lightsvec.size() + 1) * sizeof(Triangle));
    cudaMalloc(&S.BVH_d, S.bvh.size() * sizeof(BVHnode));
    cudaMalloc(&S.BVHindices_d, S.ind.size() * sizeof(int));
    cudaMemcpy(S.scene_d, S.mesh.data(), S.mesh.size() * sizeof(Triangle), cudaMemcpyHostToDevice);
    cudaMemcpy(S.lights_d, S.lightsvec.data(), S.lightsvec.size() * sizeof(Triangle), cudaMemcpyHostToDevice);
    cudaMemcpy(S.BVH_d, S.bvh.data(), S.bvh.size() * sizeof(BVHnode), cudaMemcpyHostToDevice);
    cudaMemcpy(S.BVHindices_d, S.ind.data(), S.ind.size() * sizeof(int), cudaMemcpyHostToDevice);
    return 0;
}

