// novum_main.cpp — command-line driver mirroring the reference's main()/initRender() (main.cu:235-933):
// config -> scene -> BVH -> integrator -> image. Differences: config path / spp / seed / output come from argv and
// one render is made instead of the hard-coded 75 (main.cu:925-933).
#include "novum_b200.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char** argv) {
    std::string config = "configs/config.rendertron", out = "render", gen;
    int spp = -1, renderNumber = 0, detail = 100; unsigned long long seed = 103033ull; bool raw = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto next = [&]() { return i + 1 < argc ? std::string(argv[++i]) : std::string(); };
        if (a == "--config") config = next();
        else if (a == "--spp") spp = atoi(next().c_str());
        else if (a == "--seed") seed = strtoull(next().c_str(), nullptr, 10);
        else if (a == "--out") out = next();
        else if (a == "--render-number") renderNumber = atoi(next().c_str());
        else if (a == "--raw") raw = true;
        else if (a == "--make-scene") gen = next();
        else if (a == "--detail") detail = atoi(next().c_str());
        else if (a == "--help") { printf("novum [--config f.rendertron] [--spp N] [--seed S] [--out name] [--raw] [--make-scene name --out dir --detail D]\n"); return 0; }
    }
    if (!gen.empty()) {
        int rc = novum_make_scene(gen.c_str(), out.c_str(), detail);
        if (rc) fprintf(stderr, "make-scene failed: %s\n", novum_error_string(rc));
        return rc ? 1 : 0;
    }
    NovumScene* scene = nullptr; NovumRenderConfig cfg;
    int rc = novum_scene_load_config(config.c_str(), renderNumber, &scene, &cfg);
    if (rc) { fprintf(stderr, "scene load failed: %s\n", novum_error_string(rc)); return 1; }
    if (spp > 0) cfg.sampleCount = spp;
    NovumCamera cam; novum_camera_make(&cfg, &cam);
    int nTri, nLights, nNodes; float mn[3], mx[3], radius;
    novum_scene_info(scene, &nTri, &nLights, &nNodes, mn, mx, &radius);
    printf("scene: %d triangles, %d lights, %d BVH8 nodes, radius %g\n", nTri, nLights, nNodes, radius);
    std::vector<NovumFloat4> img((size_t)cfg.width * cfg.height);
    NovumRenderStats st; NovumShard shard = {0, 1};
    rc = novum_render(scene, &cfg, &cam, 0, cfg.sampleCount, shard, seed, img.data(), &st);
    if (rc) { fprintf(stderr, "render failed: %s\n", novum_error_string(rc)); return 1; }
    double rays = (double)st.closestRays + (double)st.shadowRays;
    printf("{\"paths\": %llu, \"closest_rays\": %llu, \"shadow_rays\": %llu, \"device_ms\": %.3f, \"mrays_per_s\": %.1f, \"launches\": %llu}\n",
           (unsigned long long)st.paths, (unsigned long long)st.closestRays, (unsigned long long)st.shadowRays, st.deviceMs,
           rays / (st.deviceMs * 1e3), (unsigned long long)st.kernelLaunches);
    // main.cu:855-907: / sampleCount, NaN / Inf marking, tonemap + gamma + BGR rows — one device kernel, 3 B/pixel read back
    rc = novum_save_bmp((out + ".bmp").c_str(), img.data(), nullptr, cfg.width, cfg.height, cfg.sampleCount, NOVUM_FRAME_FINAL, cfg.postProcess);
    if (rc) fprintf(stderr, "bmp write failed: %s\n", novum_error_string(rc));
    if (raw) {
        FILE* f = fopen((out + ".raw").c_str(), "wb");
        if (f) { fwrite(img.data(), sizeof(NovumFloat4), img.size(), f); fclose(f); }
    }
    novum_scene_destroy(scene);
    return 0;
}
