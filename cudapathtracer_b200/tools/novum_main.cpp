// novum_main.cpp — command-line driver mirroring the reference's main()/initRender() (main.cu:235-933):
// config -> scene -> BVH -> integrator -> image. Differences: config path / spp / seed / output come from argv and
// one render is made instead of the hard-coded 75 (main.cu:925-933).
#include "novum_b200.h"
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

int main(int argc, char** argv) {
    std::string config = "configs/config.rendertron", out = "render", gen, checkpoint;
    int every = 16;
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
        else if (a == "--checkpoint") checkpoint = next();
        else if (a == "--every") every = atoi(next().c_str());
        else if (a == "--help") { printf("novum [--config f.rendertron] [--spp N] [--seed S] [--out name] [--raw] [--checkpoint f.novumacc [--every N]] [--make-scene name --out dir --detail D]\n"); return 0; }
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
    int finalCount = cfg.sampleCount;
    if (checkpoint.empty()) {
        rc = novum_render(scene, &cfg, &cam, 0, cfg.sampleCount, shard, seed, img.data(), &st);
        if (rc) { fprintf(stderr, "render failed: %s\n", novum_error_string(rc)); return 1; }
    } else {
        // resumable run: the sums so far are loaded if the checkpoint exists, then `every` passes at a time are rendered, added and
        // written back (with a progressive .bmp beside them); killing the process loses at most one chunk
        int done = 0, cw = 0, ch = 0, cinteg = -1; uint64_t cseed = seed;
        memset(img.data(), 0, img.size() * sizeof(NovumFloat4));
        if (novum_accum_info(checkpoint.c_str(), &cw, &ch, &done, &cseed, &cinteg) == 0) {
            if (cinteg != cfg.integrator) {          // sums of another integrator must not be continued with this one
                fprintf(stderr, "checkpoint %s was written by integrator %d, this render uses %d: refusing to resume\n", checkpoint.c_str(), cinteg, cfg.integrator);
                return 1;
            }
            rc = novum_accum_load(checkpoint.c_str(), img.data(), cfg.width, cfg.height, &done, &cseed);
            if (rc) { fprintf(stderr, "checkpoint %s does not fit this render: %s\n", checkpoint.c_str(), novum_error_string(rc)); return 1; }
            seed = cseed;
            printf("resuming at sample %d of %d (seed %llu)\n", done, cfg.sampleCount, (unsigned long long)seed);
        }
        if (every <= 0) every = 16;
        finalCount = std::max(done, cfg.sampleCount);          // a checkpoint that already holds more passes than asked for is normalised by what it holds
        std::vector<NovumFloat4> part(img.size());
        memset(&st, 0, sizeof(st));
        while (done < cfg.sampleCount) {
            const int n = std::min(every, cfg.sampleCount - done);
            NovumRenderStats ps;
            rc = novum_render(scene, &cfg, &cam, done, n, shard, seed, part.data(), &ps);
            if (rc) { fprintf(stderr, "render failed: %s\n", novum_error_string(rc)); return 1; }
            for (size_t i = 0; i < img.size(); ++i) { img[i].x += part[i].x; img[i].y += part[i].y; img[i].z += part[i].z; img[i].w += part[i].w; }
            done += n;
            st.paths += ps.paths; st.closestRays += ps.closestRays; st.shadowRays += ps.shadowRays; st.deviceMs += ps.deviceMs; st.kernelLaunches += ps.kernelLaunches;
            if ((rc = novum_accum_save(checkpoint.c_str(), img.data(), cfg.width, cfg.height, done, seed, cfg.integrator)))
                fprintf(stderr, "checkpoint write failed: %s\n", novum_error_string(rc));
            // cleanAndFormatImage's convention: currentSampleCount = passes done - 1 (deviceCode.cu:1905-1927)
            novum_save_bmp((out + "_progress.bmp").c_str(), img.data(), nullptr, cfg.width, cfg.height, done - 1, NOVUM_FRAME_PROGRESSIVE, cfg.postProcess);
        }
    }
    double rays = (double)st.closestRays + (double)st.shadowRays;
    printf("{\"paths\": %llu, \"closest_rays\": %llu, \"shadow_rays\": %llu, \"device_ms\": %.3f, \"mrays_per_s\": %.1f, \"launches\": %llu}\n",
           (unsigned long long)st.paths, (unsigned long long)st.closestRays, (unsigned long long)st.shadowRays, st.deviceMs,
           rays / (st.deviceMs * 1e3), (unsigned long long)st.kernelLaunches);
    // main.cu:855-907: / sampleCount, NaN / Inf marking, tonemap + gamma + BGR rows — one device kernel, 3 B/pixel read back
    rc = novum_save_bmp((out + ".bmp").c_str(), img.data(), nullptr, cfg.width, cfg.height, finalCount, NOVUM_FRAME_FINAL, cfg.postProcess);
    if (rc) fprintf(stderr, "bmp write failed: %s\n", novum_error_string(rc));
    if (raw) {
        FILE* f = fopen((out + ".raw").c_str(), "wb");
        if (f) { fwrite(img.data(), sizeof(NovumFloat4), img.size(), f); fclose(f); }
    }
    novum_scene_destroy(scene);
    return 0;
}
