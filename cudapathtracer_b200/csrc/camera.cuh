// camera.cuh — device-side camera: ray generation and world -> raster projection with the reference's arithmetic
// (objects.cuh:268-307 generateCameraRay, :343-374 worldToRaster).
#pragma once
#include "vecmath.cuh"
#include "../../include/novum_b200.h"

namespace nv {

// Camera::generateCameraRay (objects.cuh:268-307)
NV_D void cameraRay(const NovumCamera& c, Rng& rng, int x, int y, V3& o, V3& d) {
    float aspect = (float)c.w / (float)c.h;
    float jx = (rng.uniform() - 0.5f) * c.antiAliasJitterDist;
    float jy = (rng.uniform() - 0.5f) * c.antiAliasJitterDist;
    float u = (2.0f * ((x + jx) / (float)c.w) - 1.0f) * aspect * c.fovScale;
    float v = (2.0f * ((y + jy) / (float)c.h) - 1.0f) * c.fovScale;
    V3 org(c.cameraOrigin.x, c.cameraOrigin.y, c.cameraOrigin.z);
    V3 right(c.right.x, c.right.y, c.right.z), up(c.up.x, c.up.y, c.up.z), fwd(c.forward.x, c.forward.y, c.forward.z);
    V3 focal = org + (right * (u * c.focalDist)) + (up * (v * c.focalDist)) + (fwd * c.focalDist);
    V3 lens(0.0f);
    if (c.aperture > 0.0f) {
        float rr = rng.uniform();
        float theta = 2.0f * 3.141592f * rng.uniform();
        float radius = c.aperture * sqrtf(rr);
        float s, co; sincosf(theta, &s, &co);
        lens = (right * (radius * co)) + (up * (radius * s));
    }
    o = org + lens;
    d = normalize(focal - o);
}

// Camera::worldToRaster (objects.cuh:343-374). Returns false when the point is behind the camera or outside the frustum.
NV_D bool worldToRaster(const NovumCamera& c, V3 p, float& rx, float& ry) {
    V3 dir = p - V3(c.cameraOrigin.x, c.cameraOrigin.y, c.cameraOrigin.z);
    V3 fwd(c.forward.x, c.forward.y, c.forward.z), right(c.right.x, c.right.y, c.right.z), up(c.up.x, c.up.y, c.up.z);
    float dz = dot(dir, fwd);
    if (dz <= 0.001f) return false;
    float sx = dot(dir, right) / dz, sy = dot(dir, up) / dz;
    float aspect = (float)c.w / (float)c.h;
    float nx = sx / (aspect * c.fovScale), ny = sy / c.fovScale;
    if (nx < -1.0f || nx > 1.0f || ny < -1.0f || ny > 1.0f) return false;
    rx = (nx + 1.0f) * 0.5f * (float)c.w;
    ry = (ny + 1.0f) * 0.5f * (float)c.h;
    return true;
}

}  // namespace nv
