// novum_oracle_grid.cpp — CPU restatement of the photon hash grid, merge-radius schedule and the
// octahedral normal codec. TEST INFRASTRUCTURE ONLY (see novum_oracle.h).
#include "novum_oracle.h"
#include <math.h>
#include <string.h>
#include <vector>
#include <algorithm>
#include <numeric>

extern "C" {

// util.cuh:393-413
static bool isPrime(int n) {
    if (n <= 1) return false;
    if (n == 2 || n == 3) return true;
    if (n % 2 == 0 || n % 3 == 0) return false;
    for (int i = 5; (long long)i * i <= n; i += 6)
        if (n % i == 0 || n % (i + 2) == 0) return false;
    return true;
}
int oracle_next_prime(int n) {
    if (n <= 2) return 2;
    if (n % 2 == 0) n++;
    while (!isPrime(n)) n += 2;
    return n;
}

// util.cuh:415-418
float oracle_merge_radius(float initialRadius, float alpha, int currSample) {
    return initialRadius * sqrtf(1.0f / powf((float)(currSample + 1), alpha));
}

// integratorUtilities.cuh:382-388
void oracle_grid_index(float px, float py, float pz, const float sceneMin[3], float cell, int32_t out[3]) {
    out[0] = (int32_t)floorf((px - sceneMin[0]) / cell);
    out[1] = (int32_t)floorf((py - sceneMin[1]) / cell);
    out[2] = (int32_t)floorf((pz - sceneMin[2]) / cell);
}

// integratorUtilities.cuh:405-412 (unsigned wrap-around multiply, xor, mod table size)
uint32_t oracle_hash_grid_index(int32_t gx, int32_t gy, int32_t gz, int tableSize) {
    uint32_t n = (73856093u * (uint32_t)gx) ^ (19349663u * (uint32_t)gy) ^ (83492791u * (uint32_t)gz);
    return n % (uint32_t)tableSize;
}

// integratorUtilities.cuh:390-403 (signed multiply that wraps; same bits as the unsigned form)
uint32_t oracle_grid_hash(float px, float py, float pz, const float sceneMin[3], float mergeRadius, int tableSize) {
    int32_t g[3];
    oracle_grid_index(px, py, pz, sceneMin, mergeRadius, g);
    return oracle_hash_grid_index(g[0], g[1], g[2], tableSize);
}

// deviceCode.cu:3140-3198: hash, stable sort of (hash, index), cell_start / cell_end with 0xFFFFFFFF = empty
void oracle_hashgrid_build(const float* px, const float* py, const float* pz, int n, const float sceneMin[3],
                           float mergeRadius, int tableSize, uint32_t* sortedIndex, uint32_t* sortedHash,
                           uint32_t* cellStart, uint32_t* cellEnd) {
    std::vector<uint32_t> h(n), idx(n);
    for (int i = 0; i < n; i++) { h[i] = oracle_grid_hash(px[i], py[i], pz[i], sceneMin, mergeRadius, tableSize); idx[i] = (uint32_t)i; }
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return h[a] < h[b]; });
    for (int i = 0; i < n; i++) { sortedIndex[i] = idx[i]; sortedHash[i] = h[idx[i]]; }
    memset(cellStart, 0xFF, sizeof(uint32_t) * (size_t)tableSize);
    memset(cellEnd, 0xFF, sizeof(uint32_t) * (size_t)tableSize);
    for (int i = 0; i < n; i++) {   // deviceCode.cu:3113-3138
        uint32_t hh = sortedHash[i];
        if (i == 0 || sortedHash[i - 1] != hh) cellStart[hh] = (uint32_t)i;
        if (i == n - 1 || sortedHash[i + 1] != hh) cellEnd[hh] = (uint32_t)(i + 1);
    }
}

// deviceCode.cu:2968-2995: 27 neighbour cells, every photon of each cell distance-tested; two neighbour
// cells hashing to the same slot are both walked (photons counted twice), exactly like the reference.
void oracle_hashgrid_query_count(const float* spx, const float* spy, const float* spz, const uint32_t* cellStart,
                                 const uint32_t* cellEnd, int tableSize, const float sceneMin[3], float mergeRadius,
                                 const float* qx, const float* qy, const float* qz, int nq, uint32_t* counts) {
    float r2 = mergeRadius * mergeRadius;
    for (int q = 0; q < nq; q++) {
        int32_t c[3];
        oracle_grid_index(qx[q], qy[q], qz[q], sceneMin, mergeRadius, c);
        uint32_t cnt = 0;
        for (int z = -1; z <= 1; z++) for (int y = -1; y <= 1; y++) for (int x = -1; x <= 1; x++) {
            uint32_t h = oracle_hash_grid_index(c[0] + x, c[1] + y, c[2] + z, tableSize);
            uint32_t s = cellStart[h], e = cellEnd[h];
            if (s == 0xFFFFFFFFu) continue;
            for (int i = (int)s; i < (int)e; i++) {
                float dx = qx[q] - spx[i], dy = qy[q] - spy[i], dz = qz[q] - spz[i];
                float d2 = dx * dx + dy * dy + dz * dz;
                if (d2 <= r2) cnt++;
            }
        }
        counts[q] = cnt;
    }
}

// util.cuh:342-367
uint32_t oracle_pack_oct(float x, float y, float z) {
    float l1 = fabsf(x) + fabsf(y) + fabsf(z);
    float inv = (l1 > 0.0f) ? (1.0f / l1) : 0.0f;
    float rx = x * inv, ry = y * inv;
    if (z < 0.0f) {
        float tx = rx, ty = ry;
        rx = (1.0f - fabsf(ty)) * ((tx >= 0.0f) ? 1.0f : -1.0f);
        ry = (1.0f - fabsf(tx)) * ((ty >= 0.0f) ? 1.0f : -1.0f);
    }
    int16_t qx = (int16_t)(int)nearbyintf(fminf(fmaxf(rx, -1.0f), 1.0f) * 32767.0f);
    int16_t qy = (int16_t)(int)nearbyintf(fminf(fmaxf(ry, -1.0f), 1.0f) * 32767.0f);
    return ((uint32_t)(uint16_t)qx) | (((uint32_t)(uint16_t)qy) << 16);
}

// util.cuh:369-391 (rsqrtf replaced by 1/sqrtf: last-bit differences only)
void oracle_unpack_oct(uint32_t p, float out[3]) {
    int16_t px = (int16_t)(p & 0xFFFF), py = (int16_t)(p >> 16);
    float x = fmaxf(-1.0f, px / 32767.0f), y = fmaxf(-1.0f, py / 32767.0f);
    float z = 1.0f - (fabsf(x) + fabsf(y));
    float t = fmaxf(-z, 0.0f);
    x += (x >= 0.0f) ? -t : t;
    y += (y >= 0.0f) ? -t : t;
    float il = 1.0f / sqrtf(x * x + y * y + z * z);
    out[0] = x * il; out[1] = y * il; out[2] = z * il;
}

}  // extern "C"
