// trav_stream.cuh — persistent-warp ray streaming with per-lane refill.
//
// The first version of the traversal kernels fetched 32 rays per warp and ran them all to completion: on the mixed queues of
// a path tracer (primary rays that leave after 3 nodes next to glass-interior rays that visit 60) only 7 of 32 lanes were
// active per issued instruction (profiles/r1_wfExtend_steady.txt). Here every lane carries a resumable traversal
// (ClosestTrav / ShadowTrav); whenever fewer than kRefillBelow lanes are still busy the warp stops stepping, finished lanes
// pull new ray indices from the global work counter with ONE aggregated atomicAdd, and stepping resumes — the "while-while
// with dynamic fetch" scheme, on a compressed BVH8 whose step is one node (8 boxes) + one leaf group.
#pragma once
#include "scene_dev.cuh"

namespace nv {

constexpr int kRefillBelow = 22;      // refill as soon as this few lanes are busy (while work remains)

// load(i, o, d) reads ray i; commit(i, hit) consumes its result (called from divergent code); sync(final) is called with the
// whole warp converged before every refill and once at the end. Returns the number of rays this lane traced.
template <bool COUNT, class Load, class Commit, class Sync>
NV_D unsigned streamClosest(const Bvh8View& bvh, TravStack& ts, int n, int* fetchCounter, TravCounters cnt, Load load, Commit commit, Sync sync) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    ClosestTrav<COUNT> tr;
    bool busy = false, exhausted = false;
    int idx = 0;
    unsigned traced = 0;
    while (true) {
        sync(false);
        if (!exhausted) {
            unsigned need = __ballot_sync(0xffffffffu, !busy);
            if (need) {
                int base = 0, want = __popc(need);
                if (lane == (__ffs(need) - 1)) base = atomicAdd(fetchCounter, want);
                base = __shfl_sync(0xffffffffu, base, __ffs(need) - 1);
                if (!busy) {
                    idx = base + __popc(need & lt);
                    if (idx < n) {
                        V3 o, d;
                        load(idx, o, d);
                        tr.begin(o, d, 999999.0f, -1, ts);
                        busy = true; ++traced;
                    }
                }
                exhausted = base + want >= n;
            }
        }
        unsigned live = __ballot_sync(0xffffffffu, busy);
        if (live == 0) break;
        const int floorBusy = exhausted ? 1 : kRefillBelow;
        do {
            if (busy && tr.step(bvh, ts)) {
                commit(idx, tr.best);
                tr.flushCounters(cnt);
                busy = false;
            }
            live = __ballot_sync(0xffffffffu, busy);
        } while (__popc(live) >= floorBusy);
    }
    sync(true);
    return traced;
}

// load(i, o, d, maxT, skipTri); commit(i, throughput)
template <bool COUNT, class Load, class Commit>
NV_D unsigned streamShadow(const SceneView& sc, TravStack& ts, int n, int* fetchCounter, TravCounters cnt, Load load, Commit commit) {
    const int lane = threadIdx.x & 31;
    const unsigned lt = (1u << lane) - 1u;
    ShadowTrav<COUNT> tr;
    bool busy = false, exhausted = false;
    int idx = 0;
    unsigned traced = 0;
    while (true) {
        if (!exhausted) {
            unsigned need = __ballot_sync(0xffffffffu, !busy);
            if (need) {
                int base = 0, want = __popc(need);
                if (lane == (__ffs(need) - 1)) base = atomicAdd(fetchCounter, want);
                base = __shfl_sync(0xffffffffu, base, __ffs(need) - 1);
                if (!busy) {
                    idx = base + __popc(need & lt);
                    if (idx < n) {
                        V3 o, d; float maxT; int skip;
                        load(idx, o, d, maxT, skip);
                        tr.begin(o, d, maxT, skip, ts);
                        busy = true; ++traced;
                    }
                }
                exhausted = base + want >= n;
            }
        }
        unsigned live = __ballot_sync(0xffffffffu, busy);
        if (live == 0) break;
        const int floorBusy = exhausted ? 1 : kRefillBelow;
        do {
            if (busy && tr.step(sc, ts)) {
                commit(idx, tr.thr);
                tr.flushCounters(cnt);
                busy = false;
            }
            live = __ballot_sync(0xffffffffu, busy);
        } while (__popc(live) >= floorBusy);
    }
    return traced;
}

}  // namespace nv
