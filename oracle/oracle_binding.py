"""ctypes binding of oracle/liboracle.so — the CPU restatement of the reference's hot path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs. The product package never imports this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

TRI_DTYPE = np.dtype({"names": ["aInd", "bInd", "cInd", "naInd", "nbInd", "ncInd", "uvaInd", "uvbInd", "uvcInd", "materialID",
                                "emission", "lightInd", "triInd"],
                      "formats": ["<i4"] * 10 + [("<f4", 4), "<i4", "<i4"],
                      "offsets": [0, 4, 8, 12, 16, 20, 24, 28, 32, 36, 48, 64, 68], "itemsize": 80})
HIT_DTYPE = np.dtype([("tri", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])
BVH_DTYPE = np.dtype([("mn", "<f4", 4), ("mx", "<f4", 4), ("left", "<i4"), ("right", "<i4"), ("first", "<i4"), ("primCount", "<i4")])


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])


def lib():
    global _LIB
    if _LIB is None:
        p = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(p):
            build()
        L = C.CDLL(p)
        vp, ci, cf = C.c_void_p, C.c_int, C.c_float
        L.oracle_scene_create.restype = vp
        L.oracle_scene_create.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, ci]
        L.oracle_scene_destroy.argtypes = [vp]
        L.oracle_scene_bvh_size.argtypes = [vp, C.POINTER(ci), C.POINTER(ci)]
        L.oracle_scene_bvh_copy.argtypes = [vp, vp, vp]
        L.oracle_scene_bounds.argtypes = [vp, C.POINTER(cf * 3), C.POINTER(cf * 3), C.POINTER(cf)]
        for f in ("oracle_trace_closest", "oracle_trace_closest_lowest_id", "oracle_trace_brute"):
            getattr(L, f).argtypes = [vp, vp, ci, vp, ci]
        L.oracle_trace_shadow.argtypes = [vp, vp, vp, vp, ci, vp, ci]
        L.oracle_next_prime.argtypes = [ci]
        L.oracle_merge_radius.restype = cf
        L.oracle_merge_radius.argtypes = [cf, cf, ci]
        L.oracle_grid_hash.restype = C.c_uint32
        L.oracle_grid_hash.argtypes = [cf, cf, cf, C.POINTER(cf * 3), cf, ci]
        L.oracle_hashgrid_build.argtypes = [vp, vp, vp, ci, C.POINTER(cf * 3), cf, ci, vp, vp, vp, vp]
        L.oracle_hashgrid_query_count.argtypes = [vp, vp, vp, vp, vp, ci, C.POINTER(cf * 3), cf, vp, vp, vp, ci, vp]
        L.oracle_pack_oct.restype = C.c_uint32
        L.oracle_pack_oct.argtypes = [cf, cf, cf]
        L.oracle_unpack_oct.argtypes = [C.c_uint32, C.POINTER(cf * 3)]
        L.oracle_bsdf_eval.argtypes = [vp, ci, vp, C.POINTER(cf * 3), C.POINTER(cf * 3), cf, cf, cf, C.POINTER(cf * 3)]
        L.oracle_bsdf_pdf.restype = cf
        L.oracle_bsdf_pdf.argtypes = [vp, ci, vp, C.POINTER(cf * 3), C.POINTER(cf * 3), cf, cf, cf]
        L.oracle_bsdf_sample.argtypes = [vp, ci, vp, C.POINTER(cf * 3), cf, ci, ci, cf, cf, vp, C.POINTER(cf * 3), C.POINTER(cf * 3), C.POINTER(cf)]
        L.oracle_render_unidirectional.argtypes = [vp, vp, ci, ci, ci, ci, ci, C.c_uint64, vp, ci, vp]
        L.oracle_triangle_intersect.argtypes = [vp, vp, vp, C.POINTER(cf), C.POINTER(cf), C.POINTER(cf)]
        L.oracle_clean_and_format.argtypes = [vp, vp, vp, ci, ci, ci]
        L.oracle_final_normalise.argtypes = [vp, vp, vp, ci, ci, ci]
        L.oracle_bmp_rows.argtypes = [vp, ci, ci, ci, vp]
        L.oracle_gamma_table_check.restype = C.c_long
        L.oracle_gamma_table_check.argtypes = [vp, C.c_uint32, C.c_uint32]
        _LIB = L
    return _LIB


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleScene:
    def __init__(self, positions, normals, uvs, tris, mats, textures=None, max_leaf=2):
        self.pos = np.ascontiguousarray(positions, np.float32).reshape(-1, 4)
        self.nrm = np.ascontiguousarray(normals, np.float32).reshape(-1, 4)
        self.uv = np.ascontiguousarray(uvs, np.float32).reshape(-1, 2)
        self.tris = np.ascontiguousarray(tris)
        assert self.tris.dtype.itemsize == 80
        self.mats = np.ascontiguousarray(mats)
        assert self.mats.dtype.itemsize == 176
        self.tex = None if textures is None else np.ascontiguousarray(textures, np.float32).reshape(-1, 4)
        ntex = 0 if self.tex is None else len(self.tex)
        self.h = lib().oracle_scene_create(_p(self.pos), len(self.pos), _p(self.nrm), len(self.nrm), _p(self.uv), len(self.uv),
                                           _p(self.tris), len(self.tris), _p(self.mats), len(self.mats), _p(self.tex), ntex, max_leaf)

    def close(self):
        if self.h:
            lib().oracle_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def bounds(self):
        mn, mx, r = (C.c_float * 3)(), (C.c_float * 3)(), C.c_float()
        lib().oracle_scene_bounds(self.h, C.byref(mn), C.byref(mx), C.byref(r))
        return list(mn), list(mx), r.value

    def bvh(self):
        n, l = C.c_int(), C.c_int()
        lib().oracle_scene_bvh_size(self.h, C.byref(n), C.byref(l))
        nodes = np.zeros(n.value, BVH_DTYPE); idx = np.zeros(len(self.tris), np.int32)
        lib().oracle_scene_bvh_copy(self.h, _p(nodes), _p(idx))
        return nodes, idx, l.value

    def _trace(self, fn, rays, threads):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), HIT_DTYPE)
        getattr(lib(), fn)(self.h, _p(rays), len(rays), _p(hits), threads)
        return hits

    def trace_closest(self, rays, threads=0):
        return self._trace("oracle_trace_closest", rays, threads)

    def trace_closest_lowest_id(self, rays, threads=0):
        return self._trace("oracle_trace_closest_lowest_id", rays, threads)

    def trace_brute(self, rays, threads=0):
        return self._trace("oracle_trace_brute", rays, threads)

    def trace_shadow(self, rays, max_t, skip=None, threads=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        max_t = np.ascontiguousarray(max_t, np.float32)
        skip = None if skip is None else np.ascontiguousarray(skip, np.int32)
        out = np.zeros((len(rays), 4), np.float32)
        lib().oracle_trace_shadow(self.h, _p(rays), _p(max_t), _p(skip), len(rays), _p(out), threads)
        return out

    def render_unidirectional(self, cam_bytes, w, h, max_depth, first_sample, num_sample, seed=103033, use_mis=True, naive=False,
                              threads=0):
        colors = np.zeros((h, w, 4), np.float32)
        counts = np.zeros(2, np.uint64)
        cam = np.frombuffer(cam_bytes, np.uint8).copy()
        lib().oracle_render_unidirectional(self.h, _p(cam), max_depth, int(use_mis), int(naive), first_sample, num_sample, seed,
                                           _p(colors), threads, _p(counts))
        return colors, counts


def next_prime(n):
    return lib().oracle_next_prime(n)


def merge_radius(r0, alpha, i):
    return lib().oracle_merge_radius(r0, alpha, i)


def hashgrid_build(px, py, pz, scene_min, radius, table_size):
    px, py, pz = [np.ascontiguousarray(a, np.float32) for a in (px, py, pz)]
    n = len(px)
    idx = np.zeros(n, np.uint32); hsh = np.zeros(n, np.uint32)
    cs = np.zeros(table_size, np.uint32); ce = np.zeros(table_size, np.uint32)
    mn = (C.c_float * 3)(*scene_min)
    lib().oracle_hashgrid_build(_p(px), _p(py), _p(pz), n, C.byref(mn), radius, table_size, _p(idx), _p(hsh), _p(cs), _p(ce))
    return idx, hsh, cs, ce


def hashgrid_query_count(spx, spy, spz, cs, ce, table_size, scene_min, radius, qx, qy, qz):
    arrs = [np.ascontiguousarray(a, np.float32) for a in (spx, spy, spz, qx, qy, qz)]
    out = np.zeros(len(arrs[3]), np.uint32)
    mn = (C.c_float * 3)(*scene_min)
    lib().oracle_hashgrid_query_count(_p(arrs[0]), _p(arrs[1]), _p(arrs[2]), _p(cs), _p(ce), table_size, C.byref(mn), radius,
                                      _p(arrs[3]), _p(arrs[4]), _p(arrs[5]), len(out), _p(out))
    return out


def pack_oct(v):
    return lib().oracle_pack_oct(float(v[0]), float(v[1]), float(v[2]))


def unpack_oct(p):
    out = (C.c_float * 3)()
    lib().oracle_unpack_oct(int(p), C.byref(out))
    return np.array(list(out), np.float32)


def frame_clean(acc, overlay, n, final_mode=False):
    """cleanAndFormatImage (integratorUtilities.cuh:337-380; n = currentSampleCount) or, final_mode, main.cu:855-884 (n = sampleCount)."""
    acc = np.ascontiguousarray(acc, dtype=np.float32)
    h, w = acc.shape[:2]
    ov = None if overlay is None else np.ascontiguousarray(overlay, dtype=np.float32)
    out = np.empty((h, w, 4), dtype=np.float32)
    f = lib().oracle_final_normalise if final_mode else lib().oracle_clean_and_format
    f(_p(acc), None if ov is None else _p(ov), _p(out), w, h, int(n))
    return out


def bmp_rows(frame, post_process):
    """Image::saveImageBMP's rows (imageUtil.cu:69-100, 202-232) of a cleaned float4 frame: (h, (3w+3)&~3) uint8, B G R."""
    frame = np.ascontiguousarray(frame, dtype=np.float32)
    h, w = frame.shape[:2]
    rows = np.empty((h, (3 * w + 3) & ~3), dtype=np.uint8)
    lib().oracle_bmp_rows(_p(frame), w, h, int(bool(post_process)), _p(rows))
    return rows


def frame_finish(acc, overlay, n, final_mode, post_process):
    return bmp_rows(frame_clean(acc, overlay, n, final_mode), post_process)


def gamma_table_check(steps, first_bits=0, last_bits=0x3f800000):
    steps = np.ascontiguousarray(steps, dtype=np.float32)
    return int(lib().oracle_gamma_table_check(_p(steps), first_bits, last_bits))
