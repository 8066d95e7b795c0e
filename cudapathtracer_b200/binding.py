"""ctypes mirror of include/novum_b200.h. Every call goes through the C ABI; nothing here computes."""
import ctypes as C
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


class NovumError(RuntimeError):
    pass


def lib_path():
    return os.path.join(_HERE, "libnovum_b200.so")


def lib():
    """Loads the CUDA library. Fails loudly when it has not been built (python __graft_entry__.py build)."""
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not os.path.exists(p):
            raise NovumError(f"{p} is missing: build it with `make -C cudapathtracer_b200` (no CPU fallback exists)")
        _LIB = C.CDLL(p)
        _declare(_LIB)
    return _LIB


class F4(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("w", C.c_float)]


class Camera(C.Structure):
    _fields_ = [("cameraOrigin", F4), ("w", C.c_int32), ("h", C.c_int32), ("xRot", C.c_float), ("yRot", C.c_float),
                ("zRot", C.c_float), ("aperture", C.c_float), ("focalDist", C.c_float), ("fovScale", C.c_float),
                ("antiAliasJitterDist", C.c_float), ("_pad", C.c_int32 * 3), ("forward", F4), ("right", F4), ("up", F4)]


class RenderConfig(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("integrator", C.c_int32), ("sampleCount", C.c_int32),
                ("maxDepth", C.c_int32), ("bvhLeafSize", C.c_int32), ("sampleEnvironment", C.c_int32),
                ("postProcess", C.c_int32), ("bdptEyeDepth", C.c_int32), ("bdptLightDepth", C.c_int32),
                ("bdptLightTrace", C.c_int32), ("bdptNee", C.c_int32), ("bdptNaive", C.c_int32),
                ("bdptConnection", C.c_int32), ("bdptDrawPath", C.c_int32), ("bdptDoMis", C.c_int32),
                ("bdptPaintWeight", C.c_int32), ("vcmDoMerge", C.c_int32), ("doSPPM", C.c_int32),
                ("vcmMergeConst", C.c_float), ("vcmInitialMergeRadiusMultiplier", C.c_float),
                ("pinholeCamera", C.c_int32), ("camPos", C.c_float * 3), ("camRot", C.c_float * 3),
                ("camFov", C.c_float), ("camApeture", C.c_float), ("camFocalDist", C.c_float),
                ("numMeshes", C.c_int32), ("name", C.c_char * 64)]


class Constants(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("sampleEnvironment", "bdptLightTrace", "bdptNee", "bdptNaive", "bdptConnection",
                                         "vcmDoMerge", "doSPPM", "bdptDrawPath", "bdptDoMis", "bdptPaintWeight")]


class LaunchCacheStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("launches", "hits", "rebinds", "sceneBuilds", "enginesMoved", "progressiveSaves")]


class Shard(C.Structure):
    _fields_ = [("rank", C.c_int32), ("world", C.c_int32)]


class RenderStats(C.Structure):
    _fields_ = [("closestRays", C.c_uint64), ("shadowRays", C.c_uint64), ("paths", C.c_uint64), ("photons", C.c_uint64),
                ("lightVertices", C.c_uint64), ("kernelLaunches", C.c_uint64), ("deviceMs", C.c_float), ("_pad", C.c_float),
                ("bvhNodeVisits", C.c_uint64), ("triTests", C.c_uint64), ("stageMs", C.c_float * 4), ("stageLaunches", C.c_uint64 * 4),
                ("mergeQueries", C.c_uint64), ("mergeCandidates", C.c_uint64), ("mergeAccepted", C.c_uint64)]


# numpy dtypes with the reference's struct layouts (objects.cuh Triangle 80 B, Material 176 B)
TRI_DTYPE = np.dtype({"names": ["aInd", "bInd", "cInd", "naInd", "nbInd", "ncInd", "uvaInd", "uvbInd", "uvcInd", "materialID",
                                "emission", "lightInd", "triInd"],
                      "formats": ["<i4"] * 10 + [("<f4", 4), "<i4", "<i4"],
                      "offsets": [0, 4, 8, 12, 16, 20, 24, 28, 32, 36, 48, 64, 68], "itemsize": 80})
MAT_DTYPE = np.dtype({"names": ["hasTexture", "startInd", "width", "height", "hasTransMap", "tstartInd", "twidth", "theight", "type",
                                "albedo", "roughness", "eta", "k", "ior", "metallic", "specular", "transmission", "isSpecular",
                                "boundary", "thinWalled", "absorption", "priority"],
                      "formats": ["u1", "<i4", "<i4", "<i4", "u1", "<i4", "<i4", "<i4", "<i4", ("<f4", 4), "<f4", ("<f4", 4),
                                  ("<f4", 4), "<f4", "<f4", "<f4", "<f4", "u1", "u1", "u1", ("<f4", 4), "<i4"],
                      "offsets": [0, 4, 8, 12, 16, 20, 24, 28, 32, 48, 64, 80, 96, 112, 116, 120, 124, 128, 129, 130, 144, 160],
                      "itemsize": 176})
HIT_DTYPE = np.dtype([("tri", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])


def _header_symbols():
    hdr = os.path.join(_HERE, "..", "include", "novum_b200.h")
    with open(hdr) as f:
        text = f.read()
    return sorted(set(re.findall(r"\b(novum_[a-z0-9_]+)\s*\(", text)))


DECLARED_SYMBOLS = _header_symbols()


def _declare(L):
    vp, ci, cf = C.c_void_p, C.c_int, C.c_float
    L.novum_error_string.restype = C.c_char_p
    L.novum_error_string.argtypes = [ci]
    L.novum_scene_destroy.restype = None
    L.novum_hashgrid_destroy.restype = None
    L.novum_vcm_destroy.restype = None
    L.novum_scene_load_config.argtypes = [C.c_char_p, ci, C.POINTER(vp), C.POINTER(RenderConfig)]
    L.novum_config_load.argtypes = [C.c_char_p, C.POINTER(RenderConfig)]
    L.novum_scene_create.argtypes = [vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, vp, ci, C.POINTER(vp)]
    L.novum_scene_destroy.argtypes = [vp]
    L.novum_scene_info.argtypes = [vp, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci), C.POINTER(cf * 3), C.POINTER(cf * 3), C.POINTER(cf)]
    L.novum_scene_counts.argtypes = [vp] + [C.POINTER(ci)] * 5
    L.novum_scene_download.argtypes = [vp, vp, vp, vp, vp, vp]
    L.novum_scene_timings.argtypes = [vp, C.POINTER(cf), C.POINTER(cf), C.POINTER(ci)]
    L.novum_scene_bvh_depth.argtypes = [vp]
    L.novum_scene_set_features.argtypes = [vp, ci]
    L.novum_scene_set_environment.argtypes = [vp, C.POINTER(cf * 3)]
    L.novum_parse_double.restype = C.c_double
    L.novum_parse_double.argtypes = [C.c_char_p, C.POINTER(ci)]
    L.novum_material_table.argtypes = [vp]
    L.novum_camera_make.argtypes = [C.POINTER(RenderConfig), C.POINTER(Camera)]
    L.novum_trace_closest.argtypes = [vp, vp, ci, vp]
    L.novum_trace_shadow.argtypes = [vp, vp, vp, vp, ci, vp]
    L.novum_trace_closest_device.argtypes = [vp, vp, ci, vp, ci]
    L.novum_trace_shadow_device.argtypes = [vp, vp, vp, vp, ci, vp, ci]
    L.novum_trace_last_ms.argtypes = [vp, C.POINTER(cf)]
    L.novum_trace_counters.argtypes = [vp, ci, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.novum_set_profile.argtypes = [vp, ci]
    L.novum_update_constants.argtypes = [C.POINTER(Constants)]
    L.novum_render.argtypes = [vp, C.POINTER(RenderConfig), C.POINTER(Camera), ci, ci, Shard, C.c_uint64, vp, C.POINTER(RenderStats)]
    L.novum_render_device.argtypes = [vp, C.POINTER(RenderConfig), C.POINTER(Camera), ci, ci, Shard, C.c_uint64, vp, C.POINTER(RenderStats)]
    L.novum_make_scene.argtypes = [C.c_char_p, C.c_char_p, ci]
    L.novum_save_bmp.argtypes = [C.c_char_p, vp, vp, ci, ci, ci, ci, ci]
    L.novum_bmp_row_bytes.argtypes = [ci]
    L.novum_frame_finish_device.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp, vp]
    L.novum_frame_finish.argtypes = [vp, vp, ci, ci, ci, ci, ci, vp, vp]
    L.novum_frame_save_bmp.argtypes = [C.c_char_p, vp, vp, ci, ci, ci, ci, ci]
    L.novum_frame_save_bmp_async.argtypes = [C.c_char_p, vp, vp, ci, ci, ci, ci, ci, vp]
    L.novum_write_bmp.argtypes = [C.c_char_p, vp, ci, ci]
    L.novum_frame_last_ms.argtypes = [C.POINTER(cf)]
    L.novum_frame_gamma_steps.argtypes = [vp]
    L.novum_save_pfm.argtypes = [C.c_char_p, vp, ci, ci, cf]
    L.novum_obj_parse.argtypes = [C.c_char_p, C.POINTER(cf * 3), ci, cf, ci, ci, C.POINTER(vp)]
    L.novum_obj_free.argtypes = [vp]
    L.novum_obj_free.restype = None
    L.novum_obj_counts.argtypes = [vp, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci), C.POINTER(ci), C.POINTER(ci)]
    L.novum_obj_download.argtypes = [vp, vp, vp, vp, vp]
    L.novum_obj_device_create.argtypes = [C.POINTER(vp)]
    L.novum_obj_device_destroy.argtypes = [vp]
    L.novum_obj_device_destroy.restype = None
    L.novum_obj_device_append_file.argtypes = [vp, C.c_char_p, C.POINTER(cf * 3), ci, cf]
    L.novum_obj_device_append_text.argtypes = [vp, C.c_char_p, C.c_size_t, C.POINTER(cf * 3), ci, cf]
    L.novum_obj_device_append_device_text.argtypes = [vp, vp, C.c_size_t, C.POINTER(cf * 3), ci, cf]
    L.novum_obj_device_reset.argtypes = [vp]
    L.novum_obj_device_counts.argtypes = [vp] + [C.POINTER(ci)] * 5
    L.novum_obj_device_arrays.argtypes = [vp] + [C.POINTER(vp)] * 4
    L.novum_obj_device_download.argtypes = [vp, vp, vp, vp, vp]
    L.novum_obj_device_timings.argtypes = [vp, C.POINTER(cf), C.POINTER(cf), C.POINTER(ci), C.POINTER(C.c_longlong)]
    L.novum_scene_from_obj_device.argtypes = [vp, vp, ci, vp, ci, C.POINTER(vp)]
    L.novum_accum_save.argtypes = [C.c_char_p, vp, ci, ci, ci, C.c_uint64, ci]
    L.novum_accum_info.argtypes = [C.c_char_p, C.POINTER(ci), C.POINTER(ci), C.POINTER(ci), C.POINTER(C.c_uint64), C.POINTER(ci)]
    L.novum_accum_load.argtypes = [C.c_char_p, vp, ci, ci, C.POINTER(ci), C.POINTER(C.c_uint64)]
    L.novum_hashgrid_create.argtypes = [ci, C.POINTER(vp)]
    L.novum_hashgrid_destroy.argtypes = [vp]
    L.novum_hashgrid_table_size.argtypes = [vp]
    L.novum_hashgrid_build.argtypes = [vp, vp, vp, vp, ci, C.POINTER(cf * 3), cf, vp, vp, vp, vp]
    L.novum_hashgrid_build_device.argtypes = [vp, vp, vp, vp, ci, C.POINTER(cf * 3), cf, ci]
    L.novum_hashgrid_query_count.argtypes = [vp, vp, vp, vp, ci, vp]
    L.novum_hashgrid_last_ms.argtypes = [vp, C.POINTER(cf)]
    L.novum_launch_cache_stats.argtypes = [C.POINTER(LaunchCacheStats)]
    L.novum_launch_cache_clear.restype = None
    L.novum_launch_set_progressive.argtypes = [cf, C.c_char_p]
    L.novum_hashgrid_build_static.argtypes = [vp, vp, vp, ci, ci, C.POINTER(cf * 3), cf, vp, vp, vp, vp, vp]
    L.novum_vcm_create.argtypes = [vp, C.POINTER(RenderConfig), C.POINTER(Camera), Shard, C.c_uint64, C.POINTER(vp)]
    L.novum_vcm_destroy.argtypes = [vp]
    L.novum_vcm_light_pass.argtypes = [vp, ci, vp]
    L.novum_vcm_cache_buffer.argtypes = [vp, C.POINTER(vp), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.novum_vcm_build_grid.argtypes = [vp, ci]
    L.novum_vcm_eye_pass.argtypes = [vp, ci, vp]
    L.novum_vcm_stats.argtypes = [vp, C.POINTER(RenderStats)]
    L.novum_vcm_stream.argtypes = [vp, C.POINTER(vp)]
    L.novum_vcm_reset_stats.argtypes = [vp]
    L.novum_tiles_create.argtypes = [vp, C.POINTER(RenderConfig), C.POINTER(Camera), Shard, C.c_uint64, ci, C.POINTER(vp)]
    L.novum_tiles_destroy.argtypes = [vp]
    L.novum_tiles_destroy.restype = None
    L.novum_tiles_lanes.argtypes = [vp]
    L.novum_tiles_set_view.argtypes = [vp, C.POINTER(RenderConfig), C.POINTER(Camera), C.c_uint64]
    L.novum_tiles_ipc_export.argtypes = [vp, vp]
    L.novum_tiles_ipc_attach.argtypes = [vp, ci, vp]
    L.novum_tiles_bases.argtypes = [vp, C.POINTER(vp)]
    L.novum_tiles_attach_pointers.argtypes = [vp, ci, C.POINTER(vp)]
    L.novum_tiles_render.argtypes = [vp, ci, ci, vp, C.POINTER(RenderStats)]
    L.novum_vcm_ipc_export.argtypes = [vp, vp]
    L.novum_vcm_ipc_attach.argtypes = [vp, ci, vp]
    L.novum_vcm_peer_write.argtypes = [vp, ci]
    L.novum_vcm_attach_pointer.argtypes = [vp, ci, vp]


def _check(rc, what):
    if rc != 0:
        msg = lib().novum_error_string(rc).decode()
        raise NovumError(f"{what} failed: {msg} (code {rc})")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def device_count():
    return lib().novum_device_count()


def make_scene(name, out_dir, detail=100):
    _check(lib().novum_make_scene(name.encode(), out_dir.encode(), int(detail)), "novum_make_scene")
    return os.path.join(out_dir, "configs", "config.rendertron")


def load_config(path):
    cfg = RenderConfig()
    _check(lib().novum_config_load(path.encode(), C.byref(cfg)), "novum_config_load")
    return cfg


def material_table():
    m = np.zeros(24, dtype=MAT_DTYPE)
    _check(lib().novum_material_table(_ptr(m)), "novum_material_table")
    return m


def camera_make(cfg):
    cam = Camera()
    _check(lib().novum_camera_make(C.byref(cfg), C.byref(cam)), "novum_camera_make")
    return cam


class Scene:
    """Opaque NovumScene handle: device geometry + materials + the GPU-built BVH8."""

    def __init__(self, handle, cfg=None):
        self.h = handle
        self.cfg = cfg

    @classmethod
    def from_config(cls, path, render_number=0):
        h = C.c_void_p()
        cfg = RenderConfig()
        _check(lib().novum_scene_load_config(path.encode(), render_number, C.byref(h), C.byref(cfg)), "novum_scene_load_config")
        return cls(h, cfg)

    @classmethod
    def from_arrays(cls, positions, normals, uvs, tris, mats, textures=None):
        positions = np.ascontiguousarray(positions, np.float32).reshape(-1, 4)
        normals = np.ascontiguousarray(normals, np.float32).reshape(-1, 4)
        uvs = np.ascontiguousarray(uvs, np.float32).reshape(-1, 2)
        tris = np.ascontiguousarray(tris, TRI_DTYPE)
        mats = np.ascontiguousarray(mats, MAT_DTYPE)
        ntex = 0
        if textures is not None:
            textures = np.ascontiguousarray(textures, np.float32).reshape(-1, 4)
            ntex = len(textures)
        h = C.c_void_p()
        _check(lib().novum_scene_create(_ptr(positions), len(positions), _ptr(normals), len(normals), _ptr(uvs), len(uvs), _ptr(tris),
                                        len(tris), _ptr(mats), len(mats), _ptr(textures), ntex, C.byref(h)), "novum_scene_create")
        return cls(h)

    def close(self):
        if self.h:
            lib().novum_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        nt, nl, nn, r = C.c_int(), C.c_int(), C.c_int(), C.c_float()
        mn, mx = (C.c_float * 3)(), (C.c_float * 3)()
        _check(lib().novum_scene_info(self.h, C.byref(nt), C.byref(nl), C.byref(nn), C.byref(mn), C.byref(mx), C.byref(r)), "novum_scene_info")
        pm, bm, bl = C.c_float(), C.c_float(), C.c_int()
        _check(lib().novum_scene_timings(self.h, C.byref(pm), C.byref(bm), C.byref(bl)), "novum_scene_timings")
        return {"tris": nt.value, "lights": nl.value, "bvh8_nodes": nn.value, "min": list(mn), "max": list(mx), "radius": r.value,
                "parse_ms": pm.value, "bvh8_build_ms": bm.value, "bvh8_build_launches": bl.value,
                "bvh8_depth": lib().novum_scene_bvh_depth(self.h)}

    def set_features(self, nested_bidir=False, environment=None):
        """Opt-in features (novum_scene_set_features): nested dielectrics in the bidirectional walks; a constant environment (rgb)."""
        flags = (1 if nested_bidir else 0) | (2 if environment is not None else 0)
        if environment is not None:
            rgb = (C.c_float * 3)(*[float(x) for x in environment])
            _check(lib().novum_scene_set_environment(self.h, C.byref(rgb)), "novum_scene_set_environment")
        _check(lib().novum_scene_set_features(self.h, flags), "novum_scene_set_features")

    def download(self):
        c = [C.c_int() for _ in range(5)]
        _check(lib().novum_scene_counts(self.h, *[C.byref(x) for x in c]), "novum_scene_counts")
        npos, nnrm, nuv, ntri, nl = [x.value for x in c]
        pos = np.zeros((npos, 4), np.float32); nrm = np.zeros((nnrm, 4), np.float32); uv = np.zeros((nuv, 2), np.float32)
        tris = np.zeros(ntri, TRI_DTYPE); lights = np.zeros(nl, TRI_DTYPE)
        _check(lib().novum_scene_download(self.h, _ptr(pos), _ptr(nrm), _ptr(uv), _ptr(tris), _ptr(lights)), "novum_scene_download")
        return pos, nrm, uv, tris, lights

    def trace_closest(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.zeros(len(rays), HIT_DTYPE)
        _check(lib().novum_trace_closest(self.h, _ptr(rays), len(rays), _ptr(hits)), "novum_trace_closest")
        return hits

    def trace_shadow(self, rays, max_t, skip=None):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        max_t = np.ascontiguousarray(max_t, np.float32)
        skip = None if skip is None else np.ascontiguousarray(skip, np.int32)
        out = np.zeros((len(rays), 4), np.float32)
        _check(lib().novum_trace_shadow(self.h, _ptr(rays), _ptr(max_t), _ptr(skip), len(rays), _ptr(out)), "novum_trace_shadow")
        return out

    def trace_closest_device(self, d_rays, n, d_hits, sync=True):
        _check(lib().novum_trace_closest_device(self.h, C.c_void_p(d_rays), n, C.c_void_p(d_hits), int(sync)), "novum_trace_closest_device")

    def trace_shadow_device(self, d_rays, d_max_t, d_skip, n, d_out, sync=True):
        _check(lib().novum_trace_shadow_device(self.h, C.c_void_p(d_rays), C.c_void_p(d_max_t), C.c_void_p(d_skip), n, C.c_void_p(d_out),
                                               int(sync)), "novum_trace_shadow_device")

    def last_trace_ms(self):
        ms = C.c_float()
        _check(lib().novum_trace_last_ms(self.h, C.byref(ms)), "novum_trace_last_ms")
        return ms.value

    def counters(self, enable):
        a, b = C.c_uint64(), C.c_uint64()
        _check(lib().novum_trace_counters(self.h, int(enable), C.byref(a), C.byref(b)), "novum_trace_counters")
        return a.value, b.value

    def set_profile(self, flags):
        _check(lib().novum_set_profile(self.h, int(flags)), "novum_set_profile")

    def render(self, cfg, cam, first_sample, num_sample, seed=103033, rank=0, world=1, out=None):
        """novum_render: the sum over the passes, copied to host. `out`: optional (h, w, 4) float32 host array to fill (a
        page-locked one, e.g. torch.empty(..., pin_memory=True).numpy(), receives the frame at full PCIe rate)."""
        img = out if out is not None else np.empty((cfg.height, cfg.width, 4), np.float32)
        assert img.shape == (cfg.height, cfg.width, 4) and img.dtype == np.float32 and img.flags.c_contiguous
        st = RenderStats()
        _check(lib().novum_render(self.h, C.byref(cfg), C.byref(cam), first_sample, num_sample, Shard(rank, world), seed, _ptr(img),
                                  C.byref(st)), "novum_render")
        return img, st

    def render_device(self, cfg, cam, first_sample, num_sample, d_colors, seed=103033, rank=0, world=1):
        st = RenderStats()
        _check(lib().novum_render_device(self.h, C.byref(cfg), C.byref(cam), first_sample, num_sample, Shard(rank, world), seed,
                                         C.c_void_p(d_colors), C.byref(st)), "novum_render_device")
        return st


def obj_parse(path, emission=(0.0, 0.0, 0.0), material_id=0, y_offset=0.0, threads=0, chunk_bytes=0):
    """One OBJ file through the host front end (no GPU): -> (pos (n,4), nrm (n,4), uv (n,2), tris TRI_DTYPE, n_lights)."""
    h = C.c_void_p()
    e = (C.c_float * 3)(*[float(x) for x in emission])
    _check(lib().novum_obj_parse(os.fsencode(path), C.byref(e), int(material_id), float(y_offset), int(threads), int(chunk_bytes), C.byref(h)),
           "novum_obj_parse")
    try:
        n = [C.c_int() for _ in range(5)]
        _check(lib().novum_obj_counts(h, *[C.byref(x) for x in n]), "novum_obj_counts")
        pos = np.zeros((n[0].value, 4), np.float32); nrm = np.zeros((n[1].value, 4), np.float32)
        uv = np.zeros((n[2].value, 2), np.float32); tris = np.zeros(n[3].value, TRI_DTYPE)
        _check(lib().novum_obj_download(h, _ptr(pos), _ptr(nrm), _ptr(uv), _ptr(tris)), "novum_obj_download")
        return pos, nrm, uv, tris, n[4].value
    finally:
        lib().novum_obj_free(h)


class ObjDeviceMesh:
    """OBJ files parsed ON THE GPU (novum_obj_device_*, csrc/obj_device.cu): accumulates the files of a scene in device memory, in
    order, with the arrays of obj_parse bit for bit. NovumError(-5, unsupported) leaves the mesh unchanged: use obj_parse for that file."""

    def __init__(self):
        self.h = C.c_void_p()
        _check(lib().novum_obj_device_create(C.byref(self.h)), "novum_obj_device_create")

    def close(self):
        if self.h:
            lib().novum_obj_device_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def append_file(self, path, emission=(0.0, 0.0, 0.0), material_id=0, y_offset=0.0):
        e = (C.c_float * 3)(*[float(x) for x in emission])
        _check(lib().novum_obj_device_append_file(self.h, os.fsencode(path), C.byref(e), int(material_id), float(y_offset)), "novum_obj_device_append_file")

    def append_text(self, text, emission=(0.0, 0.0, 0.0), material_id=0, y_offset=0.0):
        e = (C.c_float * 3)(*[float(x) for x in emission])
        text = bytes(text)
        _check(lib().novum_obj_device_append_text(self.h, text, len(text), C.byref(e), int(material_id), float(y_offset)), "novum_obj_device_append_text")

    def append_device_text(self, d_text, nbytes, emission=(0.0, 0.0, 0.0), material_id=0, y_offset=0.0):
        e = (C.c_float * 3)(*[float(x) for x in emission])
        _check(lib().novum_obj_device_append_device_text(self.h, C.c_void_p(int(d_text)), int(nbytes), C.byref(e), int(material_id), float(y_offset)),
               "novum_obj_device_append_device_text")

    def reset(self):
        _check(lib().novum_obj_device_reset(self.h), "novum_obj_device_reset")

    def counts(self):
        n = [C.c_int() for _ in range(5)]
        _check(lib().novum_obj_device_counts(self.h, *[C.byref(x) for x in n]), "novum_obj_device_counts")
        return tuple(x.value for x in n)

    def device_arrays(self):
        p = [C.c_void_p() for _ in range(4)]
        _check(lib().novum_obj_device_arrays(self.h, *[C.byref(x) for x in p]), "novum_obj_device_arrays")
        return tuple(x.value for x in p)

    def download(self):
        """-> (pos (n,4), nrm (n,4), uv (n,2), tris TRI_DTYPE, n_lights), the layout of obj_parse."""
        n = self.counts()
        pos = np.zeros((n[0], 4), np.float32); nrm = np.zeros((n[1], 4), np.float32)
        uv = np.zeros((n[2], 2), np.float32); tris = np.zeros(n[3], TRI_DTYPE)
        _check(lib().novum_obj_device_download(self.h, _ptr(pos), _ptr(nrm), _ptr(uv), _ptr(tris)), "novum_obj_device_download")
        return pos, nrm, uv, tris, n[4]

    def timings(self):
        up, k, l, b = C.c_float(), C.c_float(), C.c_int(), C.c_longlong()
        _check(lib().novum_obj_device_timings(self.h, C.byref(up), C.byref(k), C.byref(l), C.byref(b)), "novum_obj_device_timings")
        return {"upload_ms": up.value, "kernel_ms": k.value, "launches": l.value, "text_bytes": b.value}

    def to_scene(self, mats, textures=None):
        """The arrays move into a Scene (BVH8 build on the device); the mesh is empty afterwards."""
        mats = np.ascontiguousarray(mats, MAT_DTYPE)
        ntex = 0
        if textures is not None:
            textures = np.ascontiguousarray(textures, np.float32).reshape(-1, 4)
            ntex = len(textures)
        h = C.c_void_p()
        _check(lib().novum_scene_from_obj_device(self.h, _ptr(mats), len(mats), _ptr(textures), ntex, C.byref(h)), "novum_scene_from_obj_device")
        return Scene(h)


def accum_save(path, colors, sample_count, seed=103033, integrator=0):
    """Checkpoint of the running sums (h, w, 4) float32 after `sample_count` passes (novum_accum_save; host only)."""
    colors = np.ascontiguousarray(colors, dtype=np.float32)
    h, w = colors.shape[:2]
    _check(lib().novum_accum_save(os.fsencode(path), _ptr(colors), w, h, int(sample_count), int(seed), int(integrator)), "novum_accum_save")


def accum_info(path):
    w, h, n, integ, seed = C.c_int(), C.c_int(), C.c_int(), C.c_int(), C.c_uint64()
    _check(lib().novum_accum_info(os.fsencode(path), C.byref(w), C.byref(h), C.byref(n), C.byref(seed), C.byref(integ)), "novum_accum_info")
    return {"width": w.value, "height": h.value, "sample_count": n.value, "seed": seed.value, "integrator": integ.value}


def accum_load(path):
    """-> (sums (h, w, 4) float32, sample_count, seed); raises on a truncated / corrupted file."""
    info = accum_info(path)
    img = np.empty((info["height"], info["width"], 4), np.float32)
    n, seed = C.c_int(), C.c_uint64()
    _check(lib().novum_accum_load(os.fsencode(path), _ptr(img), info["width"], info["height"], C.byref(n), C.byref(seed)), "novum_accum_load")
    return img, n.value, seed.value


FRAME_PROGRESSIVE, FRAME_FINAL = 0, 1


def bmp_row_bytes(w):
    return lib().novum_bmp_row_bytes(int(w))


def frame_finish(colors, overlay, sample_count, mode=FRAME_FINAL, post_process=True, want_final=False):
    """Host accumulator (h, w, 4) float32 [+ overlay] -> packed BGR rows (h, row_bytes) uint8 exactly as Image::saveImageBMP writes
    them (imageUtil.cu:82-96) [and the cleaned linear frame]. Runs novum_frame_finish (upload, one device kernel, read-back)."""
    colors = np.ascontiguousarray(colors, dtype=np.float32)
    h, w = colors.shape[:2]
    ov = None if overlay is None else np.ascontiguousarray(overlay, dtype=np.float32)
    rows = np.empty((h, bmp_row_bytes(w)), dtype=np.uint8)
    fin = np.empty((h, w, 4), dtype=np.float32) if want_final else None
    _check(lib().novum_frame_finish(_ptr(colors), None if ov is None else _ptr(ov), w, h, int(sample_count), int(mode), int(bool(post_process)),
                                    _ptr(rows), None if fin is None else _ptr(fin)), "novum_frame_finish")
    return (rows, fin) if want_final else rows


def frame_finish_device(d_colors, d_overlay, w, h, sample_count, mode, post_process, d_bgr, d_final=None, stream=None):
    """Device pointers (ints) in and out; asynchronous on `stream` (cudaStream_t as int, None = default stream)."""
    _check(lib().novum_frame_finish_device(d_colors, d_overlay, int(w), int(h), int(sample_count), int(mode), int(bool(post_process)), d_bgr,
                                           d_final, stream), "novum_frame_finish_device")


def frame_save_bmp(path, d_colors, d_overlay, w, h, sample_count, mode=FRAME_FINAL, post_process=True):
    _check(lib().novum_frame_save_bmp(os.fsencode(path), d_colors, d_overlay, int(w), int(h), int(sample_count), int(mode),
                                      int(bool(post_process))), "novum_frame_save_bmp")
    ms = C.c_float()
    lib().novum_frame_last_ms(C.byref(ms))
    return ms.value


def frame_save_bmp_async(path, d_colors, d_overlay, w, h, sample_count, mode=FRAME_PROGRESSIVE, post_process=True, stream=None):
    """Queues kernel + read-back + file write behind `stream` (a cudaStream_t value); `stream` waits for the kernel only."""
    _check(lib().novum_frame_save_bmp_async(os.fsencode(path), d_colors, d_overlay, int(w), int(h), int(sample_count), int(mode),
                                            int(bool(post_process)), stream), "novum_frame_save_bmp_async")


def frame_save_wait():
    _check(lib().novum_frame_save_wait(), "novum_frame_save_wait")


def save_bmp(path, colors, overlay, sample_count, mode=FRAME_FINAL, post_process=True):
    colors = np.ascontiguousarray(colors, dtype=np.float32)
    h, w = colors.shape[:2]
    ov = None if overlay is None else np.ascontiguousarray(overlay, dtype=np.float32)
    _check(lib().novum_save_bmp(os.fsencode(path), _ptr(colors), None if ov is None else _ptr(ov), w, h, int(sample_count), int(mode),
                                int(bool(post_process))), "novum_save_bmp")


def gamma_steps():
    s = np.empty(257, dtype=np.float32)
    _check(lib().novum_frame_gamma_steps(_ptr(s)), "novum_frame_gamma_steps")
    return s


class HashGrid:
    def __init__(self, max_photons):
        self.h = C.c_void_p()
        self.max = max_photons
        _check(lib().novum_hashgrid_create(max_photons, C.byref(self.h)), "novum_hashgrid_create")
        self.table_size = lib().novum_hashgrid_table_size(self.h)

    def close(self):
        if self.h:
            lib().novum_hashgrid_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def build(self, px, py, pz, scene_min, radius):
        px, py, pz = [np.ascontiguousarray(a, np.float32) for a in (px, py, pz)]
        n = len(px)
        idx = np.zeros(n, np.uint32); hsh = np.zeros(n, np.uint32)
        cs = np.zeros(self.table_size, np.uint32); ce = np.zeros(self.table_size, np.uint32)
        mn = (C.c_float * 3)(*scene_min)
        _check(lib().novum_hashgrid_build(self.h, _ptr(px), _ptr(py), _ptr(pz), n, C.byref(mn), radius, _ptr(idx), _ptr(hsh), _ptr(cs),
                                          _ptr(ce)), "novum_hashgrid_build")
        return idx, hsh, cs, ce

    def build_device(self, d_px, d_py, d_pz, n, scene_min, radius, sync=True):
        mn = (C.c_float * 3)(*scene_min)
        _check(lib().novum_hashgrid_build_device(self.h, C.c_void_p(d_px), C.c_void_p(d_py), C.c_void_p(d_pz), n, C.byref(mn), radius,
                                                 int(sync)), "novum_hashgrid_build_device")

    def query_count(self, qx, qy, qz):
        qx, qy, qz = [np.ascontiguousarray(a, np.float32) for a in (qx, qy, qz)]
        out = np.zeros(len(qx), np.uint32)
        _check(lib().novum_hashgrid_query_count(self.h, _ptr(qx), _ptr(qy), _ptr(qz), len(qx), _ptr(out)), "novum_hashgrid_query_count")
        return out

    def last_ms(self):
        ms = C.c_float()
        _check(lib().novum_hashgrid_last_ms(self.h, C.byref(ms)), "novum_hashgrid_last_ms")
        return ms.value


STATIC_GRID_MAX_PHOTONS = 1 << 20


def hashgrid_build_static(d_px, d_py, d_pz, n, max_photons, scene_min, radius, d_sorted_hash, d_sorted_index, d_cell_start, d_cell_end,
                          stream=None):
    """novum_hashgrid_build_static: device pointers (ints) in and out, nothing allocated by the call; asynchronous on `stream`."""
    mn = (C.c_float * 3)(*scene_min)
    vp = C.c_void_p
    _check(lib().novum_hashgrid_build_static(vp(d_px), vp(d_py), vp(d_pz), n, max_photons, C.byref(mn), radius, vp(d_sorted_hash),
                                             vp(d_sorted_index), vp(d_cell_start), vp(d_cell_end), vp(stream or 0)),
           "novum_hashgrid_build_static")


class _DevBuf:
    """Zero-copy view of a device allocation for torch.as_tensor (CUDA array interface v2)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class VcmSession:
    """Phase API of the bidirectional engine for one rank of a tile-sharded render (include/novum_b200.h, novum_vcm_*):
    light_pass -> all-gather of slabs() by the caller -> build_grid -> eye_pass."""

    def __init__(self, scene, cfg, cam, rank=0, world=1, seed=103033):
        self.h = C.c_void_p()
        self.scene, self.rank, self.world = scene, rank, world
        _check(lib().novum_vcm_create(scene.h, C.byref(cfg), C.byref(cam), Shard(rank, world), seed, C.byref(self.h)), "novum_vcm_create")
        base, slab, total = C.c_void_p(), C.c_uint64(), C.c_uint64()
        _check(lib().novum_vcm_cache_buffer(self.h, C.byref(base), C.byref(slab), C.byref(total)), "novum_vcm_cache_buffer")
        self.base, self.slab_bytes, self.total_bytes = base.value, slab.value, total.value

    def close(self):
        if self.h:
            lib().novum_vcm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def light_pass(self, it, d_colors):
        _check(lib().novum_vcm_light_pass(self.h, int(it), C.c_void_p(d_colors)), "novum_vcm_light_pass")

    def build_grid(self, it):
        _check(lib().novum_vcm_build_grid(self.h, int(it)), "novum_vcm_build_grid")

    def eye_pass(self, it, d_colors):
        _check(lib().novum_vcm_eye_pass(self.h, int(it), C.c_void_p(d_colors)), "novum_vcm_eye_pass")

    def stats(self):
        st = RenderStats()
        _check(lib().novum_vcm_stats(self.h, C.byref(st)), "novum_vcm_stats")
        return st

    def reset_stats(self):
        _check(lib().novum_vcm_reset_stats(self.h), "novum_vcm_reset_stats")

    def stream(self):
        s = C.c_void_p()
        _check(lib().novum_vcm_stream(self.h, C.byref(s)), "novum_vcm_stream")
        return s.value

    def ipc_export(self):
        """64-byte CUDA IPC handle of this rank's slab buffer (to be carried to the other ranks' processes)."""
        buf = C.create_string_buffer(64)
        _check(lib().novum_vcm_ipc_export(self.h, C.cast(buf, C.c_void_p)), "novum_vcm_ipc_export")
        return bytes(buf.raw)

    def ipc_attach(self, peer_rank, handle):
        buf = C.create_string_buffer(bytes(handle), 64)
        _check(lib().novum_vcm_ipc_attach(self.h, peer_rank, C.cast(buf, C.c_void_p)), "novum_vcm_ipc_attach")

    def attach_pointer(self, peer_rank, d_base):
        _check(lib().novum_vcm_attach_pointer(self.h, peer_rank, C.c_void_p(d_base)), "novum_vcm_attach_pointer")

    def peer_write(self, enable=True):
        _check(lib().novum_vcm_peer_write(self.h, 1 if enable else 0), "novum_vcm_peer_write")

    def slabs_tensor(self):
        """All ranks' photon slabs as one uint8 torch tensor on the current device (in-place all-gather target)."""
        import torch
        return torch.as_tensor(_DevBuf(self.base, self.total_bytes), device="cuda")

    def my_slab_tensor(self):
        return self.slabs_tensor()[self.rank * self.slab_bytes:(self.rank + 1) * self.slab_bytes]


class TileRank:
    """One rank of the tile renderer (include/novum_b200.h, novum_tiles_*): `lanes` iterations in flight, photon exchange by peer
    stores fused into the light pass, device-side rendezvous. Every rank calls render() with the same iteration range."""

    def __init__(self, scene, cfg, cam, rank=0, world=1, seed=103033, lanes=2):
        self.h = C.c_void_p()
        self.scene, self.rank, self.world, self.lanes = scene, rank, world, lanes
        _check(lib().novum_tiles_create(scene.h, C.byref(cfg), C.byref(cam), Shard(rank, world), seed, lanes, C.byref(self.h)), "novum_tiles_create")

    def close(self):
        if self.h:
            lib().novum_tiles_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def ipc_export(self):
        buf = (C.c_ubyte * (64 * self.lanes))()
        _check(lib().novum_tiles_ipc_export(self.h, buf), "novum_tiles_ipc_export")
        return bytes(buf)

    def ipc_attach(self, peer_rank, handles):
        buf = (C.c_ubyte * (64 * self.lanes)).from_buffer_copy(handles)
        _check(lib().novum_tiles_ipc_attach(self.h, peer_rank, buf), "novum_tiles_ipc_attach")

    def bases(self):
        arr = (C.c_void_p * self.lanes)()
        _check(lib().novum_tiles_bases(self.h, arr), "novum_tiles_bases")
        return list(arr)

    def attach_pointers(self, peer_rank, bases):
        arr = (C.c_void_p * self.lanes)(*bases)
        _check(lib().novum_tiles_attach_pointers(self.h, peer_rank, arr), "novum_tiles_attach_pointers")

    def render(self, first, count, d_colors):
        st = RenderStats()
        _check(lib().novum_tiles_render(self.h, first, count, C.c_void_p(d_colors), C.byref(st)), "novum_tiles_render")
        return st
