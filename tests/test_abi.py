"""The C-ABI library loads without a GPU and exports every symbol include/novum_b200.h declares; POD layouts match
the reference's structs (sizes / offsets printed from the reference's own headers: Triangle 80, Material 176,
Camera 112, BVHnode 48, Vertices 32 — SURVEY §8b)."""
import ctypes as C
import os
import subprocess

import numpy as np

import cudapathtracer_b200 as nb
from cudapathtracer_b200 import binding as B


def test_library_exports_every_declared_symbol(built):
    L = nb.lib()
    assert len(nb.DECLARED_SYMBOLS) >= 40
    missing = [s for s in nb.DECLARED_SYMBOLS if not hasattr(L, s)]
    assert not missing, f"declared in novum_b200.h but not exported: {missing}"
    assert L.novum_abi_version() == 6


def test_exports_are_plain_c_symbols(built):
    out = subprocess.check_output(["nm", "-D", "--defined-only", nb.lib_path()]).decode()
    names = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for s in nb.DECLARED_SYMBOLS:
        assert s in names, s


def test_struct_layouts_match_reference():
    assert C.sizeof(B.Camera) == 112
    assert B.Camera.forward.offset == 64 and B.Camera.right.offset == 80 and B.Camera.up.offset == 96
    assert B.Camera.aperture.offset == 36 and B.Camera.antiAliasJitterDist.offset == 48
    assert B.TRI_DTYPE.itemsize == 80 and B.TRI_DTYPE.fields["emission"][1] == 48 and B.TRI_DTYPE.fields["lightInd"][1] == 64
    assert B.MAT_DTYPE.itemsize == 176
    for name, off in (("type", 32), ("albedo", 48), ("roughness", 64), ("eta", 80), ("k", 96), ("ior", 112), ("transmission", 124),
                      ("isSpecular", 128), ("boundary", 129), ("thinWalled", 130), ("absorption", 144), ("priority", 160)):
        assert B.MAT_DTYPE.fields[name][1] == off, name
    assert C.sizeof(B.RenderStats) == 144


def test_error_strings_and_null_handles(built):
    L = nb.lib()
    assert L.novum_error_string(0) == b"ok"
    assert b"argument" in L.novum_error_string(-1)
    assert L.novum_scene_info(None, None, None, None, None, None, None) == -3
    assert L.novum_trace_closest(None, None, 0, None) == -3
    assert L.novum_config_load(None, None) == -1


def test_no_gpu_is_reported_not_faked(built):
    """Without a device the scene constructor must refuse (there is no CPU fallback to fall into)."""
    if nb.device_count() > 0:
        return
    pos = np.zeros((3, 4), np.float32); pos[1, 0] = 1; pos[2, 1] = 1
    nrm = np.zeros((1, 4), np.float32); nrm[0, 2] = 1
    uv = np.zeros((1, 2), np.float32)
    tri = np.zeros(1, B.TRI_DTYPE); tri["bInd"] = 1; tri["cInd"] = 2; tri["materialID"] = 2
    try:
        nb.Scene.from_arrays(pos, nrm, uv, tri, nb.material_table())
    except nb.NovumError as e:
        assert "CUDA" in str(e) or "device" in str(e)
    else:
        raise AssertionError("scene creation succeeded without a GPU")
