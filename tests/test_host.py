"""Host front end kept from the reference: config parser (objects.cuh:844-943), material table (main.cu:397-467),
camera basis (objects.cuh:221-325), scene generator output. No GPU needed."""
import math
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from scene_utils import parse_config


def test_config_roundtrip(scene_dirs):
    cfg = nb.load_config(os.path.join(scene_dirs["cornell"], "configs", "config.rendertron"))
    assert (cfg.width, cfg.height, cfg.sampleCount, cfg.maxDepth, cfg.bvhLeafSize) == (512, 512, 64, 40, 2)
    assert cfg.integrator == 0 and cfg.pinholeCamera == 1 and cfg.numMeshes == 6
    assert cfg.bdptEyeDepth == 16 and cfg.bdptLightDepth == 10
    assert abs(cfg.vcmMergeConst - 0.7) < 1e-7 and abs(cfg.vcmInitialMergeRadiusMultiplier - 0.01) < 1e-9
    assert list(cfg.camPos) == [0.0, 0.0, 1.0] and cfg.camFov == 60.0


def test_config_accepts_both_spellings_of_radius_key(tmp_path):
    """configs/config.rendertron:27 says 'Multipler', objects.cuh:939 matches 'Multiplier' (SURVEY F10)."""
    for key in ("VCM Initial Merge Radius Multiplier", "VCM Initial Merge Radius Multipler"):
        p = tmp_path / "c.rendertron"
        p.write_text(f"width: 8\nheight: 4\nIntegrator: SPPM\nBDPT_NEE: true\n{key}: 0.25\nMeshes (x):\na.obj; 2.0 * (1.0, 0.5, 0.0); 3\n")
        cfg = nb.load_config(str(p))
        assert cfg.vcmInitialMergeRadiusMultiplier == 0.25
        # SPPM forces the connection strategies off and merging on (main.cu:314-333)
        assert cfg.integrator == 4 and cfg.doSPPM == 1 and cfg.vcmDoMerge == 1 and cfg.bdptNee == 0 and cfg.bdptDoMis == 0
        assert cfg.numMeshes == 1


def test_config_missing_file_is_an_error():
    try:
        nb.load_config("/nonexistent/config.rendertron")
    except nb.NovumError as e:
        assert "file" in str(e)
    else:
        raise AssertionError("expected NovumError")


def test_integrator_names(tmp_path):
    for name, want in (("UNIDIRECTIONAL", 0), ("BIDIRECTIONAL", 1), ("BDPT", 1), ("NAIVE_UNIDIRECTIONAL", 2), ("VCM", 3), ("SPPM", 4),
                       ("nope", -1)):
        p = tmp_path / "c.rendertron"
        p.write_text(f"Integrator: {name}\n")
        assert nb.load_config(str(p)).integrator == want


def test_material_table_matches_reference_values():
    m = nb.material_table()
    assert len(m) == 24
    assert m["type"][0] == 2 and m["priority"][0] == 99 and m["ior"][0] == 1.0                 # air
    assert m["type"][5] == 2 and m["ior"][5] == 1.5 and m["priority"][5] == 1 and m["isSpecular"][5] == 1 and m["boundary"][5] == 1
    assert m["type"][4] == 1 and np.allclose(m["eta"][4][:3], [0.17, 0.35, 1.5]) and np.allclose(m["k"][4][:3], m["eta"][4][:3])
    assert abs(m["roughness"][4] - 0.05) < 1e-7 and abs(m["roughness"][7] - 0.15) < 1e-7
    assert np.allclose(m["absorption"][8][:3], [2.5 * 0.18, 2.5 * 1.5, 2.5 * 2.996])            # tea
    assert m["priority"][9] == 0 and abs(m["ior"][9] - 1.31) < 1e-7 and m["priority"][10] == 2
    assert m["type"][13] == 4 and abs(m["transmission"][13] - 0.15) < 1e-7 and m["thinWalled"][13] == 1
    assert m["type"][19] == 6 and m["isSpecular"][19] == 1 and m["boundary"][19] == 0
    assert np.allclose(m["albedo"][2][:3], 0.9) and np.allclose(m["albedo"][6][:3], [0.9, 0.1, 0.1]) and np.allclose(m["albedo"][23][:3], [0.1, 0.9, 0.1])
    # fields the reference leaves indeterminate are zero here
    assert not m["hasTransMap"].any() and m["isSpecular"][[1, 2, 3, 4, 6, 7]].sum() == 0


def test_camera_basis_and_pinhole_quirks(scene_dirs):
    cfg = nb.load_config(os.path.join(scene_dirs["cornell"], "configs", "config.rendertron"))
    cam = nb.camera_make(cfg)
    assert abs(cam.fovScale - math.tan(30 * 3.141592 / 180)) < 1e-6
    assert abs(cam.aperture - 1e-6) < 1e-12 and abs(cam.focalDist - 1.0 / 60.0) < 1e-9     # objects.cuh:235-236
    assert cam.antiAliasJitterDist == 2.0
    f = np.array([cam.forward.x, cam.forward.y, cam.forward.z]); r = np.array([cam.right.x, cam.right.y, cam.right.z])
    u = np.array([cam.up.x, cam.up.y, cam.up.z])
    assert np.allclose(f, [0, 0, -1]) and np.allclose(r, [1, 0, 0]) and np.allclose(u, [0, 1, 0])
    cfg.camRot[1] = 90.0
    cam = nb.camera_make(cfg)
    f = np.array([cam.forward.x, cam.forward.y, cam.forward.z])
    assert np.allclose(f, [-1, 0, 0], atol=1e-6)


def test_generated_scenes_have_normals_and_uvs(scene_dirs, py_scenes):
    for name in ("cornell", "dragon", "water"):
        cfg, pos, nrm, uv, tris, mats = py_scenes[name]
        assert len(tris) > 0 and (tris["naInd"] >= 0).all() and (tris["uvaInd"] >= 0).all()
        assert (tris["emission"][:, :3].sum(1) > 0).sum() >= 2            # a quad light = 2 triangles
        assert np.isfinite(pos).all() and np.isfinite(nrm).all()
        ln = np.linalg.norm(nrm[:, :3], axis=1)
        assert (ln > 0.99).all() and (ln < 1.01).all()
    assert len(py_scenes["cornell"][4]) == 36                              # (5 room + 12 box + 1 light) quads x 2
    cfgd, meshes = parse_config(os.path.join(scene_dirs["dragon"], "configs", "config.rendertron"))
    assert cfgd["Integrator"] == "UNIDIRECTIONAL" and cfgd["width"] == "1920" and cfgd["Sample Count"] == "1024"
    assert meshes[0][0] == "scenedata/dragon_vrip.obj" and meshes[0][2] == 5


def test_dragon_full_detail_triangle_count(tmp_path):
    """The stand-in for dragon_vrip.obj has the reference asset's scale: ~871k triangles."""
    d = str(tmp_path / "dragon")
    nb.make_scene("dragon", d, 100)
    n = sum(1 for l in open(os.path.join(d, "scenedata", "dragon_vrip.obj")) if l.startswith("f "))
    assert 860_000 <= n <= 880_000, n


def test_bench_frame_reference_arm_runs_on_cpu(built):
    """bench.py --workload frame_finish --impl reference times the reference's own host code (or the restatement when
    oracle/_ref is absent) and prints the contract's JSON line; no GPU involved."""
    import json, subprocess, sys, os
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.check_output([sys.executable, os.path.join(repo, "bench.py"), "--workload", "frame_finish", "--impl", "reference", "--steps", "1"],
                                  timeout=300).decode()
    line = json.loads([l for l in out.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["unit"] == "Mpixels/s" and line["value"] > 0
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["e2e"]["value"] == line["value"]


def test_obj_number_reader_equals_strtod(built):
    """The OBJ loader's fast number path (integer mantissa x exact power of ten) must return what strtod returns — bit for bit — and
    consume the same characters, falling back to strtod for everything outside its exact range."""
    import ctypes as C
    import struct
    import cudapathtracer_b200 as nb
    L = nb.lib()
    libc = C.CDLL(None)
    libc.strtod.restype = C.c_double
    libc.strtod.argtypes = [C.c_char_p, C.POINTER(C.c_char_p)]

    def ours(t):
        n = C.c_int()
        return L.novum_parse_double(t, C.byref(n)), n.value

    def ref(t):
        buf = C.create_string_buffer(t)
        end = C.c_char_p()
        v = libc.strtod(buf, C.byref(end))
        return v, C.cast(end, C.c_void_p).value - C.addressof(buf)

    rng = np.random.default_rng(5)
    cases = [b"0", b"-0", b"0.0", b"-0.0", b"1.", b".5", b"-.5", b"1e5", b"1E-5", b"1e", b"1e+", b"12345678901234567890", b"123456789012345",
             b"1234567890123456", b"1e22", b"1e23", b"1e-22", b"1e-23", b"3.14159 2.7", b"nan", b"inf", b"-inf", b"0x10", b"+7", b"  42", b"1.5e3x",
             b"9007199254740993", b"0.1", b"1e308", b"4.9e-324", b"", b" ", b"-", b"+.", b"1.e2", b".e2", b"\t-12.5\r", b"7/8/9"]
    for v in rng.standard_normal(20000) * 10:
        cases += [b"%.6f" % v, b"%.9g" % v, b"%.17g" % v, b"%g" % v]
    for e in rng.integers(-40, 40, 4000):
        cases.append(b"%.12ge%d" % (rng.random(), e))
    bad = []
    for t in cases:
        a, b = ours(t), ref(t)
        same = (struct.pack("d", a[0]) == struct.pack("d", b[0]) or (a[0] != a[0] and b[0] != b[0])) and a[1] == b[1]
        if not same:
            bad.append((t, a, b))
    assert not bad, bad[:5]


def test_accumulator_checkpoint_roundtrip_and_damage(built, tmp_path):
    """novum_accum_save / _info / _load: bit-exact round trip (NaN and Inf payloads included), refusal of a wrong frame size,
    of a truncated file and of a flipped payload bit; no GPU involved."""
    import ctypes as C
    import cudapathtracer_b200 as nb
    rng = np.random.default_rng(11)
    h, w = 37, 53
    acc = (rng.random((h, w, 4), dtype=np.float32) * 1000).astype(np.float32)
    acc[3, 4, 0] = np.nan; acc[5, 6, 1] = np.inf; acc[7, 8, 2] = -0.0
    p = str(tmp_path / "run.novumacc")
    nb.accum_save(p, acc, 135, seed=0xDEADBEEF12345678, integrator=3)
    assert os.path.getsize(p) == 64 + h * w * 16 and not os.path.exists(p + ".part")
    assert nb.accum_info(p) == {"width": w, "height": h, "sample_count": 135, "seed": 0xDEADBEEF12345678, "integrator": 3}
    back, n, seed = nb.accum_load(p)
    assert n == 135 and seed == 0xDEADBEEF12345678 and np.array_equal(back.view(np.uint32), acc.view(np.uint32))
    L = nb.binding.lib()
    wrong = np.empty((h, w + 1, 4), np.float32)
    assert L.novum_accum_load(os.fsencode(p), wrong.ctypes.data_as(C.c_void_p), w + 1, h, None, None) == -1      # NOVUM_E_ARG
    raw = bytearray(open(p, "rb").read())
    open(str(tmp_path / "cut.novumacc"), "wb").write(raw[: len(raw) // 2])
    with pytest.raises(nb.NovumError):
        nb.accum_load(str(tmp_path / "cut.novumacc"))
    raw[64 + 1000] ^= 0x10
    open(str(tmp_path / "flip.novumacc"), "wb").write(raw)
    with pytest.raises(nb.NovumError):
        nb.accum_load(str(tmp_path / "flip.novumacc"))
    with pytest.raises(nb.NovumError):
        nb.accum_info(str(tmp_path / "missing.novumacc"))
    with pytest.raises(nb.NovumError):
        nb.accum_save(str(tmp_path / "no_dir" / "x.novumacc"), acc, 1)


def _adversarial_obj(path, rng, n_lines=600, device_ok=False):
    """An OBJ whose result depends on everything the sequential loader decides from 'how far the file has been read'.
    device_ok: only number spellings the device parser's rules read themselves (obj_parse.cuh: at most 15 significant digits, no nan)."""
    out, nv, nvt, nvn = [], 0, 0, 0
    long_fmt = "%.15g" if device_ok else "%.17g"
    odd_vn = ["vn 0 0 0", "vn 1 2", "vn x 0 1" if device_ok else "vn nan 0 1", "vn .5 -.25e1 +1.", "vn 1e-7 1e-7 0"]
    for k in range(n_lines):
        r = rng.random()
        if r < 0.30 or nv < 4:
            out.append("v %s %s %s" % tuple(rng.choice(["%.6f", "%.3e", "%d", long_fmt]) % x for x in rng.normal(0, 3, 3))); nv += 1
        elif r < 0.38:
            out.append("vt %.5f %.5f" % tuple(rng.random(2))); nvt += 1
        elif r < 0.46:
            out.append(rng.choice(["vn %.4f %.4f %.4f" % tuple(rng.normal(0, 1, 3))] + odd_vn)); nvn += 1
        elif r < 0.50:
            out.append(rng.choice(["# comment", "s off", "", "g group", "usemtl m", "   "]))
        else:
            m = int(rng.integers(3, 6))
            hi = nv + (3 if rng.random() < 0.15 else 0)              # sometimes a vertex that is only declared LATER (must be dropped)
            idx = rng.integers(1, hi + 1, m)
            if rng.random() < 0.08:
                idx[1] = idx[0]                                       # degenerate
            if rng.random() < 0.05:
                idx[0] = -int(rng.integers(1, 3))                     # negative index: no relative addressing in the reference
            style = rng.random()
            if style < 0.3 or nvt == 0 or nvn == 0:
                toks = ["%d" % a for a in idx]                                          # generated normal + shared zero uv
            elif style < 0.5:
                toks = ["%d//%d" % (a, rng.integers(1, nvn + 1)) for a in idx]          # shared zero uv only
            elif style < 0.7:
                toks = ["%d/%d" % (a, rng.integers(1, nvt + 1)) for a in idx]           # generated normal only
            else:
                toks = ["%d/%d/%d" % (a, rng.integers(1, nvt + 1), rng.integers(1, nvn + 1)) for a in idx]
            out.append("f " + rng.choice([" ", "\t", "  "]).join(toks))
    text = "\r\n".join(out) if rng.random() < 0.3 else "\n".join(out)
    open(path, "w", newline="").write(text + ("" if rng.random() < 0.5 else "\n"))


def test_obj_chunked_parse_equals_sequential(built, tmp_path):
    """The multi-threaded OBJ front end: any chunking gives the arrays of the one-chunk (sequential) parse, bit for bit."""
    rng = np.random.default_rng(2024)
    for trial in range(12):
        p = str(tmp_path / f"adv{trial}.obj")
        _adversarial_obj(p, rng)
        e = (0.0, 0.0, 0.0) if trial % 3 else (5.0, 4.0, 3.0)
        ref = nb.obj_parse(p, e, 7, 0.25, threads=1, chunk_bytes=1 << 30)
        assert len(ref[3]) > 50 and len(ref[1]) > 10
        for threads, chunk in ((1, 48), (4, 48), (3, 333), (8, 4096), (0, 0)):
            got = nb.obj_parse(p, e, 7, 0.25, threads=threads, chunk_bytes=chunk)
            for a, b in zip(ref[:4], got[:4]):
                assert a.shape == b.shape and a.tobytes() == b.tobytes(), (trial, threads, chunk)
            assert got[4] == ref[4] == (len(ref[3]) if trial % 3 == 0 else 0)


def test_obj_parse_equals_python_loader(scene_dirs):
    """Independent restatement (tests/scene_utils.py) on a generated mesh, parsed in 64 KiB chunks by 4 threads."""
    from scene_utils import load_obj_into
    path = os.path.join(scene_dirs["dragon"], "scenedata", "dragon_vrip.obj")
    pos, nrm, uv, tris, nl = [], [], [], [], [0]
    load_obj_into(path, pos, nrm, uv, tris, (0.0, 0.0, 0.0), 5, nl)
    g = nb.obj_parse(path, (0.0, 0.0, 0.0), 5, 0.0, threads=4, chunk_bytes=65536)
    assert np.array_equal(np.array(pos, np.float32), g[0]) and np.array_equal(np.array(nrm, np.float32), g[1])
    assert np.array_equal(np.array(uv, np.float32), g[2])
    want = np.array(tris, g[3].dtype)
    for f in ("aInd", "bInd", "cInd", "naInd", "nbInd", "ncInd", "uvaInd", "uvbInd", "uvcInd", "materialID", "lightInd", "triInd"):
        assert np.array_equal(want[f], g[3][f]), f


def _obj_rules_lib(built):
    """tests/obj_rules_host.cpp: the device parser's per-line rules (csrc/obj_parse.cuh) compiled for the CPU."""
    import ctypes as C
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    out = os.path.join(here, "_build", "libobjrules.so")
    src = os.path.join(here, "obj_rules_host.cpp")
    hdr = os.path.join(here, "..", "cudapathtracer_b200", "csrc", "obj_parse.cuh")
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        os.makedirs(os.path.dirname(out), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-I" + os.path.join(here, "..", "include"),
                               "-I" + os.path.join(here, "..", "cudapathtracer_b200", "csrc"), src, "-o", out])
    lib = C.CDLL(out)
    lib.obj_rules_parse.restype = C.c_void_p
    lib.obj_rules_parse.argtypes = [C.c_char_p, C.c_size_t, C.POINTER(C.c_float), C.c_int, C.c_float]
    lib.obj_rules_free.argtypes = [C.c_void_p]
    lib.obj_rules_unsupported.argtypes = [C.c_void_p]
    lib.obj_rules_counts.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 5
    lib.obj_rules_download.argtypes = [C.c_void_p] + [C.c_void_p] * 4
    return lib


def _obj_rules_parse(lib, text, emission, material_id, y_offset):
    import ctypes as C
    from cudapathtracer_b200.binding import TRI_DTYPE as ref_tri_dtype
    e = (C.c_float * 3)(*emission)
    h = lib.obj_rules_parse(text, len(text), e, material_id, y_offset)
    try:
        n = [C.c_int() for _ in range(5)]
        lib.obj_rules_counts(h, *[C.byref(x) for x in n])
        pos = np.zeros((n[0].value, 4), np.float32); nrm = np.zeros((n[1].value, 4), np.float32)
        uv = np.zeros((n[2].value, 2), np.float32); tris = np.zeros(n[3].value, ref_tri_dtype)
        lib.obj_rules_download(h, pos.ctypes.data, nrm.ctypes.data, uv.ctypes.data, tris.ctypes.data)
        return (pos, nrm, uv, tris, n[4].value), bool(lib.obj_rules_unsupported(h))
    finally:
        lib.obj_rules_free(h)


def test_obj_device_rules_equal_host_loader(built, tmp_path):
    """The per-line rules the device OBJ parser runs (obj_parse.cuh), driven in file order on the CPU, give the host loader's arrays bit
    for bit; spellings outside the rules are reported as unsupported instead of being guessed."""
    lib = _obj_rules_lib(built)
    rng = np.random.default_rng(77)
    for trial in range(16):
        p = str(tmp_path / f"dev{trial}.obj")
        _adversarial_obj(p, rng, n_lines=700, device_ok=True)
        e = (0.0, 0.0, 0.0) if trial % 3 else (5.0, 4.0, 3.0)
        ref = nb.obj_parse(p, e, 7, 0.25, threads=1, chunk_bytes=1 << 30)
        got, unsupported = _obj_rules_parse(lib, open(p, "rb").read(), e, 7, 0.25)
        assert not unsupported, trial
        for a, b in zip(ref[:4], got[:4]):
            assert a.shape == b.shape and a.tobytes() == b.tobytes(), trial
        assert got[4] == ref[4]
    # hand-written corner cases: empty file pieces, no trailing newline, blank lines, a face before its vertices, NUL inside a line
    cases = [b"", b"\n", b"v 1 2 3", b"v 1 2 3\nv 4 5 6\nv 7 8 10\nf 1 2 3", b"f 1 2 3\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n\n\n",
             b"v 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvt 0.5 0.25\nf 1/1 2/1 3/1 4/1\nf 1 2 3\nf 1//1 2//1 3//1\nvn 0 0 1\nf 2//1 3//1 4//1\n",
             b"v 0 0 0\nv 1 0 0\nv 0 1\x00 5\nf 1 2 3\n", b"v 0 0 0\r\nv 2 0 0\r\nv 0 2 0\r\nf 1/ 2/ 3/\r\nf 1// 2// 3//\r\nf 0 1 2 3\r\n",
             b"vt 1\nvtx 3 4\nvnormal 1 0 0\nv\t1 1 1\nv 2 1 1\nv 1 2 1\nf\t1 2 3 1 2\n", b"v 1e3 -2.5E-2 +.5\nv 1. 0 0\nv 0 1e+1 0\nf 1 2 3\n"]
    for text in cases:
        p = str(tmp_path / "case.obj")
        open(p, "wb").write(text)
        ref = nb.obj_parse(p, (1.0, 0.0, 0.0), 3, -0.5, threads=1, chunk_bytes=1 << 30)
        got, unsupported = _obj_rules_parse(lib, text, (1.0, 0.0, 0.0), 3, -0.5)
        assert not unsupported, text
        for a, b in zip(ref[:4], got[:4]):
            assert a.shape == b.shape and a.tobytes() == b.tobytes(), text
        assert got[4] == ref[4], text
    for text in (b"v 0.12345678901234567 0 0\n", b"v nan 0 0\n", b"v inf 0 0\n", b"v 0x10 0 0\n", b"v 1e40 0 0\n", b"vn 1 1 1e-30\n",
                 b"v 0 0 0\nf " + b" ".join(b"1" for _ in range(17)) + b"\n", b"v 0 0 0\nf 12345678901 1 1\n"):
        _, unsupported = _obj_rules_parse(lib, text, (0.0, 0.0, 0.0), 0, 0.0)
        assert unsupported, text
