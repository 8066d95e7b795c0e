"""The OBJ front end on the device (csrc/obj_device.cu) against the host loader (novum_obj_parse = main.cu:936-1068
restated): the arrays must be identical bit for bit - positions, generated face normals, the shared zero uv, dropped triangles, light
indices - for adversarial files, for several files appended to one mesh, and for a full-size generated dragon; a scene built from the
device arrays is the scene built from the host arrays."""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from scene_utils import random_rays
from test_host import _adversarial_obj

pytestmark = pytest.mark.gpu


def _same(ref, got, what):
    for name, a, b in zip(("pos", "nrm", "uv", "tris"), ref[:4], got[:4]):
        assert a.shape == b.shape, (what, name, a.shape, b.shape)
        assert a.tobytes() == b.tobytes(), (what, name)
    assert ref[4] == got[4], what


def test_adversarial_files_equal_host_loader(built, tmp_path):
    rng = np.random.default_rng(4242)
    for trial in range(10):
        p = str(tmp_path / f"adv{trial}.obj")
        _adversarial_obj(p, rng, n_lines=int(rng.integers(300, 3000)), device_ok=True)
        e = (0.0, 0.0, 0.0) if trial % 3 else (5.0, 4.0, 3.0)
        ref = nb.obj_parse(p, e, 7, 0.25, threads=1, chunk_bytes=1 << 30)
        m = nb.ObjDeviceMesh()
        if trial % 2:
            m.append_file(p, e, 7, 0.25)
        else:
            m.append_text(open(p, "rb").read(), e, 7, 0.25)
        _same(ref, m.download(), trial)
        assert m.timings()["launches"] == 12
        m.close()


def test_corner_cases(built, tmp_path):
    cases = [b"\n", b"v 1 2 3", b"v 1 2 3\nv 4 5 6\nv 7 8 10\nf 1 2 3", b"f 1 2 3\nv 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n\n\n",
             b"v 0 0 0\nv 1 0 0\nv 0 1 0\nv 1 1 0\nvt 0.5 0.25\nf 1/1 2/1 3/1 4/1\nf 1 2 3\nf 1//1 2//1 3//1\nvn 0 0 1\nf 2//1 3//1 4//1\n",
             b"v 0 0 0\nv 1 0 0\nv 0 1\x00 5\nf 1 2 3\n", b"v 0 0 0\r\nv 2 0 0\r\nv 0 2 0\r\nf 1/ 2/ 3/\r\nf 1// 2// 3//\r\nf 0 1 2 3\r\n",
             b"vt 1\nvtx 3 4\nvnormal 1 0 0\nv\t1 1 1\nv 2 1 1\nv 1 2 1\nf\t1 2 3 1 2\n", b"v 1e3 -2.5E-2 +.5\nv 1. 0 0\nv 0 1e+1 0\nf 1 2 3\n",
             b"#" + b"x" * 5000 + b"\nv 0 0 0\nv 1 0 0\nv 0 1 0\n" + b"\n" * 4100 + b"f 1 2 3\n"]
    for k, text in enumerate(cases):
        p = str(tmp_path / "case.obj")
        open(p, "wb").write(text)
        ref = nb.obj_parse(p, (1.0, 0.0, 0.0), 3, -0.5, threads=1, chunk_bytes=1 << 30)
        m = nb.ObjDeviceMesh()
        m.append_text(text, (1.0, 0.0, 0.0), 3, -0.5)
        _same(ref, m.download(), k)
        m.close()
    m = nb.ObjDeviceMesh()
    m.append_text(b"", (0.0, 0.0, 0.0), 0, 0.0)
    assert m.counts() == (0, 0, 0, 0, 0)
    with pytest.raises(nb.NovumError):
        m.append_file(str(tmp_path / "missing.obj"))
    m.close()


def test_unsupported_spellings_are_refused_and_leave_the_mesh_unchanged(built):
    m = nb.ObjDeviceMesh()
    good = b"v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n"
    m.append_text(good)
    before = m.download()
    for text in (b"v 0.12345678901234567 0 0\n", b"v nan 0 0\n", b"v 0 0 0\nvn inf 0 0\n", b"vt 0x10 0\n", b"v 1e40 0 0\n",
                 b"v 0 0 0\nf " + b" ".join(b"1" for _ in range(17)) + b"\n"):
        with pytest.raises(nb.NovumError) as ei:
            m.append_text(text)
        assert "(code -5)" in str(ei.value), str(ei.value)
        _same(before, m.download(), text)
    m.append_text(good)            # still usable; the second file's face names ITS three positions (file-local index + positions before it)
    assert m.counts()[0] == 6 and m.counts()[3] == 2
    m.close()


def test_scene_from_config_device_path_equals_host_path(built, scene_dirs):
    """novum_scene_load_config parses on the device by default; with NOVUM_OBJ_DEVICE=0 it parses on the host. Same scene either way:
    arrays (several meshes appended, one of them emissive), light list, BVH hits."""
    for name in ("cornell", "dragon", "water"):
        cfg = os.path.join(scene_dirs[name], "configs", "config.rendertron")
        os.environ["NOVUM_OBJ_DEVICE"] = "0"
        try:
            host = nb.Scene.from_config(cfg)
        finally:
            del os.environ["NOVUM_OBJ_DEVICE"]
        dev = nb.Scene.from_config(cfg)
        a, b = host.download(), dev.download()
        for k, (x, y) in enumerate(zip(a, b)):
            assert x.shape == y.shape and x.tobytes() == y.tobytes(), (name, k)
        assert len(a[4]) > 0
        info = host.info()
        rays = random_rays(info["min"], info["max"], 20000, np.random.default_rng(5))
        ha, hb = host.trace_closest(rays), dev.trace_closest(rays)
        assert np.array_equal(ha["tri"], hb["tri"]) and np.array_equal(ha["t"].view(np.uint32), hb["t"].view(np.uint32))
        host.close(); dev.close()


def test_config_with_unsupported_file_falls_back_to_host_loader(built, tmp_path):
    d = str(tmp_path / "c")
    nb.make_scene("cornell", d, 100)
    cfg = os.path.join(d, "configs", "config.rendertron")
    ref = nb.Scene.from_config(cfg)
    a = ref.download()
    # respell one coordinate of one mesh with 17 significant digits (same value): the device rules refuse the file, the host loader reads it
    objs = [os.path.join(r, f) for r, _, fs in os.walk(d) for f in fs if f.endswith(".obj")]
    assert objs
    text = open(objs[0]).read().splitlines()
    for i, l in enumerate(text):
        if l.startswith("v ") and float(l.split()[1]) != 0.0:
            x, y, z = l.split()[1:4]
            text[i] = "v %.17f %s %s" % (float(x), y, z)          # e.g. 0.50000000000000000: 17 significant digits, the same value
            break
    else:
        raise AssertionError("no vertex with a non-zero x")
    open(objs[0], "w").write("\n".join(text) + "\n")
    got = nb.Scene.from_config(cfg)
    b = got.download()
    for k, (x, y) in enumerate(zip(a, b)):
        assert x.shape == y.shape and x.tobytes() == y.tobytes(), k
    ref.close(); got.close()


def test_full_size_dragon_equals_host_loader_and_is_faster(built, tmp_path):
    """The 871k-triangle stand-in dragon (33 MB of text): identical arrays; prints the timings the docs quote."""
    import time
    d = str(tmp_path / "dragon")
    nb.make_scene("dragon", d, 100)
    p = os.path.join(d, "scenedata", "dragon_vrip.obj")
    size = os.path.getsize(p)
    assert size > 20e6
    t0 = time.time(); ref = nb.obj_parse(p, (0.0, 0.0, 0.0), 5, 0.0); host_ms = (time.time() - t0) * 1e3
    t0 = time.time(); ref = nb.obj_parse(p, (0.0, 0.0, 0.0), 5, 0.0); host_ms = min(host_ms, (time.time() - t0) * 1e3)
    best = None
    for rep in range(3):
        m = nb.ObjDeviceMesh()
        t0 = time.time(); m.append_file(p, (0.0, 0.0, 0.0), 5, 0.0); wall = (time.time() - t0) * 1e3
        t = m.timings(); t["wall_ms"] = wall
        if best is None or wall < best["wall_ms"]:
            best = t
        if rep == 0:
            _same(ref, m.download(), "dragon")
            assert len(ref[3]) > 800000
        m.close()
    print(f"\nOBJ {size / 1e6:.1f} MB, {len(ref[3])} triangles: host loader {host_ms:.1f} ms ({os.cpu_count()} cores); device parse wall {best['wall_ms']:.1f} ms "
          f"(read+copy {best['upload_ms']:.1f} ms, kernels {best['kernel_ms']:.2f} ms in {best['launches']} launches, "
          f"{size / 1e6 / best['kernel_ms']:.1f} GB/s of text)")
    assert best["kernel_ms"] < host_ms
