"""The drop-in boundary driven by the REFERENCE'S OWN host code: oracle/_ref/novum_dropin is the reference's unmodified
main.cu + imageUtil.cu with deviceCode.cu replaced by integration/novum_shim.cu, linked against libnovum_b200.so. Every call below
goes launch_* (deviceCode.cuh:8-18) -> shim -> novum_launch_* / novum_update_constants.

  * `render`: the buffers initRender allocates (main.cu:560-780), all five integrators on the Cornell stand-in, held to the same
    reference-made goldens and the same bars as tests/test_gpu_render.py / tests/test_gpu_bidir.py;
  * `--repeat`: the second launch on the same buffers is a cache hit (no scene build);
  * `initrender`: the reference's initRender itself, twice (main.cu:925-933 renders 75 times, moving the lights each time): config,
    OBJ files, CPU BVH2, uploads, launch, final host pass and renders/<name><i>.bmp - the second render rebuilds the scene part and
    keeps the engines.
The binary is built where /root/reference exists (oracle/build_ref.sh) and travels to the GPU box."""
import json
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

import cudapathtracer_b200 as nb

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(REPO, "tests", "golden")
DROPIN = os.path.join(REPO, "oracle", "_ref", "novum_dropin")

needs_dropin = pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/novum_dropin not built (needs /root/reference at build time)")


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def block(img, k=4):
    h, w, c = img.shape
    return img.reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


def set_config(scene_dir, **kv):
    p = os.path.join(scene_dir, "configs", "config.rendertron")
    lines = open(p).read().splitlines()
    for k, v in kv.items():
        assert any(l.startswith(k + ":") for l in lines), k
        lines = [f"{k}: {v}" if l.startswith(k + ":") else l for l in lines]
    open(p, "w").write("\n".join(lines) + "\n")


def run_dropin(args, cwd=None):
    r = subprocess.run([DROPIN] + args, capture_output=True, text=True, timeout=900, cwd=cwd)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "RENDER ERROR" not in r.stderr, r.stderr[-2000:]
    info = json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", r.stdout).group(1)) if "NOVUM_REF_JSON" in r.stdout else {}
    cache = json.loads(re.search(r"NOVUM_DROPIN_CACHE (\{.*\})", r.stdout).group(1))
    return info, cache, r.stdout


@pytest.fixture()
def cornell_dir(tmp_path):
    d = str(tmp_path / "cornell")
    nb.make_scene("cornell", d, 100)
    return d


# (config name, golden file, spp rendered here, mean tolerance, block-relMSE bar) - the bars of test_gpu_render / test_gpu_bidir
CASES = [("UNIDIRECTIONAL", "unidirectional", 256, 0.02, 0.02), ("NAIVE_UNIDIRECTIONAL", "naive_unidirectional", 4096, 0.02, 0.02),
         ("BDPT", "bdpt", 1024, 0.03, 0.02), ("VCM", "vcm", 1024, 0.03, 0.02), ("SPPM", "sppm", 1024, 0.03, 0.02)]


@needs_dropin
@pytest.mark.parametrize("integ,gold,spp,mean_tol,bar", CASES)
def test_reference_main_drives_the_library(integ, gold, spp, mean_tol, bar, cornell_dir, tmp_path):
    ref = np.load(os.path.join(GOLD, f"cornell_render_{gold}.npz"))["img"]
    ok = np.isfinite(ref).all(axis=2)
    ref = np.where(ok[..., None], ref, 0.0)
    set_config(cornell_dir, width=64, height=64, Integrator=integ, **{"Sample Count": spp})
    raw = str(tmp_path / "out.raw")
    info, cache, _ = run_dropin(["render", cornell_dir, "--out", raw])
    assert info["cuda_error"] == 0 and info["w"] == 64 and info["spp"] == spp
    assert cache["launches"] == 1 and cache["scene_builds"] == 1
    img = np.fromfile(raw, np.float32).reshape(64, 64, 4)[..., :3] / spp
    assert np.isfinite(img).all() and (img >= 0).all()
    img = np.where(ok[..., None], img, 0.0)
    print(integ, "mean dropin", img.mean(), "reference", ref.mean(), "relMSE(4x4)", rel_mse(block(img), block(ref)))
    assert abs(img.mean() - ref.mean()) < mean_tol * ref.mean(), (img.mean(), ref.mean())
    assert rel_mse(block(img), block(ref)) < bar


@needs_dropin
def test_second_launch_on_the_same_buffers_builds_nothing(cornell_dir, tmp_path):
    set_config(cornell_dir, width=96, height=64, Integrator="VCM", **{"Sample Count": 8})
    raw = str(tmp_path / "out.raw")
    info, cache, _ = run_dropin(["render", cornell_dir, "--out", raw, "--repeat", "3"])
    assert cache == {"launches": 3, "hits": 2, "rebinds": 0, "scene_builds": 1, "engines_moved": 0, "progressive_saves": 0}
    each = info["launch_ms_each"]
    assert len(each) == 3 and each[1] < each[0] and each[2] < each[0]          # first launch pays BVH8 + engine allocation
    # the same frame through the scene-object API (same seed 103033, counter-based RNG): equal up to float-atomic order
    img = np.fromfile(raw, np.float32).reshape(64, 96, 4)
    scene = nb.Scene.from_config(os.path.join(cornell_dir, "configs", "config.rendertron"))
    try:
        cfg = scene.cfg
        cam = nb.camera_make(cfg)
        want, _ = scene.render(cfg, cam, 0, 8)
    finally:
        scene.close()
    assert np.allclose(img, want, rtol=2e-3, atol=1e-4)


@needs_dropin
def test_reference_initrender_twice(cornell_dir):
    """initRender itself (main.cu:235-923), renders 0 and 1: the light moves by -0.01 (main.cu:476-478), so the second call is new
    content behind the same launcher: scene part rebuilt, engines kept. The .bmp files come from the reference's own host pass."""
    set_config(cornell_dir, width=80, height=48, Integrator="BDPT", **{"Sample Count": 32})
    _, cache, out = run_dropin(["initrender", cornell_dir, "2"])
    runs = [json.loads(m) for m in re.findall(r"NOVUM_DROPIN_INITRENDER (\{.*\})", out)]
    assert [r["rc"] for r in runs] == [0, 0]
    assert cache["launches"] == 2 and cache["scene_builds"] == 2 and cache["engines_moved"] == 1 and cache["hits"] == 0
    for i in range(2):
        bmp = os.path.join(cornell_dir, "renders", f"cornell{i}.bmp")
        data = np.fromfile(bmp, np.uint8)
        assert data.size == 54 + 48 * nb.bmp_row_bytes(80) and data[54:].mean() > 20      # a lit picture, not black
    a = np.fromfile(os.path.join(cornell_dir, "renders", "cornell0.bmp"), np.uint8)[54:].astype(np.float32)
    b = np.fromfile(os.path.join(cornell_dir, "renders", "cornell1.bmp"), np.uint8)[54:].astype(np.float32)
    assert abs(a.mean() - b.mean()) < 0.05 * a.mean()           # same scene up to a light 1 cm lower


@needs_dropin
def test_draw_path_constant_is_refused_not_dropped(cornell_dir, tmp_path):
    set_config(cornell_dir, width=32, height=32, Integrator="BDPT", BDPT_DRAWPATH="true", **{"Sample Count": 2})
    r = subprocess.run([DROPIN, "render", cornell_dir, "--out", str(tmp_path / "o.raw")], capture_output=True, text=True, timeout=300)
    assert "RENDER ERROR in launch_bidirectional: unsupported" in r.stderr


def test_launch_entry_points_validate_arguments():
    L = nb.lib()
    st = nb.LaunchCacheStats()
    assert L.novum_launch_cache_stats(None) == -1
    assert L.novum_launch_cache_stats(st) == 0
    L.novum_launch_cache_clear()
