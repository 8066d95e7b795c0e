"""Image parity of the wavefront integrators (through the C ABI) against
  (a) the CPU oracle's restatement of Li_unidirectional / Li_naive_unidirectional with the SAME per-(pixel, sample)
      random streams — the two implementations then follow the same paths, so images agree far inside Monte-Carlo noise;
  (b) the reference's own high-spp renders (tests/golden/cornell_render_*.npz, produced on a B200 by oracle/_ref),
      within a stated relMSE tolerance and a per-pixel-mean t-test (different RNG: statistical parity only).
"""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


@pytest.fixture(scope="module")
def cornell(scene_dirs):
    s = nb.Scene.from_config(os.path.join(scene_dirs["cornell"], "configs", "config.rendertron"))
    yield s
    s.close()


def _small_cfg(scene, w, h, integrator=0):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(scene.cfg))
    cfg.width, cfg.height, cfg.integrator = w, h, integrator
    return cfg


@pytest.mark.parametrize("naive", [False, True])
def test_same_streams_as_oracle(naive, cornell, oracle_scenes):
    cfg = _small_cfg(cornell, 48, 48, 2 if naive else 0)
    cam = nb.camera_make(cfg)
    spp = 16
    img, st = cornell.render(cfg, cam, 0, spp)
    ref, counts = oracle_scenes["cornell"].render_unidirectional(bytes(cam), 48, 48, cfg.maxDepth, 0, spp, naive=naive)
    assert np.isfinite(img).all() and (img[..., :3] >= 0).all()
    a, b = img[..., :3] / spp, ref[..., :3] / spp
    # identical random streams: only a few paths diverge (rsqrtf / sincosf vs libm last-bit differences at a branch)
    close = np.isclose(a, b, rtol=2e-3, atol=2e-4).all(axis=2)
    assert close.mean() > 0.97, close.mean()
    assert abs(a.mean() - b.mean()) < 0.01 * b.mean()
    assert rel_mse(a, b) < 2e-3
    if not naive:
        # ray counts: the oracle traces a shadow ray for every NEE sample, we skip those with a non-positive light pdf or
        # zero contribution, and stop paths whose throughput is exactly zero — never more rays than the reference
        assert st.closestRays <= counts[0] * 1.01 and st.closestRays >= 0.9 * counts[0]
        assert st.shadowRays <= counts[1]
    assert st.paths == 48 * 48 * spp


def test_sample_ranges_and_shards_compose(cornell):
    cfg = _small_cfg(cornell, 40, 24)
    cam = nb.camera_make(cfg)
    full, _ = cornell.render(cfg, cam, 0, 8)
    a, _ = cornell.render(cfg, cam, 0, 4)
    b, _ = cornell.render(cfg, cam, 4, 4)
    assert np.allclose(full, a + b, rtol=1e-4, atol=1e-5)              # passes are keyed by (pixel, sample): order-free
    r0, _ = cornell.render(cfg, cam, 0, 8, rank=0, world=2)
    r1, _ = cornell.render(cfg, cam, 0, 8, rank=1, world=2)
    assert np.allclose(full, r0 + r1, rtol=1e-4, atol=1e-5)            # two ranks = the single-GPU image
    again, _ = cornell.render(cfg, cam, 0, 8)
    assert np.allclose(full, again, rtol=1e-5, atol=1e-6)              # only float-atomic ordering differs run to run


def test_render_resumes_from_a_checkpoint(cornell, tmp_path):
    """stop after 5 passes, write the sums, load them in a 'new process', continue at sample 5 - the result is
    the uninterrupted 12-pass render (up to float-atomic ordering) and the saved frame finishes to the same picture."""
    cfg = _small_cfg(cornell, 56, 40)
    cam = nb.camera_make(cfg)
    full, _ = cornell.render(cfg, cam, 0, 12, seed=777)
    part, _ = cornell.render(cfg, cam, 0, 5, seed=777)
    ck = str(tmp_path / "cornell.novumacc")
    nb.accum_save(ck, part, 5, seed=777, integrator=cfg.integrator)
    sums, done, seed = nb.accum_load(ck)
    assert done == 5 and seed == 777 and np.array_equal(sums.view(np.uint32), part.view(np.uint32))
    rest, _ = cornell.render(cfg, cam, done, 12 - done, seed=seed)
    assert np.allclose(full, sums + rest, rtol=1e-4, atol=1e-5)
    a = nb.frame_finish(full, None, 12, nb.FRAME_FINAL, True)
    b = nb.frame_finish(sums + rest, None, 12, nb.FRAME_FINAL, True)
    assert np.mean(a != b) < 1e-3 and np.abs(a.astype(int) - b.astype(int)).max() <= 1      # a byte may sit on a rounding step


def test_cli_checkpointed_run_resumes(scene_dirs, tmp_path):
    """The command-line driver with --checkpoint: a run cut short at 5 passes and started again reaches the same sums as an
    uninterrupted run of 8 (the tool renders `--every` passes at a time and rewrites the checkpoint after each chunk)."""
    import subprocess
    tool = os.path.join(os.path.dirname(os.path.abspath(nb.__file__)), "novum")
    if not os.path.exists(tool):
        pytest.skip("command-line driver not built")
    cfgp = os.path.join(scene_dirs["cornell"], "configs", "config.rendertron")
    cwd = scene_dirs["cornell"]

    def run(*extra):
        r = subprocess.run([tool, "--config", cfgp, "--seed", "4242", *extra], cwd=cwd, capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        return r.stdout

    ck = str(tmp_path / "cli.novumacc")
    run("--spp", "8", "--out", str(tmp_path / "full"), "--raw")
    run("--spp", "5", "--out", str(tmp_path / "cut"), "--checkpoint", ck, "--every", "3")
    assert nb.accum_info(ck)["sample_count"] == 5 and nb.accum_info(ck)["seed"] == 4242
    out = run("--spp", "8", "--out", str(tmp_path / "resumed"), "--checkpoint", ck, "--every", "3", "--raw")
    assert "resuming at sample 5 of 8" in out and nb.accum_info(ck)["sample_count"] == 8
    full = np.fromfile(str(tmp_path / "full.raw"), np.float32)
    resumed = np.fromfile(str(tmp_path / "resumed.raw"), np.float32)
    assert full.size == resumed.size and np.allclose(full, resumed, rtol=1e-4, atol=1e-5)
    sums, n, _ = nb.accum_load(ck)
    assert n == 8 and np.array_equal(sums.ravel().view(np.uint32), resumed.view(np.uint32))
    assert os.path.getsize(str(tmp_path / "resumed_progress.bmp")) == os.path.getsize(str(tmp_path / "resumed.bmp"))


def test_odd_resolution_and_zero_samples(cornell):
    cfg = _small_cfg(cornell, 37, 19)
    cam = nb.camera_make(cfg)
    img, st = cornell.render(cfg, cam, 0, 3)
    assert st.paths == 37 * 19 * 3
    # 37x19 widens the horizontal field of view past the box's side walls: only the central columns must be lit
    assert (img[:, 12:25, :3].sum(axis=2) > 0).mean() > 0.7
    img0, st0 = cornell.render(cfg, cam, 0, 0)
    assert st0.paths == 0 and not img0.any()


def test_naive_to_nee_ratio_matches_the_reference(cornell):
    """The reference's two unidirectional integrators do NOT agree with each other on this scene (pdf floor 0.01 at
    deviceCode.cu:481 only in the NEE one; tests/golden/README.md: 0.249 vs 0.274): we must reproduce that ratio."""
    ref_nee = np.load(os.path.join(GOLD, "cornell_render_unidirectional.npz"))["img"]
    ref_nai = np.load(os.path.join(GOLD, "cornell_render_naive_unidirectional.npz"))["img"]
    ref_ratio = ref_nai[np.isfinite(ref_nai).all(axis=2)].mean() / ref_nee[np.isfinite(ref_nee).all(axis=2)].mean()
    cfg = _small_cfg(cornell, 64, 64)
    cam = nb.camera_make(cfg)
    nee, _ = cornell.render(cfg, cam, 0, 256)
    cfg.integrator = 2
    nai, _ = cornell.render(cfg, cam, 0, 1024)
    ratio = (nai[..., :3].mean() / 1024) / (nee[..., :3].mean() / 256)
    assert abs(ratio - ref_ratio) < 0.03 * ref_ratio, (ratio, ref_ratio)


@pytest.mark.parametrize("integ,name", [(0, "unidirectional"), (2, "naive_unidirectional")])
def test_against_reference_render(integ, name, cornell):
    """relMSE(ours@256spp, reference@8192spp) must not exceed what 256 spp of noise explains, and the per-pixel means
    must pass a t-test at the Bonferroni-corrected level (unbiased with respect to the reference's own image)."""
    p = os.path.join(GOLD, f"cornell_render_{name}.npz")
    if not os.path.exists(p):
        pytest.skip("reference render not generated yet")
    ref = np.load(p)["img"]
    ok = np.isfinite(ref).all(axis=2)          # the reference's own render has a few non-finite pixels (it paints them magenta, main.cu:863)
    ref = np.where(ok[..., None], ref, 0.0)
    cfg = _small_cfg(cornell, 64, 64, integ)
    cam = nb.camera_make(cfg)
    batches = []
    per = 16 if integ == 0 else 256       # the naive estimator is heavy-tailed (most samples are zero): it needs far more per batch
    for b in range(16):
        img, _ = cornell.render(cfg, cam, b * per, per)
        batches.append(img[..., :3] / per)
    batches = np.array(batches)
    mean = np.where(ok[..., None], batches.mean(0), 0.0)
    assert np.isfinite(batches).all()
    assert abs(mean.mean() - ref.mean()) < 0.02 * ref.mean()
    # stated tolerance at 256 (NEE) / 4096 (naive) spp, 64x64: relMSE of 4x4 block means <= 0.02 (the naive integrator is too noisy per pixel), and per pixel <= 0.05 for NEE
    blk = lambda a: a.reshape(16, 4, 16, 4, 3).mean(axis=(1, 3))
    assert rel_mse(blk(mean), blk(ref)) < 0.02
    if integ == 0:
        assert rel_mse(mean, ref) < 0.05
    # standard error of (ours - reference): our batch spread, plus the reference's own noise at its 8192 spp
    se = batches.std(0, ddof=1) / np.sqrt(len(batches)) * np.sqrt(1.0 + 16 * per / 8192.0) + 1e-4
    tstat = (np.abs(mean - ref) / se)[ok]
    # 12288 tests; t(15) two-sided at 1e-6 is ~8.0: allow a handful of fireflies-in-the-reference outliers
    assert (tstat > 8.0).mean() < 2e-3, float((tstat > 8.0).mean())
