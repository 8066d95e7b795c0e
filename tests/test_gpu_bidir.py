"""Parity of the bidirectional wavefront engine (VCM, SPPM, BDPT; through the C ABI) against the reference's own renders
(tests/golden/cornell_render_{vcm,sppm,bdpt}.npz: launch_VCM / launch_bidirectional of the unmodified reference on a B200,
64x64, 1024 iterations) and of its two multi-GPU decompositions against the single-GPU result.

Image parity is statistical (different RNG, float atomics): stated tolerances are
  * mean radiance within 3 % of the reference image's mean,
  * relMSE of the 4x4-block-averaged images <= 0.02 (both images carry 1024-iteration noise),
and every integrator is compared with ITS OWN fixture (the reference's integrators are biased differently by their clamps,
tests/golden/README.md)."""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INTEG = {"vcm": 3, "sppm": 4, "bdpt": 1}


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def block(img, k=4):
    h, w, c = img.shape
    return img.reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


@pytest.fixture(scope="module")
def cornell(scene_dirs):
    s = nb.Scene.from_config(os.path.join(scene_dirs["cornell"], "configs", "config.rendertron"))
    yield s
    s.close()


def _cfg(scene, w, h, integrator):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(scene.cfg))
    cfg.width, cfg.height, cfg.integrator = w, h, integrator
    return cfg


@pytest.mark.parametrize("name", ["vcm", "sppm", "bdpt"])
def test_against_reference_render(name, cornell):
    ref = np.load(os.path.join(GOLD, f"cornell_render_{name}.npz"))
    spp = int(ref["spp"])
    ref = ref["img"]
    cfg = _cfg(cornell, 64, 64, INTEG[name])
    cam = nb.camera_make(cfg)
    img, st = cornell.render(cfg, cam, 0, spp)
    img = img[..., :3] / spp
    assert np.isfinite(img).all() and (img >= 0).all()
    assert st.closestRays > 0 and st.paths == 2 * 64 * 64 * spp
    if name != "bdpt":
        assert st.photons > 0
    if name != "sppm":
        assert st.shadowRays > 0
    print(name, "mean ours", img.mean(), "ref", ref.mean(), "relMSE(4x4)", rel_mse(block(img), block(ref)))
    assert abs(img.mean() - ref.mean()) < 0.03 * ref.mean(), (img.mean(), ref.mean())
    assert rel_mse(block(img), block(ref)) < 0.02


# scenes with delta materials (dielectric water slab, glass dragon): fixtures of tests/golden/make_golden_scenes.py
DELTA_CASES = [("water", 12, "unidirectional", 0), ("water", 12, "bdpt", 1), ("water", 12, "vcm", 3), ("water", 12, "sppm", 4),
               ("rainbow", 8, "vcm", 3)]


@pytest.mark.parametrize("scene_name,detail,name,integ", DELTA_CASES)
def test_delta_scenes_against_reference_render(scene_name, detail, name, integ, tmp_path):
    """Caustic paths through specular interfaces: the reference's own render of the same stand-in scene is the oracle.
    Tolerances: mean radiance within 4 %, relMSE of 4x4-block averages <= 0.05 (caustic fixtures are noisier than Cornell's,
    and the reference's firefly clamp makes single pixels heavy-tailed), non-finite reference pixels masked."""
    path = os.path.join(GOLD, f"{scene_name}_render_{name}.npz")
    assert os.path.exists(path), f"{path} missing: run tests/golden/make_golden_scenes.py on the GPU box"
    z = np.load(path)
    spp, ref = int(z["spp"]), z["img"]
    h, w = ref.shape[:2]
    d = str(tmp_path / scene_name)
    nb.make_scene(scene_name, d, int(z["detail"]))
    scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
    try:
        cfg = _cfg(scene, w, h, integ)
        cam = nb.camera_make(cfg)
        img, st = scene.render(cfg, cam, 0, spp)
    finally:
        scene.close()
    img = img[..., :3] / spp
    assert np.isfinite(img).all() and (img >= 0).all()
    ok = np.isfinite(ref).all(axis=2)
    assert ok.mean() > 0.99
    ref = np.where(ok[..., None], ref, 0.0); img = np.where(ok[..., None], img, 0.0)
    hb, wb = h - h % 4, w - w % 4
    r = rel_mse(block(img[:hb, :wb]), block(ref[:hb, :wb]))
    print(scene_name, name, "mean ours", img.mean(), "ref", ref.mean(), "relMSE(4x4)", r)
    assert abs(img.mean() - ref.mean()) < 0.04 * ref.mean(), (img.mean(), ref.mean())
    assert r < 0.05


@pytest.mark.parametrize("integ", [3, 1, 4])
def test_iteration_striping_composes(integ, cornell):
    """novum_render_device stripes whole iterations over ranks: the ranks' frames sum to the single-GPU frame."""
    cfg = _cfg(cornell, 40, 24, integ)
    cam = nb.camera_make(cfg)
    full, _ = cornell.render(cfg, cam, 0, 6)
    r0, _ = cornell.render(cfg, cam, 0, 6, rank=0, world=2)
    r1, _ = cornell.render(cfg, cam, 0, 6, rank=1, world=2)
    assert np.allclose(full, r0 + r1, rtol=2e-3, atol=1e-4)
    a, _ = cornell.render(cfg, cam, 0, 3)
    b, _ = cornell.render(cfg, cam, 3, 3)
    assert np.allclose(full, a + b, rtol=2e-3, atol=1e-4)


@pytest.mark.parametrize("integ,world", [(3, 2), (3, 3), (4, 2)])
def test_peer_write_exchange_equals_single_gpu(integ, world, cornell):
    """Peer-write form of the tile-sharded exchange: every session's light pass stores its photons (and its count) straight
    into the other sessions' slab buffers — attached by device pointer here, by CUDA IPC handle across processes — so no copy
    happens between light pass and grid build. The image must equal the single-GPU image."""
    import torch
    cfg = _cfg(cornell, 48, 30, integ)
    cam = nb.camera_make(cfg)
    iters = 4
    full, _ = cornell.render(cfg, cam, 0, iters)
    sessions = [nb.VcmSession(cornell, cfg, cam, rank=r, world=world) for r in range(world)]
    frames = [torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda") for _ in range(world)]
    sync = nb.lib().novum_sync
    try:
        with pytest.raises(nb.NovumError):
            sessions[0].peer_write(True)              # refused until every peer is attached
        for r, s in enumerate(sessions):
            for q, o in enumerate(sessions):
                if q != r:
                    s.attach_pointer(q, o.base)
            s.peer_write(True)
        for it in range(iters):
            for r, s in enumerate(sessions):
                s.light_pass(it, frames[r].data_ptr())
            sync()                                    # (across processes: a stream-ordered 4-byte all-reduce)
            counts = [int(s.slabs_tensor()[q * s.slab_bytes:q * s.slab_bytes + 4].view(torch.int32)[0]) for s in sessions for q in range(world)]
            assert counts[:world] * world == counts   # every session sees the same per-rank photon counts
            for r, s in enumerate(sessions):
                s.build_grid(it)
                s.eye_pass(it, frames[r].data_ptr())
            sync()
        total = sum(f.cpu().numpy() for f in frames)
        assert sessions[0].stats().photons > 0
    finally:
        for s in sessions:
            s.close()
    assert np.allclose(total, full, rtol=2e-3, atol=1e-4)


@pytest.mark.parametrize("integ,world", [(3, 2), (3, 3), (4, 2), (1, 2)])
def test_tile_sharded_phases_equal_single_gpu(integ, world, cornell):
    """The phase API with `world` ranks emulated on one GPU (slabs copied between the sessions where NCCL's all-gather
    would move them) reproduces the single-GPU image: walks are keyed by (pixel, iteration), photon order is free."""
    import torch
    cfg = _cfg(cornell, 48, 30, integ)
    cam = nb.camera_make(cfg)
    iters = 4
    full, st_full = cornell.render(cfg, cam, 0, iters)
    sessions = [nb.VcmSession(cornell, cfg, cam, rank=r, world=world) for r in range(world)]
    frames = [torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda") for _ in range(world)]
    sync = nb.lib().novum_sync
    try:
        for it in range(iters):
            for r, s in enumerate(sessions):
                s.light_pass(it, frames[r].data_ptr())
            sync()
            for r, s in enumerate(sessions):          # "all-gather": every session receives every other rank's slab
                for q, o in enumerate(sessions):
                    if q != r:
                        s.slabs_tensor()[q * s.slab_bytes:(q + 1) * s.slab_bytes].copy_(o.my_slab_tensor())
            torch.cuda.synchronize()
            for r, s in enumerate(sessions):
                s.build_grid(it)
                s.eye_pass(it, frames[r].data_ptr())
            sync()
        total = sum(f.cpu().numpy() for f in frames)
        photons = sessions[0].stats().photons
    finally:
        for s in sessions:
            s.close()
    if integ != 1:
        assert photons > 0
    assert np.allclose(full, total, rtol=2e-3, atol=1e-4), float(np.abs(full - total).max())


def test_sppm_uses_only_merging(cornell):
    cfg = _cfg(cornell, 32, 32, 4)
    cam = nb.camera_make(cfg)
    img, st = cornell.render(cfg, cam, 0, 8)
    assert st.shadowRays == 0 and st.photons > 0 and img[..., :3].sum() > 0


def test_strategy_toggles(cornell):
    """BDPT with every strategy off renders black; each strategy alone is non-negative and no brighter than unweighted sum."""
    cfg = _cfg(cornell, 32, 32, 1)
    cam = nb.camera_make(cfg)
    cfg.bdptLightTrace = cfg.bdptNee = cfg.bdptNaive = cfg.bdptConnection = 0
    img, _ = cornell.render(cfg, cam, 0, 4)
    assert not img[..., :3].any()
    parts = []
    for k in ("bdptLightTrace", "bdptNee", "bdptNaive", "bdptConnection"):
        setattr(cfg, k, 1)
        img, _ = cornell.render(cfg, cam, 0, 64)
        setattr(cfg, k, 0)
        assert (img[..., :3] >= 0).all() and img[..., :3].sum() > 0, k
        parts.append(img[..., :3])
    cfg.bdptLightTrace = cfg.bdptNee = cfg.bdptNaive = cfg.bdptConnection = 1
    full, _ = cornell.render(cfg, cam, 0, 64)
    # same seeds, same walks: the MIS-weighted strategies add up to the full image (NEE draws shift the eye RNG stream, so only in the mean)
    assert abs(sum(p.mean() for p in parts) - full[..., :3].mean()) < 0.05 * full[..., :3].mean()


def test_full_size_vcm_decompositions_agree(tmp_path):
    """The headline workload's scene at full tessellation (871k triangles), 1920x1080, 2 VCM iterations: the striped and the
    row-sharded (peer-write) decompositions both reproduce the single-GPU frame; photon counts and merge statistics are sane."""
    import torch
    d = str(tmp_path / "rainbow")
    nb.make_scene("rainbow", d, 100)
    scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
    try:
        cfg = _cfg(scene, 1920, 1080, 3)
        cam = nb.camera_make(cfg)
        full, st = scene.render(cfg, cam, 0, 2)
        assert st.photons > 1_000_000 and st.mergeQueries > 1_000_000 and st.mergeAccepted > st.mergeQueries
        assert st.mergeCandidates >= st.mergeAccepted
        assert np.isfinite(full).all() and full[..., :3].mean() > 0
        r0, _ = scene.render(cfg, cam, 0, 2, rank=0, world=2)
        r1, _ = scene.render(cfg, cam, 0, 2, rank=1, world=2)
        # float atomics into the same pixel from thousands of photons: compare 8x8 block sums (per-pixel order noise ~1e-3)
        def blocks(a):
            return a[..., :3].reshape(135, 8, 240, 8, 3).sum(axis=(1, 3))
        assert np.allclose(blocks(r0 + r1), blocks(full), rtol=2e-3, atol=1e-3)
        world = 2
        sessions = [nb.VcmSession(scene, cfg, cam, rank=r, world=world) for r in range(world)]
        frames = [torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda") for _ in range(world)]
        try:
            for r, s in enumerate(sessions):
                for q, o in enumerate(sessions):
                    if q != r:
                        s.attach_pointer(q, o.base)
                s.peer_write(True)
            for it in range(2):
                for r, s in enumerate(sessions):
                    s.light_pass(it, frames[r].data_ptr())
                nb.lib().novum_sync()
                for r, s in enumerate(sessions):
                    s.build_grid(it)
                    s.eye_pass(it, frames[r].data_ptr())
                nb.lib().novum_sync()
            total = sum(f.cpu().numpy() for f in frames)
        finally:
            for s in sessions:
                s.close()
        assert np.allclose(blocks(total), blocks(full), rtol=2e-3, atol=1e-3)
    finally:
        scene.close()


@pytest.mark.parametrize("integ,world,lanes", [(3, 1, 2), (3, 2, 2), (3, 3, 1), (4, 2, 2), (1, 2, 2)])
def test_tile_renderer_lanes_equal_single_gpu(integ, world, lanes, cornell):
    """novum_tiles_*: `world` ranks x `lanes` iterations in flight (ranks are threads of this process on one GPU here; across
    processes the slab buffers travel as CUDA IPC handles), photons exchanged by peer stores inside the light pass, rendezvous by
    stamps in the slab headers that a device-side wait kernel polls. The sum of the ranks' frames is the single-GPU frame."""
    import threading
    import torch
    cfg = _cfg(cornell, 48, 30, integ)
    cam = nb.camera_make(cfg)
    iters = 7
    full, st_full = cornell.render(cfg, cam, 2, iters)
    ranks = [nb.TileRank(cornell, cfg, cam, rank=r, world=world, lanes=lanes) for r in range(world)]
    frames = [torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda") for _ in range(world)]
    stats, errs = [None] * world, []
    try:
        for r, t in enumerate(ranks):
            for q, o in enumerate(ranks):
                if q != r:
                    t.attach_pointers(q, o.bases())

        def run(r, first, count):
            try:
                stats[r] = ranks[r].render(first, count, frames[r].data_ptr())
            except Exception as ex:      # noqa: BLE001
                errs.append(ex)

        for first, count in ((2, 3), (5, 4)):          # two calls: the lanes' buffer parity and stamps carry over
            th = [threading.Thread(target=run, args=(r, first, count)) for r in range(world)]
            [t.start() for t in th]
            [t.join() for t in th]
            assert not errs, errs
        total = sum(f.cpu().numpy() for f in frames)
    finally:
        for t in ranks:
            t.close()
    assert np.allclose(total, full, rtol=2e-3, atol=1e-4), float(np.abs(total - full).max())
    assert sum(s.paths for s in stats) == 2 * cfg.width * cfg.height * 4       # the last call's 4 iterations
    if integ != 1:
        assert stats[0].photons > 0
