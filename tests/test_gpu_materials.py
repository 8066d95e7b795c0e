"""GPU parity of the BSDF classes and the shadow-ray branch that Cornell / water / glass-dragon scenes never reach, against fixtures made
by the REFERENCE'S OWN device code on the `materials` stand-in (tests/golden/make_golden_materials.py): GGX metals
(reflectors.cuh:129-178), delta mirror (:59-76), leaf lobes (:420-545), bilinear texture fetch (:371-417), nested smooth dielectrics with
Beer-Lambert absorption and three priorities (deviceCode.cu:296-420), dispatch (:547-666) and the MAT_LEAF attenuation of BVHShadowRay
(integratorUtilities.cuh:215-245).

Bars: renders - mean radiance within 3 % of the reference image's mean and relMSE of the 4x4-block means <= 0.02, each integrator against
ITS OWN fixture; closest hits - id exact, t/u/v bit-equal (ties on bit-equal t reported apart); shadow throughput - bit-equal.
The headline benchmark scene (rainbow dragon, ~871k triangles, VCM) is held to a 1920x1080, 64-iteration frame of the reference:
mean within 3 %, relMSE of the 8x8-block means <= 0.05."""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from scene_utils import classify

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_mse(img, ref):
    return float(np.mean((img - ref) ** 2 / (ref ** 2 + 1e-2)))


def block(img, k=4):
    h, w, c = img.shape
    return img.reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


@pytest.fixture(scope="module")
def materials(tmp_path_factory):
    d = str(tmp_path_factory.mktemp("materials"))
    nb.make_scene("materials", d, 100)
    s = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
    yield s
    s.close()


def test_scene_holds_every_material_class(materials):
    pos, nrm, uv, tris, lights = materials.download()
    mats = nb.material_table()
    used = set(int(m) for m in np.unique(tris["materialID"]))
    assert {2, 4, 5, 6, 7, 8, 9, 12, 13, 16, 19} <= used
    assert {int(mats["type"][m]) for m in used} == {0, 1, 2, 4, 6}          # diffuse, metal, smooth dielectric, leaf, delta mirror


@pytest.mark.parametrize("name,integ,spp", [("unidirectional", 0, 4096), ("naive_unidirectional", 2, 4096), ("bdpt", 1, 1024),
                                            ("vcm", 3, 1024), ("sppm", 4, 1024)])
def test_render_against_reference(name, integ, spp, materials):
    p = os.path.join(GOLD, f"materials_render_{name}.npz")
    assert os.path.exists(p), f"{p} missing: run tests/golden/make_golden_materials.py on the GPU box"
    ref = np.load(p)["img"]
    ok = np.isfinite(ref).all(axis=2)
    assert ok.mean() > 0.99
    cfg = nb.RenderConfig.from_buffer_copy(bytes(materials.cfg))
    cfg.width, cfg.height, cfg.integrator = 64, 64, integ
    cam = nb.camera_make(cfg)
    # SPPM's fixture and our render both carry 1024-iteration photon noise (relMSE between two such renders ~0.02): four independent
    # renders (seeds) are averaged so that the comparison is limited by the fixture's noise alone
    seeds = (103033, 7, 8, 9) if name == "sppm" else (103033,)
    img = sum(materials.render(cfg, cam, 0, spp, seed=sd)[0][..., :3] for sd in seeds) / (spp * len(seeds))
    assert np.isfinite(img).all()
    if name in ("vcm",):
        # the reference's own VCM image of this scene has negative pixels (95 values, min -1.1 in the fixture): its NEE through the
        # thin-walled leaf multiplies by the signed cosine of the transmitted direction; we reproduce that and hold the image, sign included
        assert (ref[ok] < 0).any()
    else:
        assert (img >= 0).all()
    ref = np.where(ok[..., None], ref, 0.0); img = np.where(ok[..., None], img, 0.0)
    r = rel_mse(block(img), block(ref))
    print("materials", name, "mean ours", img.mean(), "ref", ref.mean(), "relMSE(4x4)", r)
    assert abs(img.mean() - ref.mean()) < 0.03 * ref.mean(), (img.mean(), ref.mean())
    assert r < 0.02


def test_closest_hit_vs_reference_golden(materials):
    g = np.load(os.path.join(GOLD, "materials_closest.npz"))
    ours = materials.trace_closest(g["rays"])
    ref = np.rec.fromarrays([g["tri"], g["t"], g["u"], g["v"]], names="tri,t,u,v")
    cls = classify(ours, ref)
    assert cls["mismatch"] == 0, cls
    assert cls["tie"] <= 1e-2 * cls["n"] and cls["superset"] <= 1e-2 * cls["n"], cls


def test_leaf_attenuated_shadow_rays_bit_equal(materials):
    g = np.load(os.path.join(GOLD, "materials_shadow.npz"))
    thr = materials.trace_shadow(g["rays"], g["maxT"], g["skip"])
    want = g["thr"]
    part = ((want[:, :3] > 0) & (want[:, :3] < 1)).any(axis=1)
    assert part.sum() > 300, int(part.sum())                     # the fixture does cross the leaves
    same = (thr[:, :3].view(np.uint32) == want[:, :3].view(np.uint32)).all(axis=1)
    print("shadow rays", len(want), "leaf-attenuated", int(part.sum()), "bit-equal", int(same.sum()))
    assert same.all(), (int((~same).sum()), thr[~same][:4], want[~same][:4])


def test_headline_scene_frame_against_reference_1080p(tmp_path):
    p = os.path.join(GOLD, "rainbow1080_vcm_blocks.npz")
    assert os.path.exists(p), f"{p} missing: run tests/golden/make_golden_materials.py on the GPU box"
    z = np.load(p)
    w, h, spp = int(z["w"]), int(z["h"]), int(z["spp"])
    d = str(tmp_path / "rainbow")
    nb.make_scene("rainbow", d, int(z["detail"]))
    scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
    try:
        cfg = scene.cfg
        cfg.width, cfg.height, cfg.integrator = w, h, 3
        cam = nb.camera_make(cfg)
        img, st = scene.render(cfg, cam, 0, spp)
    finally:
        scene.close()
    img = img[..., :3] / spp
    assert np.isfinite(img).all()
    blocks = block(img, 8)
    r = rel_mse(blocks, z["blocks"])
    print("rainbow 1080p VCM: mean ours", img.mean(), "reference", float(z["mean"]), "relMSE(8x8 blocks)", r)
    assert abs(img.mean() - float(z["mean"])) < 0.03 * float(z["mean"])
    assert r < 0.05
