"""Opt-in features of SURVEY 8f row 3 (novum_scene_set_features) - both are things the reference only half-implements, both are OFF by
default (then every integrator reproduces the reference and the goldens, which the rest of the suite holds), both change results:

  * NOVUM_FEATURE_ENVIRONMENT: a constant-radiance environment. Closed form used here: an unoccluded Lambert plane of albedo rho under
    a uniform environment L has outgoing radiance rho * L (a cosine-sampled bounce always escapes), the sky itself shows L.
  * NOVUM_FEATURE_NESTED_BIDIR: the nested-dielectric priority stack inside the BDPT / VCM / SPPM walks (the reference leaves
    `etaI = etaT = 1.0f // TEMPORARY` there, deviceCode.cu:851-852, 1177-1178, 2078-2079). Oracle: the repo's own NEE+MIS integrator,
    which honours priorities and absorption as Li_unidirectional does (deviceCode.cu:296-425) and is itself held to the reference's
    renders of the `materials` scene (tests/test_gpu_materials.py)."""
import os

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from cudapathtracer_b200.binding import TRI_DTYPE

pytestmark = pytest.mark.gpu


def _plane_scene(albedo_material=22):
    """One huge quad at z = -3 facing +z (the camera at z = +1 looks down -z): every pixel sees an unoccluded Lambert plane."""
    pos = np.array([[-500, -500, -3, 0], [500, -500, -3, 0], [500, 500, -3, 0], [-500, 500, -3, 0]], np.float32)
    nrm = np.array([[0, 0, 1, 0]], np.float32)
    uv = np.zeros((1, 2), np.float32)
    tris = np.zeros(2, TRI_DTYPE)
    tris["aInd"], tris["bInd"], tris["cInd"] = [0, 0], [1, 2], [2, 3]
    tris["materialID"] = albedo_material          # 22: Lambert 0.5 (main.cu:443-467)
    tris["lightInd"] = -51
    tris["triInd"] = [0, 1]
    return nb.Scene.from_arrays(pos, nrm, uv, tris, nb.material_table())


def _cfg(integ, w=32, h=24):
    cfg = nb.RenderConfig()
    cfg.width, cfg.height, cfg.integrator = w, h, integ
    cfg.sampleCount, cfg.maxDepth, cfg.bvhLeafSize = 1, 40, 2
    cfg.bdptEyeDepth, cfg.bdptLightDepth = 16, 10
    cfg.bdptLightTrace = cfg.bdptNee = cfg.bdptNaive = cfg.bdptConnection = cfg.bdptDoMis = cfg.vcmDoMerge = 1
    cfg.vcmMergeConst, cfg.vcmInitialMergeRadiusMultiplier = 0.7, 0.01
    cfg.pinholeCamera = 1
    cfg.camPos[0], cfg.camPos[1], cfg.camPos[2] = 0.0, 0.0, 1.0
    cfg.camFov = 60.0
    return cfg


@pytest.mark.parametrize("integ", [0, 2, 1, 3])
def test_constant_environment_closed_form(integ):
    s = _plane_scene()
    try:
        cfg = _cfg(integ)
        cam = nb.camera_make(cfg)
        spp = 64
        off, _ = s.render(cfg, cam, 0, 4)
        assert not off[..., :3].any()                      # feature off, no lights: black, as the reference's sampleSky
        s.set_features(environment=(1.0, 0.5, 2.0))
        img, _ = s.render(cfg, cam, 0, spp)
        img = img[..., :3] / spp
        want = 0.5 * np.array([1.0, 0.5, 2.0], np.float32)
        assert np.isfinite(img).all()
        got = img.reshape(-1, 3).mean(0)
        print("integrator", integ, "plane radiance", got, "closed form", want)
        assert np.allclose(got, want, rtol=0.01), (got, want)
        # a bounce off a 0.5 plane is deterministic: every pixel (the NEE+MIS walk floors its pdf at 0.01, deviceCode.cu:481: grazing samples lose a little)
        assert np.allclose(img, want[None, None, :], rtol=0.05 if integ == 0 else 0.01, atol=1e-4)
        s.set_features()                                   # off again: black again
        again, _ = s.render(cfg, cam, 0, 2)
        assert not again[..., :3].any()
    finally:
        s.close()


def test_environment_is_seen_directly_and_lights_still_work(scene_dirs):
    s = nb.Scene.from_config(os.path.join(scene_dirs["cornell"], "configs", "config.rendertron"))
    try:
        cfg = nb.RenderConfig.from_buffer_copy(bytes(s.cfg))
        cfg.width, cfg.height, cfg.integrator = 37, 19, 0     # wide frame: the side columns look past the box into the environment
        cam = nb.camera_make(cfg)
        base, _ = s.render(cfg, cam, 0, 64)
        s.set_features(environment=(0.25, 0.25, 0.25))
        img, _ = s.render(cfg, cam, 0, 64)
        sky = base[..., :3].sum(axis=2) == 0                  # pixels that saw nothing without the environment
        assert sky.any()
        assert np.allclose(img[sky][:, :3] / 64, 0.25, rtol=1e-3)
        assert (img[..., :3].sum() > base[..., :3].sum())
    finally:
        s.close()


def _quad(P, N, T, a, b, c, d, mat, emission=(0.0, 0.0, 0.0)):
    """Appends the planar quad a, b, c, d (counter-clockwise seen from its normal side) as two triangles."""
    a, b, c, d = [np.asarray(v, np.float32) for v in (a, b, c, d)]
    n = np.cross(b - a, c - a); n = n / np.linalg.norm(n)
    i0, n0 = len(P), len(N)
    P.extend([list(a) + [0], list(b) + [0], list(c) + [0], list(d) + [0]])
    N.append(list(n) + [0])
    for (x, y, z) in ((0, 1, 2), (0, 2, 3)):
        t = np.zeros((), TRI_DTYPE)
        t["aInd"], t["bInd"], t["cInd"] = i0 + x, i0 + y, i0 + z
        t["naInd"] = t["nbInd"] = t["ncInd"] = n0
        t["materialID"] = mat
        t["emission"] = list(emission) + [0.0]
        T.append(t)


def _box(P, N, T, lo, hi, mat):
    (x0, y0, z0), (x1, y1, z1) = lo, hi
    _quad(P, N, T, (x0, y0, z1), (x1, y0, z1), (x1, y1, z1), (x0, y1, z1), mat)      # +z
    _quad(P, N, T, (x1, y0, z0), (x0, y0, z0), (x0, y1, z0), (x1, y1, z0), mat)      # -z
    _quad(P, N, T, (x1, y0, z1), (x1, y0, z0), (x1, y1, z0), (x1, y1, z1), mat)      # +x
    _quad(P, N, T, (x0, y0, z0), (x0, y0, z1), (x0, y1, z1), (x0, y1, z0), mat)      # -x
    _quad(P, N, T, (x0, y1, z1), (x1, y1, z1), (x1, y1, z0), (x0, y1, z0), mat)      # +y
    _quad(P, N, T, (x0, y0, z0), (x1, y0, z0), (x1, y0, z1), (x0, y0, z1), mat)      # -y


def _room_behind_a_tea_slab():
    """A white room lit from its ceiling, seen THROUGH a 0.1-thick slab of the absorbing dielectric `tea` (material 8: ior 1.333, absorption
    2.5 * (0.18, 1.5, 2.996), main.cu:443-467) that fills the whole view. Whatever reaches the camera has crossed the slab once:
    transmittance exp(-sigma d) ~ (0.96, 0.69, 0.47) at normal incidence - a strong colour shift that only the medium logic produces."""
    P, N, T = [], [], []
    _quad(P, N, T, (-1, -1, -1), (1, -1, -1), (1, -1, -3), (-1, -1, -3), 2)              # floor
    _quad(P, N, T, (-1, 1, -3), (1, 1, -3), (1, 1, -1), (-1, 1, -1), 2)                  # ceiling
    _quad(P, N, T, (-1, -1, -3), (1, -1, -3), (1, 1, -3), (-1, 1, -3), 2)                # back wall
    _quad(P, N, T, (-1, -1, -1), (-1, -1, -3), (-1, 1, -3), (-1, 1, -1), 2)              # left
    _quad(P, N, T, (1, -1, -3), (1, -1, -1), (1, 1, -1), (1, 1, -3), 2)                  # right
    _quad(P, N, T, (-0.4, 0.99, -2.4), (0.4, 0.99, -2.4), (0.4, 0.99, -1.6), (-0.4, 0.99, -1.6), 2, emission=(12.0, 12.0, 12.0))
    _box(P, N, T, (-1.3, -1.3, -0.6), (1.3, 1.3, -0.5), 8)
    tris = np.array(T, TRI_DTYPE)
    tris["triInd"] = np.arange(len(tris))
    lit = (tris["emission"][:, :3].sum(1) > 0)
    tris["lightInd"] = np.where(lit, np.cumsum(lit) - 1, -51)
    return nb.Scene.from_arrays(np.array(P, np.float32), np.array(N, np.float32), np.zeros((1, 2), np.float32), tris, nb.material_table())


@pytest.mark.parametrize("integ,name", [(1, "bdpt"), (3, "vcm")])
def test_nested_dielectrics_in_the_bidirectional_walks(integ, name):
    s = _room_behind_a_tea_slab()
    try:
        cfg = _cfg(0, 48, 48)
        cam = nb.camera_make(cfg)
        nee, _ = s.render(cfg, cam, 0, 1024)
        nee = nee[..., :3].reshape(-1, 3).mean(0) / 1024
        cfg.integrator = integ
        plain, _ = s.render(cfg, cam, 0, 256)
        plain = plain[..., :3].reshape(-1, 3).mean(0) / 256
        s.set_features(nested_bidir=True)
        nested, _ = s.render(cfg, cam, 0, 256)
        nested = nested[..., :3].reshape(-1, 3).mean(0) / 256
        s.set_features()
        again, _ = s.render(cfg, cam, 0, 256)
        again = again[..., :3].reshape(-1, 3).mean(0) / 256
        print(name, "NEE+MIS", nee, "plain", plain, "nested", nested)
        # the colour of the picture: blue / red and green / red of the mean radiance. The unidirectional integrator is the oracle for it
        # (it carries the medium stack as the reference's does); absolute levels differ between the reference's integrators by design
        tint = lambda c: np.array([c[1] / c[0], c[2] / c[0]])
        assert np.allclose(tint(nested), tint(nee), rtol=0.05), (tint(nested), tint(nee))
        assert tint(plain)[1] > 1.5 * tint(nee)[1]                 # without the feature the slab is clear: the reference's placeholder behaviour
        assert np.allclose(again, plain, rtol=2e-3)                # off again: the default path, unchanged
    finally:
        s.close()
