"""Third set of golden fixtures, again from the REFERENCE'S OWN device code (oracle/_ref/novum_ref on a B200):

  materials_render_<integrator>.npz : the `materials` stand-in (novum_make_scene: GGX gold + steel spheres, delta mirror panel, glass box
                                      holding a tea box holding an ice box = three dielectric priorities with absorption, two thin-walled
                                      leaf quads under the light, a bilinearly textured wall; generated textures/*.bmp), 64x64:
                                      UNIDIRECTIONAL 4096 spp, NAIVE_UNIDIRECTIONAL 4096 spp, BDPT / VCM / SPPM 1024 iterations
  materials_closest.npz / materials_shadow.npz : BVHSceneIntersect / BVHShadowRay on ray batches of that scene; the shadow batch holds
                                      NEE-like rays from points below the leaves to points on the light, so many throughputs are
                                      leaf-attenuated values strictly between 0 and 1 (integratorUtilities.cuh:215-245)
  rainbow1080_vcm_blocks.npz        : the benchmark scene itself (rainbow dragon, detail 100, ~871k triangles), VCM, 1920x1080,
                                      64 iterations: 8x8-block means of the mean image (240x135x3) + the global mean - what
                                      bench.py's `relmse_vs_reference` and tests/test_gpu_materials.py hold the 4K headline config to

Run on the GPU box:  python tests/golden/make_golden_materials.py [--skip-1080]"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, run_ref, set_config, write_rays, nb  # noqa: E402
from scene_utils import load_scene_py, random_rays, camera_rays  # noqa: E402


def nee_like_rays(pos, tris, n, rng):
    """Shadow rays as nextEventEstimation shoots them: from random points on non-emissive triangles (pushed off the surface) to random
    points on emissive triangles, maxT = distance * (1 - 1e-3), skip = the origin triangle."""
    P = pos[:, :3]
    em = (tris["emission"][:, :3] ** 2).sum(1) > 0
    src, dst = np.nonzero(~em)[0], np.nonzero(em)[0]

    def sample(ids):
        t = tris[ids]
        a, b, c = P[t["aInd"]], P[t["bInd"]], P[t["cInd"]]
        r1, r2 = np.sqrt(rng.random(len(ids))).astype(np.float32), rng.random(len(ids)).astype(np.float32)
        return (a * (1 - r1)[:, None] + b * (r1 * (1 - r2))[:, None] + c * (r1 * r2)[:, None]).astype(np.float32)

    si = rng.choice(src, n)
    o, l = sample(si), sample(rng.choice(dst, n))
    d = l - o
    dist = np.linalg.norm(d, axis=1).astype(np.float32)
    rays = np.zeros((n, 8), np.float32)
    rays[:, 0:3] = o
    rays[:, 4:7] = d / dist[:, None]
    return rays, (dist * np.float32(1 - 1e-3)).astype(np.float32), si.astype(np.int32)


def main():
    if not os.path.exists(REF):
        raise SystemExit(f"{REF} missing: run oracle/build_ref.sh where /root/reference exists")
    tmp = tempfile.mkdtemp()
    d = os.path.join(tmp, "materials")
    nb.make_scene("materials", d, 100)
    cfg, pos, nrm, uv, tris, mats = load_scene_py(d)
    rng = np.random.default_rng(777)
    c = nb.load_config(os.path.join(d, "configs", "config.rendertron"))
    cam = nb.camera_make(c)
    mn, mx = pos[:, :3].min(0), pos[:, :3].max(0)
    rb, ob_ = os.path.join(tmp, "rays.bin"), os.path.join(tmp, "out.bin")
    rays = np.concatenate([camera_rays(cam, c.width, c.height, 4096, rng), random_rays(mn, mx, 4096, rng)])
    write_rays(rb, rays)
    info = run_ref(["trace", d, rb, ob_])
    res = np.fromfile(ob_, np.float32).reshape(-1, 4)
    np.savez_compressed(os.path.join(HERE, "materials_closest.npz"), rays=rays, tri=res[:, 0].view(np.int32).copy(), t=res[:, 1].copy(),
                        u=res[:, 2].copy(), v=res[:, 3].copy())
    print("materials closest", info)
    r1, t1, s1 = nee_like_rays(pos, tris, 6144, rng)
    r2 = random_rays(mn, mx, 2048, rng)
    t2 = (rng.random(2048) * 3.0 + 0.01).astype(np.float32)
    s2 = np.full(2048, -1, np.int32)
    srays, max_t, skip = np.concatenate([r1, r2]), np.concatenate([t1, t2]), np.concatenate([s1, s2])
    write_rays(rb, srays, max_t, skip)
    info = run_ref(["trace", d, rb, ob_, "--shadow"])
    thr = np.fromfile(ob_, np.float32).reshape(-1, 4)
    part = ((thr[:, :3] > 0) & (thr[:, :3] < 1)).any(axis=1)
    np.savez_compressed(os.path.join(HERE, "materials_shadow.npz"), rays=srays, maxT=max_t, skip=skip, thr=thr)
    print("materials shadow", info, "leaf-attenuated rays:", int(part.sum()), "of", len(thr))
    for integ, spp in (("UNIDIRECTIONAL", 4096), ("NAIVE_UNIDIRECTIONAL", 4096), ("BDPT", 1024), ("VCM", 1024), ("SPPM", 1024)):
        set_config(d, width=64, height=64, Integrator=integ, **{"Sample Count": spp})
        raw = os.path.join(tmp, "ref.raw")
        info = run_ref(["render", d, "--out", raw])
        img = np.fromfile(raw, np.float32).reshape(64, 64, 4)[..., :3] / spp
        np.savez_compressed(os.path.join(HERE, f"materials_render_{integ.lower()}.npz"), img=img.astype(np.float32), spp=spp, detail=100)
        print("materials", integ, info, "mean", float(np.nanmean(img)), "non-finite", int((~np.isfinite(img)).sum()))
    if "--skip-1080" in sys.argv:
        return
    d = os.path.join(tmp, "rainbow")
    nb.make_scene("rainbow", d, 100)
    w, h, spp = 1920, 1080, 64
    set_config(d, width=w, height=h, Integrator="VCM", **{"Sample Count": spp})
    raw = os.path.join(tmp, "ref1080.raw")
    info = run_ref(["render", d, "--out", raw])
    img = np.fromfile(raw, np.float32).reshape(h, w, 4)[..., :3] / spp
    bad = ~np.isfinite(img).all(axis=2)
    img = np.where(bad[..., None], 0.0, img)
    blocks = img.reshape(h // 8, 8, w // 8, 8, 3).mean(axis=(1, 3)).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "rainbow1080_vcm_blocks.npz"), blocks=blocks, mean=np.float64(img.mean()), spp=spp, detail=100, w=w, h=h,
                        non_finite=int(bad.sum()))
    print("rainbow 1080p VCM", info, "mean", float(img.mean()), "non-finite pixels", int(bad.sum()))


if __name__ == "__main__":
    main()
