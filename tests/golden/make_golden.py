"""Generates the golden fixtures in this directory by running the REFERENCE'S OWN device code
(oracle/_ref/novum_ref = DanielQ-51/cudapathtracer built unmodified for sm_100a, see oracle/build_ref.sh) on a B200:

  <scene>_closest.npz : BVHSceneIntersect + triangleIntersect on a fixed ray batch -> tri, t, u, v
  <scene>_shadow.npz  : BVHShadowRay on a fixed batch -> throughput
  cornell_render_<integrator>.npz : launch_* of the reference at 64x64, high spp -> mean image (float32 rgb)

Run on the GPU box:  python tests/golden/make_golden.py   (the reference has no CPU path, so not in the build container).
The scenes are the deterministic stand-ins written by novum_make_scene (tests use the same generator + detail).
"""
import json
import os
import re
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

import cudapathtracer_b200 as nb  # noqa: E402
from scene_utils import load_scene_py, random_rays, adversarial_rays, camera_rays  # noqa: E402

REF = os.path.join(REPO, "oracle", "_ref", "novum_ref")
SCENES = (("cornell", 100), ("dragon", 8))


def run_ref(args):
    out = subprocess.check_output([REF] + args, stderr=subprocess.STDOUT).decode()
    m = re.search(r"NOVUM_REF_JSON (\{.*\})", out)
    return json.loads(m.group(1)) if m else {}


def write_rays(path, rays, max_t=None, skip=None):
    with open(path, "wb") as f:
        f.write(np.int32(len(rays)).tobytes())
        f.write(np.ascontiguousarray(rays, np.float32).tobytes())
        if max_t is not None:
            f.write(np.ascontiguousarray(max_t, np.float32).tobytes())
            f.write(np.ascontiguousarray(skip, np.int32).tobytes())


def set_config(scene_dir, **kv):
    p = os.path.join(scene_dir, "configs", "config.rendertron")
    lines = open(p).read().splitlines()
    for k, v in kv.items():
        lines = [f"{k}: {v}" if l.startswith(k + ":") else l for l in lines]
    open(p, "w").write("\n".join(lines) + "\n")


def main():
    if not os.path.exists(REF):
        raise SystemExit(f"{REF} missing: run oracle/build_ref.sh where /root/reference exists")
    tmp = tempfile.mkdtemp()
    for name, detail in SCENES:
        d = os.path.join(tmp, name)
        nb.make_scene(name, d, detail)
        cfg, pos, nrm, uv, tris, mats = load_scene_py(d)
        rng = np.random.default_rng(2024)
        c = nb.load_config(os.path.join(d, "configs", "config.rendertron"))
        cam = nb.camera_make(c)
        mn, mx = pos[:, :3].min(0), pos[:, :3].max(0)
        rays = np.concatenate([camera_rays(cam, c.width, c.height, 3000, rng), random_rays(mn, mx, 3000, rng),
                               adversarial_rays(pos, tris, 2192, rng)])
        rb, ob_ = os.path.join(tmp, "rays.bin"), os.path.join(tmp, "out.bin")
        write_rays(rb, rays)
        info = run_ref(["trace", d, rb, ob_])
        res = np.fromfile(ob_, np.float32).reshape(-1, 4)
        np.savez_compressed(os.path.join(HERE, f"{name}_closest.npz"), rays=rays, tri=res[:, 0].view(np.int32).copy(), t=res[:, 1].copy(),
                            u=res[:, 2].copy(), v=res[:, 3].copy())
        print(name, "closest", info)
        # shadow rays: toward random points, random max distance, sometimes skipping the first-hit triangle
        srays = random_rays(mn, mx, 4096, rng)
        max_t = (rng.random(4096) * 3.0 + 0.01).astype(np.float32)
        skip = np.where(rng.random(4096) < 0.3, res[:4096, 0].view(np.int32), -1).astype(np.int32)
        write_rays(rb, srays, max_t, skip)
        info = run_ref(["trace", d, rb, ob_, "--shadow"])
        thr = np.fromfile(ob_, np.float32).reshape(-1, 4)
        np.savez_compressed(os.path.join(HERE, f"{name}_shadow.npz"), rays=srays, maxT=max_t, skip=skip, thr=thr)
        print(name, "shadow", info)
    # reference renders of the Cornell box, 64x64, for the integrators (mean image = sum / spp)
    d = os.path.join(tmp, "cornell")
    for integ, spp in (("UNIDIRECTIONAL", 8192), ("NAIVE_UNIDIRECTIONAL", 8192), ("VCM", 1024), ("SPPM", 1024), ("BDPT", 1024)):
        set_config(d, width=64, height=64, Integrator=integ, **{"Sample Count": spp})
        raw = os.path.join(tmp, "ref.raw")
        info = run_ref(["render", d, "--out", raw])
        img = np.fromfile(raw, np.float32).reshape(64, 64, 4)[..., :3] / spp
        np.savez_compressed(os.path.join(HERE, f"cornell_render_{integ.lower()}.npz"), img=img.astype(np.float32), spp=spp)
        print(integ, info, "mean", float(img.mean()))


if __name__ == "__main__":
    main()
