"""Second set of golden renders, again from the REFERENCE'S OWN device code (oracle/_ref/novum_ref on a B200): scenes with
delta materials, which the Cornell fixtures do not exercise in the bidirectional integrators.

  water_render_<integrator>.npz   : water-caustic stand-in (dielectric heightfield slab, pool, small light), detail 12, 64x64:
                                    UNIDIRECTIONAL 4096 spp; BDPT / VCM / SPPM 1024 iterations
  rainbow_render_vcm.npz          : rainbow-dragon stand-in (glass dragon over a Lambert floor), detail 8, 64x36, VCM 1024 iterations

Run on the GPU box:  python tests/golden/make_golden_scenes.py
tests/test_gpu_bidir.py::test_delta_scenes_against_reference_render compares the CUDA core with these."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_golden import REF, run_ref, set_config, nb  # noqa: E402

JOBS = (
    ("water", 12, 64, 64, (("UNIDIRECTIONAL", 4096), ("BDPT", 1024), ("VCM", 1024), ("SPPM", 1024))),
    ("rainbow", 8, 64, 36, (("VCM", 1024),)),
)


def main():
    if not os.path.exists(REF):
        raise SystemExit(f"{REF} missing: run oracle/build_ref.sh where /root/reference exists")
    tmp = tempfile.mkdtemp()
    for name, detail, w, h, runs in JOBS:
        d = os.path.join(tmp, name)
        nb.make_scene(name, d, detail)
        for integ, spp in runs:
            set_config(d, width=w, height=h, Integrator=integ, **{"Sample Count": spp})
            raw = os.path.join(tmp, "ref.raw")
            info = run_ref(["render", d, "--out", raw])
            img = np.fromfile(raw, np.float32).reshape(h, w, 4)[..., :3] / spp
            np.savez_compressed(os.path.join(HERE, f"{name}_render_{integ.lower()}.npz"), img=img.astype(np.float32), spp=spp, detail=detail)
            print(name, integ, info, "mean", float(np.nanmean(img)), "non-finite", int((~np.isfinite(img)).sum()))


if __name__ == "__main__":
    main()
