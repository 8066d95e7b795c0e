"""Renders small VCM / SPPM frames through the library currently in place and saves image + counters (A/B of two builds)."""
import os, sys, tempfile
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cudapathtracer_b200 as nb
tag = sys.argv[1]
out = {}
for name, detail, w, h, integ, iters in (("cornell", 100, 256, 256, 3, 8), ("rainbow", 8, 320, 180, 3, 8), ("water", 12, 256, 256, 4, 8), ("water", 12, 256, 256, 3, 4)):
    d = tempfile.mkdtemp()
    scene = nb.Scene.from_config(nb.make_scene(name, d, detail))
    cfg = scene.cfg; cfg.width, cfg.height, cfg.integrator = w, h, integ
    cam = nb.camera_make(cfg)
    img, st = scene.render(cfg, cam, 0, iters)
    key = f"{name}_{integ}"
    out[key + "_img"] = img
    out[key + "_stats"] = np.array([st.photons, st.mergeQueries, st.mergeCandidates, st.mergeAccepted], np.float64)
    print(tag, key, "photons", st.photons, "queries", st.mergeQueries, "cand", st.mergeCandidates, "acc", st.mergeAccepted, "mean", float(img[..., :3].mean()))
    scene.close()
np.savez(os.path.join("gpurun_out", f"ab_{tag}.npz"), **out)
