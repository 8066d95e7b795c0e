"""Where does a tile-sharded 4K VCM iteration go on ONE rank of `world`? All ranks are emulated on one GPU (sessions attached by pointer,
peer-write on), phases timed with CUDA events on the engine stream: light pass (this rank's rows), grid build (all photons), eye pass
(this rank's rows, all photons). Usage: python tools/tile_phase_timing.py [world] [first_iter] [iters]"""
import os
import sys
import tempfile

import numpy as np
import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import cudapathtracer_b200 as nb  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
first = int(sys.argv[2]) if len(sys.argv) > 2 else 5
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 4
d = os.path.join(tempfile.gettempdir(), "novum_bench_rainbow_100")
if not os.path.exists(os.path.join(d, "configs", "config.rendertron")):
    nb.make_scene("rainbow", d, 100)
scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
cfg = scene.cfg
cfg.width, cfg.height, cfg.integrator = 3840, 2160, 3
cam = nb.camera_make(cfg)
sessions = [nb.VcmSession(scene, cfg, cam, rank=r, world=world) for r in range(world)]
frame = torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda")
for r, s in enumerate(sessions):
    for q, o in enumerate(sessions):
        if q != r:
            s.attach_pointer(q, o.base)
    if world > 1:
        s.peer_write(True)
ext = torch.cuda.ExternalStream(sessions[0].stream())
ev = lambda: torch.cuda.Event(enable_timing=True)
tot = {"light": 0.0, "grid": 0.0, "eye": 0.0}
launch0 = sessions[0].stats().kernelLaunches
scene.set_profile(1)
for it in range(first - 2, first + iters):
    if it == first:
        sessions[0].reset_stats()
    e = [ev() for _ in range(4)]
    for r in range(1, world):
        sessions[r].light_pass(it, frame.data_ptr())
    sessions[0].stats()            # drains rank 0's stage events, so that the other ranks' kernels are not charged to its first launch
    e[0].record(ext)
    sessions[0].light_pass(it, frame.data_ptr())
    e[1].record(ext)
    sessions[0].build_grid(it)
    e[2].record(ext)
    sessions[0].eye_pass(it, frame.data_ptr())
    e[3].record(ext)
    torch.cuda.synchronize()
    if it >= first:
        tot["light"] += e[0].elapsed_time(e[1]); tot["grid"] += e[1].elapsed_time(e[2]); tot["eye"] += e[2].elapsed_time(e[3])
st = sessions[0].stats()
print(f"world {world}, iterations {first}..{first + iters - 1}: per iteration on rank 0: " +
      ", ".join(f"{k} {v / iters:.3f} ms" for k, v in tot.items()) + f", sum {sum(tot.values()) / iters:.3f} ms; photons {st.photons}, "
      f"launches/iteration {st.kernelLaunches / iters:.0f}")
print("  stage ms per iteration [extend, step+grid, resolve, merge]:", [round(st.stageMs[k] / iters, 3) for k in range(4)],
      "launches", [st.stageLaunches[k] // iters for k in range(4)])
for s in sessions:
    s.close()
scene.close()
