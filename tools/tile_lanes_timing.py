"""Single-GPU effect of the tile renderer's lanes (iterations in flight) on the headline workload: 4K VCM, iterations 5..24.
Usage: python tools/tile_lanes_timing.py [lanes ...]"""
import os
import sys
import tempfile
import time

import torch

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import cudapathtracer_b200 as nb  # noqa: E402

lanes_list = [int(a) for a in sys.argv[1:]] or [1, 2, 3]
d = os.path.join(tempfile.gettempdir(), "novum_bench_rainbow_100")
if not os.path.exists(os.path.join(d, "configs", "config.rendertron")):
    nb.make_scene("rainbow", d, 100)
scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
cfg = scene.cfg
cfg.width, cfg.height, cfg.integrator = 3840, 2160, 3
cam = nb.camera_make(cfg)
frame = torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda")
st = scene.render_device(cfg, cam, 0, 5, frame.data_ptr())
st = scene.render_device(cfg, cam, 5, 20, frame.data_ptr())
print(f"novum_render_device: {st.deviceMs / 20:.3f} ms/iteration, image sum {float(frame.sum()):.1f}")
for lanes in lanes_list:
    t = nb.TileRank(scene, cfg, cam, rank=0, world=1, lanes=lanes)
    frame.zero_()
    t.render(0, 5, frame.data_ptr())
    frame.zero_()
    w0 = time.time()
    st = t.render(5, 20, frame.data_ptr())
    wall = (time.time() - w0) * 1e3
    print(f"tile renderer, {lanes} lane(s): {st.deviceMs / 20:.3f} ms/iteration (wall {wall / 20:.3f}), launches {st.kernelLaunches}, image sum {float(frame.sum()):.1f}")
    t.close()
scene.close()
