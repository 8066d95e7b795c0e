"""Two tile-renderer ranks emulated on ONE GPU (threads of this process, slabs attached by pointer) at the size tests/mp_tile_check.py
uses across processes: their summed frames against the single-GPU render. Usage: python tools/tile_emul_check.py [w h iters]"""
import os, sys, tempfile, threading
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import cudapathtracer_b200 as nb
w, h, iters = (int(a) for a in (sys.argv[1:4] + ["96", "64", "9"][len(sys.argv) - 1:]))
d = os.path.join(tempfile.gettempdir(), "novum_emul_cornell")
nb.make_scene("cornell", d, 100)
scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
for name, integ in (("vcm", 3), ("sppm", 4), ("bdpt", 1)):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(scene.cfg))
    cfg.width, cfg.height, cfg.integrator = w, h, integ
    cam = nb.camera_make(cfg)
    full, _ = scene.render(cfg, cam, 0, iters)
    again, _ = scene.render(cfg, cam, 0, iters)
    for world in (2,):
        ranks = [nb.TileRank(scene, cfg, cam, rank=r, world=world, lanes=2) for r in range(world)]
        frames = [torch.zeros((h, w, 4), dtype=torch.float32, device="cuda") for _ in range(world)]
        for r, t in enumerate(ranks):
            for q, o in enumerate(ranks):
                if q != r:
                    t.attach_pointers(q, o.bases())
        for first, count in ((0, 4), (4, iters - 4)):
            th = [threading.Thread(target=lambda r=r: ranks[r].render(first, count, frames[r].data_ptr())) for r in range(world)]
            [t.start() for t in th]; [t.join() for t in th]
        total = sum(f.cpu().numpy() for f in frames)
        for t in ranks:
            t.close()
        bad = ~np.isclose(total, full, rtol=2e-3, atol=1e-4)
        print(f"{name} world {world}: max abs diff {np.abs(total - full).max():.3e}, pixels outside rtol 2e-3: {int(bad.any(axis=2).sum())} of {w * h}; "
              f"single-GPU render twice: max abs diff {np.abs(again - full).max():.3e}; mean {full[..., :3].mean() / iters:.5f}")
scene.close()
