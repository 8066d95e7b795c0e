"""Run under torchrun with N >= 2 processes, one GPU each (tests/test_gpu_multi.py launches it): the tile renderer across REAL processes -
slab buffers swapped as CUDA IPC handles, photons stored into the peers over NVLink inside the light pass, device-side rendezvous - must
reproduce the single-GPU frame; so must the NCCL all-gather exchange of the phase API. Rank 0 prints MP_TILE_CHECK_JSON {...}."""
import json
import os
import sys
import tempfile

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import numpy as np
import torch
import torch.distributed as dist

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import cudapathtracer_b200 as nb  # noqa: E402
from cudapathtracer_b200 import distributed as nd  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.cuda.set_device(local)
nb.lib().novum_set_device(local)
d = os.path.join(tempfile.gettempdir(), "novum_mp_cornell")
if rank == 0:
    nb.make_scene("cornell", d, 100)
dist.barrier()
scene = nb.Scene.from_config(os.path.join(d, "configs", "config.rendertron"))
out = {}
for name, integ in (("vcm", 3), ("sppm", 4), ("bdpt", 1)):
    cfg = nb.RenderConfig.from_buffer_copy(bytes(scene.cfg))
    cfg.width, cfg.height, cfg.integrator = 96, 64, integ
    cam = nb.camera_make(cfg)
    iters = 9
    for mode in ("p2p", "nccl"):
        colors = torch.zeros((cfg.height, cfg.width, 4), dtype=torch.float32, device="cuda")
        r = (nd.TileRenderer(scene, cfg, cam, dist, lanes=2) if mode == "p2p" else nd.PhaseTileRenderer(scene, cfg, cam, dist, exchange="nccl"))
        r.render(0, 4, colors)
        r.render(4, iters - 4, colors)
        nd.reduce_frames(colors, dist, 0)
        torch.cuda.synchronize()
        r.close()
        if rank == 0:
            full, _ = scene.render(cfg, cam, 0, iters)
            got = colors.cpu().numpy()
            out[f"{name}_{mode}"] = {"max_abs_diff": float(np.abs(got - full).max()), "allclose": bool(np.allclose(got, full, rtol=2e-3, atol=1e-4)),
                                    "mean": float(full[..., :3].mean() / iters)}
        dist.barrier()
if rank == 0:
    print("MP_TILE_CHECK_JSON " + json.dumps(out))
scene.close()
dist.barrier()
dist.destroy_process_group()
