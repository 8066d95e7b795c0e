"""Multi-GPU paths on REAL devices (skipped on a one-GPU box; run with `gpurun --gpus 2`):
  * one process per GPU under torchrun: the tile renderer over CUDA IPC + NVLink peer stores, and the NCCL all-gather exchange of the
    phase API, both against the single-GPU frame (tests/mp_tile_check.py);
  * one process, `numGpus` of the launch boundary: the reference's own main.cu (oracle/_ref/novum_dropin) with NOVUM_GPUS=2 - host threads,
    peer access, scene replicas - against the reference-made goldens."""
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import cudapathtracer_b200 as nb
from test_gpu_dropin import DROPIN, GOLD, block, rel_mse, set_config

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
two_gpus = pytest.mark.skipif(nb.device_count() < 2, reason="needs two GPUs")


@two_gpus
def test_tile_renderer_across_processes_equals_single_gpu():
    n = min(nb.device_count(), 4)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={n}", "--master-addr", "127.0.0.1", "--master-port", "29611",
           os.path.join(REPO, "tests", "mp_tile_check.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = json.loads(re.search(r"MP_TILE_CHECK_JSON (\{.*\})", r.stdout).group(1))
    assert set(res) == {f"{a}_{b}" for a in ("vcm", "sppm", "bdpt") for b in ("p2p", "nccl")}
    for k, v in res.items():
        assert v["allclose"], (k, v)


@two_gpus
@pytest.mark.skipif(not os.path.exists(DROPIN), reason="oracle/_ref/novum_dropin not built")
@pytest.mark.parametrize("integ,gold,spp", [("UNIDIRECTIONAL", "unidirectional", 256), ("VCM", "vcm", 1024), ("BDPT", "bdpt", 1024)])
def test_reference_main_on_two_gpus(integ, gold, spp, tmp_path):
    d = str(tmp_path / "cornell")
    nb.make_scene("cornell", d, 100)
    ref = np.load(os.path.join(GOLD, f"cornell_render_{gold}.npz"))["img"]
    ok = np.isfinite(ref).all(axis=2)
    ref = np.where(ok[..., None], ref, 0.0)
    set_config(d, width=64, height=64, Integrator=integ, **{"Sample Count": spp})
    raw = str(tmp_path / "out.raw")
    env = dict(os.environ, NOVUM_GPUS="2", CUDA_DEVICE_MAX_CONNECTIONS="32")
    r = subprocess.run([DROPIN, "render", d, "--out", raw, "--repeat", "2"], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0 and "RENDER ERROR" not in r.stderr, r.stdout[-2000:] + r.stderr[-2000:]
    img = np.fromfile(raw, np.float32).reshape(64, 64, 4)[..., :3] / spp
    img = np.where(ok[..., None], img, 0.0)
    assert np.isfinite(img).all()
    assert abs(img.mean() - ref.mean()) < 0.03 * ref.mean(), (img.mean(), ref.mean())
    assert rel_mse(block(img), block(ref)) < 0.02
