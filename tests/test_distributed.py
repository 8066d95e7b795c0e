"""Host-side logic of the two multi-GPU decompositions, on CPU with world_size-2 gloo process groups (no GPU needed):
row ownership of the tile-sharded phase API, pass striping of novum_render_device, and the frame reduction.
The device side of the same paths is covered by tests/test_gpu_bidir.py (ranks emulated on one GPU) and by
`bench.py --gpus N` on a multi-GPU box."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cudapathtracer_b200 import distributed as nd


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, h, w, first, count, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        assert nd.shard_of(dist) == (rank, world)
        # every rank paints its rows and its passes into a full frame, exactly as the engines accumulate
        frame = torch.zeros((h, w, 4), dtype=torch.float32)
        rows = nd.local_rows(h, rank, world)
        for y in rows:
            frame[y, :, 0] += 1.0                                # tile-sharded: one owner per row
        for s in nd.striped_samples(first, count, rank, world):
            frame[:, :, 1] += float(s)                           # striped: pass s contributes s everywhere
        frame[:, :, 2] += float(rank + 1)                        # splats: every rank may touch every pixel
        # ownership is a partition: gather the row lists and the pass lists
        gathered = [None] * world
        dist.all_gather_object(gathered, (rows, nd.striped_samples(first, count, rank, world)))
        all_rows = sorted(y for g in gathered for y in g[0])
        all_passes = sorted(s for g in gathered for s in g[1])
        assert all_rows == list(range(h))
        assert all_passes == list(range(first, first + count))
        nd.reduce_frames(frame, dist, 0)
        if rank == 0:
            np.save(os.path.join(out_dir, "sum.npy"), frame.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("h,w,first,count", [(9, 5, 0, 7), (8, 4, 3, 4), (1, 3, 10, 1)])
def test_two_rank_partition_and_reduce(tmp_path, h, w, first, count):
    world = 2
    port = _free_port()
    mp.spawn(_worker, args=(world, port, h, w, first, count, str(tmp_path)), nprocs=world, join=True)
    total = np.load(tmp_path / "sum.npy")
    assert np.array_equal(total[..., 0], np.ones((h, w), np.float32))                       # each row owned exactly once
    assert np.allclose(total[..., 1], sum(range(first, first + count)))                     # each pass rendered exactly once
    assert np.allclose(total[..., 2], sum(r + 1 for r in range(world)))                     # full-frame sum over ranks


def _exchange_worker(rank, world, port, cap, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sb = 32 + 32 * cap
        slabs = torch.full((world * sb,), 0xEE, dtype=torch.uint8)       # stale bytes everywhere
        mine = slabs[rank * sb:(rank + 1) * sb]
        n = 2 + 3 * rank                                                 # ranks store different photon counts
        mine[:32] = 0
        mine[:4].view(torch.int32)[0] = n
        mine[32:32 + 32 * n] = rank + 1
        used = nd.exchange_slabs(dist, slabs, rank, world, sb)
        assert used == 32 + 32 * (2 + 3 * (world - 1))                   # prefix of the fullest slab
        for r in range(world):
            s = slabs[r * sb:(r + 1) * sb]
            nr = 2 + 3 * r
            assert int(s[:4].view(torch.int32)[0]) == nr
            assert bool((s[32:32 + 32 * nr] == r + 1).all())
        if rank == 0:
            open(os.path.join(out_dir, "ok"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_slab_exchange(tmp_path):
    port = _free_port()
    mp.spawn(_exchange_worker, args=(2, port, 16, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok").exists()


def test_single_process_helpers():
    assert nd.shard_of(None) == (0, 1)
    assert nd.local_rows(5, 1, 2) == [1, 3]
    assert nd.local_rows(2, 2, 3) == []
    assert nd.striped_samples(4, 5, 1, 2) == [5, 7]
    assert nd.striped_samples(0, 0, 0, 2) == []
    assert nd.used_slab_bytes(0, 32 + 32 * 10) == 32
    assert nd.used_slab_bytes(3, 32 + 32 * 10) == 32 + 96
    assert nd.used_slab_bytes(99, 32 + 32 * 10) == 32 + 320          # clamped to the slab's capacity
    f = torch.ones((2, 2, 4))
    assert nd.reduce_frames(f, None) is f
