"""One process per GPU: the two multi-GPU decompositions of a render, with torch.distributed (NCCL over NVLink) as plumbing.

  render_striped  sample passes / whole VCM iterations are striped over ranks (pass i on rank i % world); no exchange during
                  the render, one framebuffer sum at the end. Works for every integrator.
  render_tiles    VCM / SPPM / BDPT with pixel rows striped over ranks for both walks of every iteration; after each light
                  pass the ranks' photon slabs are all-gathered IN PLACE (dist.all_gather_into_tensor on the slab buffer the
                  C ABI exposes), every rank builds the hash grid over all photons and runs the eye pass for its rows; the
                  full-frame accumulators (eye contributions on own rows + t=1 splats anywhere) are summed at the end.

The reference has no multi-GPU path; the single-GPU image is the oracle for both (tests/test_gpu_bidir.py,
tests/test_distributed.py)."""
import numpy as np

from . import binding as nb


def shard_of(dist):
    if dist is None or not dist.is_initialized():
        return 0, 1
    return dist.get_rank(), dist.get_world_size()


def local_rows(h, rank, world):
    """Pixel rows owned by `rank` under the row-interleaved tiling of the phase API (y % world == rank)."""
    return list(range(rank, h, world))


def striped_samples(first, count, rank, world):
    """Sample / iteration indices rendered by `rank` in novum_render_device(first, count, shard)."""
    return [first + k for k in range(rank, count, world)]


def reduce_frames(colors, dist, dst=0):
    """Sum of the ranks' full-frame accumulators on rank `dst` (NCCL reduce; gloo in the CPU tests)."""
    if dist is not None and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(colors, dst)
    return colors


def render_striped(scene, cfg, cam, first, count, colors, dist=None, seed=103033):
    """colors: torch float32 (h, w, 4) CUDA tensor, accumulated into. Returns this rank's RenderStats."""
    rank, world = shard_of(dist)
    return scene.render_device(cfg, cam, first, count, colors.data_ptr(), seed=seed, rank=rank, world=world)


SLAB_HEADER_BYTES, PHOTON_BYTES = 32, 32          # csrc/vcm.cuh: kSlabHeaderBytes, 2 x 16 B per photon record


def used_slab_bytes(max_count, slab_bytes):
    """Bytes of a slab's prefix that hold `max_count` photons (header included), clamped to the slab."""
    cap = (slab_bytes - SLAB_HEADER_BYTES) // PHOTON_BYTES
    return SLAB_HEADER_BYTES + PHOTON_BYTES * max(0, min(int(max_count), cap))


def exchange_slabs(dist, slabs, rank, world, slab_bytes):
    """In-place all-gather of the ranks' photon slabs (`slabs`: flat uint8 tensor of world x slab_bytes; rank r has filled
    slab r). Every slab is `[32 B header: count][records]` at a fixed stride and a light pass fills only a fraction of the
    worst-case capacity, so only the used prefix travels: the ranks agree on the largest count (one 4-byte all-reduce + host
    read), then gather that prefix of every slab into place. Returns the bytes gathered per slab."""
    import torch
    mine = slabs[rank * slab_bytes:(rank + 1) * slab_bytes]
    count = mine[:4].view(torch.int32).clone()
    dist.all_reduce(count, op=dist.ReduceOp.MAX)
    used = used_slab_bytes(int(count.item()), slab_bytes)
    outs = [slabs[r * slab_bytes:r * slab_bytes + used] for r in range(world)]
    dist.all_gather(outs, mine[:used])
    return used


class TileRenderer:
    """Tile-sharded bidirectional render: keeps the per-rank session (buffers, slabs) alive across calls."""

    def __init__(self, scene, cfg, cam, dist=None, seed=103033):
        import torch
        self.torch, self.dist = torch, dist
        self.rank, self.world = shard_of(dist)
        self.session = nb.VcmSession(scene, cfg, cam, rank=self.rank, world=self.world, seed=seed)
        self.slabs = self.session.slabs_tensor()
        self.mine = self.session.my_slab_tensor()
        self.ext_stream = torch.cuda.ExternalStream(self.session.stream())
        self.exchanged_bytes = 0
        self.last_exchange_bytes = 0

    def close(self):
        self.session.close()

    def exchange(self):
        """All-gather of this iteration's photons on the engine's own stream (behind the light pass, ahead of the grid build)."""
        with self.torch.cuda.stream(self.ext_stream):
            used = exchange_slabs(self.dist, self.slabs, self.rank, self.world, self.session.slab_bytes)
        self.exchanged_bytes += used * self.world
        self.last_exchange_bytes = used * self.world

    def iteration(self, it, colors):
        s = self.session
        s.light_pass(it, colors.data_ptr())
        if self.world > 1:
            # the collective runs on the engine's own stream, right behind the light pass and ahead of the grid build
            self.exchange()
        s.build_grid(it)
        s.eye_pass(it, colors.data_ptr())

    def render(self, first, count, colors):
        for it in range(first, first + count):
            self.iteration(it, colors)
        nb.lib().novum_sync()
        return self.session.stats()


def frame_to_host(colors):
    return colors.detach().cpu().numpy().astype(np.float32)
