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

    def close(self):
        self.session.close()

    def iteration(self, it, colors):
        torch, dist = self.torch, self.dist
        s = self.session
        s.light_pass(it, colors.data_ptr())
        if self.world > 1:
            # the collective runs on the engine's own stream, right behind the light pass and ahead of the grid build
            with torch.cuda.stream(self.ext_stream):
                dist.all_gather_into_tensor(self.slabs, self.mine)
            self.exchanged_bytes += self.slabs.numel()
        s.build_grid(it)
        s.eye_pass(it, colors.data_ptr())

    def render(self, first, count, colors):
        for it in range(first, first + count):
            self.iteration(it, colors)
        nb.lib().novum_sync()
        return self.session.stats()


def frame_to_host(colors):
    return colors.detach().cpu().numpy().astype(np.float32)
