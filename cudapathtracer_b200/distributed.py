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


class PhaseTileRenderer:
    """Tile-sharded bidirectional render through the PHASE API (host-ordered light pass / exchange / grid build / eye pass, one iteration
    at a time); kept for the NCCL all-gather exchange and as the fallback when CUDA IPC mapping is unavailable. Keeps the per-rank session
    (buffers, slabs) alive across calls.

    exchange="p2p" (default for world > 1): the ranks swap CUDA IPC handles of their slab buffers once; every light pass
    then stores its photons into all peers over NVLink while it appends them (the all-gather is fused into the light-pass
    kernels), and the only collectives left are two 4-byte all-reduces per iteration that order the ranks on the engine's
    stream. exchange="nccl": light pass, then an NCCL all-gather of the used slab prefixes (`exchange_slabs`). Falls back to
    "nccl" when IPC mapping is not available."""

    def __init__(self, scene, cfg, cam, dist=None, seed=103033, exchange="p2p"):
        import torch
        self.torch, self.dist = torch, dist
        self.rank, self.world = shard_of(dist)
        self.session = nb.VcmSession(scene, cfg, cam, rank=self.rank, world=self.world, seed=seed)
        self.slabs = self.session.slabs_tensor()
        self.mine = self.session.my_slab_tensor()
        self.ext_stream = torch.cuda.ExternalStream(self.session.stream())
        self.exchanged_bytes = 0
        self.last_exchange_bytes = 0
        self.exchange_mode = "none" if self.world == 1 else exchange
        self.flag = torch.zeros(1, dtype=torch.int32, device=self.slabs.device) if self.world > 1 else None
        if self.exchange_mode == "p2p":
            self.exchange_mode = "p2p" if self._attach_peers() else "nccl"

    def _attach_peers(self):
        """Swap IPC handles over the process group and map every peer's slab buffer; all ranks agree on the outcome."""
        dist = self.dist
        handles = [None] * self.world
        dist.all_gather_object(handles, self.session.ipc_export())
        ok = 1
        try:
            for r, h in enumerate(handles):
                if r != self.rank:
                    self.session.ipc_attach(r, h)
            self.session.peer_write(True)
        except nb.NovumError:
            ok = 0
        t = self.torch.tensor([ok], dtype=self.torch.int32, device=self.slabs.device)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        if int(t.item()) == 0:
            try:
                self.session.peer_write(False)
            except nb.NovumError:
                pass
            return False
        return True

    def close(self):
        self.session.close()

    def _order_ranks(self):
        """Stream-ordered rendezvous: completes on a rank only after every rank's preceding work on its engine stream."""
        with self.torch.cuda.stream(self.ext_stream):
            self.dist.all_reduce(self.flag)

    def exchange(self):
        """All-gather of this iteration's photons on the engine's own stream (behind the light pass, ahead of the grid build)."""
        with self.torch.cuda.stream(self.ext_stream):
            used = exchange_slabs(self.dist, self.slabs, self.rank, self.world, self.session.slab_bytes)
        self.exchanged_bytes += used * self.world
        self.last_exchange_bytes = used * self.world

    def iteration(self, it, colors):
        s = self.session
        s.light_pass(it, colors.data_ptr())           # p2p: also writes this rank's photons + count into every peer
        if self.exchange_mode == "p2p":
            self._order_ranks()                       # all light passes (and their peer stores) are complete
        elif self.exchange_mode == "nccl":
            self.exchange()
        s.build_grid(it)                              # reads all slabs (host reads the counts here)
        if self.exchange_mode == "p2p":
            self._order_ranks()                       # nobody overwrites a slab a peer is still compacting
        s.eye_pass(it, colors.data_ptr())

    def render(self, first, count, colors):
        """Returns the RenderStats of THIS call (the session's counters are cumulative: differences are taken here), deviceMs = CUDA-event
        time on the engine's stream."""
        torch = self.torch
        before = self.session.stats()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.ext_stream)
        for it in range(first, first + count):
            self.iteration(it, colors)
        e1.record(self.ext_stream)
        nb.lib().novum_sync()
        st = self.session.stats()
        if self.exchange_mode == "p2p":               # bytes stored into peers + received from them, last iteration
            self.last_exchange_bytes = PHOTON_BYTES * int(st.photons)
            self.exchanged_bytes += self.last_exchange_bytes * count
        for f in ("closestRays", "shadowRays", "paths", "lightVertices", "kernelLaunches", "mergeQueries", "mergeCandidates", "mergeAccepted"):
            setattr(st, f, getattr(st, f) - getattr(before, f))
        st.deviceMs = e0.elapsed_time(e1)
        return st


class TileRenderer:
    """The tile renderer of the C ABI (novum_tiles_*), one rank per process: rank r owns pixel rows y % world == r; `lanes` iterations
    are in flight per rank; every light pass stores its photons into the same lane's slab on every peer over NVLink (the slab buffers
    are swapped once as CUDA IPC handles over the process group) and the per-iteration rendezvous is a device-side wait on stamps in
    the slab headers. torch.distributed only carries the handles and the barriers around set-up and tear-down."""

    def __init__(self, scene, cfg, cam, dist=None, seed=103033, lanes=4):
        import torch
        self.torch, self.dist = torch, dist
        self.rank, self.world = shard_of(dist)
        self.lanes = lanes
        self.t = nb.TileRank(scene, cfg, cam, rank=self.rank, world=self.world, seed=seed, lanes=lanes)
        self.exchange_mode = "none" if self.world == 1 else "p2p"
        self.last_exchange_bytes = 0
        if self.world > 1:
            handles = [None] * self.world
            dist.all_gather_object(handles, self.t.ipc_export())
            for r, h in enumerate(handles):
                if r != self.rank:
                    self.t.ipc_attach(r, h)
            dist.barrier()

    def render(self, first, count, colors):
        st = self.t.render(first, count, colors.data_ptr())
        # bytes this rank stored into its peers (its own photons, once per peer), per iteration
        self.last_exchange_bytes = PHOTON_BYTES * int(st.photons) * (self.world - 1) // max(1, count)
        return st

    def close(self):
        if self.dist is not None and self.dist.is_initialized() and self.world > 1:
            self.dist.barrier()            # nobody frees slabs a peer may still be storing into
        self.t.close()


def frame_to_host(colors):
    return colors.detach().cpu().numpy().astype(np.float32)
