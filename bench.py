#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native Novum core.

A "step" is one pass of the hot path over one batch of work: `--spp-per-step` sample passes (unidirectional) or whole
VCM / BDPT / SPPM iterations of the full frame per GPU. Weak scaling: N GPUs render N x as many passes per step, striped by
pass index (`--sharding striped`, default) or — bidirectional integrators only — by pixel rows with a per-iteration NCCL
all-gather of the photon slabs (`--sharding tiles`); the full-frame accumulators are summed with one NCCL reduce.

Workloads = BASELINE.json configs on the generated stand-in scenes (the reference's OBJ files are LFS pointers):
  rainbow_vcm  configs[4]  rainbow dragon, VCM, 3840x2160         (DEFAULT: the config BASELINE.json's metric is quoted on)
  dragon_nee   configs[1]  glass dragon stand-in (~871k tris) 1920x1080, NEE+MIS          (north_star's >=5x target)
  water_bdpt   configs[2]  water caustic, BDPT, 1920x1080
  water_sppm   configs[3]  water caustic, SPPM, 1024x1024 light paths (~1M photons / iteration)
  cornell      configs[0]  Cornell box 512x512 NEE+MIS
  trace_batch  SURVEY 8d micro: closest-hit / any-hit kernels alone on identical ray batches (ours vs the reference's routines)
  hashgrid     SURVEY 8d micro / a15: photon hash grid build vs the reference's buildHashGrid on identical photon sets
  frame_finish SURVEY 8f row 2: the step after the path (clean + tonemap + gamma + BGR8 rows of a 4K frame), metric Mpixels/s
  obj_parse    SURVEY 8f row 1: the step before the path (OBJ text -> device arrays, parsed on the GPU), metric MB/s of OBJ text
metric = Mrays/s: closest-hit + shadow rays actually traced (device counters); paths/s is reported beside it.

  python bench.py --gpus N --steps K --warmup W            our arm (torchrun for N > 1)
  python bench.py --impl reference ...                      the reference's own CUDA build (oracle/_ref/novum_ref), 1 GPU
"""
import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))

UNI, BDPT, NAIVE, VCM, SPPM = 0, 1, 2, 3, 4
WORKLOADS = {
    "dragon_nee": dict(gen="dragon", integ=UNI, w=1920, h=1080, spp=32, desc="configs[1]: dragon stand-in (~871k tris, glass) 1920x1080 NEE+MIS"),
    "rainbow_vcm": dict(gen="rainbow", integ=VCM, w=3840, h=2160, spp=8, desc="configs[4]: rainbow dragon stand-in (~871k tris) 3840x2160 VCM, eye 16 / light 10"),
    "water_bdpt": dict(gen="water", integ=BDPT, w=1920, h=1080, spp=2, desc="configs[2]: water caustic stand-in 1920x1080 BDPT, all strategies + MIS"),
    "water_sppm": dict(gen="water_sppm", integ=SPPM, w=1024, h=1024, spp=4, desc="configs[3]: water caustic stand-in SPPM, 1024x1024 light paths per iteration"),
    "cornell": dict(gen="cornell", integ=UNI, w=512, h=512, spp=64, desc="configs[0]: Cornell box 512x512 NEE+MIS"),
}


def peaks():
    p = os.path.join(REPO, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        reasons = []
        for i, name in enumerate(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")):
            if any(len(r) > 3 + i and r[3 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        mx = max(int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit())
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": mx, "reasons": reasons, "samples": len(sm)}


def scene_dir_for(workload, detail, rank, dist):
    """Rank 0 generates the stand-in scene once per box with the standalone command-line tool (a subprocess: neither arm has to load
    the product library for this); other ranks wait for it."""
    gen = WORKLOADS[workload]["gen"]
    d = os.path.join(tempfile.gettempdir(), f"novum_bench_{gen}_{detail}")
    marker = os.path.join(d, ".done")
    if rank == 0 and not os.path.exists(marker):
        tool = os.path.join(REPO, "cudapathtracer_b200", "novum")
        subprocess.check_call([tool, "--make-scene", gen, "--out", d, "--detail", str(detail)], stdout=subprocess.DEVNULL)
        open(marker, "w").write("ok")
    if dist is not None:
        dist.barrier()
    return d


def cpu_oracle_baseline(sdir, workload, target_s=12.0):
    """The CPU restatement of the reference's path (oracle/, kind "port") on the host cores, on a bounded sample: the same
    scene and camera at reduced resolution through the oracle's NEE+MIS walk (closest-hit + shadow rays of the reference's
    BVHSceneIntersect / BVHShadowRay restatement), sized for ~10-20 s. NOT the headline's ray mix for the bidirectional workloads
    (no CPU VCM exists, and the reference has no CPU path at all): the figure is the host's Mrays/s on this scene's geometry."""
    import cudapathtracer_b200 as nb
    from oracle.oracle_binding import OracleScene
    scene = nb.Scene.from_config(os.path.join(sdir, "configs", "config.rendertron"))
    pos, nrm, uv, tris, lights = scene.download()
    cfg = scene.cfg
    scene.close()
    ora = OracleScene(pos, nrm, uv, tris, nb.material_table(), None, 2)
    wl = WORKLOADS[workload]
    sw, sh, spp = max(16, wl["w"] // 16), max(9, wl["h"] // 16), 1
    cfg.width, cfg.height = sw, sh
    if cfg.maxDepth <= 0:
        cfg.maxDepth = 40
    cam = nb.camera_make(cfg)
    t0 = time.time()
    img, counts = ora.render_unidirectional(bytes(cam), sw, sh, cfg.maxDepth, 0, spp)
    dt = time.time() - t0
    if dt < target_s / 4:          # grow the sample towards the target duration
        spp = int(min(256, max(1, target_s / max(dt, 1e-3))))
        t0 = time.time()
        img, counts = ora.render_unidirectional(bytes(cam), sw, sh, cfg.maxDepth, 1, spp)
        dt = time.time() - t0
    rays = float(counts[0] + counts[1])
    mix = "" if wl["integ"] == UNI else " - a DIFFERENT ray mix from this workload's bidirectional integrator, which has no CPU form"
    return {"value": rays / dt / 1e6, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "port",
            "sample": f"{sw}x{sh} px x {spp} spp of the same scene/camera through the oracle's NEE+MIS walk, OpenMP on all host cores ({dt:.1f} s){mix}",
            "paths_per_s": sw * sh * spp / dt, "rays_per_path": rays / (sw * sh * spp)}


def kernel_counters():
    """Per-kernel figures read off committed ncu --set full captures of the SHIPPED kernels (profiles/r2_kernel_counters.json, written by
    tools/profile.sh + tools/ncu_extract.py): dram bytes per launch, warp instructions per launch. None when the file is absent."""
    p = os.path.join(REPO, "profiles", "r2_kernel_counters.json")
    return json.load(open(p)) if os.path.exists(p) else None


def block_means(img, k):
    h, w, c = img.shape
    return img[:h - h % k, :w - w % k].reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


def relmse_vs_reference(scene, nb):
    """The headline scene against a frame of the REFERENCE: tests/golden/rainbow1080_vcm_blocks.npz holds the 8x8-block means of novum_ref's
    own 1920x1080, 64-iteration VCM render of the same scene files (tests/golden/make_golden_materials.py). Ours: same resolution and
    iteration count through novum_render."""
    import numpy as np
    p = os.path.join(REPO, "tests", "golden", "rainbow1080_vcm_blocks.npz")
    if not os.path.exists(p):
        return None
    z = np.load(p)
    cfg = nb.RenderConfig.from_buffer_copy(bytes(scene.cfg))
    cfg.width, cfg.height, cfg.integrator = int(z["w"]), int(z["h"]), VCM
    cam = nb.camera_make(cfg)
    spp = int(z["spp"])
    img, _ = scene.render(cfg, cam, 0, spp)
    img = img[..., :3] / spp
    b = block_means(img, 8)
    ref = z["blocks"]
    return {"relmse_8x8_blocks": float(np.mean((b - ref) ** 2 / (ref ** 2 + 1e-2))), "mean_ratio": float(img.mean() / float(z["mean"])),
            "mean_ours": float(img.mean()), "mean_reference": float(z["mean"]), "width": cfg.width, "height": cfg.height, "iterations": spp,
            "fixture": "tests/golden/rainbow1080_vcm_blocks.npz (novum_ref, same scene files)", "test_bar": "mean within 3 %, block relMSE <= 0.05 (tests/test_gpu_materials.py)"}


def run_ours(args):
    os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")      # lanes x ranks never alias a hardware queue
    import numpy as np
    import torch
    import cudapathtracer_b200 as nb
    from cudapathtracer_b200 import distributed as nd

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if not torch.cuda.is_available() or nb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    torch.cuda.set_device(local)
    nb.lib().novum_set_device(local)
    dev = torch.device("cuda", local)

    wl = WORKLOADS[args.workload]
    sdir = scene_dir_for(args.workload, args.detail, rank, dist)
    t0 = time.time()
    scene = nb.Scene.from_config(os.path.join(sdir, "configs", "config.rendertron"))
    load_s = time.time() - t0
    cfg = scene.cfg
    w, h, integ = wl["w"], wl["h"], wl["integ"]
    cfg.width, cfg.height, cfg.integrator = w, h, integ
    bidir = integ in (BDPT, VCM, SPPM)
    sharding = args.sharding or ("tiles" if (bidir and world > 1) else "striped")
    tiles = bidir and sharding == "tiles"
    cam = nb.camera_make(cfg)
    info = scene.info()
    S = args.spp_per_step or wl["spp"]
    colors = torch.zeros((h, w, 4), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)       # > L2 (126 MB)
    tiler = None
    if tiles:
        tiler = (nd.TileRenderer(scene, cfg, cam, dist, lanes=args.lanes) if args.exchange == "p2p"
                 else nd.PhaseTileRenderer(scene, cfg, cam, dist, exchange="nccl"))
    pinned = torch.empty((h, w, 4), dtype=torch.float32, pin_memory=True)     # the caller's page-locked frame, reused every step

    # Which iterations a step renders:
    #   tiles   (strong scaling): step i = iterations [i*S, (i+1)*S) of the full frame, all ranks together (rows sharded, photons exchanged)
    #   striped (weak scaling)  : step i = the SAME iterations [i*S, (i+1)*S) on every rank, rank r with seed base + r - equal work per GPU
    #                             whatever N is (the VCM merge radius shrinks with the iteration index, so striding the indices over
    #                             ranks would make an N-GPU step cheaper than a 1-GPU step); the frames are N independent estimates, summed
    def step(i, profile=0, to_host=False, do_flush=True, first=None, count=None):
        """One step (passes [i*S, (i+1)*S), or [first, first + count) when given). Returns (stats, device ms of the render)."""
        first = i * S if first is None else first
        count = S if count is None else count
        scene.set_profile(profile)
        if do_flush:
            flush.fill_(i & 0xff)      # evict L2 between steps (the working set of a step, >0.4 GB, also exceeds L2)
            torch.cuda.synchronize()
        if to_host and world == 1 and not tiles:
            _, st = scene.render(cfg, cam, first, count, out=pinned.numpy())      # novum_render: renders, copies w*h*16 B to the caller's buffer
            return st, st.deviceMs
        if to_host:
            colors.zero_()
        if tiles:
            st = tiler.render(first, count, colors)
            ms = st.deviceMs
        else:
            st = scene.render_device(cfg, cam, first, count, colors.data_ptr(), seed=103033 + rank)
            ms = st.deviceMs          # CUDA events on the library's own streams around the whole call
        if to_host:
            nd.reduce_frames(colors, dist, 0)
            if rank == 0:
                pinned.copy_(colors)
            torch.cuda.synchronize()
        return st, ms

    for i in range(args.warmup):
        step(i)
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    wall0 = time.time()
    dev_ms, rays, paths, launches, stage_ms, closest, photons = 0.0, 0, 0, 0, [0.0] * 4, 0, 0
    merge_cand, merge_acc, merge_q = 0, 0, 0
    for i in range(args.steps):
        st, ms = step(args.warmup + i, profile=1)
        dev_ms += ms
        photons = st.photons
        rays += st.closestRays + st.shadowRays; closest += st.closestRays; paths += st.paths; launches += st.kernelLaunches
        merge_cand += st.mergeCandidates; merge_acc += st.mergeAccepted; merge_q += st.mergeQueries
        for k in range(4):
            stage_ms[k] += st.stageMs[k]
    torch.cuda.synchronize()
    if dist is not None:
        dist.barrier()
    wall = time.time() - wall0
    tot = torch.tensor([float(rays), float(paths), float(launches), float(closest), float(merge_cand), float(merge_acc), float(merge_q)], dtype=torch.float64, device=dev)
    mx = torch.tensor([dev_ms, wall * 1e3], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    tot_rays, tot_paths, tot_launch, tot_closest, tot_cand, tot_acc, tot_q = [float(x) for x in tot.tolist()]
    max_dev_ms, max_wall_ms = [float(x) for x in mx.tolist()]

    # ---- end to end: the SAME steps (iteration indices, L2 flush before each) through the host-facing call: render, (N > 1: NCCL sum of
    # the frames,) framebuffer copied to the caller's page-locked buffer; wall clock around each step, the flush outside it ----
    e2e_rays, e2e_s = 0, 0.0
    for i in range(args.steps):
        flush.fill_((args.warmup + i) & 0xff)
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
        t0 = time.time()
        st, _ = step(args.warmup + i, profile=0, to_host=True, do_flush=False)
        if dist is not None:
            dist.barrier()
        e2e_s += time.time() - t0
        e2e_rays += st.closestRays + st.shadowRays
    clk = clocks.stop() if rank == 0 else None
    host_img = pinned.numpy() if rank == 0 else None
    e2e_t = torch.tensor([float(e2e_rays)], dtype=torch.float64, device=dev)
    e2e_mx = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.SUM)
        dist.all_reduce(e2e_mx, op=dist.ReduceOp.MAX)

    # ---- roofline ----
    # Top level = the closest-hit traversal (wfExtend / bdExtend), the kernel north_star's memory-bound roofline is about: per ray 32 B ray
    # + 16 B hit + 80 B x nodes + 48 B x triangles visited, nodes / triangles per ray counted live by a counted step. With
    # several lanes in flight the stages of different iterations overlap, so a stage's time is its share of the summed stage times
    # applied to the step's device time.
    # The roofline figures come from their own passes, run ONE PER CALL after the timed region (the first timed step's passes again): a
    # single pass keeps a single lane busy, so the per-stage CUDA-event times are exclusive (with several iterations in flight the
    # stages of different iterations overlap and a stage's event time is not the time the GPU spent on it).
    r_stage, r_closest, r_cand, r_acc, r_q, r_ms = [0.0] * 4, 0, 0, 0, 0, 0.0
    passes_r = S if bidir else 1
    for j in range(passes_r):
        st, ms = (step(0, profile=1, first=args.warmup * S + j, count=1) if bidir else step(args.warmup, profile=1))
        r_ms += ms; r_closest += st.closestRays; r_cand += st.mergeCandidates; r_acc += st.mergeAccepted; r_q += st.mergeQueries
        for k in range(4):
            r_stage[k] += st.stageMs[k]
    st, _ = (step(0, profile=3, first=args.warmup * S, count=1) if bidir else step(args.warmup, profile=3))      # counted (slower) kernels: nodes / triangles per ray
    vn = st.bvhNodeVisits / max(1, st.closestRays); vt = st.triTests / max(1, st.closestRays)
    bytes_per_ray = 32 + 16 + 80 * vn + 48 * vt
    peak, peak_src = peaks()
    names = ({"extend": 0, "step+grid": 1, "resolve": 2, "merge": 3} if bidir else {"extend": 0, "shade": 1, "shadow": 2, "other": 3})
    ssum = sum(stage_ms) or 1.0
    eff_ms = r_stage
    closest, merge_cand, merge_acc, merge_q = r_closest, r_cand, r_acc, r_q
    kc = kernel_counters() or {}
    ext_name = "bdExtend" if bidir else "wfExtend"
    ext_kc = kc.get(args.workload, {}).get(ext_name)
    roof = {"bound": "hbm", "kernel": f"{ext_name} (BVH8 closest hit)",
            "achieved": (closest * bytes_per_ray) / (eff_ms[0] * 1e-3) / 1e9 if eff_ms[0] > 0 else 0.0,
            "peak": peak, "unit": "GB/s", "peak_source": peak_src,
            "bytes_per_ray": bytes_per_ray, "nodes_per_ray": vn, "tris_per_ray": vt,
            "closest_rays_per_s": closest / (eff_ms[0] * 1e-3) if eff_ms[0] > 0 else None, "ms_share": stage_ms[0] / ssum,
            "kernel_ms_per_pass": eff_ms[0] / passes_r, "measured_on": f"{passes_r} pass(es) run one per call after the timed region (stages exclusive)",
            "traffic": ext_kc["dram_bytes_per_launch"] if ext_kc else None,
            "traffic_source": ext_kc["source"] if ext_kc else None,
            "stage_ms_share": {k: stage_ms[v] / ssum for k, v in names.items()},
            "note": "DRAM traffic is far below the algorithmic bytes (the BVH8 is L2-resident): the kernel is bound by dependent-load latency under "
                    "SIMT divergence, the HBM fraction is the figure north_star asks for, not its limiter (profiles/README.md)"}
    roof["frac"] = roof["achieved"] / peak
    if bidir and merge_q > 0 and eff_ms[3] > 0:
        # photon merge: L1/L2-resident photon streams (every record is re-read by ~10^3 queries), so HBM is not its bound; what it does per
        # second is (vertex, photon) pair tests, at a fixed number of warp instructions per 32 pairs (ncu of the shipped kernel)
        mkc = kc.get(args.workload, {}).get("bdMergeLambert")
        sm_hz = (clk["sm_mhz"] if clk and clk.get("sm_mhz") else 1965) * 1e6
        issue_peak = 148 * 4 * sm_hz                                         # warp instructions per second
        pairs_s = merge_cand / (eff_ms[3] * 1e-3)
        m = {"bound": "issue", "kernel": "bdMergeLambert + bdMerge (photon gather over the reachable r/2 hash-grid cells of a 4-vertex group)", "pairs_per_s": pairs_s,
             "photons_scanned_per_query": merge_cand / merge_q, "photons_merged_per_query": merge_acc / merge_q,
             "scanned_per_merged": merge_cand / max(1, merge_acc), "ms_share": stage_ms[3] / ssum,
             "issue_peak_warp_inst_per_s": issue_peak}
        if mkc:
            ipp = mkc["warp_inst_per_launch"] / max(1.0, mkc["pairs_per_launch"]) * 32.0
            m.update({"warp_inst_per_32_pairs": ipp, "achieved_warp_inst_per_s": pairs_s / 32.0 * ipp, "frac": pairs_s / 32.0 * ipp / issue_peak,
                      "traffic": mkc["dram_bytes_per_launch"], "source": mkc["source"]})
        roof["merge"] = m

    if rank == 0:
        n_iter = args.steps * S * (1 if tiles else world)
        line = {
            "metric": "Mrays/s", "value": tot_rays / (max_dev_ms * 1e-3) / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": max_dev_ms / args.steps, "higher_is_better": True,
            "scaling": "strong" if tiles else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "scene": args.workload, "tris": info["tris"], "bvh8_nodes": info["bvh8_nodes"],
                       "width": w, "height": h, "passes_per_step": S, "max_depth": cfg.maxDepth,
                       "eye_depth": cfg.bdptEyeDepth, "light_depth": cfg.bdptLightDepth,
                       "iterations_timed": f"{args.warmup * S}..{(args.warmup + args.steps) * S - 1}",
                       "sharding": (f"tiles: pixel rows y % N == rank, {getattr(tiler, 'lanes', 1)} iterations in flight per rank, photons stored into the peers' slabs "
                                    f"inside the light pass ({tiler.exchange_mode}), device-side rendezvous, frames sum-reduced (NCCL) at the end"
                                    if tiles else
                                    ("striped: every rank renders the same iteration indices with its own seed (equal work per GPU), frames sum-reduced (NCCL) at the end"
                                     if world > 1 else f"one GPU; NOVUM_LANES={os.environ.get('NOVUM_LANES', '4')} iterations in flight")),
                       "l2": "256 MB flush before every step; the per-step working set (path pool / vertex cache / queues, >0.4 GB) exceeds the 126 MB L2",
                       "scene_load_s": round(load_s, 3), "scene_parse_s": round(info["parse_ms"] * 1e-3, 3), "bvh8_build_ms": round(info["bvh8_build_ms"], 3),
                       "bvh8_build_launches": info["bvh8_build_launches"]},
            "ms_per_pass": max_dev_ms / max(1, args.steps * S),
            "paths_per_s": tot_paths / (max_dev_ms * 1e-3), "rays_per_path": tot_rays / max(1.0, tot_paths),
            "wall_ms_per_step": max_wall_ms / args.steps,
            "e2e": {"value": float(e2e_t.item()) / float(e2e_mx.item()) / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": 112 + 200,
                    "d2h_bytes_per_step": w * h * 16, "ms_per_step": float(e2e_mx.item()) * 1e3 / args.steps,
                    "note": "the same steps (same iteration indices, L2 flushed before each) through the host-facing call: camera/config in, render, "
                            + ("NCCL sum of the ranks' frames, " if world > 1 else "")
                            + "framebuffer (w*h float4) copied to page-locked host memory every step; wall clock"},
            "gpu_launches": int(tot_launch),
            "clocks": clk,
            "roofline": roof,
        }
        if bidir:
            line["photons_per_iteration"] = int(photons // max(1, S)) * (world if tiles else 1)
            line["merge"] = {"queries_per_step": int(tot_q) // max(1, args.steps), "candidates_per_step": int(tot_cand) // max(1, args.steps),
                             "merged_per_step": int(tot_acc) // max(1, args.steps)}
        if tiles:
            line["exchange_bytes_per_iteration_per_gpu"] = int(tiler.last_exchange_bytes)
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_oracle_baseline(sdir, args.workload)
        if host_img is not None:
            line["image_mean"] = float(host_img[..., :3].mean() / max(1, S * (1 if tiles else world)))
        if args.workload == "rainbow_vcm" and args.detail == 100 and not args.no_relmse:
            line["relmse_vs_reference"] = relmse_vs_reference(scene, nb)
        if world == 1 and args.workload == "rainbow_vcm" and not args.no_secondary:
            line["secondary"] = secondary_dragon_nee(args, nb)
        print(json.dumps(line))
    if tiler:
        tiler.close()
    scene.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def secondary_dragon_nee(args, nb):
    """north_star's other stated target (>= 5x the reference build's Mrays/s on the dragon NEE scene, configs[1]) measured in the same run:
    2 timed steps of 32 spp at 1920x1080 after 1 warm-up step; the reference arm prints the matching figure under the same key."""
    wl = WORKLOADS["dragon_nee"]
    sdir = scene_dir_for("dragon_nee", args.detail, 0, None)
    scene = nb.Scene.from_config(os.path.join(sdir, "configs", "config.rendertron"))
    try:
        cfg = scene.cfg
        cfg.width, cfg.height, cfg.integrator = wl["w"], wl["h"], wl["integ"]
        cam = nb.camera_make(cfg)
        import torch
        colors = torch.zeros((wl["h"], wl["w"], 4), dtype=torch.float32, device="cuda")
        S = wl["spp"]
        scene.render_device(cfg, cam, 0, S, colors.data_ptr())
        ms, rays = 0.0, 0
        for i in (1, 2):
            st = scene.render_device(cfg, cam, i * S, S, colors.data_ptr())
            ms += st.deviceMs; rays += st.closestRays + st.shadowRays
        return {"workload": wl["desc"], "metric": "Mrays/s", "value": rays / (ms * 1e-3) / 1e6, "ms_per_step": ms / 2, "passes_per_step": S,
                "steps": 2, "warmup": 1, "paths_per_s": wl["w"] * wl["h"] * S * 2 / (ms * 1e-3)}
    finally:
        scene.close()


def run_reference(args):
    """The reference's own CUDA build (unmodified sources, oracle/_ref/novum_ref) on ONE B200 of this box: it is GPU-only, so this is the
    baseline north_star names. Nothing of the product is imported here: the scene comes from the standalone generator, the work runs in
    novum_ref subprocesses. Its rays per pass come from the counting build of the same sources (novum_ref_count, one pass, never timed).
    `value` is the STOCK code path, whose BDPT / VCM launchers stop every >= 5 s to write render.bmp + an 8.3 M-value CSV
    (deviceCode.cu:1893-1934, 3447-3468); `nosave` is the same sources with that switch read as false (novum_ref_nosave)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    refc = os.path.join(REPO, "oracle", "_ref", "novum_ref_count")
    refn = os.path.join(REPO, "oracle", "_ref", "novum_ref_nosave")
    wl = WORKLOADS[args.workload]
    sdir = scene_dir_for(args.workload, args.detail, 0, None)
    w, h = wl["w"], wl["h"]
    S = args.spp_per_step or wl["spp"]
    if not os.path.exists(ref):
        base = cpu_oracle_baseline(sdir, args.workload, target_s=8.0)
        line = {"impl": "reference", "metric": "Mrays/s", "value": base["value"], "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps,
                "warmup": args.warmup, "higher_is_better": True, "config": {"workload": wl["desc"]}, "cpu_baseline": base,
                "e2e": {"value": base["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "note": "oracle/_ref/novum_ref not built on this box: CPU port of the reference's algorithm instead"}
        print(json.dumps(line))
        return

    def run(binary, scene_dir, spp):
        out = subprocess.check_output([binary, "render", scene_dir, "--spp", str(spp), "--out", os.path.join(tempfile.gettempdir(), "ref.raw")],
                                      stderr=subprocess.STDOUT).decode()
        return json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", out).group(1))

    bidir = wl["integ"] in (BDPT, VCM, SPPM)
    # bounded sample: the timed window is passes [W*S, W*S + P) with P = min(K*S, cap): the metric is a rate, and a 4K VCM pass costs
    # the reference ~2.6 s
    cap = {"rainbow_vcm": 16, "water_bdpt": 24, "water_sppm": 48, "dragon_nee": 64, "cornell": 256}[args.workload]
    passes = max(1, min(args.steps * S, cap))
    wpass = max(1, args.warmup * S)

    def timed(binary, p=None):
        p = passes if p is None else p
        a = run(binary, sdir, wpass)                          # includes curand_init, allocations: subtracted out below
        b = run(binary, sdir, wpass + p)
        return b["launch_ms"] - a["launch_ms"], b

    ms, b = timed(ref)
    paths = w * h * passes * (2 if bidir else 1)
    if os.path.exists(refc):
        nc = 1 if bidir else min(4, S)
        c = run(refc, sdir, nc)
        rays_per_pass = (c["closest_rays"] + c["shadow_rays"]) / nc
        how = f"device counters of the counting build of the same sources over {nc} pass(es)"
    else:
        base = cpu_oracle_baseline(sdir, args.workload, target_s=8.0)
        rays_per_pass = base["rays_per_path"] * w * h
        how = "rays per path of the CPU restatement"
    rays = rays_per_pass * passes
    val = rays / (ms * 1e-3) / 1e6
    line = {"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / passes * S, "ms_per_pass": ms / passes, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "scene": args.workload, "tris": b["tris"], "width": w, "height": h, "passes_per_step": S,
                       "passes_timed": f"{wpass}..{wpass + passes - 1}",
                       "timing": "launch_* CUDA-event time at W*S + P passes minus at W*S passes (drops curand_init + allocation); stock code path",
                       "rays": how},
            "paths_per_s": paths / (ms * 1e-3), "rays_per_path": rays / paths,
            "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference",
                             "sample": f"{passes} passes at {w}x{h} on one B200 (the reference has no CPU path); bounded sample of the {args.steps} x {S} passes of the other arm"},
            "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "reference_bvh_build_s": b["bvh_build_s"]}
    if bidir and os.path.exists(refn):
        pn = max(1, min(passes, 8))                            # a shorter window: the launcher has to replay the warm-up passes for every timing
        ms_n, _ = timed(refn, pn)
        line["nosave"] = {"value": rays_per_pass * pn / (ms_n * 1e-3) / 1e6, "ms_per_pass": ms_n / pn, "passes_timed": f"{wpass}..{wpass + pn - 1}",
                          "note": "same sources with DO_PROGRESSIVERENDER (util.cuh:33) read as false inside deviceCode.cu: no 5-second render.bmp / renderCSV.csv "
                                  "dump inside the timed launch; `value` above is the stock path, which the ratio is taken against"}
    if args.workload == "rainbow_vcm" and not args.no_secondary:
        wl2 = WORKLOADS["dragon_nee"]
        sdir2 = scene_dir_for("dragon_nee", args.detail, 0, None)
        S2 = wl2["spp"]
        a2, b2 = run(ref, sdir2, S2), run(ref, sdir2, 3 * S2)
        ms2 = b2["launch_ms"] - a2["launch_ms"]
        sec = {"workload": wl2["desc"], "ms_per_step": ms2 / 2, "passes_per_step": S2, "steps": 2, "warmup": 1,
               "paths_per_s": wl2["w"] * wl2["h"] * S2 * 2 / (ms2 * 1e-3)}
        if os.path.exists(refc):
            c2 = run(refc, sdir2, 2)
            sec["metric"] = "Mrays/s"; sec["value"] = (c2["closest_rays"] + c2["shadow_rays"]) / 2 * S2 * 2 / (ms2 * 1e-3) / 1e6
        line["secondary"] = sec
    print(json.dumps(line))


FRAME_DESC = "frame finish (SURVEY 8f row 2): 3840x2160 accumulator + overlay -> cleaned, ACES-tonemapped, gamma-corrected BGR8 .bmp rows"


def frame_inputs(w, h, seed=5):
    import numpy as np
    rng = np.random.default_rng(seed)
    acc = np.zeros((h, w, 4), np.float32)
    acc[..., :3] = (rng.random((h, w, 3), dtype=np.float32) ** 3) * 60.0
    acc[..., 3] = 1.0
    ov = np.zeros_like(acc)
    idx = rng.choice(h * w, h * w // 100, replace=False)
    ov.reshape(-1, 4)[idx, :3] = rng.random((idx.size, 3), dtype=np.float32)
    acc.reshape(-1, 4)[idx[:64], 0] = np.nan
    return acc, ov


def ref_bmp_ms(acc_frame, w, h, reps):
    """The reference's own host code for this row (Image::setColor + saveImageBMP, one core) through oracle/_ref/novum_ref bmp."""
    ref = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    if not os.path.exists(ref):
        return None
    tmp = tempfile.mkdtemp()
    raw = os.path.join(tmp, "f.raw")
    acc_frame.tofile(raw)
    out = subprocess.check_output([ref, "bmp", raw, str(w), str(h), "1", os.path.join(tmp, "o.bmp"), "--time", str(reps)], stderr=subprocess.STDOUT).decode()
    return json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", out).group(1))["host_ms"]


def run_frame(args):
    """--workload frame_finish: the step after the path. One step = one 4K frame finished. `value` = device-resident kernel
    throughput; `e2e` = device accumulators -> .bmp file on disk (kernel + 3 B/pixel read-back + write) per step; the reference
    arm's equivalent is two float4 frame copies + its single-core host loop + the same file write."""
    import numpy as np
    import torch
    import cudapathtracer_b200 as nb
    from oracle import oracle_binding as ob
    if not torch.cuda.is_available() or nb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    w, h, n = 3840, 2160, 64
    acc, ov = frame_inputs(w, h)
    d_acc = torch.from_numpy(acc).cuda(); d_ov = torch.from_numpy(ov).cuda()
    rowb = nb.bmp_row_bytes(w)
    d_rows = torch.empty((h, rowb), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    clocks = ClockSampler(0)
    for i in range(args.warmup):
        nb.frame_finish_device(d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, d_rows.data_ptr(), None, stream.cuda_stream)
    flush.fill_(1)
    torch.cuda.synchronize()
    clocks.start()
    # K steps back to back inside one event pair: every step streams its two inputs (265 MB, more than twice the 126 MB L2, read
    # front to back) from HBM again — ncu shows 265 MB of DRAM reads per launch in exactly this sequence (profiles/) — so no flush
    # kernel sits inside the timed region
    ev0.record(stream)
    for i in range(args.steps):
        nb.frame_finish_device(d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, d_rows.data_ptr(), None, stream.cuda_stream)
    ev1.record(stream)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / args.steps
    # parity of this very run against the restatement (sampled rows; the full frame is covered by tests/test_gpu_frame.py)
    want = ob.bmp_rows(ob.frame_clean(acc[:64], ov[:64], n, final_mode=True), True)
    parity = bool(np.array_equal(d_rows[:64].cpu().numpy(), want))
    out = os.path.join(tempfile.gettempdir(), "novum_bench_frame.bmp")
    for i in range(args.warmup):
        nb.frame_save_bmp(out, d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True)
    t0 = time.time()
    for i in range(args.steps):
        nb.frame_save_bmp(out, d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True)
    e2e_ms = (time.time() - t0) * 1e3 / args.steps
    # the asynchronous variant: how long the render stream is held per save (events around the call on that stream), and the
    # time per save when they are queued back to back (read-back and file write of save k overlap the kernel of save k+1)
    n_async = min(args.steps, 8)
    outs = [os.path.join(tempfile.gettempdir(), f"novum_bench_frame_async_{i & 1}.bmp") for i in range(n_async)]
    nb.frame_save_bmp_async(outs[0], d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, stream.cuda_stream)
    nb.frame_save_wait(); stream.synchronize()
    a0 = [torch.cuda.Event(enable_timing=True) for _ in range(n_async)]; a1 = [torch.cuda.Event(enable_timing=True) for _ in range(n_async)]
    t_queue = 0.0
    for i in range(n_async):                                   # saves spaced apart, as a progressive render issues them (one every few seconds)
        a0[i].record(stream)
        t0 = time.time()
        nb.frame_save_bmp_async(outs[i], d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, stream.cuda_stream)
        t_queue += (time.time() - t0) * 1e3 / n_async
        a1[i].record(stream)
        nb.frame_save_wait(); stream.synchronize()
    stall_ms = float(np.median([a0[i].elapsed_time(a1[i]) for i in range(n_async)]))
    t0 = time.time()
    for i in range(n_async):                                   # back to back: bounded by the file writes of the two slots
        nb.frame_save_bmp_async(outs[i], d_acc.data_ptr(), d_ov.data_ptr(), w, h, n, nb.FRAME_FINAL, True, stream.cuda_stream)
    nb.frame_save_wait(); stream.synchronize()
    async_ms = (time.time() - t0) * 1e3 / n_async
    async_same = bool(np.array_equal(np.fromfile(outs[-1], np.uint8), np.fromfile(out, np.uint8)))
    clk = clocks.stop()
    px = w * h
    algo = px * (16 + 16 + 3)
    peak, peak_src = peaks()
    line = {"metric": "Mpixels/s", "value": px / (ms * 1e-3) / 1e6, "unit": "Mpixels/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": FRAME_DESC, "width": w, "height": h, "mode": "final (main.cu:855-884), overlay present, postProcess on",
                       "l2": "inputs (2 x 133 MB) exceed the 126 MB L2 and are streamed front to back every step (265 MB of DRAM reads per launch in ncu); one flush before the timed region"},
            "e2e": {"value": px / (e2e_ms * 1e-3) / 1e6, "unit": "Mpixels/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": rowb * h,
                    "note": "novum_frame_save_bmp: device accumulators -> kernel -> 3 B/pixel read-back into page-locked memory -> .bmp on disk; the "
                            "accumulators are produced on the device by the render, so there is no H2D"},
            "save_async": {"render_stream_stall_ms": stall_ms, "host_call_ms": t_queue, "ms_per_save_back_to_back": async_ms, "saves": n_async,
                           "file_equals_blocking_variant": async_same,
                           "note": "novum_frame_save_bmp_async: the render stream waits for the finish kernel only; read-back + file write run "
                                   "behind it on the save stream / a worker thread (the blocking variant holds the caller for e2e.ms_per_step)"},
            "gpu_launches": args.steps, "clocks": clk, "parity_rows_checked": 64, "parity_ok": parity,
            "roofline": {"bound": "hbm", "kernel": "frameFinishKernel", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                         "frac": algo / (ms * 1e-3) / 1e9 / peak, "traffic": 285.0e6, "peak_source": peak_src,
                         "traffic_source": "ncu --set full of this command: dram__bytes_read 265.4 MB + dram__bytes_write 19.6 MB per launch (profiles/r1_frameFinish_final.txt); algorithmic bytes per launch 290.3 MB",
                         "bytes_per_pixel": 35, "note": "16 B accumulator + 16 B overlay read, 3 B written per pixel"}}
    if not args.no_cpu_baseline:
        fin = ob.frame_clean(acc, ov, n, final_mode=True)
        rms = ref_bmp_ms(fin, w, h, 3)
        if rms is not None:
            line["cpu_baseline"] = {"value": px / (rms * 1e-3) / 1e6, "unit": "Mpixels/s", "cores": 1, "kind": "reference", "ms_per_frame": rms,
                                    "sample": "the reference's own host code (Image::setColor + postProcessImage + saveImageBMP, single-threaded as "
                                              "shipped) on the same 3840x2160 cleaned frame, best of 3, file write included; its two 133 MB D2H copies "
                                              "and the cleaning loop (main.cu:855-884) are NOT included"}
    print(json.dumps(line))


def run_frame_reference(args):
    import numpy as np
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from oracle import oracle_binding as ob
    w, h, n = 3840, 2160, 64
    acc, ov = frame_inputs(w, h)
    fin = ob.frame_clean(acc, ov, n, final_mode=True)
    rms = ref_bmp_ms(fin, w, h, max(1, args.steps))
    kind = "reference"
    if rms is None:                                  # reference build absent: time the restatement (OpenMP) instead
        t0 = time.time(); ob.bmp_rows(fin, True); rms = (time.time() - t0) * 1e3; kind = "port"
    val = w * h / (rms * 1e-3) / 1e6
    print(json.dumps({"impl": "reference", "metric": "Mpixels/s", "value": val, "unit": "Mpixels/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": rms, "higher_is_better": True, "config": {"workload": FRAME_DESC, "width": w, "height": h},
                      "cpu_baseline": {"value": val, "unit": "Mpixels/s", "cores": 1, "kind": kind,
                                       "sample": "Image::setColor + postProcessImage + saveImageBMP on one 3840x2160 frame, best of steps"},
                      "e2e": {"value": val, "unit": "Mpixels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


OBJ_DESC = "OBJ front end (SURVEY 8f row 1): the dragon stand-in's 92 MB of OBJ text -> positions / normals / uvs / 80 B Triangle records in device memory"


def run_obj(args):
    """--workload obj_parse: the step before the path. One step = one parse of the dragon stand-in's OBJ file into device arrays.
    `value` = text already in HBM (novum_obj_device_append_device_text: 12 launches + 3 count read-backs); `e2e` = file (page cache)
    -> pinned double buffers -> HBM -> arrays (novum_obj_device_append_file). Parity of this very run: the arrays equal the host
    loader's bit for bit (novum_obj_parse, the restatement of main.cu:936-1068 that tests/test_host.py pins)."""
    import numpy as np
    import torch
    import cudapathtracer_b200 as nb
    if not torch.cuda.is_available() or nb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device - the product path has no CPU fallback")
    sdir = scene_dir_for("dragon_nee", args.detail, 0, None)
    path = os.path.join(sdir, "scenedata", "dragon_vrip.obj")
    raw = np.fromfile(path, np.uint8)
    nbytes = int(raw.size)
    d_text = torch.from_numpy(raw).cuda()
    m = nb.ObjDeviceMesh()
    clocks = ClockSampler(0)
    for i in range(max(args.warmup, 1)):
        m.reset(); m.append_device_text(d_text.data_ptr(), nbytes, (0.0, 0.0, 0.0), 5, 0.0)
    got = m.download()
    torch.cuda.synchronize()
    clocks.start()
    kernel_ms, wall = [], []
    for i in range(args.steps):
        m.reset()
        t0 = time.time()
        m.append_device_text(d_text.data_ptr(), nbytes, (0.0, 0.0, 0.0), 5, 0.0)
        wall.append((time.time() - t0) * 1e3)
        kernel_ms.append(m.timings()["kernel_ms"])
    ms = float(np.mean(kernel_ms))
    e2e, up = [], []
    for i in range(args.warmup + args.steps):
        m.reset()
        t0 = time.time()
        m.append_file(path, (0.0, 0.0, 0.0), 5, 0.0)
        if i >= args.warmup:
            e2e.append((time.time() - t0) * 1e3); up.append(m.timings()["upload_ms"])
    clk = clocks.stop()
    e2e_ms = float(np.mean(e2e))
    n = m.counts()
    t0 = time.time(); ref = nb.obj_parse(path, (0.0, 0.0, 0.0), 5, 0.0); host_ms = (time.time() - t0) * 1e3
    t0 = time.time(); ref = nb.obj_parse(path, (0.0, 0.0, 0.0), 5, 0.0); host_ms = min(host_ms, (time.time() - t0) * 1e3)
    parity = all(a.shape == b.shape and a.tobytes() == b.tobytes() for a, b in zip(ref[:4], got[:4]))
    m.close()
    out_bytes = 16 * n[0] + 16 * n[1] + 8 * n[2] + 80 * n[3]
    algo = nbytes + out_bytes                 # text read once + every record written once
    peak, peak_src = peaks()
    line = {"metric": "MB/s of OBJ text", "value": nbytes / (ms * 1e-3) / 1e6, "unit": "MB/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": OBJ_DESC, "text_bytes": nbytes, "positions": n[0], "normals": n[1], "uvs": n[2], "triangles": n[3],
                       "l2": "92 MB of text + 100 MB of records per step against a 126 MB L2; the text is re-read by three line passes",
                       "timing": "CUDA events around the 12 launches of one parse on the parser's stream (includes the three count read-backs between them)"},
            "e2e": {"value": nbytes / (e2e_ms * 1e-3) / 1e6, "unit": "MB/s", "ms_per_step": e2e_ms, "read_and_copy_ms": float(np.mean(up)),
                    "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 3 * 32,
                    "note": "novum_obj_device_append_file: pread of the file (page cache) into pinned double buffers by 4 threads, copies queued as the "
                            "pieces arrive, then the parse; the arrays stay in device memory for the BVH8 build (nothing but three 32 B count records comes back)"},
            "gpu_launches": 12 * args.steps, "clocks": clk, "parity_ok": bool(parity), "parity": "all four arrays equal novum_obj_parse (host loader) bit for bit",
            "host_call_ms": float(np.mean(wall)),
            "roofline": {"bound": "hbm", "kernel": "objEmit + objFaceCount + objPositions (one thread per line)", "achieved": algo / (ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": algo / (ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src,
                         "algorithmic_bytes": algo,
                         "note": "text read once + records written once over the whole 12-launch parse; the parse is latency-bound byte work (one thread walks "
                                 "one ~70-character line), far from the HBM roof - it is sized against the 3.6 ms it takes to get the text to the GPU"},
            "cpu_baseline": {"value": nbytes / (host_ms * 1e-3) / 1e6, "unit": "MB/s", "cores": min(16, os.cpu_count() or 1), "kind": "port", "ms": host_ms,
                             "sample": "novum_obj_parse (the product's own multi-threaded host loader = main.cu:936-1068 restated) on the same file, best of 2"}}
    print(json.dumps(line))


def run_obj_reference(args):
    """The reference's own readObjSimple (single-threaded istringstream parser, main.cu:936-1068) on the same file, timed inside novum_ref."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    ref = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    sdir = scene_dir_for("dragon_nee", args.detail, 0, None)
    path = os.path.join(sdir, "scenedata", "dragon_vrip.obj")
    nbytes = os.path.getsize(path)
    if not os.path.exists(ref):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/novum_ref not built on this box"}))
        return
    best = None
    for i in range(max(1, min(args.steps, 3))):
        out = subprocess.check_output([ref, "render", sdir, "--spp", "1", "--out", os.path.join(tempfile.gettempdir(), "ref.raw")], stderr=subprocess.STDOUT).decode()
        j = json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", out).group(1))
        if "obj_parse_s" not in j:
            print(json.dumps({"impl": "reference", "unavailable": "novum_ref on this box predates the obj_parse_s field"}))
            return
        best = j["obj_parse_s"] if best is None else min(best, j["obj_parse_s"])
    val = nbytes / best / 1e6
    print(json.dumps({"impl": "reference", "metric": "MB/s of OBJ text", "value": val, "unit": "MB/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": best * 1e3, "higher_is_better": True, "config": {"workload": OBJ_DESC, "text_bytes": nbytes},
                      "cpu_baseline": {"value": val, "unit": "MB/s", "cores": 1, "kind": "reference",
                                       "sample": "readObjSimple over the scene's meshes inside novum_ref (all but 2 KB of the text is the dragon), best of up to 3 runs"},
                      "e2e": {"value": val, "unit": "MB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


TRACE_DESC = "ray batches on the dragon stand-in's BVH (SURVEY 8d micro): primary-like, cosine-bounce secondary and shadow rays, closest-hit / any-hit"


def trace_ray_sets(scene, cfg, n_side, dev):
    """Three ray batches on the GPU (torch), deterministic: primary-like rays from the camera position through an n_side x n_side
    grid (row-major, coherent), cosine-distributed bounce rays from their hit points (incoherent), and shadow rays from the hit
    points to a jittered point light above the scene with maxT and the hit triangle as skipTri."""
    import numpy as np
    import torch
    info = scene.info()
    mn = torch.tensor(info["min"], device=dev, dtype=torch.float32); mx = torch.tensor(info["max"], device=dev, dtype=torch.float32)
    ctr = (mn + mx) * 0.5
    radius = float(torch.linalg.norm(mx - ctr))
    cam = torch.tensor([cfg.camPos[0], cfg.camPos[1], cfg.camPos[2]], device=dev, dtype=torch.float32)
    fwd = ctr - cam; fwd = fwd / torch.linalg.norm(fwd)
    up = torch.tensor([0.0, 1.0, 0.0], device=dev)
    right = torch.linalg.cross(fwd, up); right = right / torch.linalg.norm(right)
    up = torch.linalg.cross(right, fwd)
    g = torch.Generator(device=dev); g.manual_seed(1234)
    u = (torch.arange(n_side, device=dev, dtype=torch.float32) + 0.5) / n_side * 2 - 1
    yy, xx = torch.meshgrid(-u, u, indexing="ij")
    tgt = ctr[None, None, :] + 0.7 * radius * (xx[..., None] * right + yy[..., None] * up)
    d = tgt - cam
    d = d / torch.linalg.norm(d, dim=-1, keepdim=True)
    n = n_side * n_side
    prim = torch.zeros((n, 8), device=dev, dtype=torch.float32)
    prim[:, 0:3] = cam; prim[:, 4:7] = d.reshape(n, 3)
    hits = torch.zeros((n, 4), device=dev, dtype=torch.int32)
    scene.trace_closest_device(prim.data_ptr(), n, hits.data_ptr(), True)
    tri = hits[:, 0].long(); t = hits[:, 1].view(torch.float32)
    ok = tri >= 0
    pos, _, _, tris, _ = scene.download()
    P = torch.from_numpy(np.ascontiguousarray(pos[:, :3])).to(dev)
    ia = torch.from_numpy(tris["aInd"].astype(np.int64)).to(dev); ib = torch.from_numpy(tris["bInd"].astype(np.int64)).to(dev)
    ic = torch.from_numpy(tris["cInd"].astype(np.int64)).to(dev)
    tri = tri[ok]; o = prim[ok, 0:3] + t[ok, None] * prim[ok, 4:7]; din = prim[ok, 4:7]
    ng = torch.linalg.cross(P[ib[tri]] - P[ia[tri]], P[ic[tri]] - P[ia[tri]])
    ng = ng / torch.linalg.norm(ng, dim=-1, keepdim=True).clamp_min(1e-30)
    ng = torch.where((ng * din).sum(-1, keepdim=True) > 0, -ng, ng)
    m = tri.numel()
    r1 = torch.rand(m, device=dev, generator=g); r2 = torch.rand(m, device=dev, generator=g) * (2 * np.pi)
    a = torch.where(ng[:, 0:1].abs() > 0.9, torch.tensor([0.0, 1.0, 0.0], device=dev), torch.tensor([1.0, 0.0, 0.0], device=dev))
    tx = torch.linalg.cross(ng, a.expand(m, 3)); tx = tx / torch.linalg.norm(tx, dim=-1, keepdim=True)
    ty = torch.linalg.cross(ng, tx)
    sr = torch.sqrt(r1)
    dsec = sr[:, None] * torch.cos(r2)[:, None] * tx + sr[:, None] * torch.sin(r2)[:, None] * ty + torch.sqrt(1 - r1)[:, None] * ng
    sec = torch.zeros((m, 8), device=dev, dtype=torch.float32)
    sec[:, 0:3] = o + 1e-3 * ng; sec[:, 4:7] = dsec / torch.linalg.norm(dsec, dim=-1, keepdim=True)
    L = ctr + torch.tensor([0.0, 1.0, 0.0], device=dev) * (0.9 * float(mx[1] - ctr[1])) + (torch.rand((m, 3), device=dev, generator=g) - 0.5) * 0.05 * radius
    dl = L - o; dist = torch.linalg.norm(dl, dim=-1)
    sh = torch.zeros((m, 8), device=dev, dtype=torch.float32)
    sh[:, 0:3] = o; sh[:, 4:7] = dl / dist[:, None]
    return {"primary": (prim, None, None), "secondary": (sec.contiguous(), None, None),
            "shadow": (sh.contiguous(), (dist * (1 - 1e-3)).contiguous(), tri.int().contiguous())}


def ref_trace(sdir, rays, max_t, skip, tmp):
    """The reference's own BVHSceneIntersect / BVHShadowRay (unmodified, oracle/_ref/novum_ref trace) on the same ray batch on this
    box's GPU: returns (kernel ms, outputs)."""
    import numpy as np
    ref = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    rp, op = os.path.join(tmp, "rays.bin"), os.path.join(tmp, "out.bin")
    with open(rp, "wb") as f:
        np.array([len(rays)], np.int32).tofile(f); rays.tofile(f)
        if max_t is not None:
            max_t.tofile(f); skip.tofile(f)
    cmd = [ref, "trace", sdir, rp, op] + (["--shadow"] if max_t is not None else [])
    out = subprocess.check_output(cmd, stderr=subprocess.STDOUT).decode()
    js = json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", out).group(1))
    res = np.fromfile(op, np.float32).reshape(-1, 4)
    os.remove(rp); os.remove(op)
    return js["kernel_ms"], res


def run_trace(args):
    """--workload trace_batch: the intersection routines alone, on identical ray batches for both arms. One step = the three
    batches traced once. value = device-resident Mrays/s (CUDA events inside the library around each kernel); e2e = the same
    through the host-buffer entry points (rays H2D, hits D2H inside the timed region); parity of THIS run's outputs against the
    reference's own routines (bit-exact classes) is part of the line when oracle/_ref/novum_ref is on the box."""
    import numpy as np
    import torch
    import cudapathtracer_b200 as nb
    from scene_utils import classify
    if int(os.environ.get("RANK", "0")) != 0:
        return
    ref_bin = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    sdir = scene_dir_for("dragon_nee", args.detail, 0, None)
    tmp = tempfile.mkdtemp()
    if not torch.cuda.is_available() or nb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = torch.device("cuda", 0)
    scene = nb.Scene.from_config(os.path.join(sdir, "configs", "config.rendertron"))
    info = scene.info()
    sets = trace_ray_sets(scene, scene.cfg, args.rays_side, dev)
    HIT = np.dtype([("tri", "<i4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4")])
    if args.impl == "reference":
        tot_ms, tot_n, per = 0.0, 0, {}
        for name, (rays, mt, sk) in sets.items():
            ms, _ = ref_trace(sdir, rays.cpu().numpy(), None if mt is None else mt.cpu().numpy(), None if sk is None else sk.cpu().numpy(), tmp)
            per[name] = {"rays": len(rays), "kernel_ms": ms, "mrays_per_s": len(rays) / ms / 1e3}
            tot_ms += ms; tot_n += len(rays)
        val = tot_n / tot_ms / 1e3
        print(json.dumps({"impl": "reference", "metric": "Mrays/s", "value": val, "unit": "Mrays/s", "n_gpus": 1, "steps": 3, "warmup": 0,
                          "ms_per_step": tot_ms, "higher_is_better": True, "config": {"workload": TRACE_DESC, "tris": info["tris"], "rays_side": args.rays_side},
                          "sets": per, "cpu_baseline": {"value": val, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference",
                                                        "sample": "the reference's BVHSceneIntersect / BVHShadowRay kernels (one thread per ray, best of 3) on the same batches on one B200"},
                          "e2e": {"value": val, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    outs, per = {}, {}
    clocks = ClockSampler(0)

    def run_set(name):
        rays, mt, sk = sets[name]
        n = len(rays)
        if mt is None:
            out = torch.zeros((n, 4), device=dev, dtype=torch.int32)
            scene.trace_closest_device(rays.data_ptr(), n, out.data_ptr(), True)
        else:
            out = torch.zeros((n, 4), device=dev, dtype=torch.float32)
            scene.trace_shadow_device(rays.data_ptr(), mt.data_ptr(), sk.data_ptr(), n, out.data_ptr(), True)
        return out, scene.last_trace_ms()

    for i in range(args.warmup):
        for name in sets:
            run_set(name)
    clocks.start()
    tot_ms, tot_n = 0.0, 0
    for name in sets:
        per[name] = {"rays": len(sets[name][0]), "kernel_ms": 0.0}
    for i in range(args.steps):
        for name in sets:
            flush.fill_(i)                       # evict the BVH (52 MB) and the rays from L2 between launches
            torch.cuda.synchronize()
            outs[name], ms = run_set(name)
            per[name]["kernel_ms"] += ms / args.steps
    for name in sets:
        per[name]["mrays_per_s"] = per[name]["rays"] / per[name]["kernel_ms"] / 1e3
        tot_ms += per[name]["kernel_ms"]; tot_n += per[name]["rays"]
    # nodes / triangles per ray (counted launch, not timed) -> algorithmic bytes of the closest-hit batches
    scene.counters(True)
    for name in ("primary", "secondary"):
        scene.counters(True)
        run_set(name)
        nv_, nt_ = scene.counters(True)
        nr = per[name]["rays"]
        per[name].update({"nodes_per_ray": nv_ / nr, "tris_per_ray": nt_ / nr, "bytes_per_ray": 48 + 80 * nv_ / nr + 48 * nt_ / nr})
    scene.counters(False)
    # e2e: page-locked host buffers through novum_trace_closest / novum_trace_shadow (rays H2D, results D2H inside the call)
    host = {k: (r.cpu().numpy(), None if m is None else m.cpu().numpy(), None if s_ is None else s_.cpu().numpy()) for k, (r, m, s_) in sets.items()}
    pinned = {}
    for k, (r, m, s_) in sets.items():
        pr = torch.empty(r.shape, dtype=torch.float32, pin_memory=True); pr.copy_(r)
        po = torch.empty((len(r), 4), dtype=torch.float32, pin_memory=True)
        pm = ps = None
        if m is not None:
            pm = torch.empty(m.shape, dtype=torch.float32, pin_memory=True); pm.copy_(m)
            ps = torch.empty(s_.shape, dtype=torch.int32, pin_memory=True); ps.copy_(s_)
        pinned[k] = (pr, pm, ps, po)
    torch.cuda.synchronize()
    L = nb.lib()
    t0 = time.time()
    for k, (pr, pm, ps, po) in pinned.items():
        if pm is None:
            rc = L.novum_trace_closest(scene.h, pr.data_ptr(), len(pr), po.data_ptr())
        else:
            rc = L.novum_trace_shadow(scene.h, pr.data_ptr(), pm.data_ptr(), ps.data_ptr(), len(pr), po.data_ptr())
        if rc:
            raise SystemExit(f"trace failed: {rc}")
    e2e_s = time.time() - t0
    clk = clocks.stop()                 # sampled over the timed launches, the counted launches and the end-to-end calls
    peak, peak_src = peaks()
    bpr = (per["primary"]["bytes_per_ray"] * per["primary"]["rays"] + per["secondary"]["bytes_per_ray"] * per["secondary"]["rays"])
    closest_ms = per["primary"]["kernel_ms"] + per["secondary"]["kernel_ms"]
    line = {"metric": "Mrays/s", "value": tot_n / tot_ms / 1e3, "unit": "Mrays/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": tot_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": TRACE_DESC, "tris": info["tris"], "bvh8_nodes": info["bvh8_nodes"], "rays_side": args.rays_side,
                       "l2": "256 MB flush before every launch"},
            "sets": per, "gpu_launches": 3 * args.steps, "clocks": clk,
            "e2e": {"value": tot_n / e2e_s / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(sum(32 * len(r) + (8 * len(r) if m is not None else 0) for r, m, _ in host.values())),
                    "d2h_bytes_per_step": 16 * tot_n, "note": "novum_trace_closest / novum_trace_shadow with page-locked host arrays (PCIe-bound: 48-56 B per ray cross the bus)"},
            "roofline": {"bound": "hbm", "kernel": "closestBatchKernel (BVH8 closest hit, primary + secondary batches)", "achieved": bpr / (closest_ms * 1e-3) / 1e9,
                         "peak": peak, "unit": "GB/s", "frac": bpr / (closest_ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src}}
    if os.path.exists(ref_bin) and not args.no_cpu_baseline:
        parity, ref_ms, ref_n = {}, 0.0, 0
        for name, (r, m, s_) in host.items():
            ms, res = ref_trace(sdir, r, m, s_, tmp)
            ref_ms += ms; ref_n += len(r)
            per[name]["reference_kernel_ms"] = ms; per[name]["reference_mrays_per_s"] = len(r) / ms / 1e3
            per[name]["speedup"] = ms / per[name]["kernel_ms"]
            if m is None:
                ours = outs[name].cpu().numpy().view(HIT).reshape(-1)
                refh = np.zeros(len(r), HIT)
                refh["tri"] = res[:, 0].view(np.int32); refh["t"] = res[:, 1]; refh["u"] = res[:, 2]; refh["v"] = res[:, 3]
                c = classify(ours, refh); c.pop("mismatch_idx")
                parity[name] = c
            else:
                o = outs[name].cpu().numpy()
                parity[name] = {"n": len(r), "bit_equal": int((o.view(np.uint32) == res.view(np.uint32)).all(axis=1).sum())}
        line["parity_vs_reference_routines"] = parity
        line["cpu_baseline"] = {"value": ref_n / ref_ms / 1e3, "unit": "Mrays/s", "cores": os.cpu_count(), "kind": "reference",
                                "sample": "the reference's BVHSceneIntersect / BVHShadowRay kernels (one thread per ray, best of 3) on the same three batches on this B200 (the reference has no CPU path)"}
    print(json.dumps(line))
    scene.close()


GRID_DESC = "photon hash grid build (SURVEY 8d micro / row a15): hash + radix sort + cell tables on photon sets shaped like SPPM 1M and VCM 4K iterations"


def grid_photons(n, seed):
    """Photons on the six faces of a 10^3 box (uniform) with a fifth of them in two Gaussian caustic blobs on the floor."""
    import numpy as np
    rng = np.random.default_rng(seed)
    p = rng.random((n, 3), dtype=np.float32) * 10.0
    face = rng.integers(0, 6, n)
    for a in range(3):
        p[face == 2 * a, a] = 0.0
        p[face == 2 * a + 1, a] = 10.0
    k = n // 5
    blob = rng.normal(0.0, 0.15, (k, 2)).astype(np.float32)
    ctr = np.where(rng.random(k)[:, None] < 0.5, np.float32([3.0, 6.5]), np.float32([7.2, 2.8]))
    p[:k, 0] = np.clip(ctr[:, 0] + blob[:, 0], 0, 10); p[:k, 1] = 0.0; p[:k, 2] = np.clip(ctr[:, 1] + blob[:, 1], 0, 10)
    return np.ascontiguousarray(p[:, 0]), np.ascontiguousarray(p[:, 1]), np.ascontiguousarray(p[:, 2])


GRID_CASES = [("sppm_1m", 1 << 20, 10 << 20), ("vcm4k_typical", 10_500_000, 82_944_000)]
GRID_FULL = [("vcm4k_full", 82_944_000, 82_944_000)]


def ref_grid(px, py, pz, max_photons, radius, scene_min, tmp):
    import numpy as np
    ref = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    ip, op = os.path.join(tmp, "photons.bin"), os.path.join(tmp, "grid.bin")
    with open(ip, "wb") as f:
        np.array([len(px), max_photons], np.int32).tofile(f); np.array([radius] + list(scene_min), np.float32).tofile(f)
        px.tofile(f); py.tofile(f); pz.tofile(f)
    out = subprocess.check_output([ref, "grid", ip, op], stderr=subprocess.STDOUT).decode()
    js = json.loads(re.search(r"NOVUM_REF_JSON (\{.*\})", out).group(1))
    res = np.fromfile(op, np.uint32).reshape(4, -1)
    os.remove(ip); os.remove(op)
    return js, res


def run_grid(args):
    """--workload hashgrid: one step = one grid build per case. value = photons/s over the cases, device-resident (CUDA events in the
    library around hash + sort + tables); e2e = novum_hashgrid_build with host positions in and sorted order + tables out."""
    import numpy as np
    import torch
    import cudapathtracer_b200 as nb
    if int(os.environ.get("RANK", "0")) != 0:
        return
    cases = GRID_CASES + (GRID_FULL if args.grid_full else [])
    if args.grid_small:
        cases = [("small", 200_000, 1 << 20)]
    radius, scene_min = 0.0866, (0.0, 0.0, 0.0)       # 0.01 x the box's bounding radius, the reference's initial merge radius
    tmp = tempfile.mkdtemp()
    ref_bin = os.path.join(REPO, "oracle", "_ref", "novum_ref")
    if args.impl == "reference":
        per, tot_ms, tot_n = {}, 0.0, 0
        for name, n, mx in cases:
            px, py, pz = grid_photons(n, 7)
            js, _ = ref_grid(px, py, pz, mx, radius, scene_min, tmp)
            per[name] = {"photons": n, "max_photons": mx, "build_ms": js["build_ms"], "reorder_ms": js["reorder_ms"], "mphotons_per_s": n / js["build_ms"] / 1e3}
            tot_ms += js["build_ms"]; tot_n += n
        val = tot_n / tot_ms / 1e3
        print(json.dumps({"impl": "reference", "metric": "Mphotons/s", "value": val, "unit": "Mphotons/s", "n_gpus": 1, "steps": 3, "warmup": 0, "ms_per_step": tot_ms,
                          "higher_is_better": True, "config": {"workload": GRID_DESC}, "cases": per,
                          "cpu_baseline": {"value": val, "unit": "Mphotons/s", "cores": os.cpu_count(), "kind": "reference",
                                           "sample": "the reference's own buildHashGrid (computeHashes, cub::DeviceRadixSort, reorderPhotons, 2 table memsets, buildTable; best of 3) on one B200"},
                          "e2e": {"value": val, "unit": "Mphotons/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return
    if not torch.cuda.is_available() or nb.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device — the product path has no CPU fallback")
    dev = torch.device("cuda", 0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    clocks = ClockSampler(0)
    clocks.start()
    per, tot_ms, tot_n, algo = {}, 0.0, 0, 0.0
    e2e_s, e2e_h2d, e2e_d2h = 0.0, 0, 0
    for name, n, mx in cases:
        px, py, pz = grid_photons(n, 7)
        g = nb.HashGrid(mx)
        d = [torch.from_numpy(a).to(dev) for a in (px, py, pz)]
        for i in range(args.warmup):
            g.build_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, scene_min, radius, True)
        ms = 0.0
        for i in range(args.steps):
            flush.fill_(i); torch.cuda.synchronize()
            g.build_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), n, scene_min, radius, True)
            ms += g.last_ms() / args.steps
        bits = max(1, int(np.ceil(np.log2(g.table_size))))
        passes = (bits + 7) // 8
        bpp = 12 + 8 + passes * 20 + 12
        per[name] = {"photons": n, "max_photons": mx, "table_size": g.table_size, "build_ms": ms, "mphotons_per_s": n / ms / 1e3, "radix_passes": passes,
                     "bytes_per_photon": bpp, "gbs": n * bpp / (ms * 1e-3) / 1e9}
        tot_ms += ms; tot_n += n; algo += n * bpp
        t0 = time.time()
        idx, hsh, cs, ce = g.build(px, py, pz, scene_min, radius)
        e2e_s += time.time() - t0; e2e_h2d += 12 * n; e2e_d2h += 8 * n + 8 * g.table_size
        if os.path.exists(ref_bin) and not args.no_cpu_baseline:
            js, res = ref_grid(px, py, pz, mx, radius, scene_min, tmp)
            per[name].update({"reference_build_ms": js["build_ms"], "reference_reorder_ms": js["reorder_ms"], "speedup": js["build_ms"] / ms,
                              "speedup_without_reference_reorder": (js["build_ms"] - js["reorder_ms"]) / ms,
                              "parity": {"sorted_hash_equal": bool(np.array_equal(hsh, res[0])), "sorted_index_equal": bool(np.array_equal(idx, res[1])),
                                         "cell_start_equal": bool(np.array_equal(cs[hsh], res[2])), "cell_end_equal": bool(np.array_equal(ce[hsh], res[3]))}})
        g.close()
        del d
    clk = clocks.stop()
    peak, peak_src = peaks()
    line = {"metric": "Mphotons/s", "value": tot_n / tot_ms / 1e3, "unit": "Mphotons/s", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": GRID_DESC, "radius": radius, "l2": "256 MB flush before every build",
                       "scope": "hash + LSD radix sort of (hash, index) + cell tables (+ clearing the previous build's slots); the 32 B photon-record gather of the VCM engine (bdReorderPhotons) is not part of this entry point, the reference's total includes its 34 B reorder (reported separately)"},
            "cases": per, "gpu_launches": None, "clocks": clk,
            "e2e": {"value": tot_n / e2e_s / 1e6, "unit": "Mphotons/s", "h2d_bytes_per_step": e2e_h2d, "d2h_bytes_per_step": e2e_d2h,
                    "note": "novum_hashgrid_build: host positions in, sorted index + hash + both cell tables out (the tables dominate the D2H)"},
            "roofline": {"bound": "hbm", "kernel": "hashKernel + radix passes (digitHistogram, scatterPass) + tableKernel", "achieved": algo / (tot_ms * 1e-3) / 1e9, "peak": peak,
                         "unit": "GB/s", "frac": algo / (tot_ms * 1e-3) / 1e9 / peak, "traffic": None, "peak_source": peak_src}}
    if not args.no_cpu_baseline:
        from oracle import oracle_binding as ob
        px, py, pz = grid_photons(1 << 20, 7)
        t0 = time.time()
        ob.hashgrid_build(px, py, pz, scene_min, radius, ob.next_prime(2 * (10 << 20)))
        dt = time.time() - t0
        line["cpu_baseline"] = {"value": (1 << 20) / dt / 1e6, "unit": "Mphotons/s", "cores": 1, "kind": "port",
                                "sample": f"the CPU restatement (hash, stable sort, tables) on the 1M-photon case ({dt:.2f} s); the reference itself has no CPU path — its GPU build is in cases[*].reference_build_ms"}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None, help="timed steps (default 4; frame_finish: 50 — its step is a 55 us kernel)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="rainbow_vcm", choices=sorted(WORKLOADS) + ["frame_finish", "trace_batch", "hashgrid", "obj_parse"])
    ap.add_argument("--grid-small", action="store_true", help="hashgrid: one 200k-photon case (used by the GPU test suite)")
    ap.add_argument("--grid-full", action="store_true", help="hashgrid: add the 82.9 M-photon worst case of a 4K VCM iteration")
    ap.add_argument("--rays-side", type=int, default=2048, help="trace_batch: the primary batch is rays_side^2 rays (2048 -> 4.19 M)")
    ap.add_argument("--spp-per-step", type=int, default=0, help="passes per step per GPU (0 = the workload's default)")
    ap.add_argument("--sharding", default=None, choices=["striped", "tiles"],
                    help="default: tiles for the bidirectional workloads at N > 1 (strong scaling: rows sharded, photons exchanged every iteration), striped otherwise")
    ap.add_argument("--lanes", type=int, default=4, help="tiles sharding: iterations in flight per rank")
    ap.add_argument("--no-relmse", action="store_true", help="skip the 1080p frame against the reference's (rainbow_vcm only)")
    ap.add_argument("--no-secondary", action="store_true", help="skip the dragon-NEE secondary measurement (rainbow_vcm, N = 1)")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"], help="tiles sharding: photon exchange by peer stores fused into the light pass, or by an NCCL all-gather")
    ap.add_argument("--detail", type=int, default=100)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.steps is None:
        args.steps = 50 if (args.workload in ("frame_finish", "trace_batch") and args.impl == "ours") else (10 if args.workload == "obj_parse" else 4)
    if args.workload == "trace_batch":
        run_trace(args)
    elif args.workload == "hashgrid":
        run_grid(args)
    elif args.workload == "frame_finish":
        (run_frame_reference if args.impl == "reference" else run_frame)(args)
    elif args.workload == "obj_parse":
        (run_obj_reference if args.impl == "reference" else run_obj)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
