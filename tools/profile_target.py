"""The program ncu profiles (tools/profile.sh): one step of a bench.py workload between cudaProfilerStart / Stop, after warm-up, with ONE
lane so that launches come in iteration order. Prints a JSON line with the step's own counters (rays, BVH nodes / triangles per ray from a
second, counted and un-profiled run of the same step; merge pairs), which tools/ncu_extract.py joins with the ncu capture.
Usage: python tools/profile_target.py <workload> [first_pass] [passes]"""
import json
import os
import sys
import tempfile

os.environ.setdefault("NOVUM_LANES", "1")
import torch  # noqa: E402

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
import bench  # noqa: E402
import cudapathtracer_b200 as nb  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "rainbow_vcm"
wl = bench.WORKLOADS[workload]
first = int(sys.argv[2]) if len(sys.argv) > 2 else (5 if wl["integ"] != bench.UNI else 32)
passes = int(sys.argv[3]) if len(sys.argv) > 3 else (1 if wl["integ"] != bench.UNI else 32)
sdir = bench.scene_dir_for(workload, 100, 0, None)
scene = nb.Scene.from_config(os.path.join(sdir, "configs", "config.rendertron"))
cfg = scene.cfg
cfg.width, cfg.height, cfg.integrator = wl["w"], wl["h"], wl["integ"]
cam = nb.camera_make(cfg)
colors = torch.zeros((wl["h"], wl["w"], 4), dtype=torch.float32, device="cuda")
scene.render_device(cfg, cam, 0, first, colors.data_ptr())
torch.cuda.synchronize()
torch.cuda.profiler.start()
st = scene.render_device(cfg, cam, first, passes, colors.data_ptr())
torch.cuda.synchronize()
torch.cuda.profiler.stop()
scene.set_profile(3)
ct = scene.render_device(cfg, cam, first, passes, colors.data_ptr())
out = {"workload": workload, "first_pass": first, "passes": passes, "device_ms": st.deviceMs, "launches": st.kernelLaunches,
       "closest_rays": st.closestRays, "shadow_rays": st.shadowRays, "nodes_per_ray": ct.bvhNodeVisits / max(1, ct.closestRays),
       "tris_per_ray": ct.triTests / max(1, ct.closestRays), "merge_queries": st.mergeQueries, "merge_candidates": st.mergeCandidates,
       "merge_accepted": st.mergeAccepted, "photons": st.photons, "stage_ms_counted_run": [ct.stageMs[k] for k in range(4)],
       "stage_launches": [ct.stageLaunches[k] for k in range(4)]}
print("PROFILE_TARGET_JSON " + json.dumps(out))
scene.close()
