#!/usr/bin/env python
"""Turns what tools/profile.sh brought back in gpurun_out/ into the committed evidence (run here, no GPU):
     profiles/r2_<workload>_launches.txt   per-kernel totals of the launch list of the profiled step
     profiles/r2_<workload>_full.txt       per-launch counters of the ncu --set full capture (shipped kernels)
     profiles/r2_kernel_counters.json      what bench.py reads: per kernel class DRAM bytes / warp instructions per launch, pairs per launch
   python tools/ncu_extract.py <workload>"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W = sys.argv[1] if len(sys.argv) > 1 else "rainbow_vcm"
G, P = os.path.join(REPO, "gpurun_out"), os.path.join(REPO, "profiles")
KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum', 'l1tex__t_bytes.sum',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'smsp__inst_executed.sum', 'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']


def short(n):
    n = re.sub(r'\(.*', '', n); n = re.sub(r'^void ', '', n); n = n.replace('nv::', '').replace('<unnamed>::', '')
    return re.sub(r'<.*', '', n)


def num(x):
    try:
        return float(x.replace(',', ''))
    except ValueError:
        return None


target = json.loads(re.search(r"PROFILE_TARGET_JSON (\{.*\})", open(os.path.join(G, f"r2_{W}_target.log")).read()).group(1))
# ---- launch list ----
rows = [r for r in csv.reader(open(os.path.join(G, f"r2_{W}_launches.csv"))) if len(r) > 5]
hdr = rows[0]
ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
tot, cnt = collections.defaultdict(float), collections.Counter()
for r in rows[1:]:
    tot[short(r[ki])] += num(r[vi]); cnt[short(r[ki])] += 1
all_ = sum(tot.values())
with open(os.path.join(P, f"r2_{W}_launches.txt"), "w") as f:
    f.write(f"# ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none: python tools/profile_target.py {W}\n")
    f.write(f"# the profiled step: passes {target['first_pass']}..{target['first_pass'] + target['passes'] - 1}, {target['launches']} launches, "
            f"{target['device_ms']:.3f} ms un-profiled (CUDA events); per-launch times below are cold-cache and serialised: compare SHARES\n")
    f.write(f"{'kernel':44s} {'launches':>8s} {'total ms':>10s} {'share':>7s}\n")
    for n in sorted(tot, key=tot.get, reverse=True):
        f.write(f"{n[:44]:44s} {cnt[n]:8d} {tot[n] / 1e6:10.3f} {tot[n] / all_:7.3f}\n")
print(open(os.path.join(P, f"r2_{W}_launches.txt")).read())
# ---- full capture ----
rows = list(csv.reader(open(os.path.join(G, f"r2_{W}_full_raw.csv"))))       # `ncu -i <report> --page raw --csv`, exported on the GPU box (tools/profile.sh)
hdr, units, data = rows[0], rows[1], rows[2:]
kn = hdr.index('Kernel Name')
col = {k: hdr.index(k) for k in KEYS if k in hdr}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}      # -> bytes, ns


def val(r, k):
    v = num(r[col[k]])
    return None if v is None else v * SCALE.get(units[col[k]], 1.0)


per = collections.defaultdict(list)
for r in data:
    per[short(r[kn])].append({k: val(r, k) for k in col})
with open(os.path.join(P, f"r2_{W}_full.txt"), "w") as f:
    f.write(f"# ncu --profile-from-start off --set full --clock-control none --import-source on: python tools/profile_target.py {W} (SHIPPED kernels)\n")
    f.write(f"# step counters of the same program: {json.dumps(target)}\n")
    for name, launches in per.items():
        f.write(f"\n== {name}: {len(launches)} launches ==\n")
        for k in col:
            vals = [l[k] for l in launches]
            u = {"Kbyte": "byte", "Mbyte": "byte", "Gbyte": "byte", "us": "ns", "ms": "ns", "s": "ns"}.get(units[col[k]], units[col[k]])
            f.write(f"{k} [{u}]: " + " ".join("-" if v is None else (f"{v:.4g}") for v in vals) + "\n")
print(open(os.path.join(P, f"r2_{W}_full.txt")).read()[:3000])
# ---- counters for bench.py ----
jp = os.path.join(P, "r2_kernel_counters.json")
kc = json.load(open(jp)) if os.path.exists(jp) else {}
entry = {}
for name, launches in per.items():
    n = len(launches)
    dram = sum((l.get('dram__bytes_read.sum') or 0) + (l.get('dram__bytes_write.sum') or 0) for l in launches)
    inst = sum(l.get('smsp__inst_executed.sum') or 0 for l in launches)
    dur = sum(l.get('gpu__time_duration.sum') or 0 for l in launches)
    e = {"launches": n, "dram_bytes_per_launch": dram / n, "warp_inst_per_launch": inst / n, "ns_per_launch_under_ncu": dur / n,
         "registers_per_thread": launches[0].get('launch__registers_per_thread'),
         "issue_active_pct": sum(l.get('smsp__issue_active.avg.pct_of_peak_sustained_active') or 0 for l in launches) / n,
         "lanes_per_inst": sum(l.get('smsp__thread_inst_executed_per_inst_executed.ratio') or 0 for l in launches) / n,
         "source": f"profiles/r2_{W}_full.txt (ncu --set full of the shipped kernel, passes {target['first_pass']}..{target['first_pass'] + target['passes'] - 1})"}
    if name.startswith("bdMergeLambert"):
        e["pairs_per_launch"] = target["merge_candidates"] / n
    if "Extend" in name:
        e["rays_per_launch"] = target["closest_rays"] / n
        e["algorithmic_bytes_per_launch"] = target["closest_rays"] / n * (48 + 80 * target["nodes_per_ray"] + 48 * target["tris_per_ray"])
    entry[name] = e
kc[W] = entry
json.dump(kc, open(jp, "w"), indent=1)
print(json.dumps(entry, indent=1))
