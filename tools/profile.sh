#!/bin/bash
# Regenerates the raw material of profiles/r2_*: run on the GPU box (gpurun -- 'bash tools/profile.sh rainbow_vcm'), then
# `python tools/ncu_extract.py <workload>` here turns gpurun_out/ into the committed summaries + profiles/r2_kernel_counters.json.
#   1. the plain run (must exit 0 before anything is profiled)          -> gpurun_out/r2_<workload>_target.log
#   2. launch list: every launch of the profiled step with its device time (cold cache, serialised: compare SHARES) -> r2_<workload>_launches.csv
#   3. ncu --set full of the traversal / resolve / merge / shade kernels of that step                                -> r2_<workload>_full.ncu-rep
set -e
W="${1:-rainbow_vcm}"
K="${2:-bdExtend|bdResolve|bdMergeLambert|wfExtend|wfShadow}"
mkdir -p gpurun_out
python tools/profile_target.py "$W" > "gpurun_out/r2_${W}_target.log" 2>&1
grep PROFILE_TARGET_JSON "gpurun_out/r2_${W}_target.log"
ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file "gpurun_out/r2_${W}_launches.csv" \
    python tools/profile_target.py "$W" > "gpurun_out/r2_${W}_launches.log" 2>&1
ncu --profile-from-start off --set full --clock-control none --import-source on -k "regex:$K" -f -o "/tmp/r2_${W}_full" \
    python tools/profile_target.py "$W" > "gpurun_out/r2_${W}_full.log" 2>&1
# the report of ~60 launches is ~100 MB and gpurun_out/ carries 64 MiB: the counters travel as CSV (what `ncu -i ... --page raw --csv` prints)
ncu -i "/tmp/r2_${W}_full.ncu-rep" --page raw --csv > "gpurun_out/r2_${W}_full_raw.csv"
ls -la gpurun_out/r2_${W}_* /tmp/r2_${W}_full.ncu-rep
