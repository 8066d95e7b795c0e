#!/bin/bash
# usage (under gpurun --gpus N): bash tools/scale_run.sh "<gpus> <lanes>" ...   -> one bench.py line per configuration in gpurun_out/
for cfg in "$@"; do
  set -- $cfg
  if [ "$1" = "1" ]; then
    NOVUM_LANES=$2 python bench.py --steps 3 --warmup 2 --no-relmse --no-secondary --no-cpu-baseline > gpurun_out/bench_n$1_l$2.json 2> gpurun_out/bench_n$1_l$2.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus $1 --lanes $2 --steps 3 --warmup 2 --no-relmse > gpurun_out/bench_n$1_l$2.json 2> gpurun_out/bench_n$1_l$2.err
  fi
  echo "N=$1 lanes=$2 rc=$?"
  python - "$1" "$2" <<'PY'
import json, sys
try:
    l = [x for x in open(f"gpurun_out/bench_n{sys.argv[1]}_l{sys.argv[2]}.json") if x.startswith("{")]
    d = json.loads(l[-1])
    print({k: d[k] for k in ("value", "ms_per_pass", "n_gpus", "scaling")}, "e2e", round(d["e2e"]["value"], 1), "image_mean", d.get("image_mean"))
except Exception as ex:
    print("no line:", ex)
    print(open(f"gpurun_out/bench_n{sys.argv[1]}_l{sys.argv[2]}.err").read()[-600:])
PY
done
