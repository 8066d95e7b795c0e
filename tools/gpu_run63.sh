set -x
mkdir -p gpurun_out
timeout 600 python tools/ab_render.py slab2 > gpurun_out/ab_slab2.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_slab2.log 2>&1
timeout 900 python bench.py --workload water_sppm --no-cpu-baseline > gpurun_out/bench_sppm_slab2.log 2>&1
timeout 900 python -m pytest tests/test_gpu_bidir.py -m gpu -q -x > gpurun_out/pytest_slab2.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_slab2.log
python - <<'PY' > gpurun_out/ab_compare2.log 2>&1
import numpy as np
a = np.load('gpurun_out/ab_slab2.npz'); b = np.load('gpurun_out/ab_noslab.npz')
for k in a.files:
    if k.endswith('_img'):
        x, y = a[k][..., :3].astype(np.float64), b[k][..., :3].astype(np.float64)
        den = np.maximum(np.abs(y), 1e-3)
        print(k, 'max rel diff', float((np.abs(x - y) / den).max()), 'mean rel diff', float(np.abs(x - y).sum() / np.abs(y).sum()))
    else:
        print(k, a[k].tolist(), b[k].tolist())
PY
