set -x
mkdir -p gpurun_out
python - <<'PY' > gpurun_out/l2attr.txt 2>&1
import torch
p = torch.cuda.get_device_properties(0)
print(p)
import ctypes
rt = ctypes.CDLL("libcudart.so.12")
for name, a in (("MaxPersistingL2CacheSize", 108), ("MaxAccessPolicyWindowSize", 109), ("L2CacheSize", 38)):
    v = ctypes.c_int(); rt.cudaDeviceGetAttribute(ctypes.byref(v), a, 0); print(name, v.value)
PY
for t in on off; do
if [ $t = off ]; then export NOVUM_NO_L2_PERSIST=1; fi
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_nee_$t.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$t.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_$t.log 2>&1
done
unset NOVUM_NO_L2_PERSIST
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
