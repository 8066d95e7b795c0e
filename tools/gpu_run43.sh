set -x
mkdir -p gpurun_out
run() { tag=$1
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_nee_$tag.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$tag.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_$tag.log 2>&1
}
run base
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
cp cudapathtracer_b200/build_alt/libnovum_pf.so cudapathtracer_b200/libnovum_b200.so
run pf
timeout 600 python -m pytest tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/pytest_pf.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_pf.log
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
