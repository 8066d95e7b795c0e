set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
run() { tag=$1
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_nee_$tag.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$tag.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_$tag.log 2>&1
timeout 900 python bench.py --workload water_bdpt --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_bdpt_$tag.log 2>&1
}
run coop
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
cp cudapathtracer_b200/build_alt/libnovum_serial.so cudapathtracer_b200/libnovum_b200.so
run serial
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
