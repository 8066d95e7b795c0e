set -x
mkdir -p gpurun_out
run() { tag=$1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$tag.log 2>&1
timeout 900 python bench.py --workload water_bdpt --no-cpu-baseline > gpurun_out/bench_bdpt_$tag.log 2>&1
timeout 900 python bench.py --workload water_sppm --no-cpu-baseline > gpurun_out/bench_sppm_$tag.log 2>&1
}
run base71
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
for T in 5 6 8; do
cp cudapathtracer_b200/build_alt/libnovum_minb$T.so cudapathtracer_b200/libnovum_b200.so
run minb$T
done
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
