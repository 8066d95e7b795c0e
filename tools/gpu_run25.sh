set -x
mkdir -p gpurun_out
B="python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline"
timeout 900 $B > gpurun_out/bench_dragon_s5.log 2>&1
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
for mb in 6 8; do
cp cudapathtracer_b200/build_alt/libnovum_shade$mb.so cudapathtracer_b200/libnovum_b200.so
timeout 900 $B > gpurun_out/bench_dragon_s$mb.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_s$mb.log 2>&1
done
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_s5.log 2>&1
