set -x
mkdir -p gpurun_out
run() { tag=$1
timeout 900 python bench.py --workload trace_batch --steps 10 --warmup 2 --no-cpu-baseline > gpurun_out/bench_trace_$tag.log 2>&1
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_nee_$tag.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$tag.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_$tag.log 2>&1
}
run base54
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
for T in 7 12; do
cp cudapathtracer_b200/build_alt/libnovum_tp$T.so cudapathtracer_b200/libnovum_b200.so
run tp$T
timeout 600 python -m pytest tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/pytest_tp$T.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_tp$T.log
done
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
timeout 900 python bench.py --workload trace_batch > gpurun_out/bench_trace_54_full.log 2>&1; echo "rc=$?" >> gpurun_out/bench_trace_54_full.log
