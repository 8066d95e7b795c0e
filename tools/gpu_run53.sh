set -x
mkdir -p gpurun_out
timeout 900 python bench.py --workload trace_batch --steps 3 --warmup 2 > gpurun_out/bench_trace_53.log 2>&1; echo "rc=$?" >> gpurun_out/bench_trace_53.log
timeout 900 python bench.py --workload trace_batch --impl reference > gpurun_out/bench_trace_ref_53.log 2>&1; echo "rc=$?" >> gpurun_out/bench_trace_ref_53.log
