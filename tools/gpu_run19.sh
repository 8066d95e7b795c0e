set -x
mkdir -p gpurun_out
run() { tag=$1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_$tag.log 2>&1
for wl in water_bdpt water_sppm; do
timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_${wl}_$tag.log 2>&1
done
}
run call
cp cudapathtracer_b200/libnovum_b200.so /tmp/main.so
cp cudapathtracer_b200/build_alt/libnovum_alt.so cudapathtracer_b200/libnovum_b200.so
run inl
cp /tmp/main.so cudapathtracer_b200/libnovum_b200.so
run call2
