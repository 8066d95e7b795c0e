set -x
mkdir -p gpurun_out/golden
python tests/golden/make_golden_frame.py --out gpurun_out/golden > gpurun_out/make_golden_frame.log 2>&1
cp gpurun_out/golden/frame_clean.npz tests/golden/ 2>/dev/null
timeout 600 python -m pytest tests/test_gpu_frame.py tests/test_frame_oracle.py -q -x > gpurun_out/pytest_frame_45.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_frame_45.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_45.log 2>&1; echo "rc=$?" >> gpurun_out/smoke_45.log
timeout 600 python bench.py --workload frame_finish --steps 20 --warmup 5 > gpurun_out/bench_frame_45.log 2>&1; echo "rc=$?" >> gpurun_out/bench_frame_45.log
timeout 600 python bench.py --workload frame_finish --impl reference --steps 3 > gpurun_out/bench_frame_ref_45.log 2>&1
oracle/_ref/novum_ref clean /dev/null /dev/null 0 0 0 /dev/null > /dev/null 2>&1
python - > gpurun_out/ref_clean_4k.log 2>&1 <<'PY'
import sys, subprocess, tempfile, os
sys.path.insert(0, '.')
import bench
acc, ov = bench.frame_inputs(3840, 2160)
d = tempfile.mkdtemp(); acc.tofile(d + '/a.raw'); ov.tofile(d + '/o.raw')
print(subprocess.run(['oracle/_ref/novum_ref', 'clean', d + '/a.raw', d + '/o.raw', '3840', '2160', '63', d + '/c.raw'], capture_output=True, text=True).stdout)
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:frameFinish -s 5 -c 2 -o gpurun_out/frame_finish_45 python bench.py --workload frame_finish --steps 3 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_frame_45.log 2>&1
ls -la gpurun_out
