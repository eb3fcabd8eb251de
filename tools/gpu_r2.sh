# round 2, one-GPU job: GPU test suite, smoke, default bench line, reference arm
# (run with `gpurun --timeout 2400 -- 'bash tools/gpu_r2.sh'`)
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.max.mem --format=csv,noheader
( time timeout 1800 python -m pytest tests -m gpu -q --durations=15 > gpurun_out/pytest_gpu.log 2>&1 ) 2> gpurun_out/pytest_gpu.time; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
( time timeout 900 python bench.py > gpurun_out/bench_now.json 2> gpurun_out/bench_now.err ) 2> gpurun_out/bench_now.time; echo "bench rc=$?"; tail -3 gpurun_out/bench_now.err
( time timeout 600 python bench.py --impl reference > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err ) 2> gpurun_out/bench_ref.time; echo "bench ref rc=$?"
cat gpurun_out/pytest_gpu.time gpurun_out/bench_now.time gpurun_out/bench_ref.time
