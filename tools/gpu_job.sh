# one-GPU check on the box: GPU test suite, smoke, one bench line (run with `gpurun -- 'bash tools/gpu_job.sh'`)
set -x
( time timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1 ) 2> gpurun_out/pytest_gpu.time; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
( time python bench.py > gpurun_out/bench_now.json 2> gpurun_out/bench_now.err ) 2> gpurun_out/bench_now.time; echo "bench rc=$?"
cat gpurun_out/pytest_gpu.time gpurun_out/bench_now.time
