set -x
timeout 900 python -m pytest tests/test_front_gpu.py tests/test_matching_gpu.py -m gpu -x -q -s > gpurun_out/pytest_front.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/pytest_front.log
timeout 600 python bench.py --steps 40 --warmup 5 > gpurun_out/bench_front.json 2> gpurun_out/bench_front.err; echo "bench rc=$?"
cat gpurun_out/bench_front.json
