set -x
timeout 900 python -m pytest tests/test_front_gpu.py tests/test_matching_gpu.py -m gpu -q -x -s > gpurun_out/pytest_front.log 2>&1; echo "pytest rc=$?"
grep -E "agreement|passed|failed" gpurun_out/pytest_front.log | tail -12
timeout 600 python bench.py --steps 40 --warmup 5 --skip-gn > gpurun_out/bench_pad.json 2> gpurun_out/bench_pad.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_pad.err
python -c "
import json; d=json.load(open('gpurun_out/bench_pad.json')); print(d['front_end']); print(d['frame_pipeline']); print(d['e2e'])"
