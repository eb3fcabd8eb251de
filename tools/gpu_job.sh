set -x
timeout 900 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py -m gpu -q -x > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_match.log
timeout 600 python bench.py --steps 40 --warmup 5 --skip-gn > gpurun_out/bench_refine.json 2> gpurun_out/bench_refine.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_refine.json')); print(d['kernels']); print(d['value'], d['roofline']['fp16_pipe_frac'])"
timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:refine_f16_tiled -c 2 -o gpurun_out/refine_full3 -f python tools/run_match_once.py > gpurun_out/ncu_refine.log 2>&1; echo "ncu rc=$?"
