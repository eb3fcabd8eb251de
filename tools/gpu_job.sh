timeout 900 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py tests/test_streams_gpu.py -m gpu -q -x > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_match.log
python bench.py --skip-cpu --skip-refcuda > gpurun_out/bench_pool.json 2>gpurun_out/bench_pool.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_pool.json')); print(d['value'], d['kernels']['refine_ms'], d['front_end']['fused_ms'], d['front_end']['torch_glue_ms'], d['e2e']['value'], d['frame_pipeline']['ms_per_frame'])"
