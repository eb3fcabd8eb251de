timeout 900 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py tests/test_oracle.py -m gpu -q -x > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_match.log
timeout 300 python tools/kbench.py 2>&1 | grep -E "iter_proj" | head -2
python bench.py --skip-cpu --skip-refcuda > gpurun_out/bench_ip.json 2>gpurun_out/bench_ip.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_ip.json')); print(d['value'], d['kernels'], d['front_end']['fused_ms'], d['frame_pipeline']['ms_per_frame'])"
