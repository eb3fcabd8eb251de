set -x
timeout 1200 python -m pytest tests/test_gn_gpu.py -m gpu -q -x > gpurun_out/pytest_gn.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_gn.log
timeout 300 python tools/kbench.py coop > gpurun_out/coop.log 2>&1; echo "rc=$?"; tail -14 gpurun_out/coop.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_gn.json 2> gpurun_out/bench_gn.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_gn.json')); g=d['gn']; print({k:g[k] for k in ('value','ms_per_iter','linearize_ms','solve_ms','begin_ms','single_call_ms_per_iter')}, g['roofline']['whole_iter_frac'])"
