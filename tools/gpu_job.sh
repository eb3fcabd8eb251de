set -x
timeout 600 python bench.py --steps 40 --warmup 5 --skip-gn > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_t.err
python -c "
import json; d=json.load(open('gpurun_out/bench_t.json')); print(json.dumps(d['tracking'],indent=1))"
