timeout 600 python -m pytest tests/test_matching_gpu.py -m gpu -q -x -k "refine" 2>&1 | tail -2
timeout 600 python bench.py --steps 60 --warmup 5 --skip-gn --skip-cpu --skip-refcuda 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['kernels']['refine_ms'], d['value'])"
