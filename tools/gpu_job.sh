timeout 600 python bench.py --steps 40 --warmup 5 --skip-gn --skip-cpu --skip-refcuda 2> gpurun_out/b.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['frame_pipeline'])"; tail -3 gpurun_out/b.err
