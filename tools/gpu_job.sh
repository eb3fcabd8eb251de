set -x
timeout 900 python -m pytest tests/test_tracking_gpu.py -m gpu -q -x > gpurun_out/pytest_track.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/pytest_track.log
timeout 600 python bench.py --steps 40 --warmup 5 --skip-gn > gpurun_out/bench_track.json 2> gpurun_out/bench_track.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/bench_track.json')); print(json.dumps(d['tracking'],indent=1))"
timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -k regex:track_pose -o gpurun_out/track_full2 -f python tools/run_track_once.py > gpurun_out/ncu_track.log 2>&1; echo "ncu rc=$?"
