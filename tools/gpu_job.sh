set -x
timeout 900 python -m pytest tests/test_tracking_gpu.py tests/test_front_gpu.py -m gpu -q -s > gpurun_out/pytest_track.log 2>&1; echo "pytest rc=$?"
grep -E "rel err|iterations|passed|failed|Error|error" gpurun_out/pytest_track.log | head -60
