set -x
timeout 1200 python -m pytest tests/test_tracking_gpu.py -m gpu -q -x > gpurun_out/pytest_track.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_track.log
