set -x
timeout 900 python -m pytest tests/test_front_gpu.py -m gpu -q -x > gpurun_out/pytest_front.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_front.log
