set -x
timeout 900 python -m pytest tests/test_pointmap_gpu.py -m gpu -q > gpurun_out/pytest_pm.log 2>&1; echo "pytest rc=$?"
tail -25 gpurun_out/pytest_pm.log
