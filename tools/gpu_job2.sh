set -x
timeout 900 python -m pytest tests/test_multigpu_gpu.py -m gpu -q -x > gpurun_out/pytest_mg.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mg.log
NGPU=2 bash tools/gpu_job8.sh
