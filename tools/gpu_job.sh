set -x
timeout 900 python -m pytest tests/test_gn_gpu.py -m gpu -x -q > gpurun_out/pytest_gn.log 2>&1; echo "pytest rc=$?"
timeout 300 python tools/kbench.py coop > gpurun_out/coop.log 2>&1; echo "rc=$?"
