set -x
timeout 600 python -m pytest tests/test_gn_gpu.py -m gpu -x -q > gpurun_out/pytest_gn.log 2>&1; echo "pytest rc=$?"
M3R_GN_SOLVER=coop timeout 600 python -m pytest tests/test_gn_gpu.py -m gpu -x -q > gpurun_out/pytest_gn_coop.log 2>&1; echo "pytest coop rc=$?"
timeout 300 python tools/run_gn_once.py > gpurun_out/gn_once.log 2>&1; echo "gn_once rc=$?"
