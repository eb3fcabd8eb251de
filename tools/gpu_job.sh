set -x
timeout 1500 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py -m gpu -q -x -s -k "random_init or oracle_twin" > gpurun_out/pytest_rand.log 2>&1; echo "pytest rc=$?"
grep -E "bit-identical|agreement|passed|failed|Error|^E " gpurun_out/pytest_rand.log | head -60
