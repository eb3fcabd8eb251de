set -x
timeout 1200 python -m pytest tests/test_gn_gpu.py -m gpu -q -x > gpurun_out/pytest_gn.log 2>&1; echo "pytest rc=$?"
tail -2 gpurun_out/pytest_gn.log
timeout 300 python tools/kbench.py coop > gpurun_out/coop.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/coop.log
timeout 600 python tools/gn_scale.py rays:64 points:256 2>&1 | tail -2 | cut -c1-330
