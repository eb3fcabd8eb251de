timeout 900 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py tests/test_streams_gpu.py -m gpu -q -x > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_match.log
timeout 300 python tools/kbench.py 2>&1 | grep refine | head -2
timeout 300 python tools/kbench.py 2>&1 | grep refine | head -2
