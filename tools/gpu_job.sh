timeout 600 python -m pytest tests/test_matching_gpu.py -m gpu -q -x -k refine > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_match.log
timeout 300 python tools/kbench.py 2>&1 | grep refine | head -2
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:refine_plan -c 4 python tools/refine_var.py 2>&1 | grep -E "gpu__time_duration" | head -12
