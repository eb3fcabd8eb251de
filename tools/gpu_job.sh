timeout 900 python -m pytest tests/test_matching_gpu.py tests/test_front_gpu.py -m gpu -q -x > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_match.log
for g in 0 148 296 472 768; do
  echo "NSPLIT=$g"; M3R_REFINE_NSPLIT=$g timeout 300 python tools/refine_var.py 2>&1 | tail -1
done
