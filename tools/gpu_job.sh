set -x
python -m pytest tests/test_gn_gpu.py -m gpu -x -q > gpurun_out/pytest_gn.log 2>&1; echo "pytest rc=$?"
python tools/run_gn_once.py > gpurun_out/gn_once.log 2>&1; echo "gn_once rc=$?"
G="python tools/run_gn_once.py"
$G > gpurun_out/gn_plain.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k 'regex:gn_accumulate|gn_compact|gn_pack' -c 4 -o gpurun_out/prof_gn_r1c $G > gpurun_out/ncu_gn.log 2>&1
echo "ncu rc=$?"
