set -x
python -m pytest tests/test_matching_gpu.py -m gpu -x -q > gpurun_out/pytest_match.log 2>&1; echo "pytest rc=$?"
python tools/kbench.py > gpurun_out/kbench.log 2>&1; echo "kbench rc=$?"
G="python tools/run_match_once.py"
$G > gpurun_out/m_plain.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k 'regex:refine_f16_tiled|iter_proj_kernel' -c 4 -o gpurun_out/prof_match_r1b $G > gpurun_out/ncu_m.log 2>&1
echo "ncu rc=$?"
