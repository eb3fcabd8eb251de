set -x
timeout 900 python -m pytest tests/test_front_gpu.py -m gpu -q -x > gpurun_out/pytest_front.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_front.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/frame_launches.csv python tools/run_frame_once.py > gpurun_out/ncu_frame2.log 2>&1; echo "ncu rc=$?"
grep -E "match_confidence|prep_rays|match_proj|update_pointmap|track_pose|refine" gpurun_out/frame_launches.csv | awk -F'","' '{print $5, $NF}' | head -14
