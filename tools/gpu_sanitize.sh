# compute-sanitizer over a reduced set of GPU tests that exercise every hand-rolled
# synchronisation in the library: mbarrier + cp.async.bulk (refine), per-warp cp.async rings (GN
# accumulation), the grid barrier carried across launches (tracking), the cooperative Cholesky,
# the cross-rank release/acquire flags (loop-back exchange), programmatic dependent launch
# (refine plan -> main).  Summaries go to gpurun_out/sanitize_<tool>.txt.
set -x
mkdir -p gpurun_out
SEL='test_golden_matching_gpu or test_refine_edge_cases or test_golden_gn_gpu or test_failed_factorisation or test_solver_by_system_size or test_loopback_exchange or test_early_stop_then_second or test_golden_tracking_gpu or test_golden_front_end_gpu'
for tool in memcheck racecheck synccheck; do
  ( time timeout ${SAN_TIMEOUT:-900} compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 \
      --log-file gpurun_out/sanitize_${tool}.raw \
      python -m pytest tests -m gpu -q -x -k "$SEL" ${SAN_EXTRA:-} > gpurun_out/sanitize_${tool}.pytest 2>&1 ) 2> gpurun_out/sanitize_${tool}.time
  echo "$tool rc=$?"
  tail -3 gpurun_out/sanitize_${tool}.pytest
  grep -c "=========" gpurun_out/sanitize_${tool}.raw
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Error:" gpurun_out/sanitize_${tool}.raw | sort | uniq -c | head -20
done
