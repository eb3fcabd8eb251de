set -x
( time python bench.py > gpurun_out/bench_r1.json 2> gpurun_out/bench_r1.err ) 2> gpurun_out/bench_r1.time; echo "bench rc=$?"
( time python bench.py --impl reference > gpurun_out/bench_r1_ref.json 2> gpurun_out/bench_r1_ref.err ) 2> gpurun_out/bench_r1_ref.time; echo "ref rc=$?"
B="python bench.py --steps 3 --warmup 3 --skip-cpu --skip-refcuda --profile"
$B > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches_r1.csv $B > gpurun_out/ncu_l.log 2>&1
echo "launchlist rc=$?"
$B > gpurun_out/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on --profile-from-start off -k 'regex:refine_f16|refine_plan|iter_proj_kernel|gn_accumulate|gn_chol_solve|gn_assemble' -c 14 -o gpurun_out/prof_r1o -f $B > gpurun_out/ncu_f.log 2>&1
echo "full rc=$?"
cat gpurun_out/bench_r1.time gpurun_out/bench_r1_ref.time
