# round 2 evidence (one GPU): the default bench line + reference arm, then -- each only after its own command has
# exited 0 without ncu -- the ncu launch list of the bench command and --set full captures of the top kernels
set -x
mkdir -p gpurun_out
( time python bench.py > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err ); echo "bench rc=$?"
( time python bench.py --impl reference > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err ); echo "ref rc=$?"
python bench.py --steps 3 --warmup 3 --profile --skip-cpu --skip-refcuda > /dev/null 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 600 --csv \
    --log-file gpurun_out/r2_launches.csv python bench.py --steps 3 --warmup 3 --profile --skip-cpu --skip-refcuda > gpurun_out/r2_ncu_launches.log 2>&1
echo "launch list rc=$?"
python tools/run_match_once.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'refine|iter_proj' -c 6 \
    -o gpurun_out/r2_match -f python tools/run_match_once.py > gpurun_out/r2_ncu_match.log 2>&1
echo "match full rc=$?"
python tools/run_gn_once.py > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gn_' -c 14 \
    -o gpurun_out/r2_gn -f python tools/run_gn_once.py > gpurun_out/r2_ncu_gn.log 2>&1
echo "gn full rc=$?"
ls -la gpurun_out/*.ncu-rep
