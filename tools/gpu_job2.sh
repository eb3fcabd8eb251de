set -x
timeout 900 python -m pytest tests/test_multigpu_gpu.py -m gpu -q -x > gpurun_out/pytest_mg.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_mg.log
N=2
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench$N rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err; echo "benchref$N rc=$?"
cut -c1-400 gpurun_out/bench_n$N.json; cut -c1-200 gpurun_out/bench_ref_n$N.json
