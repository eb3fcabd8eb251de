# round 2, N-GPU job (gpurun --gpus N -- 'bash tools/gpu_r2_n.sh N'): multi-GPU tests, then the bench as the driver launches it
N=${1:-2}
set -x
mkdir -p gpurun_out
nvidia-smi topo -m | head -12
( time timeout 600 python -m pytest tests/test_multigpu_gpu.py tests/test_multi_single_process_gpu.py -q > gpurun_out/pytest_multigpu_n$N.log 2>&1 ); echo "pytest rc=$?"; tail -4 gpurun_out/pytest_multigpu_n$N.log
( time timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err ); echo "bench rc=$?"; tail -3 gpurun_out/bench_n$N.err
( time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_ref_n$N.json 2> gpurun_out/bench_ref_n$N.err ); echo "ref rc=$?"
