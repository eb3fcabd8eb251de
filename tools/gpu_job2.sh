set -x
nvidia-smi -L
python -m pytest tests/test_multigpu_gpu.py tests/test_gn_gpu.py -m gpu -x -q > gpurun_out/pytest_mg.log 2>&1; echo "pytest rc=$?"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 50 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench2 rc=$?"
python bench.py --steps 50 --skip-cpu --skip-refcuda > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench1 rc=$?"
