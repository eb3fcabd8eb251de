set -x
N=${NGPU:-8}
nvidia-smi -L | wc -l
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N tools/gn_scale.py rays:64 calib:64 points:256 > gpurun_out/gn_scale_n$N.jsonl 2> gpurun_out/gn_scale_n$N.err; echo "scale$N rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 50 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench$N rc=$?"
cat gpurun_out/gn_scale_n$N.jsonl
tail -2 gpurun_out/gn_scale_n$N.err
