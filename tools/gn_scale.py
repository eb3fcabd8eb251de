"""Sharded Gauss-Newton sweep for BASELINE configs 3 and 4 (and the headline 64-KF rays graph):
run under torchrun with N ranks (or plainly for N = 1).  One JSON line per (kind, keyframes)
on rank 0: iterations/s of the whole job, per-stage times (max over ranks for the total),
shard sizes.  Usage:
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
      --master-port P tools/gn_scale.py rays:64 calib:64 points:256"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
for p in ("lightweight-mast3r-slam_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch  # noqa: E402
import bench  # noqa: E402
import synth  # noqa: E402


def main():
    rank, local_rank, world = bench.dist_env()
    torch.cuda.set_device(local_rank)
    if world > 1:
        import torch.distributed as dist

        saved = os.dup(1)
        os.dup2(2, 1)  # NCCL's banner goes to stderr
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        dist.barrier()
        os.dup2(saved, 1)
    D = dict(synth.GN_DEFAULTS)
    for spec in sys.argv[1:]:
        kind, nkf = spec.split(":")
        g = bench.make_gn_graph(int(nkf), "cuda", seed=0)
        r = bench.time_gn(g, D, kind, rank, world, iters=10, reps=2, exchange="peer")
        rn = bench.time_gn(g, D, kind, rank, world, iters=10, reps=1, exchange="nccl") if world > 1 else None
        modes = {}
        if world > 1:
            from mast3r_slam_backends import _lib

            for name, mode in (("oneshot", 1), ("twophase", 2)):
                _lib.debug_set_exchange_mode(mode)
                rm = bench.time_gn(g, D, kind, rank, world, iters=10, reps=1, exchange="peer")
                modes[name + "_exchange_ms"] = rm["allreduce_ms"]
                modes[name + "_ms_per_iter"] = rm["total_ms"] / 10
                modes[name + "_pose_equal"] = bool(torch.equal(rm["Twc"], r["Twc"]))
            _lib.debug_set_exchange_mode(0)
        if rank == 0:
            E, n, N = g["E"], g["n"], g["N"]
            print(json.dumps({
                "kind": kind, "keyframes": N, "directed_edges": E, "points": n, "n_gpus": world,
                "iters_per_s": 10 / (r["total_ms"] * 1e-3), "ms_per_iter": r["total_ms"] / 10,
                "begin_ms": r["begin_ms"], "linearize_ms": r["linearize_ms"], "exchange_ms": r["allreduce_ms"],
                "solve_ms": r["solve_ms"], "iters_done": r["iters_done"], "system_dim": 7 * (N - 1),
                "exchange_bytes": (7 * (N - 1)) ** 2 * 8 + 7 * (N - 1) * 8,
                "pose_err_vs_truth": float((r["Twc"] - g["Twc_true"]).abs().max()),
                "pose_err_init": float((g["Twc_init"] - g["Twc_true"]).abs().max()),
                "nccl_exchange_ms": rn["allreduce_ms"] if rn else None,
                "nccl_ms_per_iter": rn["total_ms"] / 10 if rn else None, **modes}), flush=True)
        del g
        torch.cuda.empty_cache()
    if world > 1:
        import torch.distributed as dist

        dist.destroy_process_group()


if __name__ == "__main__":
    main()
