"""World-size-2 test of the SPMD Gauss-Newton driver (mast3r_slam_backends.staged.run_spmd,
shard_range, TorchAllReduce) on CPU with the gloo backend.

The CUDA stage (StagedGN) cannot run here, so the ranks plug in a stage built from the CPU
oracle with the same interface: linearise the own edge shard into a dense fp64 system, sum the
systems over the ranks, solve and retract the replica.  Checks: the sharded solve reproduces
the oracle's single-process Gauss-Newton, and the replicas stay bitwise identical."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import synth
from test_oracle import gn_params


class OracleStage:
    """begin / linearize(into) / solve_retract / system, backed by oracle/ (CPU)."""

    def __init__(self, oracle, P, g, rank, world):
        from mast3r_slam_backends.staged import shard_range

        self.o, self.P, self.g = oracle, P, g
        self.N = g["Xs"].shape[0]
        self.D = 7 * (self.N - 1)
        self.begin_, self.end_ = shard_range(g["ii"].shape[0], rank, world)
        # the shard: edges outside it contribute nothing on this rank
        vm = torch.zeros_like(g["valid_match"])
        vm[self.begin_:self.end_] = g["valid_match"][self.begin_:self.end_]
        self.valid = vm
        self.Twc = g["Twc_init"].clone()
        self.system = torch.zeros(self.D * self.D + self.D, dtype=torch.float64)
        self.dx = torch.zeros((self.N - 1, 7), dtype=torch.float32)
        self.done, self.iters = False, 0

    def begin(self):
        self.done, self.iters = False, 0
        self.ie, self.je, _ = self.o.edge_indices(self.g["ii"], self.g["jj"])

    def linearize(self, into=None):
        assert into is None
        if self.done:
            return
        g = self.g
        Hs, gs = self.o.gn_accumulate(self.P, self.Twc, g["Xs"], g["Cs"], g["ii"], g["jj"],
                                      g["idx_ii2jj"], self.valid, g["Q"])
        H = torch.zeros((self.D, self.D), dtype=torch.float64)
        b = torch.zeros(self.D, dtype=torch.float64)
        Hs, gs = Hs.double(), gs.double()
        for e in range(self.begin_, self.end_):
            io, jo = int(self.ie[e]) - 1, int(self.je[e]) - 1  # pose 0 is fixed
            for blk, (r, c) in enumerate([(io, io), (io, jo), (jo, io), (jo, jo)]):
                if r >= 0 and c >= 0:
                    H[7 * r:7 * r + 7, 7 * c:7 * c + 7] += Hs[blk, e]
            if io >= 0:
                b[7 * io:7 * io + 7] += gs[0, e]
            if jo >= 0:
                b[7 * jo:7 * jo + 7] += gs[1, e]
        self.system[: self.D * self.D] = H.flatten()
        self.system[self.D * self.D:] = b

    def solve_retract(self, delta_thresh):
        if self.done:
            return
        H = self.system[: self.D * self.D].view(self.D, self.D)
        b = self.system[self.D * self.D:]
        L, info = torch.linalg.cholesky_ex(H)
        x = torch.zeros_like(b) if int(info) != 0 else -torch.cholesky_solve(b[:, None], L)[:, 0]
        self.dx = x.float().view(self.N - 1, 7)
        self.Twc = self.o.pose_retr(self.Twc, self.dx)
        self.iters += 1
        if float(self.dx.norm()) < delta_thresh:
            self.done = True


def _worker(rank, world, port, kind, out_dir):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import m3r_oracle as oracle
        from mast3r_slam_backends.staged import TorchAllReduce, run_spmd

        torch.set_num_threads(2)
        g = synth.make_graph(5, 24, 32, seed=3)
        P = gn_params(oracle, kind, g)
        stage = OracleStage(oracle, P, g, rank, world)
        run_spmd(stage, TorchAllReduce(), 4, 0.0)
        torch.save({"Twc": stage.Twc, "dx": stage.dx, "iters": stage.iters,
                    "shard": (stage.begin_, stage.end_)}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("kind", ["rays", "points"])
def test_sharded_gn_two_ranks_gloo(oracle, kind, tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), kind, str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(os.path.join(tmp_path, "rank0.pt"))
    r1 = torch.load(os.path.join(tmp_path, "rank1.pt"))
    # the two shards partition the edge list
    g = synth.make_graph(5, 24, 32, seed=3)
    E = g["ii"].shape[0]
    assert r0["shard"][0] == 0 and r0["shard"][1] == r1["shard"][0] and r1["shard"][1] == E
    # replicas are bitwise identical (same all-reduce result, same solve)
    assert torch.equal(r0["Twc"], r1["Twc"]) and torch.equal(r0["dx"], r1["dx"])
    assert r0["iters"] == 4
    # and equal to the single-process solve of the whole graph
    P = gn_params(oracle, kind, g)
    To, dxo, _ = oracle.gauss_newton(P, g["Twc_init"], g["Xs"], g["Cs"], g["ii"], g["jj"],
                                     g["idx_ii2jj"], g["valid_match"], g["Q"], 4, 0.0)
    assert (r0["Twc"] - To).abs().max() < 1e-5
    assert torch.equal(r0["Twc"][0], g["Twc_init"][0])  # pose 0 stays fixed


def test_shard_range_partitions():
    from mast3r_slam_backends.staged import shard_range

    for E in (0, 1, 7, 108, 492, 2030):
        for world in (1, 2, 3, 4, 8):
            ranges = [shard_range(E, r, world) for r in range(world)]
            assert ranges[0][0] == 0 and ranges[-1][1] == E
            assert all(a[1] == b[0] for a, b in zip(ranges, ranges[1:]))
            sizes = [b - a for a, b in ranges]
            assert max(sizes) - min(sizes) <= 1
