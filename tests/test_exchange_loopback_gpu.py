"""The multi-GPU exchange kernels (m3r_gn_comm_*, csrc/gn_comm.cuh) on ONE device: the ranks
are streams of device 0, each with its own workspace and exchange area, connected with
m3r_gn_comm_connect_local (same-device peers need no peer enabling).  The kernels, flags,
round parities and reduction orders are exactly those of the N-GPU run; only the links are
missing.  This is what the driver's 1-GPU lease can execute; tests/test_multigpu_gpu.py runs
the same protocol over NVLink when two GPUs are visible.

Forward progress on one device: CUDA does not guarantee that kernels of two streams of ONE device
run side by side, so a reduce kernel must never have to spin for a flag here.  The host (one thread
drives all ranks) therefore orders the streams with events exactly as m3r_gn_multi_solve does:
signal -> event; phase 1 of the reduce behind every rank's event -> event; phase 2 behind those
(m3r_gn_comm_reduce_phase).  The flag protocol itself still runs in every kernel and is what the
assertions check; the spin path is exercised where ranks own their device (tests/test_multigpu_gpu.py,
bench.py under torchrun)."""
import ctypes

import pytest
import torch

import synth
from test_gn_gpu import D, KINDS, _call

pytestmark = pytest.mark.gpu


class _Rank:
    def __init__(self, kind, g, rank, world):
        from mast3r_slam_backends import _lib
        from mast3r_slam_backends.staged import StagedGN, shard_range

        self.stream = torch.cuda.Stream()
        self.ev_sig, self.ev_p1 = torch.cuda.Event(), torch.cuda.Event()
        b0, b1 = shard_range(g["E"], rank, world)
        t = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in g.items()}
        self.Twc = t["Twc_init"].clone()
        torch.cuda.synchronize()
        with torch.cuda.stream(self.stream):  # the workspace is keyed by the stream
            self.stage = StagedGN(kind, self.Twc, t["Xs"], t["Cs"], t["ii"], t["jj"],
                                  t["idx_ii2jj"][b0:b1].contiguous(), t["valid_match"][b0:b1].contiguous(),
                                  t["Q"][b0:b1].contiguous(), edge_begin=b0, edge_end=b1, K=t["K"],
                                  height=g["h"], width=g["w"], pixel_border=D["pixel_border"], z_eps=D["z_eps"],
                                  sigma_ray=D["sigma_ray"], sigma_dist=D["sigma_dist"],
                                  sigma_pixel=D["sigma_pixel"], sigma_depth=D["sigma_depth"],
                                  sigma_point=D["sigma_point"], C_thresh=D["C_thresh"], Q_thresh=D["Q_thresh"])
            self.comm = ctypes.c_void_p()
            _lib.check(_lib.gn_comm_create(ctypes.byref(self.comm), rank, world, self.stage.system.numel()),
                       "comm_create")
        torch.cuda.synchronize()

    def sptr(self):
        return ctypes.c_void_p(self.stream.cuda_stream)


def _connect(ranks):
    from mast3r_slam_backends import _lib

    arr = (ctypes.c_void_p * len(ranks))(*[r.comm.value for r in ranks])
    _lib.check(_lib.gn_comm_connect_local(arr, len(ranks)), "connect_local")


def _solve(ranks, iters, delta_thresh, delay_rank=None):
    """One sharded call of `iters` rounds on every rank (the order run_spmd issues per rank,
    interleaved over the ranks)."""
    from mast3r_slam_backends import _lib

    for k, r in enumerate(ranks):
        with torch.cuda.stream(r.stream):
            if delay_rank == k:
                torch.cuda._sleep(int(2e8))  # ~0.1 s: this rank arrives late at every handshake
            r.stage.begin()
    for _ in range(iters):
        for r in ranks:
            with torch.cuda.stream(r.stream):
                buf = ctypes.c_void_p()
                _lib.check(_lib.gn_comm_next_round(r.comm, ctypes.byref(buf)), "next_round")
                r.stage.linearize(into=buf.value)
        for r in ranks:
            _lib.check(_lib.gn_comm_signal(r.comm, r.stage.ws.handle, r.sptr()), "signal")
            r.ev_sig.record(r.stream)
        for phase, before, after in ((1, "ev_sig", "ev_p1"), (2, "ev_p1", None)):
            for r in ranks:
                for q in ranks:
                    if q is not r:
                        r.stream.wait_event(getattr(q, before))
                _lib.check(_lib.gn_comm_reduce_phase(r.comm, r.stage.ws.handle,
                                                     ctypes.c_void_p(r.stage.system.data_ptr()), r.stage.N, phase,
                                                     r.sptr()), "reduce_phase")
                if after:
                    getattr(r, after).record(r.stream)
        for r in ranks:
            with torch.cuda.stream(r.stream):
                r.stage.solve_retract(delta_thresh)
    done = []
    for r in ranks:
        with torch.cuda.stream(r.stream):
            done.append(r.stage.iterations_done())  # raises on an exchange timeout
    torch.cuda.synchronize()
    return done


def _destroy(ranks):
    from mast3r_slam_backends import _lib

    torch.cuda.synchronize()
    for r in ranks:
        _lib.gn_comm_destroy(r.comm)
    _lib.debug_set_exchange_mode(0)


@pytest.mark.parametrize("mode", [1, 2], ids=["oneshot", "twophase"])
@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("kind", KINDS)
def test_loopback_exchange(backends, kind, world, mode):
    """Replicas bitwise identical, equal to the unsharded solve to 1e-5."""
    from mast3r_slam_backends import _lib

    _lib.debug_set_exchange_mode(mode)
    iters = 4
    g = synth.make_graph(7, 96, 128, seed=5)
    ranks = [_Rank(kind, g, r, world) for r in range(world)]
    _connect(ranks)
    done = _solve(ranks, iters, 0.0)
    assert done == [iters] * world
    for r in ranks[1:]:
        assert torch.equal(r.Twc, ranks[0].Twc) and torch.equal(r.stage.dx, ranks[0].stage.dx)
    Twc1 = g["Twc_init"].cuda().clone()
    _call(backends, kind, g, Twc1, iters, 0.0)
    err = (ranks[0].Twc - Twc1).abs().max().item()
    print(f"{kind} world={world} mode={mode}: sharded vs unsharded poses, max abs diff {err:.2e}")
    assert err < 1e-5
    _destroy(ranks)


@pytest.mark.parametrize("mode", [1, 2], ids=["oneshot", "twophase"])
@pytest.mark.parametrize("max_iter", [4, 5])
def test_early_stop_then_second_call_with_a_late_rank(backends, mode, max_iter):
    """A call that converges after its first iteration leaves max_iter - 1 rounds in which nothing
    is exchanged (3: odd, 4: even); the handshake has to keep running through them, otherwise the
    next call on the same communicators starts on a buffer parity a slower peer may still be
    reading.  The second call runs with rank 1 arriving 0.1 s late at every handshake."""
    from mast3r_slam_backends import _lib

    _lib.debug_set_exchange_mode(mode)
    world = 2
    g = synth.make_graph(6, 64, 96, seed=11)
    ranks = [_Rank("rays", g, r, world) for r in range(world)]
    _connect(ranks)
    done = _solve(ranks, max_iter, 1e9)  # the first step is below the threshold: stop
    assert done == [1] * world
    assert torch.equal(ranks[0].Twc, ranks[1].Twc)
    first = ranks[0].Twc.clone()
    done = _solve(ranks, 3, 0.0, delay_rank=1)
    assert done == [3] * world
    assert torch.equal(ranks[0].Twc, ranks[1].Twc) and torch.equal(ranks[0].stage.dx, ranks[1].stage.dx)
    # the same two calls on one GPU, unsharded
    Twc1 = g["Twc_init"].cuda().clone()
    _call(backends, "rays", g, Twc1, max_iter, 1e9)
    assert (first - Twc1).abs().max().item() < 1e-5
    _call(backends, "rays", g, Twc1, 3, 0.0)
    err = (ranks[0].Twc - Twc1).abs().max().item()
    print(f"early stop ({max_iter - 1} idle rounds), second call with a late rank: max abs diff {err:.2e}")
    assert err < 1e-5
    _destroy(ranks)
