"""Two-GPU test of the NVLink peer-memory exchange (m3r_gn_comm_*): both ranks live in this
process (one per device, m3r_gn_comm_connect_local), each linearises its edge shard into its
exchange buffer, signals, reduces over peer memory, solves its replica.  Skipped on boxes with
a single GPU; the one-process-per-GPU form (CUDA IPC) is what bench.py runs under torchrun."""
import ctypes

import pytest
import torch

import synth
from test_gn_gpu import D, KINDS, _call

pytestmark = pytest.mark.gpu


def _stage(kind, g, dev, rank, world):
    from mast3r_slam_backends.staged import StagedGN, shard_range

    b0, b1 = shard_range(g["E"], rank, world)
    t = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in g.items()}
    Twc = t["Twc_init"].clone()
    with torch.cuda.device(dev):
        st = StagedGN(kind, Twc, t["Xs"], t["Cs"], t["ii"], t["jj"], t["idx_ii2jj"][b0:b1].contiguous(),
                      t["valid_match"][b0:b1].contiguous(), t["Q"][b0:b1].contiguous(), edge_begin=b0,
                      edge_end=b1, K=t["K"], height=g["h"], width=g["w"], pixel_border=D["pixel_border"],
                      z_eps=D["z_eps"], sigma_ray=D["sigma_ray"], sigma_dist=D["sigma_dist"],
                      sigma_pixel=D["sigma_pixel"], sigma_depth=D["sigma_depth"],
                      sigma_point=D["sigma_point"], C_thresh=D["C_thresh"], Q_thresh=D["Q_thresh"])
    return st, Twc


@pytest.mark.parametrize("mode", [1, 2], ids=["oneshot", "twophase"])
@pytest.mark.parametrize("kind", KINDS)
def test_peer_exchange_two_gpus(backends, kind, mode):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from mast3r_slam_backends import _lib

    _lib.debug_set_exchange_mode(mode)  # reduce in one shot / reduce-scatter + all-gather
    world, iters = 2, 4
    g = synth.make_graph(7, 96, 128, seed=5)
    devs = [torch.device("cuda", r) for r in range(world)]
    stages, poses, comms = [], [], []
    for r, dev in enumerate(devs):
        st, Twc = _stage(kind, g, dev, r, world)
        stages.append(st), poses.append(Twc)
        h = ctypes.c_void_p()
        with torch.cuda.device(dev):
            _lib.check(_lib.gn_comm_create(ctypes.byref(h), r, world, st.system.numel()), "comm_create")
        comms.append(h)
    arr = (ctypes.c_void_p * world)(*[c.value for c in comms])
    _lib.check(_lib.gn_comm_connect_local(arr, world), "connect_local")

    def stream(dev):
        return ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    for st in stages:
        st.begin()
    for _ in range(iters):
        for r, st in enumerate(stages):
            buf = ctypes.c_void_p()
            _lib.check(_lib.gn_comm_next_round(comms[r], ctypes.byref(buf)), "next_round")
            st.linearize(into=buf.value)
        for r, st in enumerate(stages):
            with torch.cuda.device(devs[r]):
                _lib.check(_lib.gn_comm_signal(comms[r], st.ws.handle, stream(devs[r])), "signal")
        for r, st in enumerate(stages):
            with torch.cuda.device(devs[r]):
                _lib.check(_lib.gn_comm_reduce(comms[r], st.ws.handle, ctypes.c_void_p(st.system.data_ptr()),
                                               st.N, stream(devs[r])), "reduce")
        for st in stages:
            st.solve_retract(0.0)
    for dev in devs:
        torch.cuda.synchronize(dev)
    assert all(st.iterations_done() == iters for st in stages)
    # replicas bitwise identical; equal to the single-GPU solve of the whole graph
    assert torch.equal(poses[0].cpu(), poses[1].cpu())
    assert torch.equal(stages[0].dx.cpu(), stages[1].dx.cpu())
    Twc1 = g["Twc_init"].cuda().clone()
    _call(backends, kind, g, Twc1, iters, 0.0)
    err = (poses[0].cpu() - Twc1.cpu()).abs().max().item()
    print(f"{kind}: 2-GPU sharded vs 1-GPU poses, max abs diff {err:.2e}")
    assert err < 1e-5
    for c in comms:
        _lib.gn_comm_destroy(c)
    _lib.debug_set_exchange_mode(0)
