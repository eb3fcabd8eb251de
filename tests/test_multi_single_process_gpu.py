"""Multi-GPU Gauss-Newton behind the unchanged single-process call (m3r_gn_multi_*, SURVEY.md
7.3-9): `gauss_newton_rays / _calib / _points` with M3R_GN_DEVICES / set_gn_devices.

On a 1-GPU lease the device list repeats device 0 (loop-back ranks: own workspace, exchange area
and stream each, the caller's tensors read "remotely" through the same code path); with two or
more GPUs visible the second parametrisation spreads the ranks over real devices and the other
ranks read the caller's tensors over NVLink peer access."""
import pytest
import torch

import synth
from test_gn_gpu import D, KINDS, _call

pytestmark = pytest.mark.gpu


def _device_lists():
    out = [("loopback2", [0, 0]), ("loopback3", [0, 0, 0])]
    if torch.cuda.is_available() and torch.cuda.device_count() >= 2:
        out.append(("two_gpus", [0, 1]))
    if torch.cuda.is_available() and torch.cuda.device_count() >= 4:
        out.append(("four_gpus", [0, 1, 2, 3]))
    return out


@pytest.mark.parametrize("name,ids", _device_lists(), ids=[n for n, _ in _device_lists()])
@pytest.mark.parametrize("kind", KINDS)
def test_drop_in_call_on_several_devices(backends, kind, name, ids):
    g = synth.make_graph(9, 96, 128, seed=4, device="cuda")
    T1 = g["Twc_init"].clone()
    backends.set_gn_devices(None)
    (dx1,) = _call(backends, kind, g, T1, 6, 0.0)
    try:
        backends.set_gn_devices(ids)
        for rep in range(2):  # the second call reuses the cached handle (exchange rounds carry on)
            TG = g["Twc_init"].clone()
            (dxG,) = _call(backends, kind, g, TG, 6, 0.0)
            rep_ = backends.last_multi_report(TG)
            err = (TG - T1).abs().max().item()
            print(f"{kind} {name} call {rep}: vs 1 GPU max abs diff {err:.2e}, begin ms per rank "
                  f"{[round(x, 3) for x in rep_['begin_ms']]}")
            assert rep_["iterations"] == 6 and rep_["replicas_bitwise_equal"]
            assert err < 1e-5 and (dxG - dx1).abs().max().item() < 1e-5
            assert torch.equal(TG[0], g["Twc_init"][0])
    finally:
        backends.set_gn_devices(None)


@pytest.mark.parametrize("mode", [1, 2], ids=["oneshot", "twophase"])
def test_exchange_forms_four_ranks(backends, mode):
    """Both forms of the exchange under the single-process driver (one-shot reduce; reduce-scatter +
    all-gather, which has a second signal and a second event fence), four ranks, an early stop in
    the middle of the first call."""
    from mast3r_slam_backends import _lib

    g = synth.make_graph(10, 64, 96, seed=7, device="cuda")
    try:
        _lib.debug_set_exchange_mode(mode)
        for delta, iters in ((1e9, 1), (0.0, 4)):
            backends.set_gn_devices(None)
            T1 = g["Twc_init"].clone()
            _call(backends, "calib", g, T1, 4, delta)
            backends.set_gn_devices([0, 0, 0, 0])
            TG = g["Twc_init"].clone()
            _call(backends, "calib", g, TG, 4, delta)
            r = backends.last_multi_report(TG)
            assert r["iterations"] == iters and r["replicas_bitwise_equal"], r
            assert (TG - T1).abs().max().item() < 1e-5
    finally:
        _lib.debug_set_exchange_mode(0)
        backends.set_gn_devices(None)


def test_growing_graph_and_early_stop(backends):
    """Keyframes arrive one by one (the exchange areas are re-allocated when the graph outgrows
    them) and a call may stop early; the next call on the same handle still agrees with one GPU."""
    try:
        for N in (4, 7, 12, 30):
            g = synth.make_graph(N, 48, 64, seed=N, device="cuda")
            for delta in (1e9, 0.0):
                T1 = g["Twc_init"].clone()
                backends.set_gn_devices(None)
                _call(backends, "rays", g, T1, 5, delta)
                backends.set_gn_devices([0, 0])
                TG = g["Twc_init"].clone()
                _call(backends, "rays", g, TG, 5, delta)
                r = backends.last_multi_report(TG)
                assert r["iterations"] == (1 if delta > 0 else 5) and r["replicas_bitwise_equal"], (N, delta, r)
                assert (TG - T1).abs().max().item() < 1e-5
    finally:
        backends.set_gn_devices(None)


def test_environment_variable_form(backends, monkeypatch):
    g = synth.make_graph(5, 48, 64, seed=2, device="cuda")
    T1 = g["Twc_init"].clone()
    backends.set_gn_devices(None)
    _call(backends, "points", g, T1, 3, 0.0)
    try:
        backends.set_gn_devices("env")
        monkeypatch.setenv("M3R_GN_DEVICES", "0,0,0,0")
        TG = g["Twc_init"].clone()
        _call(backends, "points", g, TG, 3, 0.0)
        assert backends.last_multi_report(TG)["devices"] == [0, 0, 0, 0]
        assert (TG - T1).abs().max().item() < 1e-5
        monkeypatch.setenv("M3R_GN_DEVICES", "1,0")
        with pytest.raises(RuntimeError, match="must be the tensors' device"):
            _call(backends, "points", g, g["Twc_init"].clone(), 1, 0.0)
        monkeypatch.setenv("M3R_GN_DEVICES", "64")
        with pytest.raises(RuntimeError, match="CUDA devices are visible"):
            _call(backends, "points", g, g["Twc_init"].clone(), 1, 0.0)
    finally:
        backends.set_gn_devices(None)
