"""CPU twin of tests/test_reference_callers_gpu.py: the reference's own matching.py and
global_opt.FactorGraph, unmodified (tests/refpy.py), with `mast3r_slam_backends` = the CPU oracle.
Checks the plumbing every GPU run of that test depends on (stubs, SharedKeyframes, the in-place
pose view) and that the oracle is a complete stand-in for the five exports."""
import types

import pytest
import torch

import refpy
import synth


def _oracle_backend(oracle):
    mod = types.ModuleType("mast3r_slam_backends")
    mod.iter_proj = oracle.iter_proj
    mod.refine_matches = oracle.refine_matches

    def rays(Twc, Xs, Cs, ii, jj, idx, vm, Q, sr, sd, Ct, Qt, mi, dt):
        P = oracle.make_params("rays", sigma_ray=sr, sigma_dist=sd, C_thresh=Ct, Q_thresh=Qt)
        To, dx, _ = oracle.gauss_newton(P, Twc, Xs, Cs, ii, jj, idx, vm, Q, mi, dt)
        Twc.copy_(To)  # the export updates the caller's view in place (gn_kernels.cu:1212)
        return [dx]

    def calib(Twc, Xs, Cs, K, ii, jj, idx, vm, Q, h, w, pb, ze, sp, sd, Ct, Qt, mi, dt):
        P = oracle.make_params("calib", K=K, height=h, width=w, pixel_border=pb, z_eps=ze, sigma_pixel=sp,
                               sigma_depth=sd, C_thresh=Ct, Q_thresh=Qt)
        To, dx, _ = oracle.gauss_newton(P, Twc, Xs, Cs, ii, jj, idx, vm, Q, mi, dt)
        Twc.copy_(To)
        return [dx]

    mod.gauss_newton_rays, mod.gauss_newton_calib = rays, calib
    return mod


@pytest.mark.parametrize("calib", [False, True], ids=["rays", "calib"])
def test_reference_factor_graph_over_the_oracle(oracle, monkeypatch, calib):
    if refpy.reference_root() is None:
        pytest.skip("neither baseline/_ref nor /root/reference present")
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)  # profiler.py:18 (no GPU here)
    ns = refpy.load(_oracle_backend(oracle))
    N, h, w, seed = 5, 48, 64, 0
    si, sj = synth.make_graph_edges(N, seed)
    ii, jj = si.tolist(), sj.tolist()
    pairs = {(a, b) for a, b in zip(ii, jj)} | {(b, a) for a, b in zip(ii, jj)}
    scene = refpy.make_scene(N, h, w, seed, "cpu", sorted(pairs))
    ns.cfg.config["use_calib"] = calib
    ns.mu.mast3r_match_symmetric = refpy.SyntheticDecoder(ns, scene)
    ns.global_opt.mast3r_match_symmetric = ns.mu.mast3r_match_symmetric
    frames = refpy.make_keyframes(ns, scene, "cpu")
    fg = ns.global_opt.FactorGraph(None, frames, scene["K"] if calib else None, "cpu")
    for k in range(0, len(ii), 4):
        assert fg.add_factors(ii[k:k + 4], jj[k:k + 4], 0.1)
    assert fg.ii.tolist() == ii and fg.jj.tolist() == jj
    (fg.solve_GN_calib if calib else fg.solve_GN_rays)()
    T = frames.T_WC[:N, 0]
    err0 = (scene["T_init"] - scene["T_true"]).abs().max().item()
    err1 = (T - scene["T_true"]).abs().max().item()
    assert torch.equal(T[0], scene["T_init"][0]) and err1 < err0
