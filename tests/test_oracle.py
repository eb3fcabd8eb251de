"""CPU tests of the oracle itself: known-answer tests the math admits and the
committed golden vectors produced by the reference's own CUDA kernels (tests/golden/)."""
import glob
import math
import os

import numpy as np
import pytest
import torch

import synth

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def gn_params(oracle, kind, g, **over):
    D = dict(synth.GN_DEFAULTS)
    D.update(over)
    return oracle.make_params(kind, sigma_ray=D["sigma_ray"], sigma_dist=D["sigma_dist"], K=g["K"],
                              height=g["h"], width=g["w"], pixel_border=D["pixel_border"],
                              z_eps=D["z_eps"], sigma_pixel=D["sigma_pixel"],
                              sigma_depth=D["sigma_depth"], sigma_point=D["sigma_point"],
                              C_thresh=D["C_thresh"], Q_thresh=D["Q_thresh"])


# ---------------------------------------------------------------------------- half rounding
def test_half_conversions_match_torch(oracle):
    L = oracle.lib()
    allh = torch.arange(0, 65536, dtype=torch.int32).to(torch.uint16).view(torch.float16)
    ref = allh.float()
    for bits in list(range(0, 65536, 7)) + [0x7BFF, 0x7C00, 0xFC00, 0x0001, 0x03FF, 0x0400, 0x8000]:
        a = L.m3r_oracle_h2f(bits)
        r = ref[bits].item()
        assert (math.isnan(a) and math.isnan(r)) or a == r
    gen = torch.Generator().manual_seed(0)
    x = torch.cat([torch.randn(20000, generator=gen) * 3, torch.randn(2000, generator=gen) * 1e-5,
                   torch.randn(2000, generator=gen) * 3e4,
                   torch.tensor([65504.0, 65519.9, 65520.0, 1e9, -1e9, 6e-8, 2.98e-8, 2.99e-8, 0.0, -0.0])])
    want = x.half().view(torch.int16).to(torch.int32) & 0xFFFF
    for v, wv in zip(x.tolist(), want.tolist()):
        assert L.m3r_oracle_f2h(v) == wv, v


def test_refine_half_sw_equals_hw(oracle):
    m = synth.make_match_inputs(24, 32, seed=3)
    p1 = torch.stack(torch.meshgrid(torch.arange(32), torch.arange(24), indexing="xy"), -1).reshape(1, -1, 2)
    D11, D21 = m["D11"].half(), m["D21"].view(1, -1, 24).half()
    (a,) = oracle.refine_matches(D11, D21, p1, 3, 5)
    oracle.force_sw_half(True)
    (b,) = oracle.refine_matches(D11, D21, p1, 3, 5)
    oracle.force_sw_half(False)
    assert torch.equal(a, b)


# ---------------------------------------------------------------------------- iter_proj
def test_iter_proj_identity(oracle):
    """pts = normalize(X11) with identity init returns the identity map, all converged."""
    m = synth.make_match_inputs(32, 48, seed=1)
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X11"])
    p, conv = oracle.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    inner = (p0[..., 0] >= 1) & (p0[..., 0] <= 46) & (p0[..., 1] >= 1) & (p0[..., 1] <= 30)
    assert (p - p0).abs()[inner].max() < 1e-3
    assert conv[inner].all()


def test_iter_proj_recovers_projection(oracle):
    m = synth.make_match_inputs(48, 64, seed=2, noise=0.0)
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"])
    p, conv = oracle.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    tu = m["true_uv"]
    inner = (tu[..., 0] > 2) & (tu[..., 0] < 61) & (tu[..., 1] > 2) & (tu[..., 1] < 45)
    err = (p - tu).norm(dim=-1)[inner & conv]
    assert conv[inner].float().mean() > 0.95
    assert err.median() < 0.15  # bilinear ray interpolation error, sub-pixel


def test_iter_proj_zero_iters(oracle):
    m = synth.make_match_inputs(16, 16, seed=0)
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"])
    p, conv = oracle.iter_proj(rays, pts, p0, 0, 1e-8, 1e-6)
    assert not conv.any()
    assert torch.equal(p[..., 0], p0[..., 0].clamp(1, 14)) and torch.equal(p[..., 1], p0[..., 1].clamp(1, 14))


# ---------------------------------------------------------------------------- refine_matches
def _brute_force_refine(D11, D21, p1, radius, dilation_max):
    """float64 brute force of the coarse-to-fine window argmax (well-separated maxima)."""
    b, h, w, f = D11.shape
    out = p1.clone()
    D11d, D21d = D11.double(), D21.double()
    for bi in range(b):
        for k in range(p1.shape[1]):
            u0, v0 = int(p1[bi, k, 0]), int(p1[bi, k, 1])
            best, bu, bv = 0.0, u0, v0
            for d in range(dilation_max, 0, -1):
                for i in range(-radius, radius + 1):
                    for j in range(-radius, radius + 1):
                        u, v = u0 + i * d, v0 + j * d
                        if 0 <= u < w and 0 <= v < h:
                            s = float((D21d[bi, k] * D11d[bi, v, u]).sum())
                            if s > best + 1e-9:
                                best, bu, bv = s, u, v
                u0, v0 = bu, bv
            out[bi, k, 0], out[bi, k, 1] = bu, bv
    return out


def test_refine_f64_matches_brute_force(oracle):
    m = synth.make_match_inputs(16, 20, seed=4, desc_noise=0.0)
    gen = torch.Generator().manual_seed(0)
    p1 = torch.stack([torch.randint(0, 20, (1, 320), generator=gen),
                      torch.randint(0, 16, (1, 320), generator=gen)], -1)
    D11, D21 = m["D11"].double(), m["D21"].view(1, -1, 24).double()
    (got,) = oracle.refine_matches(D11, D21, p1, 2, 3)
    want = _brute_force_refine(D11, D21, p1, 2, 3)
    assert (got == want).all(-1).float().mean() > 0.995  # exact ties aside


def test_refine_nonpositive_scores_do_not_move(oracle):
    """max_score starts at +0 for half (numeric_limits<c10::Half>::min() is Half()), so a
    window whose best score is <= 0 leaves the pixel where it was."""
    D11 = torch.ones(1, 8, 8, 24).half()
    D21 = -torch.ones(1, 64, 24).half()
    p1 = torch.stack(torch.meshgrid(torch.arange(8), torch.arange(8), indexing="xy"), -1).reshape(1, -1, 2)
    (out,) = oracle.refine_matches(D11, D21, p1, 3, 5)
    assert torch.equal(out, p1)


def test_refine_radius0_is_identity(oracle):
    m = synth.make_match_inputs(12, 12, seed=5)
    p1 = torch.stack(torch.meshgrid(torch.arange(12), torch.arange(12), indexing="xy"), -1).reshape(1, -1, 2)
    (out,) = oracle.refine_matches(m["D11"].half(), m["D21"].view(1, -1, 24).half(), p1, 0, 5)
    assert torch.equal(out, p1)


# ---------------------------------------------------------------------------- Gauss-Newton
@pytest.mark.parametrize("kind", ["rays", "calib", "points"])
def test_gn_recovers_known_poses(oracle, kind):
    # matches are rounded to whole pixels, so the optimum is biased by O(1/focal); the bias
    # shrinks with resolution (checked at 96x128, where it is well below the perturbation)
    g = synth.make_graph(5, 96, 128, seed=1, point_noise=0.0)
    P = gn_params(oracle, kind, g)
    T, dx, it = oracle.gauss_newton(P, g["Twc_init"], g["Xs"], g["Cs"], g["ii"], g["jj"],
                                    g["idx_ii2jj"], g["valid_match"], g["Q"], 10, 1e-8)
    e0 = (g["Twc_init"] - g["Twc_true"]).abs().max()
    e1 = (T - g["Twc_true"]).abs().max()
    assert torch.equal(T[0], g["Twc_init"][0])  # pose 0 is fixed
    assert e1 < 0.3 * e0
    assert dx.norm() < 1e-4


def test_gn_all_invalid_gives_zero_step(oracle):
    """Zero Hessian -> Cholesky failure -> dx = 0, Twc untouched (gn_kernels.cu:147-150)."""
    g = synth.make_graph(4, 16, 16, seed=2)
    P = gn_params(oracle, "rays", g)
    vm = torch.zeros_like(g["valid_match"])
    T, dx, it = oracle.gauss_newton(P, g["Twc_init"], g["Xs"], g["Cs"], g["ii"], g["jj"],
                                    g["idx_ii2jj"], vm, g["Q"], 3, 1e-8)
    assert torch.equal(T, g["Twc_init"]) and (dx == 0).all() and it == 1


def test_retraction_identity_and_branches(oracle):
    T = synth.make_poses(4, 0).float()
    out = oracle.pose_retr(T, torch.zeros(3, 7))
    assert torch.allclose(out, T, atol=0, rtol=0)
    # The small-angle branches (EPS = 1e-6, gn_kernels.cu:34,306,350-373) must join the
    # general formulas continuously: evaluate just below and just above each threshold.
    # (sigma is kept at 0 or 0.3: for |sigma| barely above EPS the reference's fp32
    # (scale-1)/sigma cancels catastrophically, which the oracle reproduces on purpose.)
    def retr(xi):
        return oracle.pose_retr(T, torch.tensor([xi] * 3))[1:]

    for sg in (0.0, 0.3):
        lo = retr([0.1, -0.2, 0.05, 0.99e-6, 0.0, 0.0, sg])   # |theta| < EPS branch
        hi = retr([0.1, -0.2, 0.05, 1.01e-6, 0.0, 0.0, sg])   # general branch
        assert (lo - hi).abs().max() < 1e-5
        lo = retr([0.1, -0.2, 0.05, 0.99e-3, 0.0, 0.0, sg])   # theta^2 < EPS in expSO3
        hi = retr([0.1, -0.2, 0.05, 1.01e-3, 0.0, 0.0, sg])
        assert (lo - hi).abs().max() < 1e-5
    # and a plain rotation + scale agrees with the float64 group composition
    xi = torch.tensor([[0.0, 0.0, 0.0, 0.02, -0.03, 0.01, 0.1]] * 3)
    a = oracle.pose_retr(T, xi)[1:]
    b = synth.sim3_perturb(T.double()[1:], xi.double()).float()
    assert (a - b).abs().max() < 1e-6


def test_edge_indices(oracle):
    ii = torch.tensor([10, 30, 30, 50, 7])
    jj = torch.tensor([30, 10, 50, 7, 10])
    ie, je, nu = oracle.edge_indices(ii, jj)
    assert nu == 4 and ie.tolist() == [1, 2, 2, 3, 0] and je.tolist() == [2, 1, 3, 0, 1]


def test_gn_blocks_identities(oracle):
    """H_ii = H_jj, H_ij = H_ji^T = -H_jj up to the per-term rounding of w*Ja*Jc, g_i = -g_j."""
    g = synth.make_graph(3, 16, 24, seed=3)
    P = gn_params(oracle, "rays", g)
    Hs, gs = oracle.gn_accumulate(P, g["Twc_init"], g["Xs"], g["Cs"], g["ii"], g["jj"],
                                  g["idx_ii2jj"], g["valid_match"], g["Q"])
    assert torch.equal(Hs[0], Hs[3]) and torch.equal(gs[0], -gs[1])
    assert torch.equal(Hs[1], Hs[2].transpose(-1, -2))
    assert torch.allclose(Hs[1], -Hs[3], rtol=1e-4, atol=1e-3 * Hs[3].abs().max().item())


# ---------------------------------------------------------------------------- golden vectors
def _golden(prefix):
    return sorted(glob.glob(os.path.join(GOLD, prefix + "*.npz")))


@pytest.mark.parametrize("path", _golden("match_") or [None])
def test_golden_matching(oracle, path):
    """Oracle vs outputs of the reference's CUDA kernels (generated on a B200 by
    tests/golden/make_golden.py)."""
    if path is None:
        pytest.skip("no golden matching vectors committed yet")
    z = np.load(path)
    rays, pts, p0 = (torch.from_numpy(z[k]) for k in ("rays", "pts", "p_init"))
    p, conv = oracle.iter_proj(rays, pts, p0, int(z["max_iter"]), float(z["lambda_init"]),
                               float(z["cost_thresh"]))
    ref_p, ref_conv = torch.from_numpy(z["ref_p_new"]), torch.from_numpy(z["ref_converged"])
    agree = (p.long() == ref_p.long()).all(-1)
    assert agree.float().mean() >= 0.999, f"iter_proj index agreement {agree.float().mean()}"
    assert (conv == ref_conv).float().mean() >= 0.999
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16)
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16)
    p1 = torch.from_numpy(z["ref_p1"])
    (out,) = oracle.refine_matches(D11, D21, p1, int(z["radius"]), int(z["dilation_max"]))
    assert torch.equal(out, torch.from_numpy(z["ref_p1_new"])), "refine_matches must be bit-exact"
    if "ref_p1_new_f32" in z:
        (o32,) = oracle.refine_matches(D11.float(), D21.float(), p1, int(z["radius"]), int(z["dilation_max"]))
        assert (o32 == torch.from_numpy(z["ref_p1_new_f32"])).all(-1).float().mean() >= 0.9999


@pytest.mark.parametrize("path", _golden("gn_") or [None])
def test_golden_gn(oracle, path):
    if path is None:
        pytest.skip("no golden GN vectors committed yet")
    z = np.load(path)
    kind = str(z["kind"])
    t = {k: torch.from_numpy(z[k]) for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q", "K")}
    g = dict(K=t["K"], h=int(z["height"]), w=int(z["width"]))
    P = gn_params(oracle, kind, g)
    for iters, key in ((1, "ref_1"), (int(z["max_iter"]), "ref_full")):
        T, dx, it = oracle.gauss_newton(P, t["Twc"], t["Xs"], t["Cs"], t["ii"], t["jj"], t["idx"],
                                        t["valid"].bool(), t["Q"], iters, float(z["delta_thresh"]))
        rT, rdx = torch.from_numpy(z[key + "_Twc"]), torch.from_numpy(z[key + "_dx"])
        assert (T - rT).abs().max() <= 1e-4 * max(1.0, rT.abs().max().item()), (kind, key)
        if iters == 1:
            assert (dx - rdx).norm() <= 1e-4 * rdx.norm() + 1e-7, (kind, (dx - rdx).norm() / rdx.norm())


@pytest.mark.parametrize("path", _golden("front_") or [None])
def test_golden_front_end(oracle, path):
    """Oracle restatement of the Python glue around the kernels (oracle.match,
    oracle.prep_for_iter_proj) vs the reference's own matching.py / image.py run on the same
    inputs (tests/golden/make_golden_front.py)."""
    if path is None:
        pytest.skip("no golden front-end vectors committed yet")
    z = np.load(path)
    X11, X21 = torch.from_numpy(z["X11"]), torch.from_numpy(z["X21"])
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16)
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16)
    idx_init = torch.from_numpy(z["idx_init"]) if "idx_init" in z else None
    rays, pts, p_init = oracle.prep_for_iter_proj(X11, X21, idx_init)
    assert torch.equal(rays, torch.from_numpy(z["ref_rays"]))
    assert torch.equal(pts, torch.from_numpy(z["ref_pts"]))
    assert torch.equal(p_init, torch.from_numpy(z["ref_p_init"]))
    cfg = {k: z[k].item() for k in ("max_iter", "lambda_init", "convergence_thresh", "dist_thresh",
                                    "radius", "dilation_max")}
    idx, valid = oracle.match(X11, X21, D11, D21, idx_init, **cfg)
    assert idx.dtype == torch.int64 and valid.dtype == torch.bool and valid.shape == (*idx.shape, 1)
    assert torch.equal(idx, torch.from_numpy(z["ref_idx"]))
    assert torch.equal(valid, torch.from_numpy(z["ref_valid"]))


def test_golden_random_init_network(oracle):
    """The production network with seeded random weights as the input distribution
    (tests/golden/make_golden_mast3r.py): near-degenerate rays, 4 % of the LM runs converge,
    descriptor scores without a clear maximum.  The oracle restatement must reproduce the
    reference's matching.py on these values exactly as on the smooth scenes."""
    path = os.path.join(GOLD, "mast3r_rand.npz")
    if not os.path.exists(path):
        pytest.skip("no random-init network vectors committed")
    z = np.load(path)
    X11, X21 = torch.from_numpy(z["X11"]), torch.from_numpy(z["X21"])
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16)
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16)
    # the head ranges the fixture is supposed to carry (catmlp_dpt_head.py:17-38, postprocess.py:49-58)
    assert z["C11"].min() >= 1.0 and z["C21"].min() >= 1.0
    assert z["Q11"].min() >= 0.0 and z["Q21"].min() >= 0.0
    assert abs(float(D11.float().norm(dim=-1).mean()) - 1.0) < 1e-3
    rays, pts, p_init = oracle.prep_for_iter_proj(X11, X21, None)
    assert torch.equal(rays, torch.from_numpy(z["ref_rays"]))
    cfg = {k: z[k].item() for k in ("max_iter", "lambda_init", "convergence_thresh", "dist_thresh",
                                    "radius", "dilation_max")}
    p, conv = oracle.iter_proj(rays, pts, p_init, cfg["max_iter"], cfg["lambda_init"], cfg["convergence_thresh"])
    assert torch.equal(p, torch.from_numpy(z["ref_p_proj"]))
    assert torch.equal(conv, torch.from_numpy(z["ref_converged"]))
    idx, valid = oracle.match(X11, X21, D11, D21, None, **cfg)
    assert torch.equal(idx, torch.from_numpy(z["ref_idx"]))
    assert torch.equal(valid, torch.from_numpy(z["ref_valid"]))


TRACK_KEYS = ("max_iters", "rel_error", "delta_norm", "huber", "sigma_ray", "sigma_dist", "sigma_pixel",
              "sigma_depth", "pixel_border", "depth_eps")


def load_track_golden(path):
    z = np.load(path)
    t = {k: torch.from_numpy(z[k]) for k in ("Xf", "Xk", "Qk", "valid_opt", "meas_k", "valid_meas_k", "K",
                                             "T_WCk", "T_WCf_init", "T_WCf_true")}
    cfg = {k: z[k].item() for k in TRACK_KEYS}
    ref = {k: torch.from_numpy(z[k]) for k in z.files if k.startswith("ref_")}
    return t, cfg, ref, (int(z["h"]), int(z["w"]))


@pytest.mark.parametrize("path", _golden("track_") or [None])
def test_golden_tracking(oracle, path):
    """oracle.track_pose_rays / track_pose_calib vs the reference's own tracker.py loop run on
    the same inputs (tests/golden/make_golden_track.py)."""
    if path is None:
        pytest.skip("no golden tracking vectors committed yet")
    t, cfg, ref, img = load_track_golden(path)
    Twf, Tkf, it, cost = oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                t["valid_opt"], cfg)
    assert (Twf - ref["ref_rays_T_WCf"]).abs().max() < 1e-5
    assert (Tkf - ref["ref_rays_T_CkCf"]).abs().max() < 1e-5
    Twf, Tkf, it, cost = oracle.track_pose_calib(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                 t["valid_opt"], t["meas_k"], t["valid_meas_k"], t["K"],
                                                 img, cfg)
    assert (Twf - ref["ref_calib_T_WCf"]).abs().max() < 1e-5
    assert (Tkf - ref["ref_calib_T_CkCf"]).abs().max() < 1e-5


def test_sim3_stand_in_group_laws(oracle):
    S = oracle.Sim3
    g = torch.Generator().manual_seed(0)
    d = torch.randn(4, 8, generator=g)
    d[:, 3:7] /= d[:, 3:7].norm(dim=-1, keepdim=True)
    d[:, 7] = torch.exp(0.3 * d[:, 7])
    A, B = S(d[:2]), S(d[2:])
    p = torch.randn(2, 3, generator=g)
    assert ((A * B).act(p) - A.act(B.act(p))).abs().max() < 1e-5
    assert (A.inv().act(A.act(p)) - p).abs().max() < 1e-5
    I = (A.inv() * A).data
    assert (I - S.Identity(2).data).abs().max() < 1e-6
    # retr(0) is the identity; retr is a LEFT update: a pure translation tau adds to t
    assert torch.equal(A.retr(torch.zeros(2, 7)).data, A.data)
    tau = torch.zeros(2, 7)
    tau[:, :3] = torch.tensor([0.1, -0.2, 0.3])
    assert (A.retr(tau).data[:, :3] - (A.data[:, :3] + tau[:, :3])).abs().max() < 1e-6
