"""GPU parity tests of the fused matching front end: m3r_prep_rays
and m3r_match through mast3r_slam_backends.matching, against the committed golden vectors
(the reference's own matching.py / image.py run on CPU, tests/golden/make_golden_front.py) and
against the CPU oracle on seeded inputs.

Tolerances: the ray image is fp32 arithmetic in a different summation order than
torch's conv2d -> 2e-6 absolute (values are O(1)); match indices and the valid mask agree on
>= 99.9 % of pixels (BASELINE.json north_star, same bar as iter_proj: a sub-pixel position
that lands within rounding of an integer may truncate to the neighbouring pixel)."""
import glob
import os

import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu
M = synth.MATCH_DEFAULTS
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RAY_TOL = 2e-6


@pytest.fixture(scope="module")
def front(backends):
    from mast3r_slam_backends import matching

    matching.set_config({"matching": dict(M)})
    return matching


def _case(h, w, b, seed, init):
    m = synth.make_match_inputs(h, w, b=b, seed=seed, device="cuda")
    idx_init = None
    if init != "identity":
        tu = m["true_uv"].round().long()
        if init == "shifted":
            tu = tu + torch.tensor([3, -2], device="cuda")
        idx_init = (tu[..., 1].clamp(0, h - 1) * w + tu[..., 0].clamp(0, w - 1)).contiguous()
    return m, idx_init


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "front_*.npz"))) or [None])
def test_golden_front_end_gpu(front, path):
    if path is None:
        pytest.skip("no golden front-end vectors committed")
    z = np.load(path)
    X11, X21 = torch.from_numpy(z["X11"]).cuda(), torch.from_numpy(z["X21"]).cuda()
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16).cuda()
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16).cuda()
    idx_init = torch.from_numpy(z["idx_init"]).cuda() if "idx_init" in z else None
    rays = front.ray_image_with_grad(X11)
    err = (rays.cpu() - torch.from_numpy(z["ref_rays"])).abs().max().item()
    print(f"{os.path.basename(path)}: ray image max abs err {err:.3e}")
    assert err <= RAY_TOL
    r2, pts, p0 = front.prep_for_iter_proj(X11, X21, idx_init)
    assert torch.equal(r2, rays)
    assert (pts.cpu() - torch.from_numpy(z["ref_pts"])).abs().max() <= 1e-6
    assert torch.equal(p0.cpu(), torch.from_numpy(z["ref_p_init"]))
    idx, valid = front.match(X11, X21, D11, D21, idx_init)
    assert idx.dtype == torch.int64 and valid.dtype == torch.bool
    assert idx.shape == z["ref_idx"].shape and valid.shape == z["ref_valid"].shape
    fi = (idx.cpu() == torch.from_numpy(z["ref_idx"])).float().mean().item()
    fv = (valid.cpu() == torch.from_numpy(z["ref_valid"])).float().mean().item()
    print(f"{os.path.basename(path)}: idx agreement {fi:.6f}, valid agreement {fv:.6f}")
    assert fi >= 0.999 and fv >= 0.999


@pytest.mark.parametrize("h,w,b", [(24, 32, 1), (37, 53, 2), (8, 33, 1), (3, 3, 1), (2, 2, 1), (384, 512, 1)])
def test_ray_image_vs_oracle(front, oracle, h, w, b):
    """Partial tiles, one-tile and minimum-size images (reflect padding needs >= 2)."""
    g = torch.Generator().manual_seed(h * 1000 + w)
    X11 = torch.randn(b, h, w, 3, generator=g) * torch.tensor([1.0, 1.0, 0.3]) + torch.tensor([0.0, 0.0, 2.0])
    X11[0, 0, 0] = 0.0  # zero vector: F.normalize's eps clamp, result 0
    ref, _, _ = oracle.prep_for_iter_proj(X11, X11, None)
    rays = front.ray_image_with_grad(X11.cuda())
    assert rays.shape == (b, h, w, 9)
    assert torch.isfinite(rays).all()
    assert (rays.cpu() - ref).abs().max().item() <= RAY_TOL
    assert torch.equal(rays[0, 0, 0, :3].cpu(), torch.zeros(3))


@pytest.mark.parametrize("h,w,b,seed,init", [(48, 64, 1, 0, "identity"), (96, 160, 3, 2, "true"),
                                              (61, 83, 2, 3, "shifted"), (224, 224, 1, 4, "identity"),
                                              (384, 512, 1, 5, "true")])
def test_match_vs_oracle(front, oracle, h, w, b, seed, init):
    m, idx_init = _case(h, w, b, seed, init)
    idx, valid = front.match(m["X11"], m["X21"], m["D11"], m["D21"], idx_init)
    io, vo = oracle.match(m["X11"], m["X21"], m["D11"], m["D21"], idx_init, **M)
    fi = (idx.cpu() == io).float().mean().item()
    fv = (valid.cpu() == vo).float().mean().item()
    print(f"match {h}x{w} b={b} init={init}: idx agreement {fi:.6f}, valid agreement {fv:.6f}, "
          f"valid fraction {valid.float().mean().item():.3f}")
    assert idx.shape == (b, h * w) and valid.shape == (b, h * w, 1)
    assert fi >= 0.999 and fv >= 0.999
    assert (idx >= 0).all() and (idx < h * w).all()


def test_match_equals_unfused_calls(front, backends):
    """The fused call and the reference-shaped sequence over the same kernels (ray image from
    m3r_prep_rays, then iter_proj, the torch epilogue of matching.py:66-85, refine_matches)
    agree; differences can only come from F.normalize vs the in-kernel normalisation of X21."""
    h, w, b = 120, 168, 2
    m, idx_init = _case(h, w, b, 9, "true")
    idx, valid = front.match(m["X11"], m["X21"], m["D11"], m["D21"], idx_init)
    rays, pts, p0 = front.prep_for_iter_proj(m["X11"], m["X21"], idx_init)
    p, conv = backends.iter_proj(rays, pts.contiguous(), p0.contiguous(), M["max_iter"], M["lambda_init"],
                                 M["convergence_thresh"])
    p1 = p.long()
    bi = torch.arange(b, device="cuda")[:, None].repeat(1, h * w)
    d = torch.linalg.norm(m["X11"][bi, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - m["X21"], dim=-1)
    v2 = conv & (d < M["dist_thresh"]).view(b, -1)
    (p1n,) = backends.refine_matches(m["D11"].half().contiguous(),
                                     m["D21"].view(b, h * w, -1).half().contiguous(), p1,
                                     M["radius"], M["dilation_max"])
    idx2 = front.pixel_to_lin(p1n, w)
    assert (idx == idx2).float().mean() >= 0.9995
    assert (valid[..., 0] == v2).float().mean() >= 0.9995


def test_match_radius0_and_generic_refine(front, oracle):
    """radius == 0 skips refinement (matching.py:78); fdim != 24 takes the generic kernels."""
    h, w, b = 40, 56, 1
    m, idx_init = _case(h, w, b, 12, "true")
    try:
        front.set_config({"matching": dict(M, radius=0)})
        idx, valid = front.match(m["X11"], m["X21"], None, None, idx_init)
        io, vo = oracle.match(m["X11"], m["X21"], m["D11"], m["D21"], idx_init, **dict(M, radius=0))
        assert (idx.cpu() == io).float().mean() >= 0.999 and (valid.cpu() == vo).float().mean() >= 0.999
        front.set_config({"matching": dict(M, radius=2, dilation_max=3)})
        m16 = synth.make_match_inputs(h, w, b=b, seed=12, fdim=16, device="cuda")
        idx, valid = front.match(m16["X11"], m16["X21"], m16["D11"], m16["D21"], idx_init)
        io, vo = oracle.match(m16["X11"], m16["X21"], m16["D11"], m16["D21"], idx_init,
                              **dict(M, radius=2, dilation_max=3))
        assert (idx.cpu() == io).float().mean() >= 0.999 and (valid.cpu() == vo).float().mean() >= 0.999
    finally:
        front.set_config({"matching": dict(M)})


def test_match_errors_and_empty(front):
    X = torch.rand(1, 16, 16, 3, device="cuda") + 1.0
    D = torch.rand(1, 16, 16, 24, device="cuda")
    with pytest.raises(RuntimeError):
        front.match(X.cpu(), X, D, D)
    with pytest.raises(RuntimeError):
        front.match(X.double(), X.double(), D, D)
    with pytest.raises(RuntimeError):
        front.match(X, X, D, D, torch.zeros(1, 256, dtype=torch.int32, device="cuda"))
    with pytest.raises(RuntimeError):
        front.match(X, X[:, :8], D, D)
    idx, valid = front.match(X[:0], X[:0], D[:0], D[:0])
    assert idx.shape == (0, 256) and valid.shape == (0, 256, 1)
    # identical views of a smooth surface: every interior pixel matches itself and is valid
    m = synth.make_match_inputs(32, 48, b=1, seed=3, device="cuda")
    idx, valid = front.match(m["X11"], m["X11"].clone(), m["D11"], m["D11"].clone())
    own = torch.arange(32 * 48, device="cuda").view(32, 48)
    assert valid.view(32, 48)[2:-2, 2:-2].all()
    assert torch.equal(idx.view(32, 48)[2:-2, 2:-2], own[2:-2, 2:-2])


# ------------------------------------------------------------------------- match confidences
def _confidence_reference(idx, Qs, Qd, vm, Q_conf, Cs=None, Cd=None, C_conf=0.0):
    """The reference's torch lines, verbatim in structure: tracker.py:41,59-67,104-110 and
    global_opt.py:52-68 (batched gather through batch_inds)."""
    b, n = idx.shape
    batch_inds = torch.arange(b, device=idx.device)[:, None].repeat(1, n)
    Q = torch.sqrt(Qs[batch_inds, idx] * Qd)
    valid_q = vm & (Q > Q_conf)
    valid_opt = None
    if Cs is not None:
        valid_opt = vm & (Cs[batch_inds, idx] > C_conf) & (Cd > C_conf) & (Q > Q_conf)
    uniq = [torch.unique(idx[i][vm[i, :, 0]]).shape[0] for i in range(b)]
    return Q, valid_q, valid_opt, uniq


@pytest.mark.parametrize("b,n,n_src,with_c", [(1, 196608, 196608, True), (4, 50176, 50176, False),
                                              (3, 1000, 777, True), (1, 5, 9, True)])
def test_match_confidence_bit_exact(front, b, n, n_src, with_c):
    g = torch.Generator(device="cuda").manual_seed(b * 131 + n)
    idx = torch.randint(0, n_src, (b, n), generator=g, device="cuda")
    idx[:, : n // 3] = idx[:, :1]  # many duplicates
    Qs = torch.exp(1.0 + 0.5 * torch.randn(b, n_src, 1, generator=g, device="cuda"))
    Qd = torch.exp(1.0 + 0.5 * torch.randn(b, n, 1, generator=g, device="cuda"))
    vm = torch.rand(b, n, 1, generator=g, device="cuda") < 0.8
    Cs = Cd = None
    if with_c:
        Cs = torch.randn(b, n_src, 1, generator=g, device="cuda") + 0.5
        Cd = torch.randn(b, n, 1, generator=g, device="cuda") + 0.5
    Q, vq, vo, counts = front.match_confidence(idx, Qs, Qd, vm, 1.5, Cs, Cd, 0.0, count_unique=True)
    Qr, vqr, vor, uniq = _confidence_reference(idx, Qs, Qd, vm, 1.5, Cs, Cd, 0.0)
    assert torch.equal(Q, Qr) and torch.equal(vq, vqr)  # float path is bit-identical
    assert counts[:, 0].tolist() == vqr.sum(dim=(1, 2)).tolist()
    assert counts[:, 2].tolist() == uniq
    if with_c:
        assert torch.equal(vo, vor) and counts[:, 1].tolist() == vor.sum(dim=(1, 2)).tolist()
    else:
        assert vo is None and counts[:, 1].tolist() == [0] * b
    # without the distinct count
    _, _, _, c2 = front.match_confidence(idx, Qs, Qd, vm, 1.5, Cs, Cd, 0.0)
    assert c2[:, 2].tolist() == [0] * b and torch.equal(c2[:, :2], counts[:, :2])


def test_match_confidence_tracker_usage(front):
    """b = 1 exactly as FrameTracker.track uses it, on the synthetic tracking inputs."""
    t = synth.make_track_inputs(96, 128, seed=4, device="cuda")
    T = synth.TRACK_DEFAULTS
    Qk, valid_kf, valid_opt, c = front.match_confidence(
        t["idx_f2k"][None], t["Qff"][None], t["Qkf"][None], t["valid_match_k"][None], T["Q_conf"],
        t["Cf_canon"][None], t["Ck"][None], T["C_conf"], count_unique=True)
    assert torch.equal(Qk[0], t["Qk"]) and torch.equal(valid_opt[0], t["valid_opt"])
    assert torch.equal(valid_kf[0], t["valid_match_k"] & (t["Qk"] > T["Q_conf"]))
    n = t["n"]
    assert c[0, 1].item() == int(t["valid_opt"].sum()) and c[0, 0].item() == int(valid_kf.sum())
    assert c[0, 2].item() == torch.unique(t["idx_f2k"][t["valid_match_k"][:, 0]]).shape[0]
    assert 0 < c[0, 1].item() / n < 1


def test_match_confidence_errors(front):
    idx = torch.zeros(1, 8, dtype=torch.int64, device="cuda")
    Q = torch.ones(1, 8, 1, device="cuda")
    vm = torch.ones(1, 8, 1, dtype=torch.bool, device="cuda")
    with pytest.raises(RuntimeError):
        front.match_confidence(idx.int(), Q, Q, vm, 1.5)
    with pytest.raises(RuntimeError):
        front.match_confidence(idx, Q, Q, vm, 1.5, C_src=Q)
    with pytest.raises(RuntimeError):
        front.match_confidence(idx, Q, Q[:, :4], vm, 1.5)
    # out-of-range indices do not fault and never count as matches of a distinct source point
    bad = idx.clone()
    bad[0, 0], bad[0, 1] = -5, 99
    _, _, _, c = front.match_confidence(bad, Q, Q, vm, 0.5, count_unique=True)
    torch.cuda.synchronize()
    assert c[0, 2].item() == 1


def test_random_init_network_distribution(front, backends, oracle):
    """Head outputs of the production network with seeded random weights
    (tests/golden/make_golden_mast3r.py; BASELINE.json: "random-init MASt3R weights").  The
    pointmaps are a blob around the origin and the LM iterations converge on 4 % of the
    pixels, so the bars are stated per stage, each on the reference's own intermediate:
    ray image to RAY_TOL; iter_proj on the reference's ray image bit-identical on every pixel;
    refine_matches bit-exact on the reference's projected pixels; the fused match (its own
    ray image, 1e-7 away from torch's) on >= 99.9 % of the valid pixels and of the valid
    mask, with the agreement on converged and on all pixels reported -- an iterate that does
    not converge is a chaotic orbit of the damped step that a 1-ulp input change redirects,
    and nobody reads it (matching.py:68-69 masks it); confidences bit-exact."""
    path = os.path.join(GOLD, "mast3r_rand.npz")
    if not os.path.exists(path):
        pytest.skip("no random-init network vectors committed")
    z = np.load(path)
    X11, X21 = torch.from_numpy(z["X11"]).cuda(), torch.from_numpy(z["X21"]).cuda()
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16).cuda()
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16).cuda()
    b, h, w, _ = X11.shape
    ref_rays = torch.from_numpy(z["ref_rays"])
    rays = front.ray_image_with_grad(X11)
    err = (rays.cpu() - ref_rays).abs().max().item()
    print(f"ray image max abs err {err:.3e}")
    assert err <= RAY_TOL
    # iter_proj on the reference's own ray image
    pts = torch.from_numpy(z["ref_pts"]).cuda()  # F.normalize differs by an ulp between CPU and GPU
    p0 = front.lin_to_pixel(torch.arange(h * w, device="cuda")[None].repeat(b, 1), w).float()
    p, conv = backends.iter_proj(ref_rays.cuda(), pts, p0, int(z["max_iter"]), float(z["lambda_init"]),
                                 float(z["convergence_thresh"]))
    rp, rc = torch.from_numpy(z["ref_p_proj"]), torch.from_numpy(z["ref_converged"]).bool()
    rv = torch.from_numpy(z["ref_valid"])[..., 0]
    exact = (p.cpu() == rp).all(-1).float().mean().item()
    print(f"iter_proj on the reference's ray image: bit-identical (u,v) on {exact:.6f} of all pixels "
          f"({int(rc.sum())} converged, {int(rv.sum())} valid)")
    # same inputs -> the same orbit on every pixel, converged or not (the oracle carries the
    # reference binary's FMA placement, test_matching_gpu.py::test_oracle_twin_of_reference_cuda)
    assert torch.equal(p.cpu(), rp) and torch.equal(conv.cpu().bool(), rc)
    # refine_matches: bit-exact on identical input pixels
    p1 = rp.long().cuda()
    (got,) = backends.refine_matches(D11, D21.view(b, h * w, -1), p1, int(z["radius"]), int(z["dilation_max"]))
    (want,) = oracle.refine_matches(D11.cpu(), D21.view(b, h * w, -1).cpu(), p1.cpu(), int(z["radius"]),
                                    int(z["dilation_max"]))
    assert torch.equal(got.cpu(), want)
    # the fused front end against the reference's matching.match
    idx, valid = front.match(X11, X21, D11, D21)
    ri, rv = torch.from_numpy(z["ref_idx"]), torch.from_numpy(z["ref_valid"])
    fi_conv = (idx.cpu() == ri)[rc].float().mean().item()
    fv = (valid.cpu() == rv).float().mean().item()
    print(f"fused match: idx agreement on converged {fi_conv:.5f}, on all {(idx.cpu() == ri).float().mean().item():.5f}, "
          f"valid agreement {fv:.5f}")
    fi_valid = (idx.cpu() == ri)[rv[..., 0]].float().mean().item()
    print(f"fused match: idx agreement on the {int(rv.sum())} valid pixels {fi_valid:.5f}")
    # the bar is stated on valid pixels (BASELINE.json); 482 converged pixels resolve 0.2 % steps,
    # so one redirected orbit among them is reported above, not failed
    assert fi_valid >= 0.999 and fv >= 0.999 and fi_conv >= 0.995
    # confidences of the match (tracker.py:41 form), bit-exact against torch
    Q11 = torch.from_numpy(z["Q11"]).cuda().view(b, -1, 1)
    Q21 = torch.from_numpy(z["Q21"]).cuda().view(b, -1, 1)
    Q, vq, _, counts = front.match_confidence(idx, Q11, Q21, valid, 1.5, count_unique=True)
    Qr = torch.sqrt(torch.gather(Q11[..., 0], 1, idx)[..., None] * Q21)
    assert torch.equal(Q, Qr) and torch.equal(vq, valid & (Qr > 1.5))
    assert counts[0, 2].item() == torch.unique(idx[0][valid[0, :, 0]]).numel()
