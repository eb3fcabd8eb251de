"""GPU parity tests of iter_proj / refine_matches through the drop-in module (C ABI ->
sm_100a kernels) against the CPU oracle, the committed golden vectors and, when
oracle/_ref is present, the reference's own CUDA kernels on the same device.

Tolerances (BASELINE.json north_star): refine_matches indices bit-exact given identical
input; iter_proj integer match indices agree on >= 99.9 % of pixels."""
import glob
import os

import numpy as np
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu
M = synth.MATCH_DEFAULTS
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _inputs(h, w, b=1, seed=0, init="identity"):
    m = synth.make_match_inputs(h, w, b=b, seed=seed, device="cuda")
    idx_init = None
    if init == "true":  # steady state: previous frame's matches
        tu = m["true_uv"].round().long()
        idx_init = (tu[..., 1].clamp(0, h - 1) * w + tu[..., 0].clamp(0, w - 1))
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"], idx_init)
    if init == "random":  # arbitrary initial guesses: exercises the unstaged fallback
        g = torch.Generator(device="cuda").manual_seed(seed)
        p0 = torch.rand(p0.shape, generator=g, device="cuda") * torch.tensor([w + 20.0, h + 20.0], device="cuda") - 10
    return m, rays, pts, p0.contiguous()


def _agree(p_a, p_b):
    return (p_a.long() == p_b.long()).all(-1).float().mean().item()


@pytest.mark.parametrize("h,w,b,seed,init", [(48, 64, 1, 0, "identity"), (224, 224, 1, 1, "identity"),
                                              (96, 160, 3, 2, "true"), (60, 52, 2, 3, "random"),
                                              (384, 512, 1, 4, "identity"), (384, 512, 1, 5, "true")])
def test_iter_proj_vs_oracle(backends, oracle, h, w, b, seed, init):
    m, rays, pts, p0 = _inputs(h, w, b, seed, init)
    p, conv = backends.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    po, co = oracle.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    assert p.dtype == torch.float32 and conv.dtype == torch.bool and p.shape == (b, h * w, 2)
    frac = _agree(p.cpu(), po)
    print(f"iter_proj {h}x{w} b={b} init={init}: index agreement {frac:.6f}, "
          f"max |dp| {(p.cpu() - po).abs().max().item():.3e}, converged agreement "
          f"{(conv.cpu() == co).float().mean().item():.6f}")
    assert frac >= 0.999
    assert (conv.cpu() == co).float().mean() >= 0.999


@pytest.mark.parametrize("init", ["identity", "random"])
def test_iter_proj_variants_bit_identical(backends, init):
    """The memoised kernel and the re-evaluating kernel compute the same bits."""
    m, rays, pts, p0 = _inputs(120, 168, 2, 7, init)
    a = backends.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, _variant=1)
    c = backends.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6, _variant=0)
    assert torch.equal(a[0], c[0]) and torch.equal(a[1], c[1])


def test_iter_proj_ragged_and_empty(backends, oracle):
    # n not a multiple of the block size, and n != h*w (a subset of query points)
    m, rays, pts, p0 = _inputs(40, 56, 1, 8)
    sub = 1000 + 37
    p, conv = backends.iter_proj(rays, pts[:, :sub].contiguous(), p0[:, :sub].contiguous(), 10, 1e-8, 1e-6)
    po, co = oracle.iter_proj(rays, pts[:, :sub], p0[:, :sub], 10, 1e-8, 1e-6)
    assert _agree(p.cpu(), po) >= 0.999
    p, conv = backends.iter_proj(rays, pts[:, :0].contiguous(), p0[:, :0].contiguous(), 10, 1e-8, 1e-6)
    assert p.shape == (1, 0, 2) and conv.shape == (1, 0)
    p, conv = backends.iter_proj(rays, pts, p0, 0, 1e-8, 1e-6)  # zero iterations: clamp only
    assert not conv.any() and torch.equal(p[..., 0], p0[..., 0].clamp(1, 54))


def test_iter_proj_identity_fullsize(backends):
    """Size-independent property at BASELINE size: matching a pointmap against itself from
    the identity init returns the identity, all converged (interior pixels)."""
    m, rays, pts, p0 = _inputs(384, 512, 1, 9)
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X11"])
    p, conv = backends.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    inner = (p0[..., 0] >= 1) & (p0[..., 0] <= 510) & (p0[..., 1] >= 1) & (p0[..., 1] <= 382)
    assert (p - p0).abs()[inner].max() < 1e-3 and conv[inner].all()


def test_iter_proj_errors(backends):
    m, rays, pts, p0 = _inputs(16, 16)
    with pytest.raises(RuntimeError, match="must be contiguous"):
        backends.iter_proj(rays.permute(0, 2, 1, 3), pts, p0, 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError):
        backends.iter_proj(rays.double(), pts, p0, 10, 1e-8, 1e-6)


# ------------------------------------------------------------------------------ refine
def _p1(h, w, b, n, seed, spread=0):
    g = torch.Generator().manual_seed(seed)
    u = torch.randint(-spread, w + spread, (b, n), generator=g)
    v = torch.randint(-spread, h + spread, (b, n), generator=g)
    return torch.stack([u, v], -1).cuda()


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32, torch.float64])
@pytest.mark.parametrize("h,w,b,fdim,radius,dil", [(48, 64, 2, 24, 3, 5), (40, 40, 1, 16, 2, 3),
                                                     (33, 47, 1, 7, 1, 2), (64, 64, 1, 24, 4, 1)])
def test_refine_bit_exact_vs_oracle(backends, oracle, dtype, h, w, b, fdim, radius, dil):
    m = synth.make_match_inputs(h, w, b=b, seed=11, fdim=fdim, device="cuda")
    D11 = m["D11"].to(dtype).contiguous()
    D21 = m["D21"].view(b, h * w, fdim).to(dtype).contiguous()
    p1 = _p1(h, w, b, h * w, 1, spread=6)  # includes starts outside the image
    (out,) = backends.refine_matches(D11, D21, p1, radius, dil)
    (ref,) = oracle.refine_matches(D11, D21, p1, radius, dil)
    assert out.dtype == torch.int64 and torch.equal(out.cpu(), ref)


@pytest.mark.parametrize("h,w,b,kind", [(96, 128, 1, "random"), (96, 128, 2, "smooth"), (77, 101, 1, "smooth"),
                                        (384, 512, 1, "smooth"), (50, 70, 1, "jumpy"), (16, 16, 1, "smooth"),
                                        (200, 328, 2, "collapsed"), (384, 512, 1, "collapsed")])
def test_refine_kernels_agree(backends, oracle, h, w, b, kind):
    """Tiled (TMA-staged shared-memory box; planned heavy-first order by default, variant 3 = plain
    grid of tiles), direct-pair and generic kernels: same bits.
    `smooth` keeps every window inside the staged box (unrolled shared-memory path, image
    borders masked); `jumpy` and `random` push windows out of the box (checked global path)."""
    m = synth.make_match_inputs(h, w, b=b, seed=12, device="cuda")
    D11, D21 = m["D11"].half(), m["D21"].view(b, -1, 24).half()
    if kind == "random":
        p1 = _p1(h, w, b, h * w, 2, spread=4)
    else:
        p1 = m["true_uv"].round().long().clone()
        g = torch.Generator().manual_seed(5)
        p1 += torch.randint(-2, 3, p1.shape, generator=g).cuda()
        if kind == "jumpy":  # a tenth of the matches are far from their neighbours'
            sel = (torch.rand(p1.shape[:2], generator=g) < 0.1).cuda()
            p1[sel] += torch.tensor([23, -19], device="cuda")
        if kind == "collapsed":  # the right part of view 2 sees nothing of view 1: iter_proj's clamp
            p1[..., 0] = (p1[..., 0] + w // 3).clamp(1, w - 2)  # puts those matches on the border column
            p1[..., 1] = p1[..., 1].clamp(1, h - 2)  # ("heavy" tiles of the plan pass, taken first)
    p1 = p1.contiguous()
    outs = [backends.refine_matches(D11, D21, p1, 3, 5, _variant=v)[0] for v in (0, 1, 2, 3)]
    assert all(torch.equal(outs[0], o) for o in outs[1:])
    if h * w <= 128 * 128:
        (ref,) = oracle.refine_matches(D11, D21, p1, 3, 5)
        assert torch.equal(outs[0].cpu(), ref)
    for dil in (1, 2, 7, 9):  # 9: windows wider than the staged box
        a = backends.refine_matches(D11, D21, p1, 3, dil, _variant=0)[0]
        c = backends.refine_matches(D11, D21, p1, 3, dil, _variant=1)[0]
        assert torch.equal(a, c), f"dilation_max={dil}"


def test_refine_many_tiles(backends):
    """More tiles than the plan pass orders: the unplanned grid of tiles; and tile counts that end
    inside a thread's group of 16 flags."""
    for h, w, b in [(640, 656, 3), (16, 272, 1), (48, 528, 1), (384, 512, 3)]:  # 4920, 17, 99 and 2304 tiles
        g = torch.Generator().manual_seed(h + w)
        D11 = torch.randn(b, h, w, 24, generator=g).half().cuda()
        D21 = torch.randn(b, h * w, 24, generator=g).half().cuda()
        vv, uu = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
        p1 = torch.stack([uu, vv], -1).view(1, h * w, 2).repeat(b, 1, 1)
        p1 = (p1 + torch.randint(-3, 4, p1.shape, generator=g)).cuda().contiguous()
        a = backends.refine_matches(D11, D21, p1, 3, 5, _variant=0)[0]
        c = backends.refine_matches(D11, D21, p1, 3, 5, _variant=2)[0]
        assert torch.equal(a, c), (h, w, b)


def test_refine_edge_cases(backends, oracle):
    m = synth.make_match_inputs(24, 32, seed=13, device="cuda")
    D11, D21 = m["D11"].half(), m["D21"].view(1, -1, 24).half()
    p1 = _p1(24, 32, 1, 24 * 32, 3)
    for radius, dil in [(0, 5), (3, 0), (0, 0)]:
        (out,) = backends.refine_matches(D11, D21, p1, radius, dil)
        (ref,) = oracle.refine_matches(D11, D21, p1, radius, dil)
        assert torch.equal(out.cpu(), ref)
    far = p1.clone()
    far[0, :10] = torch.tensor([10 ** 12, -10 ** 12], device="cuda")  # int64 range, never inside
    (out,) = backends.refine_matches(D11, D21, far, 3, 5)
    (ref,) = oracle.refine_matches(D11, D21, far, 3, 5)
    assert torch.equal(out.cpu(), ref) and torch.equal(out[0, :10], far[0, :10])
    # non-positive scores never win (max_score starts at +0 for half)
    (out,) = backends.refine_matches(torch.ones_like(D11), -torch.ones_like(D21), p1, 3, 5)
    assert torch.equal(out, p1)
    (out,) = backends.refine_matches(D11, D21[:, :0].contiguous(), p1[:, :0].contiguous(), 3, 5)
    assert out.shape == (1, 0, 2)
    with pytest.raises(RuntimeError):
        backends.refine_matches(D11, D21, p1.int(), 3, 5)


def test_match_pipeline_fullsize(backends, oracle):
    """BASELINE config 2 end to end at 512x384 exactly as matching.py:60-85 issues it."""
    h, w = 384, 512
    m, rays, pts, p0 = _inputs(h, w, 1, 21, "identity")
    p, conv = backends.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    p1 = p.long()
    D11, D21 = m["D11"].half(), m["D21"].view(1, h * w, -1).half()
    (p1n,) = backends.refine_matches(D11, D21, p1, M["radius"], M["dilation_max"])
    (ref,) = oracle.refine_matches(D11, D21, p1, M["radius"], M["dilation_max"])
    assert torch.equal(p1n.cpu(), ref)  # bit-exact at full size, given identical iter_proj output
    reach = M["radius"] * sum(range(1, M["dilation_max"] + 1))
    assert (p1n - p1).abs().max() <= reach
    assert (p1n[..., 0] >= 0).all() and (p1n[..., 0] < w).all() and (p1n[..., 1] >= 0).all() and (p1n[..., 1] < h).all()
    tu = m["true_uv"].round().long()
    ok = conv & (tu[..., 0] > 20) & (tu[..., 0] < w - 20) & (tu[..., 1] > 20) & (tu[..., 1] < h - 20)
    hit = ((p1n - tu).abs().max(-1).values <= 1)[ok].float().mean().item()
    print(f"refined matches within 1 px of the true projection: {hit:.4f}")
    assert hit > 0.9


# ------------------------------------------------------------------------------ reference CUDA
@pytest.mark.parametrize("h,w,b,seed,init", [(224, 224, 1, 0, "identity"), (384, 512, 1, 1, "identity"),
                                              (384, 512, 2, 2, "true"), (72, 88, 2, 3, "random")])
def test_vs_reference_cuda(backends, refcuda, h, w, b, seed, init):
    if refcuda is None:
        pytest.skip("oracle/_ref/mast3r_slam_backends_ref.so not built")
    if (h * w) % 16:
        pytest.skip("the reference kernels have no bounds guard (h*w must be a multiple of 16)")
    m, rays, pts, p0 = _inputs(h, w, b, seed, init)
    p, conv = backends.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    rp, rconv = refcuda.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    torch.cuda.synchronize()
    frac = _agree(p, rp)
    exact = (p == rp).all(-1).float().mean().item()
    print(f"vs reference CUDA {h}x{w} b={b} init={init}: index agreement {frac:.6f}, "
          f"bit-identical (u,v) {exact:.6f}, converged agreement {(conv == rconv).float().mean().item():.6f}")
    assert frac >= 0.999 and (conv == rconv).float().mean() >= 0.999
    rp1 = rp.long()
    for dt in (torch.float16, torch.float32):
        D11 = m["D11"].to(dt).contiguous()
        D21 = m["D21"].view(b, h * w, -1).to(dt).contiguous()
        (mine,) = backends.refine_matches(D11, D21, rp1, M["radius"], M["dilation_max"])
        (ref,) = refcuda.refine_matches(D11, D21, rp1, M["radius"], M["dilation_max"])
        torch.cuda.synchronize()
        assert torch.equal(mine, ref), f"refine_matches {dt} not bit-exact vs reference CUDA"


def _chaotic_inputs(kind):
    """iter_proj inputs on which most LM runs do not converge, so that every rounding of the
    arithmetic decides which orbit a pixel follows."""
    if kind == "network":
        z = np.load(os.path.join(GOLD, "mast3r_rand.npz"))
        rays = torch.from_numpy(z["ref_rays"])
        b, h, w, _ = rays.shape
        pts = torch.from_numpy(z["ref_pts"])
    else:  # white noise ray image with its own finite-difference "gradients"
        g = torch.Generator().manual_seed(11)
        b, h, w = 2, 64, 80
        rays = torch.randn(b, h, w, 9, generator=g)
        rays[..., :3] = torch.nn.functional.normalize(rays[..., :3], dim=-1)
        pts = torch.nn.functional.normalize(torch.randn(b, h * w, 3, generator=g), dim=-1)
    idx = torch.arange(h * w)[None].repeat(b, 1)
    p0 = torch.stack((idx % w, idx // w), -1).float()
    return rays.contiguous(), pts.contiguous(), p0.contiguous()


@pytest.mark.parametrize("kind", ["network", "noise"])
def test_oracle_twin_of_reference_cuda(backends, oracle, refcuda, kind):
    """The CPU oracle, this build's kernel and the reference's own CUDA kernel follow the same
    orbit on EVERY pixel, converged or not: (u,v) bit-identical.  This is what pins the
    oracle's explicit FMA placement (oracle/m3r_oracle.c, iter_proj_point) to the reference
    binary; the smooth golden scenes cannot see it."""
    if refcuda is None:
        pytest.skip("oracle/_ref/mast3r_slam_backends_ref.so not built")
    if kind == "network" and not os.path.exists(os.path.join(GOLD, "mast3r_rand.npz")):
        pytest.skip("no random-init network vectors committed")
    rays, pts, p0 = _chaotic_inputs(kind)
    args = (M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    op, oc = oracle.iter_proj(rays, pts, p0, *args)
    rp, rc = refcuda.iter_proj(rays.cuda(), pts.cuda(), p0.cuda(), *args)
    kp, kc = backends.iter_proj(rays.cuda(), pts.cuda(), p0.cuda(), *args)
    torch.cuda.synchronize()
    o_r = (op == rp.cpu()).all(-1).float().mean().item()
    k_r = (kp == rp).all(-1).float().mean().item()
    print(f"{kind}: converged {rc.float().mean().item():.3f}; bit-identical (u,v): oracle vs reference CUDA {o_r:.6f}, "
          f"this kernel vs reference CUDA {k_r:.6f}")
    assert torch.equal(op, rp.cpu()) and torch.equal(oc.bool(), rc.cpu().bool())
    assert torch.equal(kp, rp) and torch.equal(kc.bool(), rc.bool())


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "match_*.npz"))) or [None])
def test_golden_matching_gpu(backends, path):
    if path is None:
        pytest.skip("no golden matching vectors committed yet")
    z = np.load(path)
    rays, pts, p0 = (torch.from_numpy(z[k]).cuda() for k in ("rays", "pts", "p_init"))
    p, conv = backends.iter_proj(rays, pts, p0, int(z["max_iter"]), float(z["lambda_init"]), float(z["cost_thresh"]))
    assert _agree(p.cpu(), torch.from_numpy(z["ref_p_new"])) >= 0.999
    D11 = torch.from_numpy(z["D11_f16"]).view(torch.float16).cuda()
    D21 = torch.from_numpy(z["D21_f16"]).view(torch.float16).cuda()
    p1 = torch.from_numpy(z["ref_p1"]).cuda()
    (out,) = backends.refine_matches(D11, D21, p1, int(z["radius"]), int(z["dilation_max"]))
    assert torch.equal(out.cpu(), torch.from_numpy(z["ref_p1_new"]))
