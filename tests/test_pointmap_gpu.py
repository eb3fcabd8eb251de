"""GPU parity tests of the pointmap producers against the reference's
torch lines (mast3r_slam/frame.py:69-101, geometry.py:37-42,107-123) evaluated on the same
device.  weighted_pointmap, indep_conf and constrain_points_to_ray are bit-exact; the
spherical mode goes through atan2/acos/sin/cos and is held to 1e-5 relative."""
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pm(backends):
    from mast3r_slam_backends import pointmaps

    return pointmaps


def _obs(n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    Xc = torch.randn(n, 3, generator=g, device="cuda") + torch.tensor([0.0, 0.0, 3.0], device="cuda")
    X = Xc + 0.01 * torch.randn(n, 3, generator=g, device="cuda")
    Cc = 1.0 + torch.rand(n, 1, generator=g, device="cuda") * 5
    C = 1.0 + torch.rand(n, 1, generator=g, device="cuda") * 5
    return Xc, Cc, X, C


def _ref_weighted(Xc, Cc, X, C):  # frame.py:74-77
    return ((Cc * Xc) + (C * X)) / (Cc + C), Cc + C


def _ref_indep(Xc, Cc, X, C):  # frame.py:69-73
    Xc, Cc = Xc.clone(), Cc.clone()
    m = C > Cc
    Xc[m.repeat(1, 3)] = X[m.repeat(1, 3)]
    Cc[m] = C[m]
    return Xc, Cc


def _ref_spherical(Xc, Cc, X, C):  # frame.py:78-101
    def c2s(P):
        r = torch.linalg.norm(P, dim=-1, keepdim=True)
        x, y, z = torch.tensor_split(P, 3, dim=-1)
        return torch.cat((r, torch.atan2(y, x), torch.acos(z / r)), dim=-1)

    def s2c(S):
        r, phi, theta = torch.tensor_split(S, 3, dim=-1)
        return torch.cat((r * torch.sin(theta) * torch.cos(phi), r * torch.sin(theta) * torch.sin(phi),
                          r * torch.cos(theta)), dim=-1)

    s = ((Cc * c2s(Xc)) + (C * c2s(X))) / (Cc + C)
    return s2c(s), Cc + C


@pytest.mark.parametrize("n", [1, 1000, 196608])
def test_weighted_pointmap_and_indep_conf_bit_exact(pm, n):
    Xc, Cc, X, C = _obs(n, n)
    a, b = Xc.clone(), Cc.clone()
    pm.update_pointmap(a, b, X, C, "weighted_pointmap")
    rX, rC = _ref_weighted(Xc, Cc, X, C)
    assert torch.equal(a, rX) and torch.equal(b, rC)
    a, b = Xc.clone(), Cc.clone()
    pm.update_pointmap(a, b, X, C, "indep_conf")
    rX, rC = _ref_indep(Xc, Cc, X, C)
    assert torch.equal(a, rX) and torch.equal(b, rC)


def test_weighted_spherical(pm):
    Xc, Cc, X, C = _obs(50000, 5)
    a, b = Xc.clone(), Cc.clone()
    pm.update_pointmap(a, b, X, C, "weighted_spherical")
    rX, rC = _ref_spherical(Xc, Cc, X, C)
    assert torch.equal(b, rC)
    assert ((a - rX).abs().max() / rX.abs().max()).item() < 1e-5


def test_update_with_transport(pm, oracle):
    """X <- T.act(X) folded into the update (tracker.py:97-98) vs the oracle's Sim3 stand-in."""
    Xc, Cc, X, C = _obs(20000, 6)
    T = torch.tensor([[0.1, -0.2, 0.05, 0.02, -0.03, 0.01, 0.0, 1.1]])
    T[0, 6] = (1 - (T[0, 3:6] ** 2).sum()).sqrt()
    a, b = Xc.clone(), Cc.clone()
    pm.update_pointmap(a, b, X, C, "weighted_pointmap", T=T.cuda())
    Xt = oracle.Sim3(T).act(X.cpu())
    rX, rC = _ref_weighted(Xc.cpu(), Cc.cpu(), Xt, C.cpu())
    assert torch.equal(b.cpu(), rC)
    assert ((a.cpu() - rX).abs().max() / rX.abs().max()).item() < 1e-6


def test_constrain_points_to_ray_bit_exact(pm):
    h, w, b = 48, 64, 2
    K = synth.default_K(h, w).float().cuda()
    g = torch.Generator(device="cuda").manual_seed(1)
    Xs = torch.randn(b, h * w, 3, generator=g, device="cuda") + 2.0
    out = pm.constrain_points_to_ray((h, w), Xs, K)
    # geometry.py:37-42 + 107-123
    u, v = torch.meshgrid(torch.arange(w), torch.arange(h), indexing="xy")
    uv = torch.stack((u, v), dim=-1).unsqueeze(0).repeat(b, 1, 1, 1).to(device="cuda", dtype=Xs.dtype)
    uv = uv.view(*Xs.shape[:-1], 2)
    z = Xs[..., 2:3]
    tmp1 = (uv[..., 0] - K[0, 2]) / K[0, 0]
    tmp2 = (uv[..., 1] - K[1, 2]) / K[1, 1]
    dP = torch.empty(uv.shape[:-1] + (3, 1), device="cuda", dtype=K.dtype)
    dP[..., 0, 0], dP[..., 1, 0], dP[..., 2, 0] = tmp1, tmp2, 1.0
    ref = torch.squeeze(z[..., None, :] * dP, dim=-1)
    assert torch.equal(out, ref)


def test_packed_keyframes_and_errors(pm):
    n = 500
    store = pm.PackedKeyframes(4, n)
    Xc, Cc, X, C = _obs(n, 9)
    k = store.append(Xc, Cc)
    store.update(k, X, C)
    Xs, Cs = store.views()
    rX, rC = _ref_weighted(Xc, Cc, X, C)
    assert Xs.shape == (1, n, 3) and torch.equal(Xs[0], rX) and torch.equal(Cs[0], rC / 2)
    assert Xs.data_ptr() == store.X.data_ptr()  # a view: nothing is re-stacked
    with pytest.raises(ValueError):
        pm.update_pointmap(Xc, Cc, X, C, "best_score")
    with pytest.raises(RuntimeError):
        pm.update_pointmap(Xc.cpu(), Cc, X, C)
    with pytest.raises(RuntimeError):
        pm.update_pointmap(Xc, Cc, X[:-1].contiguous(), C)
