"""Loader for the reference's OWN Python callers of the hot path -- mast3r_slam/matching.py and
mast3r_slam/global_opt.py (FactorGraph), with frame.py, geometry.py, image.py, config.py --
imported UNMODIFIED from baseline/_ref (installed by oracle/install_reference_py.py; travels to
the GPU box) or, in the build container, from /root/reference.  Test infrastructure only.

What cannot be imported is replaced, nothing else:
  * `lietorch` (unpinned git dependency, not vendored) -> `Sim3` below: a tensor
    holder with the members these callers touch (.data, indexing, Identity, embedded_dim);
    none of the code under test composes or retracts poses with it (the solver does that on
    the raw [N,8] view, global_opt.py:142);
  * `mast3r_slam.mast3r_utils` (pulls the MASt3R network) -> a stub whose
    `mast3r_match_symmetric` is the reference's own post-decoder part (mast3r_utils.py:154-187,
    restated) fed from a synthetic scene instead of the decoder;
  * `mast3r_slam_backends` -> whichever implementation the test selects: this library's drop-in
    module, the reference's CUDA extension (oracle/_ref), or the CPU oracle.
"""
import os
import sys
import threading
import types

import torch

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CANDIDATES = [os.path.join(_ROOT, "baseline", "_ref"), "/root/reference"]


def reference_root():
    for c in CANDIDATES:
        if os.path.exists(os.path.join(c, "mast3r_slam", "global_opt.py")):
            return c
    return None


class Sim3:
    """Minimal lietorch.Sim3 stand-in: data [..., 8] = (t, q_xyzw, s)."""

    embedded_dim = 8

    def __init__(self, data):
        self.data = data

    @staticmethod
    def Identity(n, device="cpu", dtype=torch.float32):
        d = torch.zeros((n, 8), dtype=dtype, device=device)
        d[:, 6] = 1.0
        d[:, 7] = 1.0
        return Sim3(d)

    def __getitem__(self, k):
        return Sim3(self.data[k])


class _Value:
    def __init__(self, v):
        self.value = v


class FakeManager:
    """What SharedKeyframes asks of multiprocessing.Manager (frame.py:223-225), in-process."""

    def RLock(self):
        return threading.RLock()

    def Value(self, typecode, v):
        return _Value(v)

    def list(self):
        return []


_loaded = {}


def load(backend_module):
    """Import (once) the reference's matching / global_opt / frame modules and point their
    `mast3r_slam_backends` at `backend_module`.  Returns a namespace."""
    root = reference_root()
    if root is None:
        return None
    if not _loaded:
        lt = types.ModuleType("lietorch")
        lt.Sim3 = Sim3
        sys.modules["lietorch"] = lt
        mu = types.ModuleType("mast3r_slam.mast3r_utils")
        mu.resize_img = None
        mu.mast3r_match_symmetric = None  # set per scene, see SyntheticDecoder
        # the reference modules bind `mast3r_slam_backends` at import (matching.py:5, global_opt.py:9);
        # the name is registered only for the duration of those imports and restored afterwards, so
        # that the rest of the test session keeps seeing the real package under that name
        prev_backends = sys.modules.get("mast3r_slam_backends")
        sys.modules["mast3r_slam_backends"] = backend_module
        if root not in sys.path:
            sys.path.insert(0, root)
        import mast3r_slam  # noqa: F401

        sys.modules["mast3r_slam.mast3r_utils"] = mu
        import mast3r_slam.config as cfg

        cwd = os.getcwd()
        try:
            os.chdir(root)  # load_config resolves "config/base.yaml" relative to the cwd
            cfg.load_config("config/base.yaml")
        finally:
            os.chdir(cwd)
        import mast3r_slam.frame as frame
        import mast3r_slam.global_opt as global_opt
        import mast3r_slam.matching as matching

        if prev_backends is not None:
            sys.modules["mast3r_slam_backends"] = prev_backends
        else:
            del sys.modules["mast3r_slam_backends"]
        _loaded.update(cfg=cfg, frame=frame, global_opt=global_opt, matching=matching, mu=mu)
    ns = types.SimpleNamespace(**_loaded)
    ns.matching.mast3r_slam_backends = backend_module
    ns.global_opt.mast3r_slam_backends = backend_module
    return ns


class SyntheticDecoder:
    """Stands where the MASt3R decoder stands in mast3r_match_symmetric (mast3r_utils.py:149-187):
    produces, for a batch of keyframe pairs (i, j), the four pointmaps / descriptor maps /
    confidences the decoder would, from a ray-cast synthetic scene, then runs the reference's own
    `matching.match` on them exactly as mast3r_utils.py:162-187 does.  Keyframe k is recognised by
    feat[k, 0, 0, 0] == k (the test writes that into SharedKeyframes.feat)."""

    def __init__(self, ns, scene):
        self.ns, self.s = ns, scene

    def __call__(self, model, feat_i, pos_i, feat_j, pos_j, shape_i, shape_j):
        s = self.s
        ki = feat_i[:, 0, 0].round().long().tolist()
        kj = feat_j[:, 0, 0].round().long().tolist()
        Xii = torch.stack([s["X_self"][k] for k in ki])           # view i in frame i
        Xjj = torch.stack([s["X_self"][k] for k in kj])
        Xji = torch.stack([s["X_in"][(j, i)] for i, j in zip(ki, kj)])  # view j's points in frame i
        Xij = torch.stack([s["X_in"][(i, j)] for i, j in zip(ki, kj)])
        Dii = torch.stack([s["D"][k] for k in ki])
        Djj = torch.stack([s["D"][k] for k in kj])
        Dji = torch.stack([s["D_noisy"][(j, i)] for i, j in zip(ki, kj)])
        Dij = torch.stack([s["D_noisy"][(i, j)] for i, j in zip(ki, kj)])
        Qii = torch.stack([s["Q"][(i, i)] for i in ki])
        Qjj = torch.stack([s["Q"][(j, j)] for j in kj])
        Qji = torch.stack([s["Q"][(j, i)] for i, j in zip(ki, kj)])
        Qij = torch.stack([s["Q"][(i, j)] for i, j in zip(ki, kj)])
        b = Xii.shape[0]
        X11 = torch.cat((Xii, Xjj), dim=0)
        X21 = torch.cat((Xji, Xij), dim=0)
        D11 = torch.cat((Dii, Djj), dim=0)
        D21 = torch.cat((Dji, Dij), dim=0)
        idx_1_to_2, valid_match_2 = self.ns.matching.match(X11, X21, D11, D21)
        match_b = X11.shape[0] // 2
        return (idx_1_to_2[:match_b], idx_1_to_2[match_b:], valid_match_2[:match_b], valid_match_2[match_b:],
                Qii.view(b, -1, 1), Qjj.view(b, -1, 1), Qji.view(b, -1, 1), Qij.view(b, -1, 1))


def make_scene(N, h, w, seed, device, pairs):
    """Pointmaps, descriptors and confidences of N keyframes of the synth scene and of the ordered
    pairs in `pairs` ((a, b): view a's points expressed in frame b)."""
    import synth

    K = synth.default_K(h, w).to(device)
    T = synth.make_poses(N, seed, device)
    gen = torch.Generator().manual_seed(9000 + seed)
    f32 = torch.float32
    X_self, Pw, D = {}, {}, {}
    for k in range(N):
        X, P = synth.render_pointmap(T[k], K, h, w)
        X_self[k] = X.view(h, w, 3).to(f32)
        Pw[k] = P
        D[k] = synth.descriptor_field(P, 24, seed).view(h, w, 24).to(f32)
    X_in, D_noisy, Q = {}, {}, {}
    for (a, b) in pairs:
        Xa_in_b = synth.sim3_inv_act(T[b], Pw[a])
        nz = torch.randn(h * w, 3, generator=gen, dtype=torch.float64).to(device) * 1e-3
        X_in[(a, b)] = (Xa_in_b + nz).view(h, w, 3).to(f32)
        Dn = synth.descriptor_field(Pw[a], 24, seed)
        Dn = Dn + torch.randn(h * w, 24, generator=gen, dtype=torch.float64).to(device) * 0.05
        D_noisy[(a, b)] = (Dn / Dn.norm(dim=-1, keepdim=True)).view(h, w, 24).to(f32)
        Q[(a, b)] = torch.exp(1.0 + 0.5 * torch.randn(h, w, generator=gen)).to(device)
    for k in range(N):
        Q[(k, k)] = torch.exp(1.0 + 0.5 * torch.randn(h, w, generator=gen)).to(device)
    C = {k: (1.0 + torch.exp(torch.randn(h * w, 1, generator=gen))).to(device) for k in range(N)}
    xi = torch.randn(N, 7, generator=gen, dtype=torch.float64).to(device) * 0.02
    xi[0] = 0
    T_init = synth.sim3_perturb(T, xi).to(f32)
    return dict(N=N, h=h, w=w, K=K.to(f32), T_true=T.to(f32), T_init=T_init, X_self=X_self, X_in=X_in, D=D,
                D_noisy=D_noisy, Q=Q, C=C)


def make_keyframes(ns, scene, device):
    """A SharedKeyframes buffer (the reference's own class, frame.py:221-330) holding the scene."""
    s = scene
    N, h, w = s["N"], s["h"], s["w"]
    frames = ns.frame.SharedKeyframes(FakeManager(), h, w, buffer=N + 1, device=device)
    for k in range(N):
        frames.X[k] = s["X_self"][k].view(-1, 3)
        frames.C[k] = s["C"][k]
        frames.N[k] = 1
        frames.N_updates[k] = 1
        frames.T_WC[k] = s["T_init"][k][None]
        frames.feat[k] = float(k)
        frames.dataset_idx[k] = k
    frames.n_size.value = N
    frames.K[:] = s["K"]
    return frames
