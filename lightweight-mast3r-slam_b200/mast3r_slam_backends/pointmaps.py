"""Host-side mirror of the pointmap producers around the Gauss-Newton:
`Frame.update_pointmap` (mast3r_slam/frame.py:41-102), `constrain_points_to_ray`
(mast3r_slam/geometry.py:37-42) and the keyframe stacking of `FactorGraph.get_poses_points`
(mast3r_slam/global_opt.py:114-121).  No CPU path.
"""
import ctypes

import torch

from . import _lib


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32c(name, t):
    if not t.is_cuda or t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be a CUDA float32 tensor (mast3r_slam_backends has no CPU implementation)")
    if not t.is_contiguous():
        raise RuntimeError(f"{name} must be contiguous")
    return t


def update_pointmap(X_canon, C_canon, X, C, filtering_mode="weighted_pointmap", T=None):
    """In-place fusion of a new observation (X [n,3], C [n,1]) into the running canonical
    pointmap (X_canon [n,3], C_canon [n,1] = running confidence sum), frame.py:69-101, for
    filtering_mode in {weighted_pointmap, weighted_spherical, indep_conf}.  With T ([1,8]
    Sim3 data on the device, e.g. the T_CkCf the tracking solve returned) the observation is
    first transported, X <- T.act(X) (tracker.py:97-98), in the same pass.
    The caller keeps the counters (`N`, `N_updates`) exactly as frame.py does."""
    if filtering_mode not in _lib.FILTER_MODES:
        raise ValueError(f"update_pointmap: filtering_mode {filtering_mode!r} has no device kernel "
                         "(first / recent / best_score are plain copies, frame.py:54-68)")
    Xc, Cc, X, C = _f32c("X_canon", X_canon), _f32c("C_canon", C_canon), _f32c("X", X), _f32c("C", C)
    n = Cc.numel()
    if Xc.numel() != 3 * n or X.numel() != 3 * n or C.numel() != n:
        raise RuntimeError("update_pointmap: X_canon/X must be [n,3] and C_canon/C [n,1]")
    if T is not None:
        T = _f32c("T", T)
        if T.numel() != 8:
            raise RuntimeError("update_pointmap: T must hold one Sim3 (8 floats)")
    with torch.cuda.device(Xc.device):
        rc = _lib.update_pointmap(_lib.FILTER_MODES[filtering_mode], _ptr(Xc), _ptr(Cc), _ptr(X), _ptr(C),
                                  _ptr(T), n, _stream())
    _lib.check(rc, "update_pointmap")
    return X_canon, C_canon


def constrain_points_to_ray(img_size, Xs, K):
    """geometry.py:37-42: Xs [b,h*w,3] (or [b,h,w,3]) -> points snapped onto their pixel rays."""
    Xs, K = _f32c("Xs", Xs), _f32c("K", K)
    h, w = int(img_size[0]), int(img_size[1])
    if Xs.shape[-1] != 3 or Xs.numel() % (3 * h * w) != 0 or K.numel() != 9:
        raise RuntimeError("constrain_points_to_ray: expected Xs [b,h*w,3] and K [3,3]")
    b = Xs.numel() // (3 * h * w)
    out = torch.empty_like(Xs)
    with torch.cuda.device(Xs.device):
        rc = _lib.constrain_points_to_ray(_ptr(Xs), _ptr(out), _ptr(K), b, h, w, _stream())
    _lib.check(rc, "constrain_points_to_ray")
    return out


class PackedKeyframes:
    """Persistent packed store of the keyframes' canonical pointmaps and confidences.

    `FactorGraph.get_poses_points` re-stacks every keyframe on every solve
    (global_opt.py:114-121: torch.stack over all unique keyframes, 3.1 MB x N copied per call
    at 512x384).  Here the rows live in two preallocated device buffers, updates go in place
    through `update_pointmap`, and `views(N)` hands the Gauss-Newton contiguous [N,n,3] /
    [N,n,1] views without copying anything."""

    def __init__(self, capacity, num_points, device="cuda"):
        self.capacity, self.n = int(capacity), int(num_points)
        self.X = torch.zeros((self.capacity, self.n, 3), dtype=torch.float32, device=device)
        self.C = torch.zeros((self.capacity, self.n, 1), dtype=torch.float32, device=device)
        self.N = torch.zeros(self.capacity, dtype=torch.int32)  # per-row update counts (host)
        self.size = 0

    def append(self, X, C):
        if self.size == self.capacity:
            raise RuntimeError("PackedKeyframes: capacity exhausted")
        k = self.size
        self.X[k].copy_(X.reshape(self.n, 3))
        self.C[k].copy_(C.reshape(self.n, 1))
        self.N[k] = 1
        self.size += 1
        return k

    def update(self, k, X, C, filtering_mode="weighted_pointmap", T=None):
        if not 0 <= k < self.size:
            raise IndexError(k)
        update_pointmap(self.X[k], self.C[k], X.reshape(self.n, 3).contiguous(),
                        C.reshape(self.n, 1).contiguous(), filtering_mode, T)
        if filtering_mode == "indep_conf":
            self.N[k] = 1  # frame.py:73
        else:
            self.N[k] += 1  # frame.py:77,101

    def views(self, N=None):
        """(Xs [N,n,3], Cs [N,n,1] average confidences as Frame.get_average_conf, frame.py:106-107)."""
        N = self.size if N is None else N
        counts = self.N[:N].to(self.C.device, torch.float32).view(N, 1, 1)
        return self.X[:N], self.C[:N] / counts
