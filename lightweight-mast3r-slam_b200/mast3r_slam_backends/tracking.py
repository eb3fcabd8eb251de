"""Host-side mirror of the pose solve in the reference's `mast3r_slam/tracker.py`
(SURVEY.md 8f row N2).

`opt_pose_ray_dist_sim3` and `opt_pose_calib_sim3` keep the argument order and return
convention of the FrameTracker methods of the same name (tracker.py:173, :217), with poses
passed as their `lietorch.Sim3.data` tensors ([1,8]: t xyz, q xyzw, s):

    T_WCf, T_CkCf = opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid)

Each call is ONE kernel launch (m3r_track_pose, include/m3r_backends.h) that runs every
Gauss-Newton iteration on the device, where the reference issues ~40 torch kernels and two
host synchronisations per iteration.  When the 7x7 factorisation fails the reference's
`torch.linalg.cholesky` raises (caught in FrameTracker.track, tracker.py:90-92); here
`torch.linalg.LinAlgError` is raised from the status word the kernel leaves behind.

`config` mirrors `mast3r_slam.config.config["tracking"]` (config/base.yaml:16-31).
"""
import ctypes

import torch

from . import _lib

config = {
    "tracking": {
        "max_iters": 50, "C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3, "delta_norm": 1e-3,
        "huber": 1.345, "sigma_ray": 0.003, "sigma_dist": 10.0, "sigma_pixel": 1.0,
        "sigma_depth": 10.0, "pixel_border": -10, "depth_eps": 1e-6, "min_match_frac": 0.05,
        "match_frac_thresh": 0.333,
    }
}


def set_config(cfg):
    """Use the reference's loaded config dict (mast3r_slam/config.py) for the parameters."""
    config["tracking"] = dict(cfg["tracking"])


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _f32(name, t, shape_tail):
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (mast3r_slam_backends has no CPU implementation)")
    if t.dtype != torch.float32:
        raise RuntimeError(f"expected {name} to have dtype torch.float32, but got {t.dtype}")
    if tuple(t.shape[1:]) not in shape_tail:
        raise RuntimeError(f"{name}: unexpected shape {tuple(t.shape)}")
    return t.contiguous()


def _mask(name, t, n):
    if not t.is_cuda or t.dtype not in (torch.bool, torch.uint8) or t.numel() != n:
        raise RuntimeError(f"{name} must be a CUDA bool tensor with {n} elements")
    return t.contiguous()


class TrackInfo:
    """Result record of the last solve: iterations executed, final cost, status."""

    def __init__(self, status, iterations, cost):
        self.status, self.iterations, self.cost = status, iterations, cost


last_info = None
_scratch = {}


def _solve(kind, Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, idx_f2k, check):
    global last_info
    cfg = config["tracking"]
    Qk = _f32("Qk", Qk, ((1,), ()))
    n = Qk.shape[0]
    dev = Qk.device
    Xf = _f32("Xf", Xf, ((3,),))
    valid = _mask("valid", valid, n)
    T_WCf = _f32("T_WCf", T_WCf.reshape(1, 8), ((8,),))
    T_WCk = _f32("T_WCk", T_WCk.reshape(1, 8), ((8,),))
    if idx_f2k is None:
        if Xf.shape[0] != n:
            raise RuntimeError("Xf and Qk must have the same number of points")
    else:
        if not idx_f2k.is_cuda or idx_f2k.dtype != torch.int64 or idx_f2k.numel() != n:
            raise RuntimeError(f"idx_f2k must be a CUDA int64 tensor with {n} elements")
        idx_f2k = idx_f2k.contiguous()
    prm = _lib.TrackParams()
    prm.kind = kind
    prm.max_iters = int(cfg["max_iters"])
    prm.rel_error, prm.delta_norm, prm.huber_k = float(cfg["rel_error"]), float(cfg["delta_norm"]), float(cfg["huber"])
    if kind == _lib.GN_RAYS:
        K = None
        Xk = _f32("Xk", Xk, ((3,),))
        if Xk.shape[0] != n:
            raise RuntimeError("Xk and Qk must have the same number of points")
        prm.sigma_a, prm.sigma_b = float(cfg["sigma_ray"]), float(cfg["sigma_dist"])
    else:
        meas_k = _f32("meas_k", meas_k, ((3,),))
        valid_meas_k = _mask("valid_meas_k", valid_meas_k, n)
        if meas_k.shape[0] != n:
            raise RuntimeError("meas_k and Qk must have the same number of points")
        prm.sigma_a, prm.sigma_b = float(cfg["sigma_pixel"]), float(cfg["sigma_depth"])
        K = _f32("K", K.reshape(1, 9), ((9,),))
        prm.height, prm.width = int(img_size[0]), int(img_size[1])
        prm.pixel_border, prm.z_eps = float(cfg["pixel_border"]), float(cfg["depth_eps"])
        Xk = None
    key = (dev.index, torch.cuda.current_stream(dev).cuda_stream)
    if key not in _scratch:
        _scratch[key] = torch.zeros(_lib.track_scratch_doubles(), dtype=torch.float64, device=dev)
    result = torch.empty(ctypes.sizeof(_lib.TrackResult), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.track_pose(
            ctypes.byref(prm), _ptr(Xf), Xf.shape[0], _ptr(idx_f2k), _ptr(Xk), _ptr(Qk), _ptr(valid),
            _ptr(meas_k), _ptr(valid_meas_k), _ptr(K), n, _ptr(T_WCf), _ptr(T_WCk), _ptr(result),
            _ptr(_scratch[key]), _stream(),
        )
    _lib.check(rc, "track_pose")
    poses = result[:64].view(torch.float32)
    T_WCf_new, T_CkCf = poses[0:8].view(1, 8), poses[8:16].view(1, 8)
    if check:
        # the one host read of the call (the reference reads the cost every iteration)
        tail = result[64:80].cpu()
        status, iterations = tail[:8].view(torch.int32).tolist()
        last_info = TrackInfo(status, iterations, float(tail[8:16].view(torch.float64)[0]))
        if status == _lib.TRACK_CHOLESKY_FAILED:
            raise torch.linalg.LinAlgError(
                "linalg.cholesky: The factorization could not be completed because the input is not "
                f"positive-definite (tracking pose solve, iteration {iterations})")
    return T_WCf_new, T_CkCf


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, idx_f2k=None, check=True):
    """tracker.py:173-215.  Xf [n,3] (already gathered) or, with `idx_f2k` [n], the canonical
    frame pointmap that the kernel gathers itself; Xk [n,3]; Qk [n,1]; valid [n,1] bool;
    T_WCf, T_WCk [1,8].  Returns (T_WCf [1,8], T_CkCf [1,8]) on the device.  check=False skips
    the status read (no host sync; the caller inspects `status` later)."""
    return _solve(_lib.GN_RAYS, Xf, Xk, T_WCf, T_WCk, Qk, valid, None, None, None, None, idx_f2k, check)


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                        idx_f2k=None, check=True):
    """tracker.py:217-266.  meas_k [n,3] = (u, v, log z), valid_meas_k [n,1] as built by
    FrameTracker.get_points_poses (tracker.py:143-153); K [3,3]; img_size (h, w)."""
    return _solve(_lib.GN_CALIB, Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size,
                  idx_f2k, check)
