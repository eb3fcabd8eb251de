"""Drop-in replacement for the reference's native extension `mast3r_slam_backends`.

Same import name, same five exports, same positional signatures and return conventions as
the pybind11 module of the reference (mast3r_slam/backend/src/gn.cpp:116-123, declarations in
backend/include/gn.h), so `mast3r_slam/matching.py` and `mast3r_slam/global_opt.py` call it
unchanged.  Underneath, every call goes through the C ABI of libm3r_b200.so
(include/m3r_backends.h) to hand-written sm_100a kernels on torch's current CUDA stream.
There is no CPU path: CPU tensors raise, and a missing library fails the import.

Differences from the reference that a caller can observe:
  * kernels run on torch's *current* stream, not the legacy default stream;
  * the Gauss-Newton calls never block the host (the reference synchronises three times per
    iteration, gn_kernels.cu:1199-1222); `Twc` is still updated in place and `[dx]` returned;
  * dtype/device are validated (the reference only checks contiguity, gn.h:5);
  * with M3R_GN_DEVICES=G (or `set_gn_devices`) the same three Gauss-Newton calls spread the
    directed edges over G GPUs of the box from this one process (m3r_gn_multi_*): the other
    devices read the caller's tensors over NVLink peer access while they prepare their shard,
    iterate on local memory and exchange the assembled systems over peer memory; `Twc` on the
    caller's device is updated in place as before.
"""
import ctypes
import os
import threading

import torch

from . import _lib
from ._lib import GnProblem

__all__ = [
    "gauss_newton_points",
    "gauss_newton_rays",
    "gauss_newton_calib",
    "iter_proj",
    "refine_matches",
]


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check_contiguous(**tensors):
    # same message as the reference's CHECK_CONTIGUOUS (gn.h:5)
    for name, t in tensors.items():
        if not t.is_contiguous():
            raise RuntimeError(f"{name} must be contiguous")


def _check_cuda(dtype_by_name, **tensors):
    dev = None
    for name, t in tensors.items():
        if not t.is_cuda:
            raise RuntimeError(
                f"{name} must be a CUDA tensor (mast3r_slam_backends has no CPU implementation)"
            )
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"{name} must be on the same device as the other arguments")
        want = dtype_by_name.get(name)
        if want is not None and t.dtype != want:
            raise RuntimeError(f"expected {name} to have dtype {want}, but got {t.dtype}")
    return dev


_REFINE_DTYPES = {
    torch.float16: _lib.DTYPE_F16,
    torch.float32: _lib.DTYPE_F32,
    torch.float64: _lib.DTYPE_F64,
}


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh,
              _variant=0):
    """iter_proj(rays_img_with_grad[b,h,w,9], pts_3d_norm[b,n,3], p_init[b,n,2], max_iter,
    lambda_init, cost_thresh) -> [p_new[b,n,2] f32, converged[b,n] bool]   (gn.h:89-95)"""
    _check_contiguous(rays_img_with_grad=rays_img_with_grad, pts_3d_norm=pts_3d_norm, p_init=p_init)
    f32 = torch.float32
    dev = _check_cuda(
        {"rays_img_with_grad": f32, "pts_3d_norm": f32, "p_init": f32},
        rays_img_with_grad=rays_img_with_grad, pts_3d_norm=pts_3d_norm, p_init=p_init,
    )
    if rays_img_with_grad.dim() != 4 or rays_img_with_grad.shape[3] != 9:
        raise RuntimeError("rays_img_with_grad must have shape [b,h,w,9]")
    b, n = p_init.shape[0], p_init.shape[1]
    h, w = rays_img_with_grad.shape[1], rays_img_with_grad.shape[2]
    if pts_3d_norm.shape[0] != b or pts_3d_norm.shape[1] != n or rays_img_with_grad.shape[0] != b:
        raise RuntimeError("iter_proj: inconsistent batch / point counts")
    p_new = torch.empty((b, n, 2), dtype=f32, device=dev)
    converged = torch.empty((b, n), dtype=torch.bool, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.iter_proj_v(
            _ptr(rays_img_with_grad), _ptr(pts_3d_norm), _ptr(p_init), _ptr(p_new), _ptr(converged),
            b, h, w, n, int(max_iter), float(lambda_init), float(cost_thresh), int(_variant),
            _stream(),
        )
    _lib.check(rc, "iter_proj")
    return [p_new, converged]


def refine_matches(D11, D21, p1, radius, dilation_max, _variant=0):
    """refine_matches(D11[b,h,w,F], D21[b,n,F], p1[b,n,2] int64, radius, dilation_max)
    -> [p1_new[b,n,2] int64]   (gn.h:112-117)"""
    _check_contiguous(D11=D11, D21=D21, p1=p1)
    dev = _check_cuda({"p1": torch.int64}, D11=D11, D21=D21, p1=p1)
    if D11.dtype not in _REFINE_DTYPES:
        # AT_DISPATCH_FLOATING_TYPES_AND_HALF (matching_kernels.cu:103)
        raise RuntimeError(f'"refine_matches_kernel" not implemented for \'{D11.dtype}\'')
    if D21.dtype != D11.dtype:
        raise RuntimeError(f"expected D21 to have dtype {D11.dtype}, but got {D21.dtype}")
    if D11.dim() != 4 or D21.dim() != 3 or p1.dim() != 3 or p1.shape[2] != 2:
        raise RuntimeError("refine_matches: expected D11[b,h,w,F], D21[b,n,F], p1[b,n,2]")
    b, n = p1.shape[0], p1.shape[1]
    h, w, fdim = D11.shape[1], D11.shape[2], D11.shape[3]
    if D21.shape[0] != b or D21.shape[1] != n or D21.shape[2] != fdim or D11.shape[0] != b:
        raise RuntimeError("refine_matches: inconsistent shapes")
    p1_new = torch.empty((b, n, 2), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.refine_matches_v(
            _ptr(D11), _ptr(D21), _ptr(p1), _ptr(p1_new), b, h, w, n, fdim,
            _REFINE_DTYPES[D11.dtype], int(radius), int(dilation_max), int(_variant), _stream(),
        )
    _lib.check(rc, "refine_matches")
    return [p1_new]


# ------------------------------------------------------------------------------------------
# Gauss-Newton
# ------------------------------------------------------------------------------------------
class _Workspace:
    """Owns one `m3r_gn_workspace` (device scratch cached across calls)."""

    def __init__(self, device):
        self.device = device
        self.handle = ctypes.c_void_p()
        with torch.cuda.device(device):
            _lib.check(_lib.gn_workspace_create(ctypes.byref(self.handle)), "gn_workspace_create")

    def __del__(self):
        try:
            if self.handle:
                _lib.gn_workspace_destroy(self.handle)
        except Exception:
            pass


_workspaces = {}
_ws_lock = threading.Lock()


def _workspace(device):
    """One workspace per (device, host thread, CUDA stream): the scratch (partials, per-edge blocks,
    system, flags, compacted lists) is ordered only by the stream the calls are enqueued on, so two
    streams must never share it (the reference serialised everything on the legacy default
    stream).  `release_workspaces()` frees them."""
    index = device.index if device.index is not None else torch.cuda.current_device()
    key = (index, threading.get_ident(), torch.cuda.current_stream(index).cuda_stream)
    with _ws_lock:
        ws = _workspaces.get(key)
        if ws is None:
            ws = _workspaces[key] = _Workspace(device)
    return ws


def release_workspaces():
    """Free every cached Gauss-Newton workspace (device scratch grows to the largest graph seen).
    The caller must have synchronised the streams that used them."""
    global _last_multi
    with _ws_lock:
        _workspaces.clear()
        _multis.clear()
        _last_multi = None


# ---- single-process multi-GPU (include/m3r_backends.h, m3r_gn_multi_*) -------------------------
_gn_devices_override = None
_multis = {}
_last_multi = None


def set_gn_devices(devices):
    """Use several GPUs for the Gauss-Newton exports from this process.  `devices`: None (one
    GPU), an int G (the caller's device and the next G-1), or an explicit list whose first entry
    is the device the caller's tensors live on (an index may repeat: loop-back ranks on one
    GPU), or "env" (default: follow the environment variable M3R_GN_DEVICES, which takes the same
    forms, "4" or "0,1,2,3").  Anything but "env" overrides the environment variable."""
    global _gn_devices_override
    if isinstance(devices, str):
        if devices != "env":
            raise ValueError('set_gn_devices: expected None, an int, a list of device indices or "env"')
        _gn_devices_override = None
    else:
        _gn_devices_override = devices if devices is not None else ()


def _gn_device_list(dev):
    spec = _gn_devices_override
    if spec is None:
        env = os.environ.get("M3R_GN_DEVICES", "").strip()
        if not env:
            return None
        spec = [int(x) for x in env.split(",")] if "," in env else int(env)
    if spec == () or spec is None:
        return None
    index = dev.index if dev.index is not None else torch.cuda.current_device()
    if isinstance(spec, int):
        if spec <= 1:
            return None
        count = torch.cuda.device_count()
        if spec > count:
            raise RuntimeError(f"M3R_GN_DEVICES={spec} but only {count} CUDA devices are visible")
        ids = [(index + k) % count for k in range(spec)]
    else:
        ids = [int(x) for x in spec]
        if len(ids) <= 1:
            return None
        if ids[0] != index:
            raise RuntimeError(f"the first Gauss-Newton device ({ids[0]}) must be the tensors' device ({index})")
    return ids


class _Multi:
    def __init__(self, ids):
        self.ids = ids
        self.handle = ctypes.c_void_p()
        arr = (ctypes.c_int * len(ids))(*ids)
        _lib.check(_lib.gn_multi_create(ctypes.byref(self.handle), arr, len(ids)), "gn_multi_create")

    def __del__(self):
        try:
            if self.handle:
                _lib.gn_multi_destroy(self.handle)
        except Exception:
            pass


def _multi(ids):
    key = (tuple(ids), threading.get_ident(), torch.cuda.current_stream(ids[0]).cuda_stream)
    with _ws_lock:
        m = _multis.get(key)
        if m is None:
            m = _multis[key] = _Multi(ids)
    return m


def last_multi_report(Twc=None):
    """After a multi-GPU Gauss-Newton call: {"devices", "begin_ms" per rank (for ranks > 0 the
    scatter: remote reads + packing + compaction), "iterations", "replicas_bitwise_equal"}.
    Synchronises every rank; for tests and benchmarks."""
    m = _last_multi
    if m is None:
        return None
    G = len(m.ids)
    begin = (ctypes.c_float * G)()
    iters, same = ctypes.c_int(0), ctypes.c_int(-1)
    _lib.check(_lib.gn_multi_report(m.handle, _ptr(Twc) if Twc is not None else None,
                                    Twc.shape[0] if Twc is not None else 0, begin, ctypes.byref(iters),
                                    ctypes.byref(same) if Twc is not None else None), "gn_multi_report")
    return {"devices": list(m.ids), "begin_ms": list(begin), "iterations": iters.value,
            "replicas_bitwise_equal": bool(same.value) if Twc is not None else None}


def _gn_multi_call(ids, kind, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, dx, N, n, E, max_iter,
                   delta_thresh, K=None, **params):
    global _last_multi
    p = GnProblem()
    p.kind = kind
    p.num_poses, p.num_points, p.num_edges = N, n, E
    p.edge_begin, p.edge_end = 0, E
    p.Twc, p.Xs, p.Cs = Twc.data_ptr(), Xs.data_ptr(), Cs.data_ptr()
    p.ii, p.jj = ii.data_ptr(), jj.data_ptr()
    p.idx_ii2jj, p.valid_match, p.Q = idx_ii2jj.data_ptr(), valid_match.data_ptr(), Q.data_ptr()
    p.K = K.data_ptr() if K is not None else None
    for k, v in params.items():
        setattr(p, k, v)
    m = _multi(ids)
    with torch.cuda.device(ids[0]):
        rc = _lib.gn_multi_solve(m.handle, ctypes.byref(p), _ptr(dx), int(max_iter), float(delta_thresh),
                                 _stream())
    _lib.check(rc, "gn_multi_solve")
    _last_multi = m


def _gn_common_checks(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, extra=None):
    tensors = dict(Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx_ii2jj=idx_ii2jj,
                   valid_match=valid_match, Q=Q)
    if extra:
        tensors.update(extra)
    _check_contiguous(**tensors)
    f32, i64 = torch.float32, torch.int64
    dtypes = {"Twc": f32, "Xs": f32, "Cs": f32, "ii": i64, "jj": i64, "idx_ii2jj": i64,
              "valid_match": torch.bool, "Q": f32, "K": f32}
    dev = _check_cuda(dtypes, **tensors)
    N, n = Xs.shape[0], Xs.shape[1]
    E = ii.shape[0]
    if Twc.shape[0] != N or Twc.shape[-1] != 8 or Twc.numel() != N * 8:
        raise RuntimeError("Twc must have shape [num_poses, 8] matching Xs")
    if Cs.numel() != N * n or Xs.numel() != N * n * 3:
        raise RuntimeError("Xs must be [N,n,3] and Cs [N,n,1]")
    if jj.shape[0] != E or idx_ii2jj.numel() != E * n or valid_match.numel() != E * n \
            or Q.numel() != E * n:
        raise RuntimeError("edge tensors must be ii[E], jj[E], idx_ii2jj[E,n], valid_match[E,n,1], Q[E,n,1]")
    return dev, N, n, E


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray, sigma_dist,
                      C_thresh, Q_thresh, max_iter, delta_thresh):
    """gn.h:45-55.  Updates Twc in place, returns [dx[N-1,7]] of the last executed iteration."""
    dev, N, n, E = _gn_common_checks(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q)
    dx = torch.zeros((max(N - 1, 0), 7), dtype=torch.float32, device=dev)
    if N < 2 or max_iter <= 0:
        return [dx]
    ids = _gn_device_list(dev)
    if ids is not None:
        _gn_multi_call(ids, _lib.GN_RAYS, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, dx, N, n, E, max_iter,
                       delta_thresh, sigma_ray=float(sigma_ray), sigma_dist=float(sigma_dist),
                       C_thresh=float(C_thresh), Q_thresh=float(Q_thresh))
        return [dx]
    ws = _workspace(dev)
    with torch.cuda.device(dev):
        rc = _lib.gauss_newton_rays(
            ws.handle, _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj), _ptr(idx_ii2jj),
            _ptr(valid_match), _ptr(Q), _ptr(dx), N, n, E, float(sigma_ray), float(sigma_dist),
            float(C_thresh), float(Q_thresh), int(max_iter), float(delta_thresh), _stream(),
        )
    _lib.check(rc, "gauss_newton_rays")
    return [dx]


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, height, width,
                       pixel_border, z_eps, sigma_pixel, sigma_depth, C_thresh, Q_thresh,
                       max_iter, delta_thresh):
    """gn.h:74-87.  Updates Twc in place, returns [dx[N-1,7]]."""
    dev, N, n, E = _gn_common_checks(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q,
                                     extra={"K": K})
    if K.numel() != 9:
        raise RuntimeError("K must be [3,3]")
    dx = torch.zeros((max(N - 1, 0), 7), dtype=torch.float32, device=dev)
    if N < 2 or max_iter <= 0:
        return [dx]
    ids = _gn_device_list(dev)
    if ids is not None:
        _gn_multi_call(ids, _lib.GN_CALIB, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, dx, N, n, E, max_iter,
                       delta_thresh, K=K, height=int(height), width=int(width), pixel_border=int(pixel_border),
                       z_eps=float(z_eps), sigma_pixel=float(sigma_pixel), sigma_depth=float(sigma_depth),
                       C_thresh=float(C_thresh), Q_thresh=float(Q_thresh))
        return [dx]
    ws = _workspace(dev)
    with torch.cuda.device(dev):
        rc = _lib.gauss_newton_calib(
            ws.handle, _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(K), _ptr(ii), _ptr(jj),
            _ptr(idx_ii2jj), _ptr(valid_match), _ptr(Q), _ptr(dx), N, n, E, int(height),
            int(width), int(pixel_border), float(z_eps), float(sigma_pixel), float(sigma_depth),
            float(C_thresh), float(Q_thresh), int(max_iter), float(delta_thresh), _stream(),
        )
    _lib.check(rc, "gauss_newton_calib")
    return [dx]


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point, C_thresh,
                        Q_thresh, max_iter, delta_thresh):
    """gn.h:22-31.  Updates Twc in place, returns [dx[N-1,7]]."""
    dev, N, n, E = _gn_common_checks(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q)
    dx = torch.zeros((max(N - 1, 0), 7), dtype=torch.float32, device=dev)
    if N < 2 or max_iter <= 0:
        return [dx]
    ids = _gn_device_list(dev)
    if ids is not None:
        _gn_multi_call(ids, _lib.GN_POINTS, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, dx, N, n, E, max_iter,
                       delta_thresh, sigma_point=float(sigma_point), C_thresh=float(C_thresh),
                       Q_thresh=float(Q_thresh))
        return [dx]
    ws = _workspace(dev)
    with torch.cuda.device(dev):
        rc = _lib.gauss_newton_points(
            ws.handle, _ptr(Twc), _ptr(Xs), _ptr(Cs), _ptr(ii), _ptr(jj), _ptr(idx_ii2jj),
            _ptr(valid_match), _ptr(Q), _ptr(dx), N, n, E, float(sigma_point), float(C_thresh),
            float(Q_thresh), int(max_iter), float(delta_thresh), _stream(),
        )
    _lib.check(rc, "gauss_newton_points")
    return [dx]
