"""Host-side mirror of the reference's `mast3r_slam/matching.py`.

Same function names, arguments and return conventions as the reference module, so its callers
(`tracker.py`, `global_opt.py`, `main.py` relocalisation) can import this one instead:

    match(X11, X21, D11, D21, idx_1_to_2_init=None) -> (idx_1_to_2 [b,n] i64, valid [b,n,1] bool)

The difference is underneath: where the reference runs ~25 torch kernels around its two
extension calls (normalise, permute, pad, two depthwise convolutions, concat, arange/repeat,
fancy-index gather, norm, compare, casts, pixel_to_lin), `match` here is one C-ABI call,
`m3r_match` (include/m3r_backends.h), that launches three kernels and allocates two scratch
buffers.  There is no CPU path.

`config` mirrors `mast3r_slam.config.config["matching"]` (config/base.yaml:8-14); a host that
already loaded the reference's config can pass its dict through `set_config`.
"""
import ctypes

import torch

from . import _lib

config = {
    "matching": {
        "max_iter": 10,
        "lambda_init": 1e-8,
        "convergence_thresh": 1e-6,
        "dist_thresh": 1e-1,
        "radius": 3,
        "dilation_max": 5,
    }
}


def set_config(cfg):
    """Use the reference's loaded config dict (mast3r_slam/config.py) for the thresholds."""
    config["matching"] = dict(cfg["matching"])


_DTYPES = {torch.float16: _lib.DTYPE_F16, torch.float32: _lib.DTYPE_F32, torch.float64: _lib.DTYPE_F64}


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda_f32(**tensors):
    for name, t in tensors.items():
        if not t.is_cuda:
            raise RuntimeError(f"{name} must be a CUDA tensor (mast3r_slam_backends has no CPU implementation)")
        if t.dtype != torch.float32:
            raise RuntimeError(f"expected {name} to have dtype torch.float32, but got {t.dtype}")


def pixel_to_lin(p1, w):
    """matching.py:13-15"""
    return p1[..., 0] + (w * p1[..., 1])


def lin_to_pixel(idx_1_to_2, w):
    """matching.py:18-22"""
    return torch.stack((idx_1_to_2 % w, idx_1_to_2 // w), dim=-1)


def ray_image_with_grad(X11):
    """The ray-image half of prep_for_iter_proj (matching.py:29-36 + image.py:5-38):
    X11 [b,h,w,3] f32 -> rays_with_grad_img [b,h,w,9] f32, one kernel."""
    _need_cuda_f32(X11=X11)
    if X11.dim() != 4 or X11.shape[3] != 3:
        raise RuntimeError("X11 must have shape [b,h,w,3]")
    X11 = X11.contiguous()
    b, h, w, _ = X11.shape
    rays = torch.empty((b, h, w, 9), dtype=torch.float32, device=X11.device)
    with torch.cuda.device(X11.device):
        rc = _lib.prep_rays(_ptr(X11), _ptr(rays), b, h, w, _stream())
    _lib.check(rc, "prep_rays")
    return rays


def prep_for_iter_proj(X11, X21, idx_1_to_2_init):
    """matching.py:25-49 -> (rays_with_grad_img, pts3d_norm, p_init).  Kept for callers that
    drive iter_proj themselves; `match` does not go through it."""
    b, h, w, _ = X11.shape
    rays_with_grad_img = ray_image_with_grad(X11)
    pts3d_norm = torch.nn.functional.normalize(X21.view(b, -1, 3), dim=-1)
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = lin_to_pixel(idx_1_to_2_init, w).float()
    return rays_with_grad_img, pts3d_norm, p_init


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None):
    """matching.py:52-88, fused."""
    cfg = config["matching"]
    _need_cuda_f32(X11=X11, X21=X21)
    if X11.dim() != 4 or X11.shape[3] != 3 or X21.shape != X11.shape:
        raise RuntimeError("match: expected X11 and X21 of shape [b,h,w,3]")
    dev = X11.device
    b, h, w = X21.shape[:3]
    n = h * w
    if b == 0:
        return (torch.empty((0, n), dtype=torch.int64, device=dev),
                torch.empty((0, n, 1), dtype=torch.bool, device=dev))
    X11c, X21c = X11.contiguous(), X21.contiguous()
    radius = int(cfg["radius"])
    D11h = D21h = None
    fdim, dt = 0, _lib.DTYPE_F16
    if radius > 0:
        # the reference refines in half (matching.py:79-80); .half() is a no-op when it already is
        D11h = D11.half().contiguous()
        D21h = D21.reshape(b, n, -1).half().contiguous()
        if D11h.device != dev or D21h.device != dev:
            raise RuntimeError("match: descriptors must be on the same device as the pointmaps")
        if D11h.dim() != 4 or tuple(D11h.shape[:3]) != (b, h, w) or D21h.shape[2] != D11h.shape[3]:
            raise RuntimeError("match: expected D11 [b,h,w,F] and D21 [b,h,w,F]")
        fdim = D11h.shape[3]
    idx0 = None
    if idx_1_to_2_init is not None:
        idx0 = idx_1_to_2_init
        if not idx0.is_cuda or idx0.dtype != torch.int64 or tuple(idx0.shape) != (b, n):
            raise RuntimeError("match: idx_1_to_2_init must be a CUDA int64 tensor of shape [b,h*w]")
        idx0 = idx0.contiguous()
    idx_1_to_2 = torch.empty((b, n), dtype=torch.int64, device=dev)
    valid = torch.empty((b, n), dtype=torch.bool, device=dev)
    rays_ws = torch.empty((b, h, w, 12), dtype=torch.float32, device=dev)
    p1_ws = torch.empty((2, b, n, 2), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.match(
            _ptr(X11c), _ptr(X21c), _ptr(D11h), _ptr(D21h), fdim, dt, _ptr(idx0), _ptr(idx_1_to_2),
            _ptr(valid), _ptr(rays_ws), _ptr(p1_ws), b, h, w, int(cfg["max_iter"]),
            float(cfg["lambda_init"]), float(cfg["convergence_thresh"]), float(cfg["dist_thresh"]),
            radius, int(cfg["dilation_max"]), _stream(),
        )
    _lib.check(rc, "match")
    return idx_1_to_2, valid.unsqueeze(-1)


def match(X11, X21, D11, D21, idx_1_to_2_init=None):
    """matching.py:8-10"""
    return match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init)


def match_confidence(idx, Q_src, Q_dst, valid_match, Q_conf, C_src=None, C_dst=None, C_conf=0.0,
                     count_unique=False):
    """The gathers and reductions that follow a match, one kernel.

    idx [b,n] i64; Q_src [b,n_src,(1)], Q_dst [b,n,(1)] f32; valid_match [b,n,(1)] bool;
    optional averaged confidences C_src [b,n_src,(1)], C_dst [b,n,(1)].
    Returns (Q [b,n,1], valid_q [b,n,1] bool, valid_opt [b,n,1] bool or None, counts [b,3] i32)
    with counts = (sum valid_q, sum valid_opt, number of distinct idx among valid_match).

    tracker.py:41,59-67,104-110 (b = 1):
        Qk, valid_kf, valid_opt, c = match_confidence(idx_f2k[None], Qff[None], Qkf[None], ...)
        match_frac = c[0,1] / n;  match_frac_k = c[0,0] / n;  unique_frac_f = c[0,2] / n
    global_opt.py:55-68:
        Qj, valid_j, _, c = match_confidence(idx_i2j, Qii, Qji, valid_match_j, Q_conf)
        match_frac_j = c[:,0] / n
    """
    if not idx.is_cuda or idx.dtype != torch.int64 or idx.dim() != 2:
        raise RuntimeError("match_confidence: idx must be a CUDA int64 tensor of shape [b,n]")
    b, n = idx.shape
    dev = idx.device
    Qs, Qd = Q_src.reshape(b, -1), Q_dst.reshape(b, -1)
    _need_cuda_f32(Q_src=Qs, Q_dst=Qd)
    n_src = Qs.shape[1]
    vm = valid_match.reshape(b, -1)
    if not vm.is_cuda or vm.dtype not in (torch.bool, torch.uint8) or vm.shape[1] != n or Qd.shape[1] != n:
        raise RuntimeError("match_confidence: valid_match / Q_dst must be CUDA tensors with n entries per row")
    Cs = Cd = None
    if C_src is not None or C_dst is not None:
        if C_src is None or C_dst is None:
            raise RuntimeError("match_confidence: C_src and C_dst go together")
        Cs, Cd = C_src.reshape(b, -1).contiguous(), C_dst.reshape(b, -1).contiguous()
        _need_cuda_f32(C_src=Cs, C_dst=Cd)
        if Cs.shape[1] != n_src or Cd.shape[1] != n:
            raise RuntimeError("match_confidence: C_src / C_dst have the wrong number of points")
    idx, Qs, Qd, vm = idx.contiguous(), Qs.contiguous(), Qd.contiguous(), vm.contiguous()
    Q = torch.empty((b, n, 1), dtype=torch.float32, device=dev)
    valid_q = torch.empty((b, n, 1), dtype=torch.bool, device=dev)
    valid_opt = torch.empty((b, n, 1), dtype=torch.bool, device=dev) if Cs is not None else None
    counts = torch.empty((b, 3), dtype=torch.int32, device=dev)
    scratch = None
    if count_unique and b > 0:
        scratch = torch.empty(max(_lib.match_confidence_scratch_bytes(b, n_src), 4), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = _lib.match_confidence(
            _ptr(idx), _ptr(Qs), _ptr(Qd), _ptr(vm), _ptr(Cs), _ptr(Cd), float(Q_conf), float(C_conf),
            _ptr(Q), _ptr(valid_q), _ptr(valid_opt), _ptr(counts), _ptr(scratch), b, n, n_src, _stream(),
        )
    _lib.check(rc, "match_confidence")
    return Q, valid_q, valid_opt, counts
