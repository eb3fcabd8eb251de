"""ctypes wrapper of the CPU oracle (oracle/libm3r_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this.  Functions take and return CPU torch tensors with the reference's layouts.
"""
import ctypes
import os
import subprocess

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libm3r_oracle.so")


def build(force=False):
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(
            os.path.join(_HERE, "m3r_oracle.c")):
        subprocess.check_call(["make", "-C", _HERE, "-B" if force else "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.m3r_oracle_h2f.restype = ctypes.c_float
        _lib.m3r_oracle_h2f.argtypes = [ctypes.c_uint16]
        _lib.m3r_oracle_f2h.restype = ctypes.c_uint16
        _lib.m3r_oracle_f2h.argtypes = [ctypes.c_float]
        _lib.m3r_oracle_num_threads.restype = ctypes.c_int
        _lib.m3r_oracle_gauss_newton.restype = ctypes.c_int
        _lib.m3r_oracle_gn_solve.restype = ctypes.c_int
        _lib.m3r_oracle_edge_indices.restype = ctypes.c_int
    return _lib


class GnParams(ctypes.Structure):
    _fields_ = [
        ("kind", ctypes.c_int),
        ("sigma_ray", ctypes.c_float), ("sigma_dist", ctypes.c_float),
        ("fx", ctypes.c_float), ("fy", ctypes.c_float), ("cx", ctypes.c_float), ("cy", ctypes.c_float),
        ("height", ctypes.c_int), ("width", ctypes.c_int), ("pixel_border", ctypes.c_int),
        ("z_eps", ctypes.c_float), ("sigma_pixel", ctypes.c_float), ("sigma_depth", ctypes.c_float),
        ("sigma_point", ctypes.c_float),
        ("C_thresh", ctypes.c_float), ("Q_thresh", ctypes.c_float),
    ]


KINDS = {"rays": 0, "calib": 1, "points": 2}


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _cpu(t, dtype=None):
    t = t.detach().cpu()
    if dtype is not None:
        t = t.to(dtype)
    return t.contiguous()


def num_threads():
    return lib().m3r_oracle_num_threads()


def set_num_threads(n):
    lib().m3r_oracle_set_num_threads(ctypes.c_int(n))


def force_sw_half(on):
    lib().m3r_oracle_force_sw_half(ctypes.c_int(1 if on else 0))


def iter_proj(rays_img_with_grad, pts_3d_norm, p_init, max_iter, lambda_init, cost_thresh):
    rays = _cpu(rays_img_with_grad, torch.float32)
    pts = _cpu(pts_3d_norm, torch.float32)
    p0 = _cpu(p_init, torch.float32)
    b, h, w, _ = rays.shape
    n = p0.shape[1]
    p_new = torch.zeros((b, n, 2), dtype=torch.float32)
    conv = torch.zeros((b, n), dtype=torch.uint8)
    lib().m3r_oracle_iter_proj(_p(rays), _p(pts), _p(p0), _p(p_new), _p(conv), b, h, w, n,
                               int(max_iter), ctypes.c_float(lambda_init),
                               ctypes.c_float(cost_thresh))
    return [p_new, conv.bool()]


_DT = {torch.float16: 0, torch.float32: 1, torch.float64: 2}


def refine_matches(D11, D21, p1, radius, dilation_max):
    D11c, D21c = _cpu(D11), _cpu(D21)
    assert D11c.dtype == D21c.dtype and D11c.dtype in _DT
    p = _cpu(p1, torch.int64)
    b, h, w, f = D11c.shape
    n = p.shape[1]
    out = torch.zeros((b, n, 2), dtype=torch.int64)
    lib().m3r_oracle_refine_matches(_p(D11c), _p(D21c), _p(p), _p(out), b, h, w, n, f,
                                    _DT[D11c.dtype], int(radius), int(dilation_max))
    return [out]


def make_params(kind, *, sigma_ray=1.0, sigma_dist=1.0, K=None, height=0, width=0, pixel_border=0,
                z_eps=0.0, sigma_pixel=1.0, sigma_depth=1.0, sigma_point=1.0, C_thresh=0.0,
                Q_thresh=0.0):
    p = GnParams()
    p.kind = KINDS[kind]
    p.sigma_ray, p.sigma_dist = sigma_ray, sigma_dist
    if K is not None:
        Kc = _cpu(K, torch.float32)
        p.fx, p.fy, p.cx, p.cy = float(Kc[0, 0]), float(Kc[1, 1]), float(Kc[0, 2]), float(Kc[1, 2])
    p.height, p.width, p.pixel_border = int(height), int(width), int(pixel_border)
    p.z_eps, p.sigma_pixel, p.sigma_depth = z_eps, sigma_pixel, sigma_depth
    p.sigma_point = sigma_point
    p.C_thresh, p.Q_thresh = C_thresh, Q_thresh
    return p


def edge_indices(ii, jj):
    ii, jj = _cpu(ii, torch.int64), _cpu(jj, torch.int64)
    E = ii.shape[0]
    ie = torch.zeros(E, dtype=torch.int32)
    je = torch.zeros(E, dtype=torch.int32)
    nu = lib().m3r_oracle_edge_indices(_p(ii), _p(jj), E, _p(ie), _p(je))
    return ie, je, nu


def gn_accumulate(params, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q):
    """Returns (Hs [4,E,7,7] f32, gs [2,E,7] f32) of one accumulation launch."""
    Twc, Xs, Cs, Q = (_cpu(t, torch.float32) for t in (Twc, Xs, Cs, Q))
    idx = _cpu(idx_ii2jj, torch.int64)
    vm = _cpu(valid_match).to(torch.uint8).contiguous()
    ie, je, _ = edge_indices(ii, jj)
    N, n, E = Xs.shape[0], Xs.shape[1], ie.shape[0]
    Hs = torch.zeros((4, E, 7, 7), dtype=torch.float32)
    gs = torch.zeros((2, E, 7), dtype=torch.float32)
    lib().m3r_oracle_gn_accumulate(ctypes.byref(params), _p(Twc), _p(Xs), _p(Cs), _p(ie), _p(je),
                                   _p(idx), _p(vm), _p(Q), N, n, E, _p(Hs), _p(gs))
    return Hs, gs


def gn_solve(Hs, gs, ii, jj, num_poses):
    """Returns (fail, x[(N-1)*7] f64): the reference's SparseBlock assemble + LLT solve."""
    ie, je, _ = edge_indices(ii, jj)
    Hs, gs = _cpu(Hs, torch.float32), _cpu(gs, torch.float32)
    x = torch.zeros(7 * (num_poses - 1), dtype=torch.float64)
    fail = lib().m3r_oracle_gn_solve(_p(Hs), _p(gs), _p(ie), _p(je), int(num_poses), ie.shape[0],
                                     _p(x))
    return fail, x


def pose_retr(Twc, dx, num_fix=1):
    T = _cpu(Twc, torch.float32).clone()
    d = _cpu(dx, torch.float32)
    lib().m3r_oracle_pose_retr(_p(T), _p(d), T.shape[0], int(num_fix))
    return T


def gauss_newton(params, Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, max_iter, delta_thresh):
    """Full driver loop.  Returns (Twc_new [N,8], dx [N-1,7], iterations executed)."""
    T = _cpu(Twc, torch.float32).clone()
    Xs, Cs, Q = (_cpu(t, torch.float32) for t in (Xs, Cs, Q))
    ii, jj, idx = (_cpu(t, torch.int64) for t in (ii, jj, idx_ii2jj))
    vm = _cpu(valid_match).to(torch.uint8).contiguous()
    N, n, E = Xs.shape[0], Xs.shape[1], ii.shape[0]
    dx = torch.zeros((N - 1, 7), dtype=torch.float32)
    iters = lib().m3r_oracle_gauss_newton(ctypes.byref(params), _p(T), _p(Xs), _p(Cs), _p(ii),
                                          _p(jj), _p(idx), _p(vm), _p(Q), N, n, E, int(max_iter),
                                          ctypes.c_float(delta_thresh), _p(dx))
    return T, dx, iters


# --------------------------------------------------------------------------------------------
# matching front end: CPU restatement of mast3r_slam/matching.py and
# mast3r_slam/image.py around the two kernels above.  Pinned by tests/golden/front_*.npz,
# which tests/golden/make_golden_front.py writes by running the reference's own matching.py
# (imported from /root/reference, its extension module replaced by this oracle's kernels).
# --------------------------------------------------------------------------------------------
def img_gradient(img):
    """image.py:5-38: depthwise 3x3 taps (1/32)[-3 0 3; -10 0 10; -3 0 3] and the transpose,
    reflect padding.  img is [b,c,h,w]."""
    import torch.nn.functional as F

    c = img.shape[1]
    gxk = (1.0 / 32.0) * torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]],
                                      dtype=img.dtype)
    pad = F.pad(img, (1, 1, 1, 1), mode="reflect")
    gx = F.conv2d(pad, gxk.repeat(c, 1, 1, 1), groups=c)
    gy = F.conv2d(pad, gxk.T.contiguous().repeat(c, 1, 1, 1), groups=c)
    return gx, gy


def prep_for_iter_proj(X11, X21, idx_1_to_2_init=None):
    """matching.py:25-49.  Returns (rays_with_grad_img [b,h,w,9], pts3d_norm [b,n,3], p_init)."""
    import torch.nn.functional as F

    X11, X21 = _cpu(X11, torch.float32), _cpu(X21, torch.float32)
    b, h, w, _ = X11.shape
    rays = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    rays_with_grad = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts = F.normalize(X21.reshape(b, -1, 3), dim=-1)
    if idx_1_to_2_init is None:
        idx_1_to_2_init = torch.arange(h * w)[None, :].repeat(b, 1)
    idx0 = _cpu(idx_1_to_2_init, torch.int64)
    p_init = torch.stack((idx0 % w, idx0 // w), dim=-1).float()  # lin_to_pixel, matching.py:18-22
    return rays_with_grad, pts, p_init


def match(X11, X21, D11, D21, idx_1_to_2_init=None, *, max_iter=10, lambda_init=1e-8,
          convergence_thresh=1e-6, dist_thresh=0.1, radius=3, dilation_max=5):
    """matching.py:8-88 (match -> match_iterative_proj).  Returns (idx_1_to_2 [b,n] i64,
    valid_match2 [b,n,1] bool)."""
    X11, X21 = _cpu(X11, torch.float32), _cpu(X21, torch.float32)
    b, h, w = X21.shape[:3]
    rays, pts, p_init = prep_for_iter_proj(X11, X21, idx_1_to_2_init)
    p1, valid_proj2 = iter_proj(rays, pts, p_init, max_iter, lambda_init, convergence_thresh)
    p1 = p1.long()
    batch_inds = torch.arange(b)[:, None].repeat(1, h * w)
    dists2 = torch.linalg.norm(X11[batch_inds, p1[..., 1], p1[..., 0], :].reshape(b, h, w, 3) - X21,
                               dim=-1)
    valid_proj2 = valid_proj2 & (dists2 < dist_thresh).view(b, -1)
    if radius > 0:
        (p1,) = refine_matches(_cpu(D11).half(), _cpu(D21).reshape(b, h * w, -1).half(), p1, radius,
                               dilation_max)
    idx_1_to_2 = p1[..., 0] + w * p1[..., 1]  # pixel_to_lin, matching.py:13-15
    return idx_1_to_2, valid_proj2.unsqueeze(-1)
