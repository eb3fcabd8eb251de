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


# --------------------------------------------------------------------------------------------
# tracking pose solve: CPU restatement of FrameTracker.opt_pose_ray_dist_sim3
# / opt_pose_calib_sim3 / solve (mast3r_slam/tracker.py:157-266), the geometry helpers they call
# (mast3r_slam/geometry.py:12-33 point_to_ray_dist, :45-53 act_Sim3, :64-107 project_calib) and
# nonlinear_optimizer.py:5-35 (check_convergence, huber).  Pinned by tests/golden/track_*.npz,
# which tests/golden/make_golden_track.py writes by running the reference's own tracker.py /
# geometry.py / nonlinear_optimizer.py on CPU with `lietorch` replaced by the Sim3 class below.
# --------------------------------------------------------------------------------------------
class Sim3:
    """CPU stand-in for lietorch.Sim3: the reference's pip dependency `lietorch @ git+https://
    github.com/princeton-vl/lietorch.git` (pyproject.toml:15, unpinned, not vendored, CUDA-only
    build).  data [..., 8] = (t xyz, q xyzw, s).  Restates the published group operations of
    lietorch/include/{sim3,rxso3,so3}.h: act p -> s R p + t; inv; composition; retr(a) =
    Exp(a) * X (left), with Exp through this oracle's C restatement of the reference's own copy
    of it (gn_kernels.cu:323-413, itself pinned by the Gauss-Newton golden vectors)."""

    def __init__(self, data):
        self.data = data

    @staticmethod
    def Identity(n, **kw):
        d = torch.zeros((n, 8), dtype=torch.float32)
        d[:, 6] = 1.0
        d[:, 7] = 1.0
        return Sim3(d)

    @staticmethod
    def _rot(q, p):
        # Eigen::Quaternion::_transformVector: uv = 2 qv x p; p + w uv + qv x uv
        qv, qw = q[..., :3], q[..., 3:4]
        uv = 2.0 * torch.linalg.cross(qv.expand_as(p), p)
        return p + qw * uv + torch.linalg.cross(qv.expand_as(p), uv)

    @staticmethod
    def _qmul(a, b):
        ax, ay, az, aw = a.unbind(-1)
        bx, by, bz, bw = b.unbind(-1)
        return torch.stack((aw * bx + ax * bw + ay * bz - az * by,
                            aw * by - ax * bz + ay * bw + az * bx,
                            aw * bz + ax * by - ay * bx + az * bw,
                            aw * bw - ax * bx - ay * by - az * bz), dim=-1)

    def act(self, p):
        t, q, s = self.data[..., :3], self.data[..., 3:7], self.data[..., 7:8]
        return s * Sim3._rot(q, p) + t

    def inv(self):
        t, q, s = self.data[..., :3], self.data[..., 3:7], self.data[..., 7:8]
        qi = q * torch.tensor([-1.0, -1.0, -1.0, 1.0])
        si = 1.0 / s
        return Sim3(torch.cat((-(si * Sim3._rot(qi, t)), qi, si), dim=-1))

    def __mul__(self, other):
        t, q, s = self.data[..., :3], self.data[..., 3:7], self.data[..., 7:8]
        t2, q2, s2 = other.data[..., :3], other.data[..., 3:7], other.data[..., 7:8]
        return Sim3(torch.cat((s * Sim3._rot(q, t2) + t, Sim3._qmul(q, q2), s * s2), dim=-1))

    def retr(self, a):
        d = self.data.reshape(-1, 8).contiguous().clone()
        return Sim3(pose_retr(d, a.reshape(-1, 7).contiguous(), num_fix=0).view(self.data.shape))


TRACK_DEFAULTS = dict(max_iters=50, C_conf=0.0, Q_conf=1.5, rel_error=1e-3, delta_norm=1e-3,
                      huber=1.345, sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0,
                      sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6, min_match_frac=0.05,
                      match_frac_thresh=0.333)  # config/base.yaml:16-31


def _huber(r, k):  # nonlinear_optimizer.py:29-35
    r_abs = torch.abs(r)
    return torch.where(r_abs < k, torch.ones(1, dtype=r.dtype), k / r_abs)


def _track_solve(sqrt_info, r, J, k):  # tracker.py:157-171
    whitened_r = sqrt_info * r
    robust_sqrt_info = sqrt_info * torch.sqrt(_huber(whitened_r, k))
    mdim = J.shape[-1]
    A = (robust_sqrt_info[..., None] * J).view(-1, mdim)
    b = (robust_sqrt_info * r).view(-1, 1)
    H = A.T @ A
    g = -A.T @ b
    cost = 0.5 * (b.T @ b).item()
    L = torch.linalg.cholesky(H, upper=False)  # raises when H is not positive definite
    tau = torch.cholesky_solve(g, L, upper=False).view(1, -1)
    return tau, cost, H, g


def _converged(rel_error, delta_norm, old_cost, new_cost, delta):  # nonlinear_optimizer.py:5-26
    import math

    rel_dec = math.fabs((old_cost - new_cost) / old_cost) if old_cost != 0 else float("nan")
    return rel_dec < rel_error or float(torch.linalg.norm(delta)) < delta_norm


def _ray_dist(X):  # geometry.py:12-33
    d = torch.linalg.norm(X, dim=-1, keepdim=True)
    d_inv = 1.0 / d
    r = d_inv * X
    rd = torch.cat((r, d), dim=-1)
    eye = torch.eye(3, dtype=X.dtype).expand(*X.shape[:-1], 3, 3)
    dr_dX = d_inv.unsqueeze(-1) * (eye - (d_inv ** 2).unsqueeze(-1) * (X.unsqueeze(-1) @ X.unsqueeze(-2)))
    return rd, torch.cat((dr_dX, r.unsqueeze(-2)), dim=-2)


def _act_jac(T, p):  # geometry.py:45-53
    pW = T.act(p)
    x, y, z = pW.unbind(-1)
    o = torch.zeros_like(x)
    skew = torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).view(*pW.shape[:-1], 3, 3)
    eye = torch.eye(3, dtype=pW.dtype).expand(*pW.shape[:-1], 3, 3)
    return pW, torch.cat([eye, -skew, pW.unsqueeze(-1)], dim=-1)


def track_pose_rays(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, trace=None):
    """tracker.py:173-215.  Xf [n,3] (already gathered by idx_f2k), Xk [n,3], poses [1,8],
    Qk [n,1], valid [n,1] bool.  Returns (T_WCf [1,8], T_CkCf [1,8], iterations, last cost)."""
    c = dict(TRACK_DEFAULTS, **(cfg or {}))
    Xf, Xk, Qk = _cpu(Xf, torch.float32), _cpu(Xk, torch.float32), _cpu(Qk, torch.float32)
    valid = _cpu(valid).bool()
    Twf, Twk = Sim3(_cpu(T_WCf, torch.float32)), Sim3(_cpu(T_WCk, torch.float32))
    si_ray = 1 / c["sigma_ray"] * valid * torch.sqrt(Qk)
    si_dist = 1 / c["sigma_dist"] * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((si_ray.repeat(1, 3), si_dist), dim=1)
    T = Twk.inv() * Twf
    d = torch.linalg.norm(Xk, dim=-1, keepdim=True)
    rd_k = torch.cat((Xk / d, d), dim=-1)
    old_cost, cost, it = float("inf"), 0.0, 0
    for step in range(c["max_iters"]):
        P, dP = _act_jac(T, Xf)
        rd_f, drd = _ray_dist(P)
        r = rd_k - rd_f
        J = -drd @ dP
        tau, cost, H, g = _track_solve(sqrt_info, r, J, c["huber"])
        if trace is not None:
            trace.append(dict(H=H, g=g, cost=cost, tau=tau, T=T.data.clone()))
        T = T.retr(tau)
        it = step + 1
        if _converged(c["rel_error"], c["delta_norm"], old_cost, cost, tau):
            break
        old_cost = cost
    return (Twk * T).data, T.data, it, cost


def track_pose_calib(Xf, Xk, T_WCf, T_WCk, Qk, valid, meas_k, valid_meas_k, K, img_size, cfg=None,
                     trace=None):
    """tracker.py:217-266 with geometry.project_calib (geometry.py:64-107).  meas_k [n,3] =
    (u, v, log z) of the keyframe points, valid_meas_k [n,1]; K [3,3]; img_size (h, w)."""
    c = dict(TRACK_DEFAULTS, **(cfg or {}))
    Xf, Xk, Qk = _cpu(Xf, torch.float32), _cpu(Xk, torch.float32), _cpu(Qk, torch.float32)
    meas_k, K = _cpu(meas_k, torch.float32), _cpu(K, torch.float32)
    valid, valid_meas_k = _cpu(valid).bool(), _cpu(valid_meas_k).bool()
    Twf, Twk = Sim3(_cpu(T_WCf, torch.float32)), Sim3(_cpu(T_WCk, torch.float32))
    si_pix = 1 / c["sigma_pixel"] * valid * torch.sqrt(Qk)
    si_dep = 1 / c["sigma_depth"] * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((si_pix.repeat(1, 2), si_dep), dim=1)
    T = Twk.inv() * Twf
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    border, z_eps = c["pixel_border"], c["depth_eps"]
    old_cost, cost, it = float("inf"), 0.0, 0
    for step in range(c["max_iters"]):
        P, dP = _act_jac(T, Xf)
        p = (K @ P[..., None]).squeeze(-1)
        p = (p / p[..., 2:3])[..., :2]
        u, v = p[..., 0:1], p[..., 1:2]
        x, y, z = P[..., 0:1], P[..., 1:2], P[..., 2:3]
        valid_z = z > z_eps
        valid_proj = ((u > border) & (u < img_size[1] - 1 - border) & (v > border) &
                      (v < img_size[0] - 1 - border) & valid_z)
        logz = torch.log(z)
        logz[~valid_z] = 0.0
        pz = torch.cat((p, logz), dim=-1)
        z_inv = 1.0 / z[..., 0]
        dpz = torch.zeros(*P.shape[:-1], 3, 3)
        dpz[..., 0, 0] = fx
        dpz[..., 1, 1] = fy
        dpz[..., 0, 2] = -fx * x[..., 0] * z_inv
        dpz[..., 1, 2] = -fy * y[..., 0] * z_inv
        dpz *= z_inv[..., None, None]
        dpz[..., 2, 2] = z_inv
        sqrt_info2 = (valid_proj & valid_meas_k) * sqrt_info
        r = meas_k - pz
        J = -dpz @ dP
        tau, cost, H, g = _track_solve(sqrt_info2, r, J, c["huber"])
        if trace is not None:
            trace.append(dict(H=H, g=g, cost=cost, tau=tau, T=T.data.clone()))
        T = T.retr(tau)
        it = step + 1
        if _converged(c["rel_error"], c["delta_norm"], old_cost, cost, tau):
            break
        old_cost = cost
    return (Twk * T).data, T.data, it, cost
