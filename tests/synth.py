"""Seeded synthetic inputs for the hot path, torch, any device.

One ray-cast scene serves every export: a smooth height-field z = f(x, y) seen by pinhole
cameras with known Sim(3) poses T_WC = (t, q_xyzw, s).  Per keyframe the pointmap lies
exactly on the pixel rays (what `constrain_points_to_ray` produces for the calibrated
solver, geometry.py:37-42); matches between keyframes are the projection of one keyframe's
points into the other, rounded to the nearest pixel, so residuals look like real data
(sub-pixel mismatch + noise) and the true poses are a known answer.
"""
import math

import torch
import torch.nn.functional as F


# ------------------------------------------------------------------------------------------
# Sim(3) helpers (float64 torch, used to BUILD inputs, not as an oracle)
# ------------------------------------------------------------------------------------------
def quat_to_R(q):
    x, y, z, w = q.unbind(-1)
    return torch.stack([
        1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w),
        2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w),
        2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y),
    ], dim=-1).view(*q.shape[:-1], 3, 3)


def axis_angle_to_quat(v):
    th = v.norm(dim=-1, keepdim=True).clamp_min(1e-12)
    return torch.cat([v / th * torch.sin(0.5 * th), torch.cos(0.5 * th)], dim=-1)


def sim3_act(T, X):
    """T [...,8] acting on X [...,n,3]: s R X + t."""
    t, q, s = T[..., :3], T[..., 3:7], T[..., 7:8]
    R = quat_to_R(q)
    return s.unsqueeze(-2) * (X @ R.transpose(-1, -2)) + t.unsqueeze(-2)


def sim3_inv_act(T, P):
    t, q, s = T[..., :3], T[..., 3:7], T[..., 7:8]
    R = quat_to_R(q)
    return ((P - t.unsqueeze(-2)) @ R) / s.unsqueeze(-2)


def quat_mul(a, b):
    ax, ay, az, aw = a.unbind(-1)
    bx, by, bz, bw = b.unbind(-1)
    return torch.stack([
        aw * bx + ax * bw + ay * bz - az * by,
        aw * by - ax * bz + ay * bw + az * bx,
        aw * bz + ax * by - ay * bx + az * bw,
        aw * bw - ax * bx - ay * by - az * bz,
    ], dim=-1)


def sim3_perturb(T, xi):
    """Left-compose a small (tau, omega, sigma) onto T (first-order exact enough for tests)."""
    tau, om, sg = xi[..., :3], xi[..., 3:6], xi[..., 6:7]
    dq = axis_angle_to_quat(om)
    ds = torch.exp(sg)
    dR = quat_to_R(dq)
    t = ds * (T[..., :3].unsqueeze(-2) @ dR.transpose(-1, -2)).squeeze(-2) + tau
    q = quat_mul(dq, T[..., 3:7])
    return torch.cat([t, q, ds * T[..., 7:8]], dim=-1)


# ------------------------------------------------------------------------------------------
# scene
# ------------------------------------------------------------------------------------------
def default_K(h, w):
    """TUM-fr1-like intrinsics scaled to the image (dataloader.py:79-81 at 512x384)."""
    sx, sy = w / 512.0, h / 384.0
    return torch.tensor([[413.8 * sx, 0.0, 254.9 * sx], [0.0, 413.2 * sy, 204.2 * sy],
                         [0.0, 0.0, 1.0]], dtype=torch.float64)


def _surface(x, y):
    f = 2.5 + 0.30 * torch.sin(1.3 * x) * torch.cos(1.1 * y) + 0.08 * torch.sin(3.1 * x + 0.5) \
        + 0.06 * torch.cos(2.7 * y - 0.3)
    fx = 0.30 * 1.3 * torch.cos(1.3 * x) * torch.cos(1.1 * y) + 0.08 * 3.1 * torch.cos(3.1 * x + 0.5)
    fy = -0.30 * 1.1 * torch.sin(1.3 * x) * torch.sin(1.1 * y) - 0.06 * 2.7 * torch.sin(2.7 * y - 0.3)
    return f, fx, fy


def make_poses(N, seed, device="cpu"):
    """Cameras on a small circle looking down +z, rotations of a few degrees, scales in
    [0.8, 1.25]; pose 0 is the identity (it is the fixed pose of the solver)."""
    g = torch.Generator().manual_seed(1000 + seed)
    ang = torch.arange(N, dtype=torch.float64) * (2 * math.pi / max(N, 8)) * 1.0
    t = torch.stack([0.25 * torch.cos(ang) - 0.25, 0.25 * torch.sin(ang),
                     0.05 * torch.sin(2 * ang)], dim=-1)
    om = (torch.rand(N, 3, generator=g, dtype=torch.float64) - 0.5) * 2 * math.radians(4.0)
    q = axis_angle_to_quat(om)
    s = 0.8 + 0.45 * torch.rand(N, 1, generator=g, dtype=torch.float64)
    T = torch.cat([t, q, s], dim=-1)
    T[0] = torch.tensor([0, 0, 0, 0, 0, 0, 1, 1], dtype=torch.float64)
    return T.to(device)


def render_pointmap(T, K, h, w):
    """Ray-cast the height-field from camera T [8] (float64).  Returns (X [h*w,3] in the
    camera frame, on the pixel rays; P_w [h*w,3] world points)."""
    dev = T.device
    v, u = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=dev),
                          torch.arange(w, dtype=torch.float64, device=dev), indexing="ij")
    K = K.to(dev)
    dirs = torch.stack([(u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1], torch.ones_like(u)],
                       dim=-1).view(-1, 3)
    R = quat_to_R(T[3:7])
    d = dirs @ R.T
    c = T[:3]
    lam = torch.full((h * w,), 2.5, dtype=torch.float64, device=dev)
    for _ in range(12):
        P = c + lam[:, None] * d
        f, fx, fy = _surface(P[:, 0], P[:, 1])
        gval = P[:, 2] - f
        gder = d[:, 2] - fx * d[:, 0] - fy * d[:, 1]
        lam = lam - gval / gder
    P = c + lam[:, None] * d
    X = dirs * (lam / T[7])[:, None]
    return X, P


def descriptor_field(P_w, fdim, seed):
    """World-anchored unit-norm descriptor: 6 low-frequency + (fdim-6) band-limited channels."""
    g = torch.Generator().manual_seed(7000 + seed)
    nlow = min(6, fdim)
    fr_lo = (torch.rand(nlow, 2, generator=g, dtype=torch.float64) - 0.5) * 4.0
    fr_hi = (torch.rand(fdim - nlow, 2, generator=g, dtype=torch.float64) - 0.5) * 60.0
    fr = torch.cat([fr_lo, fr_hi], 0).to(P_w.device)
    ph = (torch.rand(fdim, generator=g, dtype=torch.float64) * 2 * math.pi).to(P_w.device)
    D = torch.sin(P_w[:, :2] @ fr.T + ph)
    return D / D.norm(dim=-1, keepdim=True)


def img_gradient(img):
    """Same taps as the reference's img_gradient (image.py:5-38): 1/32*[[-3,0,3],[-10,0,10],
    [-3,0,3]] and its transpose, reflect padding, depthwise."""
    c = img.shape[1]
    gxk = (1.0 / 32.0) * torch.tensor([[-3.0, 0.0, 3.0], [-10.0, 0.0, 10.0], [-3.0, 0.0, 3.0]],
                                      device=img.device, dtype=img.dtype)
    gyk = gxk.T.contiguous()
    pad = F.pad(img, (1, 1, 1, 1), mode="reflect")
    gx = F.conv2d(pad, gxk.repeat(c, 1, 1, 1), groups=c)
    gy = F.conv2d(pad, gyk.repeat(c, 1, 1, 1), groups=c)
    return gx, gy


def prep_for_iter_proj(X11, X21, idx_init=None):
    """What matching.prep_for_iter_proj (matching.py:25-49) hands to iter_proj."""
    b, h, w, _ = X11.shape
    rays = F.normalize(X11, dim=-1).permute(0, 3, 1, 2)
    gx, gy = img_gradient(rays)
    rays_with_grad = torch.cat((rays, gx, gy), dim=1).permute(0, 2, 3, 1).contiguous()
    pts = F.normalize(X21.reshape(b, -1, 3), dim=-1)
    if idx_init is None:
        idx_init = torch.arange(h * w, device=X11.device)[None, :].repeat(b, 1)
    p_init = torch.stack((idx_init % w, idx_init // w), dim=-1).float()
    return rays_with_grad, pts.contiguous(), p_init.contiguous()


def make_match_inputs(h, w, b=1, seed=0, fdim=24, device="cpu", noise=1e-3, desc_noise=0.05):
    """BASELINE config 2: frame -> keyframe matching inputs.
    Returns dict with X11 [b,h,w,3], X21 [b,h,w,3] (view-2 points in view-1's frame),
    D11 [b,h,w,F] f32, D21 [b,h,w,F] f32, true_uv [b,h*w,2] (float projection)."""
    K = default_K(h, w).to(device)
    g = torch.Generator().manual_seed(100 + seed)
    X11s, X21s, D11s, D21s, uvs = [], [], [], [], []
    for bi in range(b):
        T = make_poses(2 + bi, seed * 17 + bi, device)
        Ti = T[0]
        Tj = T[-1].clone()
        Tj[:3] = Tj[:3] * 0.3 + torch.tensor([0.04, -0.03, 0.01], dtype=torch.float64, device=device)
        Xi, Pi = render_pointmap(Ti, K, h, w)
        Xj, Pj = render_pointmap(Tj, K, h, w)
        Xj_in_i = sim3_inv_act(Ti, sim3_act(Tj, Xj))
        uv = torch.stack([K[0, 0] * Xj_in_i[:, 0] / Xj_in_i[:, 2] + K[0, 2],
                          K[1, 1] * Xj_in_i[:, 1] / Xj_in_i[:, 2] + K[1, 2]], dim=-1)
        nz = torch.randn(h * w, 3, generator=g, dtype=torch.float64).to(device) * noise
        D1 = descriptor_field(Pi, fdim, seed)
        D2 = descriptor_field(Pj, fdim, seed)
        D2 = D2 + torch.randn(h * w, fdim, generator=g, dtype=torch.float64).to(device) * desc_noise
        D2 = D2 / D2.norm(dim=-1, keepdim=True)
        X11s.append(Xi.view(h, w, 3))
        X21s.append((Xj_in_i + nz).view(h, w, 3))
        D11s.append(D1.view(h, w, fdim))
        D21s.append(D2.view(h, w, fdim))
        uvs.append(uv)
    f32 = torch.float32
    return dict(X11=torch.stack(X11s).to(f32), X21=torch.stack(X21s).to(f32),
                D11=torch.stack(D11s).to(f32).contiguous(), D21=torch.stack(D21s).to(f32).contiguous(),
                true_uv=torch.stack(uvs).to(f32), K=K.to(f32), h=h, w=w)


def make_graph_edges(N, seed, extra_per_kf=3):
    """(k-1,k) for all k plus up to `extra_per_kf` earlier keyframes per k (mirrors
    n_consec=1 + retrieval k=3, main.py:118-128, config/base.yaml:53).  Stored (undirected)."""
    g = torch.Generator().manual_seed(3000 + seed)
    ii, jj = [], []
    for k in range(1, N):
        ii.append(k - 1), jj.append(k)
        cand = list(range(0, k - 1))
        if cand:
            perm = torch.randperm(len(cand), generator=g).tolist()[:extra_per_kf]
            for c in sorted(perm):
                ii.append(cand[c]), jj.append(k)
    return torch.tensor(ii, dtype=torch.long), torch.tensor(jj, dtype=torch.long)


def make_graph(N, h, w, seed=0, device="cpu", pose_noise=0.02, point_noise=2e-3, kf_id_stride=1,
               extra_per_kf=3, on_rays=True):
    """BASELINE configs 3-5: an N-keyframe graph with two-way edges as
    FactorGraph.prep_two_way_edges builds them (global_opt.py:106-112).

    Returns dict: Twc_true/Twc_init [N,8] f32, Xs [N,n,3], Cs [N,n,1], K [3,3],
    ii/jj [E] int64 (global keyframe ids = kf_id_stride * packed index), idx_ii2jj [E,n] int64,
    valid_match [E,n,1] bool, Q [E,n,1] f32."""
    n = h * w
    K = default_K(h, w).to(device)
    T_true = make_poses(N, seed, device)
    g = torch.Generator().manual_seed(5000 + seed)
    Xs64, Pws = [], []
    for k in range(N):
        X, P = render_pointmap(T_true[k], K, h, w)
        if point_noise > 0:
            # noise along the ray keeps the points on their pixel rays (calibrated solver)
            dz = 1.0 + torch.randn(n, 1, generator=g, dtype=torch.float64).to(device) * point_noise
            X = X * dz if on_rays else X + torch.randn(n, 3, generator=g, dtype=torch.float64).to(device) * point_noise
        Xs64.append(X)
        Pws.append(P)
    Xs64 = torch.stack(Xs64)
    si, sj = make_graph_edges(N, seed, extra_per_kf)
    ii = torch.cat([si, sj]).to(device)  # two-way
    jj = torch.cat([sj, si]).to(device)
    E = ii.shape[0]
    idx = torch.empty((E, n), dtype=torch.long, device=device)
    valid = torch.empty((E, n, 1), dtype=torch.bool, device=device)
    for e in range(E):
        i, j = int(ii[e]), int(jj[e])
        # point k of j, seen from i (true poses, noise-free world points)
        Xin_i = sim3_inv_act(T_true[i], Pws[j])
        u = K[0, 0] * Xin_i[:, 0] / Xin_i[:, 2] + K[0, 2]
        v = K[1, 1] * Xin_i[:, 1] / Xin_i[:, 2] + K[1, 2]
        ur, vr = torch.round(u).long(), torch.round(v).long()
        ok = (ur >= 0) & (ur < w) & (vr >= 0) & (vr < h) & (Xin_i[:, 2] > 0)
        idx[e] = torch.where(ok, vr * w + ur, torch.zeros_like(ur))
        valid[e, :, 0] = ok
    # confidences: Q ~ exp(N(1, 0.5^2)) (about 75 % above Q_conf = 1.5), C = 1 + exp(N(0,1))
    Q = torch.exp(1.0 + 0.5 * torch.randn(E, n, 1, generator=g)).to(device)
    valid &= (torch.rand(E, n, 1, generator=g) < 0.9).to(device)
    Cs = (1.0 + torch.exp(torch.randn(N, n, 1, generator=g))).to(device)
    xi = torch.randn(N, 7, generator=g, dtype=torch.float64).to(device) * pose_noise
    xi[0] = 0
    T_init = sim3_perturb(T_true, xi)
    f32 = torch.float32
    return dict(Twc_true=T_true.to(f32).contiguous(), Twc_init=T_init.to(f32).contiguous(),
                Xs=Xs64.to(f32).contiguous(), Cs=Cs.to(f32).contiguous(), K=K.to(f32).contiguous(),
                ii=(ii * kf_id_stride).contiguous(), jj=(jj * kf_id_stride).contiguous(),
                idx_ii2jj=idx.contiguous(), valid_match=valid.contiguous(),
                Q=Q.to(f32).contiguous(), h=h, w=w, N=N, E=E, n=n)


# parameters from config/base.yaml:35-49 (local_opt) and :8-14 (matching)
GN_DEFAULTS = dict(sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0, sigma_depth=10.0,
                   sigma_point=0.05, C_thresh=0.0, Q_thresh=1.5, pixel_border=-10, z_eps=1e-6,
                   max_iter=10, delta_thresh=1e-8)
MATCH_DEFAULTS = dict(max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6, dist_thresh=0.1,
                      radius=3, dilation_max=5)


# parameters from config/base.yaml:16-31 (tracking)
TRACK_DEFAULTS = dict(max_iters=50, C_conf=0.0, Q_conf=1.5, rel_error=1e-3, delta_norm=1e-3,
                      huber=1.345, sigma_ray=0.003, sigma_dist=10.0, sigma_pixel=1.0,
                      sigma_depth=10.0, pixel_border=-10, depth_eps=1e-6, min_match_frac=0.05,
                      match_frac_thresh=0.333)


def make_track_inputs(h, w, seed=0, device="cpu", pose_noise=0.03, point_noise=2e-3):
    """BASELINE config 2's pose solve: one frame f against the last keyframe k, as
    FrameTracker.track hands it to opt_pose_* (tracker.py:29-90, 129-155).

    Returns dict: Xf_canon/Xk [n,3], Cf/Ck [n,1], idx_f2k [n] (for keyframe point k, its match in
    the frame), valid_match_k [n,1] bool, Qff/Qkf [n,1] (match confidences of the two views),
    T_WCf_init/T_WCf_true/T_WCk [1,8], K [3,3], and the derived arguments of the pose solve:
    Xf [n,3] = Xf_canon[idx_f2k], Qk, valid_opt, meas_k [n,3], valid_meas_k [n,1]."""
    g = make_graph(2, h, w, seed=seed, device=device, pose_noise=pose_noise, point_noise=point_noise)
    n = h * w
    e = int(((g["ii"] == 1) & (g["jj"] == 0)).nonzero()[0, 0])
    idx = g["idx_ii2jj"][e].contiguous()
    valid_match = g["valid_match"][e].contiguous()
    gen = torch.Generator().manual_seed(7000 + seed)
    Qff = torch.exp(1.0 + 0.5 * torch.randn(n, 1, generator=gen)).to(device)
    Qkf = torch.exp(1.0 + 0.5 * torch.randn(n, 1, generator=gen)).to(device)
    Xf_canon, Xk = g["Xs"][1].contiguous(), g["Xs"][0].contiguous()
    Cf_canon, Ck = g["Cs"][1].contiguous(), g["Cs"][0].contiguous()
    Qk = torch.sqrt(Qff[idx] * Qkf)  # tracker.py:41
    Cf = Cf_canon[idx]
    T = TRACK_DEFAULTS
    valid_opt = valid_match & (Cf > T["C_conf"]) & (Ck > T["C_conf"]) & (Qk > T["Q_conf"])  # :62-66
    vv, uu = torch.meshgrid(torch.arange(h, device=device), torch.arange(w, device=device), indexing="ij")
    uv = torch.stack((uu, vv), dim=-1).view(-1, 2).float()
    valid_meas = Xk[..., 2:3] > T["depth_eps"]
    meas = torch.cat((uv, torch.log(Xk[..., 2:3].clamp_min(1e-30))), dim=-1)
    meas[~valid_meas.repeat(1, 3)] = 0.0
    return dict(Xf_canon=Xf_canon, Xk=Xk, Cf_canon=Cf_canon, Ck=Ck, idx_f2k=idx, valid_match_k=valid_match,
                Qff=Qff, Qkf=Qkf, T_WCk=g["Twc_init"][0:1].contiguous(),
                T_WCf_init=g["Twc_init"][1:2].contiguous(), T_WCf_true=g["Twc_true"][1:2].contiguous(),
                K=g["K"], h=h, w=w, n=n, Xf=Xf_canon[idx].contiguous(), Qk=Qk.contiguous(), Cf=Cf.contiguous(),
                valid_opt=valid_opt.contiguous(), meas_k=meas.contiguous(), valid_meas_k=valid_meas.contiguous())


# ------------------------------------------------------------------------------------------
# the reference's tracking loop as plain torch on any device (timing baseline for bench.py:
# what FrameTracker.opt_pose_ray_dist_sim3 executes on the GPU -- ~40 torch kernels and two
# host syncs per iteration -- with lietorch's Sim3 replaced by the torch ops below)
# ------------------------------------------------------------------------------------------
def sim3_exp(xi):
    """Exp of (tau, omega, sigma) [...,7] -> Sim3 data [...,8] (lietorch rxso3/sim3 formulas,
    general branch with guards; gn_kernels.cu:323-390)."""
    tau, phi, sigma = xi[..., :3], xi[..., 3:6], xi[..., 6:7]
    scale = torch.exp(sigma)
    th2 = (phi * phi).sum(-1, keepdim=True)
    th = torch.sqrt(th2)
    small_t, small_s = th < 1e-6, sigma.abs() < 1e-6
    th_s = torch.where(small_t, torch.ones_like(th), th)
    th2_s = th_s * th_s
    sg_s = torch.where(small_s, torch.ones_like(sigma), sigma)
    q = torch.cat([phi * torch.where(small_t, 0.5 - th2 / 48, torch.sin(0.5 * th_s) / th_s),
                   torch.where(small_t, 1 - th2 / 8, torch.cos(0.5 * th_s))], dim=-1)
    C = torch.where(small_s, torch.ones_like(sigma), (scale - 1) / sg_s)
    a, b, c = scale * torch.sin(th_s), scale * torch.cos(th_s), th2_s + sg_s * sg_s
    A_gen = (a * sg_s + (1 - b) * th_s) / (th_s * c)
    B_gen = (C - ((b - 1) * sg_s + a * th_s) / c) / th2_s
    A_s0 = (1 - torch.cos(th_s)) / th2_s
    B_s0 = (th_s - torch.sin(th_s)) / (th2_s * th_s)
    A_t0 = ((sg_s - 1) * scale + 1) / (sg_s * sg_s)
    B_t0 = (scale * 0.5 * sg_s * sg_s + scale - 1 - sg_s * scale) / (sg_s * sg_s * sg_s)
    A = torch.where(small_s, torch.where(small_t, torch.full_like(th, 0.5), A_s0),
                    torch.where(small_t, A_t0, A_gen))
    B = torch.where(small_s, torch.where(small_t, torch.full_like(th, 1 / 6), B_s0),
                    torch.where(small_t, B_t0, B_gen))
    pt = torch.linalg.cross(phi, tau)
    t = C * tau + A * pt + B * torch.linalg.cross(phi, pt)
    return torch.cat([t, q, scale], dim=-1)


def sim3_mul(A, B):
    """A * B on Sim3 data [...,8]."""
    t = A[..., 7:8] * (quat_to_R(A[..., 3:7]) @ B[..., :3].unsqueeze(-1)).squeeze(-1) + A[..., :3]
    return torch.cat([t, quat_mul(A[..., 3:7], B[..., 3:7]), A[..., 7:8] * B[..., 7:8]], dim=-1)


def sim3_inv(A):
    qi = A[..., 3:7] * torch.tensor([-1.0, -1.0, -1.0, 1.0], device=A.device)
    si = 1.0 / A[..., 7:8]
    t = -si * (quat_to_R(qi) @ A[..., :3].unsqueeze(-1)).squeeze(-1)
    return torch.cat([t, qi, si], dim=-1)


def torch_track_pose_rays(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None):
    """tracker.py:157-215 op for op in torch, on the tensors' device; returns (T_WCf, T_CkCf,
    iterations)."""
    import math

    c = dict(TRACK_DEFAULTS, **(cfg or {}))
    dev = Xf.device
    si_ray = 1 / c["sigma_ray"] * valid * torch.sqrt(Qk)
    si_dist = 1 / c["sigma_dist"] * valid * torch.sqrt(Qk)
    sqrt_info = torch.cat((si_ray.repeat(1, 3), si_dist), dim=1)
    T = sim3_mul(sim3_inv(T_WCk), T_WCf)
    d = torch.linalg.norm(Xk, dim=-1, keepdim=True)
    rd_k = torch.cat(((1.0 / d) * Xk, d), dim=-1)
    old_cost, it = float("inf"), 0
    for step in range(c["max_iters"]):
        pW = sim3_act(T, Xf.unsqueeze(0)).squeeze(0)
        x, y, z = pW.unbind(-1)
        o = torch.zeros_like(x)
        skew = torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).view(-1, 3, 3)
        dP = torch.cat([torch.eye(3, device=dev).repeat(pW.shape[0], 1, 1), -skew, pW.unsqueeze(-1)], dim=-1)
        dd = torch.linalg.norm(pW, dim=-1, keepdim=True)
        d_inv = 1.0 / dd
        r_ = d_inv * pW
        rd_f = torch.cat((r_, dd), dim=-1)
        eye = torch.eye(3, device=dev).repeat(pW.shape[0], 1, 1)
        dr = d_inv.unsqueeze(-1) * (eye - (d_inv ** 2).unsqueeze(-1) * (pW.unsqueeze(-1) @ pW.unsqueeze(-2)))
        drd = torch.cat((dr, r_.unsqueeze(-2)), dim=-2)
        r = rd_k - rd_f
        J = -drd @ dP
        wr = sqrt_info * r
        r_abs = torch.abs(wr)
        hub = torch.where(r_abs < c["huber"], torch.ones(1, device=dev), c["huber"] / r_abs)
        rsi = sqrt_info * torch.sqrt(hub)
        A = (rsi[..., None] * J).view(-1, 7)
        b = (rsi * r).view(-1, 1)
        H = A.T @ A
        g = -A.T @ b
        cost = 0.5 * (b.T @ b).item()
        L = torch.linalg.cholesky(H, upper=False)
        tau = torch.cholesky_solve(g, L, upper=False).view(1, -1)
        T = sim3_mul(sim3_exp(tau), T)
        it = step + 1
        rel_dec = math.fabs((old_cost - cost) / old_cost) if old_cost != 0 else float("nan")
        if rel_dec < c["rel_error"] or torch.linalg.norm(tau) < c["delta_norm"]:
            break
        old_cost = cost
    return sim3_mul(T_WCk, T), T, it
