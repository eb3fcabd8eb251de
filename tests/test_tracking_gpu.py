"""GPU parity tests of the tracking pose solve: m3r_track_pose through
mast3r_slam_backends.tracking, against the committed golden vectors (the reference's own
tracker.py loop run on CPU, tests/golden/make_golden_track.py) and the CPU oracle.

Tolerance (BASELINE.json north_star, Gauss-Newton): poses within 1e-4 relative.  The solve is
an iteration with a data-dependent stop test, so the comparison is made (a) with the stop
tests off and a fixed iteration count, where only rounding differs, and (b) with the default
stop tests, where both sides must stop after the same number of iterations."""
import glob
import os

import pytest
import torch

import synth
from test_oracle import load_track_golden

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TOL = 1e-4


@pytest.fixture(scope="module")
def tracking(backends):
    from mast3r_slam_backends import tracking as tr

    yield tr
    tr.set_config({"tracking": dict(synth.TRACK_DEFAULTS)})


def _rel(a, b):
    return ((a.cpu() - b).abs().max() / b.abs().max()).item()


def _cuda(t):
    return {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in t.items()}


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLD, "track_*.npz"))) or [None])
def test_golden_tracking_gpu(tracking, path):
    if path is None:
        pytest.skip("no golden tracking vectors committed")
    t, cfg, ref, img = load_track_golden(path)
    tracking.set_config({"tracking": dict(synth.TRACK_DEFAULTS, **cfg)})
    d = _cuda(t)
    Twf, Tkf = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                               d["valid_opt"])
    e1, e2 = _rel(Twf, ref["ref_rays_T_WCf"]), _rel(Tkf, ref["ref_rays_T_CkCf"])
    print(f"{os.path.basename(path)} rays: rel err T_WCf {e1:.2e}, T_CkCf {e2:.2e}, "
          f"{tracking.last_info.iterations} iterations")
    assert Twf.shape == (1, 8) and Tkf.shape == (1, 8) and e1 < TOL and e2 < TOL
    Twf, Tkf = tracking.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                            d["valid_opt"], d["meas_k"], d["valid_meas_k"], d["K"], img)
    e1, e2 = _rel(Twf, ref["ref_calib_T_WCf"]), _rel(Tkf, ref["ref_calib_T_CkCf"])
    print(f"{os.path.basename(path)} calib: rel err T_WCf {e1:.2e}, T_CkCf {e2:.2e}, "
          f"{tracking.last_info.iterations} iterations")
    assert e1 < TOL and e2 < TOL


@pytest.mark.parametrize("kind", ["rays", "calib"])
@pytest.mark.parametrize("h,w,seed,iters", [(48, 64, 3, 1), (48, 64, 3, 4), (120, 160, 4, 8), (384, 512, 5, 5)])
def test_fixed_iterations_vs_oracle(tracking, oracle, kind, h, w, seed, iters):
    """Stop tests off: after exactly `iters` iterations only rounding separates the two."""
    cfg = dict(synth.TRACK_DEFAULTS, rel_error=0.0, delta_norm=0.0, max_iters=iters)
    tracking.set_config({"tracking": cfg})
    t = synth.make_track_inputs(h, w, seed=seed, pose_noise=0.04)
    d = _cuda(t)
    if kind == "rays":
        To, Tko, ito, _ = oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                 t["valid_opt"], cfg)
        Tw, Tk = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                                 d["valid_opt"])
    else:
        To, Tko, ito, _ = oracle.track_pose_calib(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                  t["valid_opt"], t["meas_k"], t["valid_meas_k"], t["K"],
                                                  (h, w), cfg)
        Tw, Tk = tracking.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                              d["valid_opt"], d["meas_k"], d["valid_meas_k"], d["K"], (h, w))
    assert tracking.last_info.iterations == iters == ito
    e1, e2 = _rel(Tw, To), _rel(Tk, Tko)
    print(f"{kind} {h}x{w} {iters} it: rel err T_WCf {e1:.2e}, T_CkCf {e2:.2e}")
    assert e1 < TOL and e2 < TOL


@pytest.mark.parametrize("kind", ["rays", "calib"])
@pytest.mark.parametrize("h,w,seed,noise", [(96, 128, 6, 0.02), (224, 224, 7, 0.05), (384, 512, 8, 0.03)])
def test_default_stop_tests_vs_oracle(tracking, oracle, kind, h, w, seed, noise):
    tracking.set_config({"tracking": dict(synth.TRACK_DEFAULTS)})
    t = synth.make_track_inputs(h, w, seed=seed, pose_noise=noise)
    d = _cuda(t)
    if kind == "rays":
        To, Tko, ito, costo = oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                     t["valid_opt"])
        Tw, Tk = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                                 d["valid_opt"])
    else:
        To, Tko, ito, costo = oracle.track_pose_calib(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                      t["valid_opt"], t["meas_k"], t["valid_meas_k"], t["K"],
                                                      (h, w))
        Tw, Tk = tracking.opt_pose_calib_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                              d["valid_opt"], d["meas_k"], d["valid_meas_k"], d["K"], (h, w))
    info = tracking.last_info
    e1 = _rel(Tw, To)
    err_true = (Tw.cpu() - t["T_WCf_true"]).abs().max().item()
    print(f"{kind} {h}x{w}: {info.iterations} iterations (oracle {ito}), cost {info.cost:.6e} (oracle "
          f"{costo:.6e}), rel err {e1:.2e}, |T - T_true| {err_true:.2e}")
    assert info.iterations == ito
    assert abs(info.cost - costo) <= 1e-4 * abs(costo)
    assert e1 < TOL and _rel(Tk, Tko) < TOL


def test_in_kernel_gather_equals_pregathered(tracking):
    """idx_f2k handed to the kernel (tracker.py:155 folded in) == Xf gathered by the caller, bitwise."""
    tracking.set_config({"tracking": dict(synth.TRACK_DEFAULTS)})
    d = _cuda(synth.make_track_inputs(120, 160, seed=9))
    a = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    b = tracking.opt_pose_ray_dist_sim3(d["Xf_canon"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                        d["valid_opt"], idx_f2k=d["idx_f2k"])
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    # and the run is reproducible bit for bit (fixed-order reductions, no atomics)
    c = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    assert torch.equal(a[0], c[0]) and torch.equal(a[1], c[1])


def test_unstaged_path_large_n(tracking, oracle):
    """More points than the shared-memory staging holds (148 x 6144): the kernel re-reads HBM."""
    cfg = dict(synth.TRACK_DEFAULTS, rel_error=0.0, delta_norm=0.0, max_iters=3)
    tracking.set_config({"tracking": cfg})
    t = synth.make_track_inputs(192, 256, seed=10)
    rep = 20  # 983 040 points
    big = {k: (t[k].repeat(rep, 1) if k in ("Xf", "Xk", "Qk", "valid_opt") else t[k]) for k in t}
    d = _cuda(big)
    Tw, Tk = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    # replicating every point scales H and g alike: same steps as the small problem
    To, Tko, _, _ = oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                           t["valid_opt"], cfg)
    assert _rel(Tw, To) < TOL and _rel(Tk, Tko) < TOL


@pytest.mark.parametrize("n_target", [148 * 5500, 148 * 5700, 148 * 6200])
def test_staging_capacity_boundary(tracking, oracle, n_target):
    """Point counts just below and above what the shared-memory staging holds (about 5.6k points
    per CTA next to the 47 KB of reduction buffers): both sides must launch and agree."""
    cfg = dict(synth.TRACK_DEFAULTS, rel_error=0.0, delta_norm=0.0, max_iters=2)
    tracking.set_config({"tracking": cfg})
    t = synth.make_track_inputs(96, 128, seed=12)
    rep = -(-n_target // t["n"])
    d = _cuda({k: (t[k].repeat(rep, 1)[:n_target].contiguous() if k in ("Xf", "Xk", "Qk", "valid_opt") else t[k])
               for k in t})
    Tw, Tk = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    assert tracking.last_info.status == 0 and tracking.last_info.iterations == 2
    c = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in d.items()}
    To, Tko, _, _ = oracle.track_pose_rays(c["Xf"], c["Xk"], c["T_WCf_init"], c["T_WCk"], c["Qk"], c["valid_opt"], cfg)
    assert _rel(Tw, To) < TOL and _rel(Tk, Tko) < TOL


def test_degenerate_and_errors(tracking):
    tracking.set_config({"tracking": dict(synth.TRACK_DEFAULTS)})
    d = _cuda(synth.make_track_inputs(24, 32, seed=11))
    none = torch.zeros_like(d["valid_opt"])
    with pytest.raises(torch.linalg.LinAlgError):  # H = 0: the reference's cholesky raises too
        tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], none)
    assert tracking.last_info.status == 1 and tracking.last_info.iterations == 1
    # check=False: no host sync, no exception; the pose comes back unchanged
    Tw, Tk = tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], none,
                                             check=False)
    assert (Tw - d["T_WCf_init"]).abs().max() < 1e-6
    with pytest.raises(RuntimeError):
        tracking.opt_pose_ray_dist_sim3(d["Xf"].cpu(), d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    with pytest.raises(RuntimeError):
        tracking.opt_pose_ray_dist_sim3(d["Xf"][:-1], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], d["valid_opt"])
    with pytest.raises(RuntimeError):
        tracking.opt_pose_ray_dist_sim3(d["Xf"].double(), d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"],
                                        d["valid_opt"])
    # a single valid point cannot constrain 7 degrees of freedom
    one = none.clone()
    one[5] = True
    with pytest.raises(torch.linalg.LinAlgError):
        tracking.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf_init"], d["T_WCk"], d["Qk"], one)
