#!/usr/bin/env python
"""Golden vectors for the tracking pose solve.

Runs the reference's OWN Python code for this path -- FrameTracker.opt_pose_ray_dist_sim3,
opt_pose_calib_sim3 and solve (mast3r_slam/tracker.py:157-266), with geometry.py and
nonlinear_optimizer.py, imported unmodified from /root/reference -- on CPU.  What cannot be
imported here is replaced: `lietorch` (CUDA-only pip dependency, not vendored) by
oracle.Sim3, and the two modules tracker.py imports but this path never calls
(mast3r_slam.frame, mast3r_slam.mast3r_utils: they pull the MASt3R network) by empty stubs.
The vectors pin oracle.track_pose_rays / track_pose_calib and the CUDA kernels.

Run in the build container (needs /root/reference):   python tests/golden/make_golden_track.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import m3r_oracle as oracle  # noqa: E402
import synth  # noqa: E402


def import_reference_tracker():
    lt = types.ModuleType("lietorch")
    lt.Sim3 = oracle.Sim3
    sys.modules["lietorch"] = lt
    fr = types.ModuleType("mast3r_slam.frame")
    fr.Frame = object
    mu = types.ModuleType("mast3r_slam.mast3r_utils")
    mu.mast3r_match_asymmetric = None
    sys.path.insert(0, "/root/reference")
    import mast3r_slam  # noqa: F401

    sys.modules["mast3r_slam.frame"] = fr
    sys.modules["mast3r_slam.mast3r_utils"] = mu
    import mast3r_slam.config as cfg
    import mast3r_slam.tracker as tracker

    cfg.config.update({"tracking": dict(synth.TRACK_DEFAULTS), "use_calib": False})
    return tracker


def main():
    tracker = import_reference_tracker()
    torch.set_num_threads(4)
    cases = {"a": (48, 64, 0, 0.03, {}), "b": (96, 128, 1, 0.06, {}),
             # fixed iteration count: both convergence tests off
             "c": (48, 64, 2, 0.03, dict(rel_error=0.0, delta_norm=0.0, max_iters=6))}
    for name, (h, w, seed, noise, over) in cases.items():
        t = synth.make_track_inputs(h, w, seed=seed, pose_noise=noise)
        cfg = dict(synth.TRACK_DEFAULTS, **over)
        self = types.SimpleNamespace(cfg=cfg)
        self.solve = types.MethodType(tracker.FrameTracker.solve, self)
        S = oracle.Sim3
        out = dict(h=h, w=w, **{k: np.asarray(v) for k, v in cfg.items()})
        for k in ("Xf", "Xk", "Qk", "valid_opt", "meas_k", "valid_meas_k", "K", "T_WCk", "T_WCf_init",
                  "T_WCf_true"):
            out[k] = t[k].numpy()
        Twf, Tkf = tracker.FrameTracker.opt_pose_ray_dist_sim3(
            self, t["Xf"], t["Xk"], S(t["T_WCf_init"].clone()), S(t["T_WCk"].clone()), t["Qk"], t["valid_opt"])
        out["ref_rays_T_WCf"], out["ref_rays_T_CkCf"] = Twf.data.numpy(), Tkf.data.numpy()
        Twf, Tkf = tracker.FrameTracker.opt_pose_calib_sim3(
            self, t["Xf"], t["Xk"], S(t["T_WCf_init"].clone()), S(t["T_WCk"].clone()), t["Qk"], t["valid_opt"],
            t["meas_k"], t["valid_meas_k"], t["K"], (h, w))
        out["ref_calib_T_WCf"], out["ref_calib_T_CkCf"] = Twf.data.numpy(), Tkf.data.numpy()
        np.savez_compressed(os.path.join(HERE, f"track_{name}.npz"), **out)
        err0 = (t["T_WCf_init"] - t["T_WCf_true"]).abs().max().item()
        err1 = (torch.from_numpy(out["ref_rays_T_WCf"]) - t["T_WCf_true"]).abs().max().item()
        err2 = (torch.from_numpy(out["ref_calib_T_WCf"]) - t["T_WCf_true"]).abs().max().item()
        print(f"wrote track_{name}: pose error {err0:.3e} -> rays {err1:.3e}, calib {err2:.3e}")


if __name__ == "__main__":
    main()
