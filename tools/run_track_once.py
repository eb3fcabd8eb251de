"""One tracking pose solve at 512x384 (ITERS fixed iterations) for profiling."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch, synth
from mast3r_slam_backends import tracking as tr
kind = os.environ.get("KIND", "rays")
tr.set_config({"tracking": dict(synth.TRACK_DEFAULTS, rel_error=0.0, delta_norm=0.0,
                                max_iters=int(os.environ.get("ITERS", 20)))})
t = synth.make_track_inputs(384, 512, seed=3, device="cuda")
def run():
    if kind == "rays":
        return tr.opt_pose_ray_dist_sim3(t["Xf_canon"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                         t["valid_opt"], idx_f2k=t["idx_f2k"], check=False)
    return tr.opt_pose_calib_sim3(t["Xf_canon"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"],
                                  t["meas_k"], t["valid_meas_k"], t["K"], (384, 512), idx_f2k=t["idx_f2k"], check=False)
for _ in range(3):
    run()
torch.cuda.synchronize(); torch.cuda.profiler.start()
run()
torch.cuda.synchronize(); torch.cuda.profiler.stop()
