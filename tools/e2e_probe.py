"""Probe: streamed e2e of the five-export path vs the fused match, more steps, CPU-side loop time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, torch
import synth
import mast3r_slam_backends as be
M = synth.MATCH_DEFAULTS
sets = bench.make_match_sets("cuda", 8)
for steps in (20, 100):
    a = bench.time_matching_e2e(be, sets, M, steps, 3, 1)
    t0 = time.perf_counter()
    f = bench.time_front_e2e(sets, M, steps, 3, 1)
    print(steps, "five-export streamed ms/step", a[0] / steps, "serial", a[1] / steps, "fused streamed", f["ms_per_step"],
          "wall per step incl warmup", (time.perf_counter() - t0) / (steps + 3) * 1e3)
