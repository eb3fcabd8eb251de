#!/usr/bin/env python
"""Time refine_matches (tiled f16 path) with the library named by M3R_B200_LIB: median / mean / min."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch
import bench
import mast3r_slam_backends as be

sets = bench.make_match_sets("cuda", 8)
for s in sets:
    p, _ = be.iter_proj(s["rays"], s["pts"], s["p_init"], 10, 1e-8, 1e-6)
    s["p1"] = p.long()
for i in range(10):
    be.refine_matches(sets[i % 8]["D11"], sets[i % 8]["D21"], sets[i % 8]["p1"], 3, 5)
torch.cuda.synchronize()
reps = 200
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
for i in range(reps):
    s = sets[i % 8]
    ev[i][0].record(); be.refine_matches(s["D11"], s["D21"], s["p1"], 3, 5); ev[i][1].record()
torch.cuda.synchronize()
t = [a.elapsed_time(b) * 1e3 for a, b in ev]
# back-to-back (the way bench.py's value sees it)
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for i in range(reps):
    s = sets[i % 8]
    be.refine_matches(s["D11"], s["D21"], s["p1"], 3, 5)
b.record(); torch.cuda.synchronize()
print(f"{os.environ.get('M3R_B200_LIB', 'default')}: median {statistics.median(t):.1f} mean {statistics.mean(t):.1f} "
      f"min {min(t):.1f} us; back-to-back {a.elapsed_time(b) * 1e3 / reps:.1f} us per call")
