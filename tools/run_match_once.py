"""A few 512x384 matching passes (iter_proj -> long -> refine_matches) for profiling."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch, bench, synth
import mast3r_slam_backends as be
sets = bench.make_match_sets("cuda", 4)
M = synth.MATCH_DEFAULTS
for s in sets:
    bench.match_step(be, s, M)
torch.cuda.synchronize(); torch.cuda.profiler.start()
for s in sets:
    bench.match_step(be, s, M)
torch.cuda.synchronize(); torch.cuda.profiler.stop()
print("ok")
