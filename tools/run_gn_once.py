"""One 64-keyframe GN solve (2 iterations) for profiling."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch, bench, synth
g = bench.make_gn_graph(int(os.environ.get("KF", 64)), "cuda")
torch.cuda.synchronize(); torch.cuda.profiler.start()
r = bench.time_gn(g, synth.GN_DEFAULTS, os.environ.get("KIND", "rays"), 0, 1, iters=2, reps=1)
torch.cuda.profiler.stop()
print(r["linearize_ms"], r["solve_ms"])
