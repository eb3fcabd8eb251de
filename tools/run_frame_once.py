"""Two passes of the whole per-frame geometric pipeline at 512x384 (fused match, confidences,
tracking pose solve, keyframe update) for profiling."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch, bench, synth
sets = bench.make_match_sets("cuda", 4)
M = synth.MATCH_DEFAULTS
bench.time_frame_pipeline(sets, M, 4, 2)
torch.cuda.synchronize(); torch.cuda.profiler.start()
r = bench.time_frame_pipeline(sets, M, 2, 0)
torch.cuda.synchronize(); torch.cuda.profiler.stop()
print(r)
