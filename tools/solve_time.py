"""GN phase times (begin / linearize / solve per iteration) for the three graph sizes of BASELINE's configs."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests", "."):
    sys.path.insert(0, os.path.join(ROOT, p))
import torch, bench, synth
D = synth.GN_DEFAULTS
for N, kind in ((16, "rays"), (64, "rays"), (256, "points")):
    if len(sys.argv) > 1 and str(N) not in sys.argv[1:]:
        continue
    g = bench.make_gn_graph(N, "cuda")
    r = bench.time_gn(g, D, kind, 0, 1, reps=3)
    print(f"N={N} {kind}: ms/iter {r['total_ms']/10:.4f} begin {r['begin_ms']:.3f} linearize {r['linearize_ms']:.4f} solve {r['solve_ms']:.4f}"
          f" pose_err_vs_truth {(r['Twc'] - g['Twc_true']).abs().max().item():.3e}")
    del g; torch.cuda.empty_cache()
