"""64-keyframe GN timing (bench.time_gn) with the library named by M3R_B200_LIB."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch, bench, synth
for kind in os.environ.get("KINDS", "rays").split(","):
    g = bench.make_gn_graph(int(os.environ.get("KF", 64)), "cuda")
    bench.time_gn(g, synth.GN_DEFAULTS, kind, 0, 1)
    r = bench.time_gn(g, synth.GN_DEFAULTS, kind, 0, 1)
    print(f"{os.path.basename(os.environ.get('M3R_B200_LIB', 'default'))} {kind}: linearize {r['linearize_ms']:.4f} solve {r['solve_ms']:.4f} "
          f"total/iter {r['total_ms'] / 10:.4f} ms")
