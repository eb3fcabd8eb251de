#!/usr/bin/env python
"""Kernel micro-timings on the B200 (CUDA events, L2-cold rotation): variants side by side."""
import os, sys, statistics
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("lightweight-mast3r-slam_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
sys.path.insert(0, ROOT)
import torch
import bench, synth
import mast3r_slam_backends as be

M = synth.MATCH_DEFAULTS


def timeit(fn, sets, reps=40, warm=5):
    for i in range(warm):
        fn(sets[i % len(sets)])
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps)]
    for i in range(reps):
        ev[i][0].record(); fn(sets[i % len(sets)]); ev[i][1].record()
    torch.cuda.synchronize()
    t = [a.elapsed_time(b) for a, b in ev]
    return statistics.median(t) * 1e3, min(t) * 1e3


def main():
    sets = bench.make_match_sets("cuda", 8)
    for s in sets:
        p, _ = be.iter_proj(s["rays"], s["pts"], s["p_init"], 10, 1e-8, 1e-6)
        s["p1"] = p.long()
    for v in (0, 1):
        med, mn = timeit(lambda s: be.iter_proj(s["rays"], s["pts"], s["p_init"], 10, 1e-8, 1e-6, _variant=v), sets)
        print(f"iter_proj variant {v}: median {med:.1f} us, min {mn:.1f} us")
    for v in (0, 3, 2, 1):
        med, mn = timeit(lambda s: be.refine_matches(s["D11"], s["D21"], s["p1"], 3, 5, _variant=v), sets)
        print(f"refine_matches variant {v}: median {med:.1f} us, min {mn:.1f} us")
    for extra in sys.argv[1:]:
        pass


if __name__ == "__main__" and len(sys.argv) == 1:
    main()


def coop_profile():
    """Per-phase cycles of the cooperative Cholesky/solve kernel (CTA 0)."""
    import ctypes
    from mast3r_slam_backends import _lib
    D = synth.GN_DEFAULTS
    g = bench.make_gn_graph(int(os.environ.get("KF", 64)), "cuda")
    out = (ctypes.c_longlong * 16)()
    st = (ctypes.c_longlong * 16)()
    _lib.lib.m3r_debug_coop_profile(out)
    _lib.lib.m3r_debug_coop_stamps(st)
    r = bench.time_gn(g, D, "rays", 0, 1, reps=1)
    _lib.lib.m3r_debug_coop_profile(out)
    _lib.lib.m3r_debug_coop_stamps(st)
    n = max(out[6], 1)
    names = ["phaseA", "sync1", "phaseB", "sync2", "backward", "finish", "launches", "-", "A:loads", "A:diag update",
             "A:potrf", "T:row loads + fold", "T:wait for L11", "A:wait for the TRSM warps", "T:trsm + row stores",
             "T:wait y_k + rhs update"]
    print(f"coop solver: {out[6]} launches, solve_ms {r['solve_ms']:.3f}, linearize_ms {r['linearize_ms']:.3f}")
    for k, nm in enumerate(names):
        if nm not in ("launches", "-"):
            print(f"  {nm:18s} {out[k] / n / 1.965e3:8.1f} us per launch")


    panels = max(out[6], 1) * ((7 * (int(os.environ.get("KF", 64)) - 1) + 31) // 32)
    stamps = ["chain starts", "chain done", "scaling warp starts", "scaling warp done", "TRSM warp: fold done",
              "TRSM warp: TRSM done", "TRSM warp: rows staged", "y_k done", "inverse done", "CTA barrier passed",
              "store warp: rows are staged", "store warp: rows stored", "TRSM warp: rhs updated"]
    print("  timeline of one panel step in CTA 0 (us since the step began, mean over the panels):")
    for k, nm in enumerate(stamps):
        print(f"    {nm:26s} {st[k] / panels / 1.965e3:6.2f}")


if __name__ == "__main__" and "coop" in sys.argv:
    coop_profile()


def cluster_profile():
    """Per-phase cycles of the cluster Cholesky/solve kernel (CTA 0)."""
    import ctypes
    from mast3r_slam_backends import _lib
    D = synth.GN_DEFAULTS
    g = bench.make_gn_graph(int(os.environ.get("KF", 64)), "cuda")
    out = (ctypes.c_longlong * 9)()
    _lib.lib.m3r_debug_cluster_profile(out)
    r = bench.time_gn(g, D, "rays", 0, 1, reps=1)
    _lib.lib.m3r_debug_cluster_profile(out)
    n = max(out[8], 1)
    names = ["load", "potrf+wait", "trsm", "barrier2", "syrk", "barrier3", "backward", "finish"]
    print(f"cluster solver: {out[8]} launches, solve_ms {r['solve_ms']:.3f}, linearize_ms {r['linearize_ms']:.3f}")
    for k, nm in enumerate(names):
        print(f"  {nm:11s} {out[k] / n / 1.965e3:8.1f} us per launch")


if __name__ == "__main__" and "cluster" in sys.argv:
    cluster_profile()
