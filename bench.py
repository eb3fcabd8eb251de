#!/usr/bin/env python
"""Benchmark of the geometric hot path (BASELINE.json metric: matched px/s + GN iters/s at
512x384 with a 64-keyframe graph).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One JSON line on stdout (rank 0).  A "step" is one matching pass over one synthetic 512x384
frame exactly as mast3r_slam/matching.py:60-85 issues it: iter_proj -> p.long() ->
refine_matches (BASELINE config 2, b = 1).  `value` is matched px/s with inputs resident in
HBM; `e2e` is the same pass through the drop-in module with pinned HOST buffers (H2D of the
five inputs and D2H of the matches inside the timed region).  The Gauss-Newton half of the
metric is reported in the `gn` object of the same line (64-keyframe gauss_newton_rays,
max_iter = 10, delta_thresh = 0 so that all 10 iterations run); at N > 1 its edges are
sharded over the ranks with one all-reduce per iteration, matching runs as replicas.

`--impl reference` times the CPU restatement of the reference path (oracle/, all host
threads) on the same workload, rank 0 only.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (os.path.join(ROOT, "lightweight-mast3r-slam_b200"), os.path.join(ROOT, "oracle"),
           os.path.join(ROOT, "tests")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import torch  # noqa: E402

H, W, FDIM = 384, 512, 24
N_PTS = H * W
NUM_SETS = 8  # input sets rotated through the timed region: 8 x 38 MB > 126 MB of L2
# algorithmic bytes per point
B_ITER_PROJ, B_REFINE, B_GN = 65, 128, 45


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons sampled DURING the timed region."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.path = index, None, None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1]))
            except ValueError:
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        os.unlink(self.path)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm))
        return out


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), \
        int(os.environ.get("WORLD_SIZE", 1))


def barrier(world):
    if world > 1:
        torch.distributed.barrier()


def max_over_ranks(x, world):
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return t.item()


# --------------------------------------------------------------------------------------------
# matching (BASELINE config 2)
# --------------------------------------------------------------------------------------------
def make_match_sets(device, num_sets, seed0=0):
    import synth

    sets = []
    for s in range(num_sets):
        m = synth.make_match_inputs(H, W, b=1, seed=seed0 + s, fdim=FDIM, device=device)
        rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"])
        sets.append(dict(rays=rays, pts=pts, p_init=p0, D11=m["D11"].half().contiguous(),
                         D21=m["D21"].view(1, N_PTS, FDIM).half().contiguous()))
    return sets


def match_step(be, s, M):
    p, conv = be.iter_proj(s["rays"], s["pts"], s["p_init"], M["max_iter"], M["lambda_init"],
                           M["convergence_thresh"])
    p1 = p.long()
    (p1n,) = be.refine_matches(s["D11"], s["D21"], p1, M["radius"], M["dilation_max"])
    return p1n, conv


def time_matching(be, sets, M, steps, warmup, world, per_kernel=True):
    """K steps back to back between one pair of CUDA events on the launching stream; extra
    events around each kernel give the per-kernel durations over the same timed region."""
    for i in range(warmup):
        match_step(be, sets[i % len(sets)], M)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(steps)]
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(steps):
        s = sets[i % len(sets)]
        if per_kernel:
            ev[i][0].record()
        p, conv = be.iter_proj(s["rays"], s["pts"], s["p_init"], M["max_iter"], M["lambda_init"],
                               M["convergence_thresh"])
        if per_kernel:
            ev[i][1].record()
        p1 = p.long()
        if per_kernel:
            ev[i][2].record()
        (p1n,) = be.refine_matches(s["D11"], s["D21"], p1, M["radius"], M["dilation_max"])
        if per_kernel:
            ev[i][3].record()
    t1.record()
    torch.cuda.synchronize()
    barrier(world)
    total_ms = max_over_ranks(t0.elapsed_time(t1), world)
    out = {"total_ms": total_ms}
    if per_kernel:
        out["iter_proj_ms"] = statistics.mean(e[0].elapsed_time(e[1]) for e in ev)
        out["cast_ms"] = statistics.mean(e[1].elapsed_time(e[2]) for e in ev)
        out["refine_ms"] = statistics.mean(e[2].elapsed_time(e[3]) for e in ev)
    return out


def time_matching_e2e(be, sets, M, steps, warmup, world):
    """Same pass through the public module with pinned HOST buffers: H2D of the five inputs
    and D2H of the matches (p1_new, converged) inside the timed region, every step."""
    host = [{k: v.cpu().pin_memory() for k, v in s.items()} for s in sets[:4]]
    out_p = torch.empty((1, N_PTS, 2), dtype=torch.int64).pin_memory()
    out_c = torch.empty((1, N_PTS), dtype=torch.bool).pin_memory()
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())
    d2h = out_p.numel() * 8 + out_c.numel()

    def step(hs):
        dev = {k: v.cuda(non_blocking=True) for k, v in hs.items()}
        p1n, conv = match_step(be, dev, M)
        out_p.copy_(p1n, non_blocking=True)
        out_c.copy_(conv, non_blocking=True)

    for i in range(warmup):
        step(host[i % len(host)])
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(steps):
        step(host[i % len(host)])
    t1.record()
    torch.cuda.synchronize()
    barrier(world)
    ms = max_over_ranks(t0.elapsed_time(t1), world)
    return ms, h2d, d2h
