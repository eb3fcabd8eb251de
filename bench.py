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


def traffic(kernel):
    """Measured DRAM bytes per launch of `kernel` (ncu --set full, profiles/traffic.json)."""
    try:
        return float(json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[kernel]["bytes"])
    except Exception:
        return None


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
        if s < 4:  # raw pointmaps for the whole-front-end timing (time_front_end)
            FRONT_INPUTS.append((m["X11"].contiguous(), m["X21"].contiguous()))
    return sets


FRONT_INPUTS = []


def time_front_end(be, sets, M, steps, warmup):
    """mast3r_slam/matching.py::match as a whole (SURVEY.md 8f rows N1+N3), inputs resident:
    the fused C-ABI call (3 launches) next to the reference-shaped sequence -- the torch glue of
    matching.py:25-88 around this library's two kernels."""
    import synth
    from mast3r_slam_backends import matching as front

    front.set_config({"matching": dict(M)})

    def fused(i):
        X11, X21 = FRONT_INPUTS[i % len(FRONT_INPUTS)]
        s = sets[i % len(FRONT_INPUTS)]
        return front.match(X11, X21, s["D11"], s["D21"])

    def unfused(i):
        X11, X21 = FRONT_INPUTS[i % len(FRONT_INPUTS)]
        s = sets[i % len(FRONT_INPUTS)]
        rays, pts, p0 = synth.prep_for_iter_proj(X11, X21)
        p, conv = be.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
        p1 = p.long()
        bi = torch.arange(1, device=X11.device)[:, None].repeat(1, N_PTS)
        d = torch.linalg.norm(X11[bi, p1[..., 1], p1[..., 0], :].reshape(1, H, W, 3) - X21, dim=-1)
        valid = conv & (d < M["dist_thresh"]).view(1, -1)
        (p1n,) = be.refine_matches(s["D11"].half(), s["D21"].half(), p1, M["radius"], M["dilation_max"])
        return p1n[..., 0] + W * p1n[..., 1], valid.unsqueeze(-1)

    out = {}
    for name, fn in (("fused_ms", fused), ("torch_glue_ms", unfused)):
        for i in range(warmup):
            fn(i)
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for i in range(steps):
            fn(i)
        t1.record()
        torch.cuda.synchronize()
        out[name] = t0.elapsed_time(t1) / steps
    out["fused_launches"] = 4  # prep_rays, match_proj, refine_plan, refine
    out["fused_px_per_s"] = N_PTS / (out["fused_ms"] * 1e-3)
    return out


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
    if not per_kernel:
        # The device idles ~1 ms in FRONT of the start event while the host enqueues the first steps:
        # the timed region (between the two events, on the device) then starts with work queued, as
        # in a streaming caller, instead of exposing the host's first launches -- with 8 ranks on one
        # box a single descheduled rank otherwise sets the max over ranks of a 3 ms region.
        torch.cuda._sleep(2_000_000)
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
    mine = t0.elapsed_time(t1)
    total_ms = max_over_ranks(mine, world)
    out = {"total_ms": total_ms, "rank_ms": mine}
    if per_kernel:
        out["iter_proj_ms"] = statistics.mean(e[0].elapsed_time(e[1]) for e in ev)
        out["cast_ms"] = statistics.mean(e[1].elapsed_time(e[2]) for e in ev)
        out["refine_ms"] = statistics.mean(e[2].elapsed_time(e[3]) for e in ev)
    return out


def _pipelined(host, outs, run, steps, warmup, world):
    """Streamed frames: a copy-in stream uploads step i+1 from pinned host memory into the
    other half of a double buffer while the compute stream works on step i and a copy-out
    stream downloads step i-1's result.  Every H2D and D2H byte is inside the timed region;
    the events are recorded on the compute stream, which joins the two copy streams first."""
    cur = torch.cuda.current_stream()
    s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
    dev = [{k: torch.empty_like(v, device="cuda") for k, v in host[0].items()} for _ in range(2)]
    dev_out = [[torch.empty_like(o, device="cuda") for o in outs] for _ in range(2)]
    ev_in = [torch.cuda.Event() for _ in range(2)]
    ev_done = [torch.cuda.Event() for _ in range(2)]
    ev_out = [torch.cuda.Event() for _ in range(2)]

    def loop(count):
        for i in range(count):
            k = i & 1
            with torch.cuda.stream(s_in):
                if i >= 2:
                    s_in.wait_event(ev_done[k])  # the kernels that read this half are finished
                for name, v in host[i % len(host)].items():
                    dev[k][name].copy_(v, non_blocking=True)
                ev_in[k].record(s_in)
            cur.wait_event(ev_in[k])
            if i >= 2:
                cur.wait_event(ev_out[k])  # the result half is free again
            res = run(dev[k])
            for d, r in zip(dev_out[k], res):
                d.copy_(r.view(d.shape), non_blocking=True)
            ev_done[k].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_done[k])
                for o, d in zip(outs, dev_out[k]):
                    o.copy_(d, non_blocking=True)
                ev_out[k].record(s_out)
        cur.wait_stream(s_in)
        cur.wait_stream(s_out)

    loop(warmup)
    torch.cuda.synchronize()
    barrier(world)
    torch.cuda.synchronize()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    loop(steps)
    t1.record()
    torch.cuda.synchronize()
    barrier(world)
    return max_over_ranks(t0.elapsed_time(t1), world)


def time_matching_e2e(be, sets, M, steps, warmup, world):
    """Same pass through the public module with pinned HOST buffers: H2D of the five inputs
    and D2H of the matches (p1_new, converged) inside the timed region, every step.  Returns
    (pipelined ms, serial ms, h2d bytes, d2h bytes): `serial` issues copy-in, kernels and
    copy-out of a step back to back on one stream; `pipelined` streams the frames."""
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
    serial_ms = max_over_ranks(t0.elapsed_time(t1), world)
    pipe_ms = _pipelined(host, [out_p, out_c], lambda d: match_step(be, d, M), steps, warmup, world)
    return pipe_ms, serial_ms, h2d, d2h


def time_front_e2e(sets, M, steps, warmup, world):
    """The fused `matching.match` from pinned host buffers (X11, X21, D11, D21 in; idx, valid
    out), streamed like time_matching_e2e: fewer bytes cross PCIe than through the five-export
    boundary because the ray image and the normalised points are produced on the device."""
    from mast3r_slam_backends import matching as front

    front.set_config({"matching": dict(M)})
    host = []
    for i in range(len(FRONT_INPUTS)):
        X11, X21 = FRONT_INPUTS[i]
        host.append({"X11": X11.cpu().pin_memory(), "X21": X21.cpu().pin_memory(),
                     "D11": sets[i]["D11"].cpu().pin_memory(), "D21": sets[i]["D21"].cpu().pin_memory()})
    out_i = torch.empty((1, N_PTS), dtype=torch.int64).pin_memory()
    out_v = torch.empty((1, N_PTS, 1), dtype=torch.bool).pin_memory()
    h2d = sum(v.numel() * v.element_size() for v in host[0].values())
    d2h = out_i.numel() * 8 + out_v.numel()
    ms = _pipelined(host, [out_i, out_v], lambda d: front.match(d["X11"], d["X21"], d["D11"], d["D21"]),
                    steps, warmup, world)
    return {"value": world * N_PTS * steps / (ms * 1e-3), "unit": "px/s", "ms_per_step": ms / steps,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h}


def time_frame_pipeline(sets, M, steps, warmup):
    """BASELINE config 2 as a whole, the geometric part of FrameTracker.track (tracker.py:29-112)
    for one 512x384 frame: match -> confidences/fractions -> pose solve -> keyframe pointmap
    update with the solved relative pose.  Six launches, no host synchronisation in between
    (the pose travels from the solve to the update in device memory)."""
    import synth
    from mast3r_slam_backends import _lib, matching as front, pointmaps, tracking as tr

    T = synth.TRACK_DEFAULTS
    front.set_config({"matching": dict(M)})
    tr.set_config({"tracking": dict(T)})
    t = synth.make_track_inputs(H, W, seed=3, device="cuda", pose_noise=0.03)
    Xk_canon, Ck_sum = t["Xk"].clone(), t["Ck"].clone()
    Xkf, Ckf = t["Xk"].clone(), t["Ck"].clone()

    def frame(i):
        X11, X21 = FRONT_INPUTS[i % len(FRONT_INPUTS)]
        s = sets[i % len(FRONT_INPUTS)]
        idx, valid = front.match(X11, X21, s["D11"], s["D21"])
        Qk, valid_kf, valid_opt, counts = front.match_confidence(
            idx, t["Qff"][None], t["Qkf"][None], valid, T["Q_conf"], t["Cf_canon"][None], t["Ck"][None],
            T["C_conf"], count_unique=True)
        T_WCf, T_CkCf = tr.opt_pose_ray_dist_sim3(t["Xf_canon"], t["Xk"], t["T_WCf_init"], t["T_WCk"], Qk[0],
                                                  valid_opt[0], idx_f2k=idx[0], check=False)
        pointmaps.update_pointmap(Xk_canon, Ck_sum, Xkf, Ckf, "weighted_pointmap", T=T_CkCf)
        return counts

    def torch_epilogues(idx, valid):
        """tracker.py:41,59-67,104-110 and frame.py:74-77 as the reference issues them (torch)."""
        i1 = idx[0]
        Qk = torch.sqrt(t["Qff"][i1] * t["Qkf"])
        vm = valid[0]
        valid_opt = vm & (t["Cf_canon"][i1] > T["C_conf"]) & (t["Ck"] > T["C_conf"]) & (Qk > T["Q_conf"])
        valid_kf = vm & (Qk > T["Q_conf"])
        fr = (valid_opt.sum() / valid_opt.numel(), valid_kf.sum() / valid_kf.numel())
        uniq = torch.unique(i1[vm[:, 0]]).shape[0]  # host sync, as in the reference
        Xn = ((Ck_sum * Xk_canon) + (Ckf * Xkf)) / (Ck_sum + Ckf)
        return fr, uniq, Xn, Ck_sum + Ckf

    idx_w, valid_w = front.match(*FRONT_INPUTS[0], sets[0]["D11"], sets[0]["D21"])
    for _ in range(3):
        torch_epilogues(idx_w, valid_w)
    torch.cuda.synchronize()
    tt0 = time.perf_counter()
    for _ in range(10):
        torch_epilogues(idx_w, valid_w)
    torch.cuda.synchronize()
    torch_epi_ms = (time.perf_counter() - tt0) / 10 * 1e3

    for i in range(warmup):
        frame(i)
    torch.cuda.synchronize()
    l0 = _lib.launch_count()
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for i in range(steps):
        frame(i)
    t1.record()
    torch.cuda.synchronize()
    ms = t0.elapsed_time(t1) / steps
    return {"ms_per_frame": ms, "frames_per_s": 1e3 / ms, "gpu_launches_per_frame": (_lib.launch_count() - l0) / steps,
            "stages": "match(4) + match_confidence(1) + track_pose(1) + update_pointmap(1); memsets not counted",
            "torch_confidence_and_update_ms": torch_epi_ms,
            "torch_note": "the reference's torch lines for the confidence/fraction/unique epilogue and the weighted "
                          "keyframe update alone (wall clock, includes torch.unique's host sync); here they are two "
                          "of the six launches"}


def time_tracking(steps, warmup, cpu=True):
    """The per-frame pose solve of BASELINE config 2 (tracker.py:173-266) at
    512x384: one cooperative launch per call.  `us_per_iter` comes from a fixed-iteration run
    (stop tests off) so launch + staging and the per-iteration cost separate."""
    import synth
    from mast3r_slam_backends import tracking as tr

    t = synth.make_track_inputs(H, W, seed=3, device="cuda", pose_noise=0.03)
    out = {}

    def run(kind, cfg, reps):
        tr.set_config({"tracking": cfg})
        if kind == "rays":
            fn = lambda: tr.opt_pose_ray_dist_sim3(t["Xf_canon"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                   t["valid_opt"], idx_f2k=t["idx_f2k"], check=False)
        else:
            fn = lambda: tr.opt_pose_calib_sim3(t["Xf_canon"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                t["valid_opt"], t["meas_k"], t["valid_meas_k"], t["K"], (H, W),
                                                idx_f2k=t["idx_f2k"], check=False)
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(reps):
            fn()
        t1.record()
        torch.cuda.synchronize()
        return t0.elapsed_time(t1) / reps

    base = dict(synth.TRACK_DEFAULTS)
    for kind in ("rays", "calib"):
        ms = run(kind, base, steps)
        # iteration count of the default configuration (one checked call)
        tr.set_config({"tracking": base})
        if kind == "rays":
            tr.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
        else:
            tr.opt_pose_calib_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"],
                                   t["meas_k"], t["valid_meas_k"], t["K"], (H, W))
        iters = tr.last_info.iterations
        fixed = dict(base, rel_error=0.0, delta_norm=0.0)
        ms5 = run(kind, dict(fixed, max_iters=5), steps)
        ms45 = run(kind, dict(fixed, max_iters=45), steps)
        out[kind] = {"ms_per_call": ms, "iterations": iters, "us_per_iter": (ms45 - ms5) / 40 * 1e3,
                     "fixed_us": (ms5 - 5 * (ms45 - ms5) / 40) * 1e3, "gpu_launches_per_call": 1}
    tr.set_config({"tracking": base})
    # what the reference itself runs on a GPU: its torch loop (tracker.py:173-215) op for op,
    # two host syncs per iteration (synth.torch_track_pose_rays)
    for _ in range(2):
        synth.torch_track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        Tw_t, _, it_t = synth.torch_track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"],
                                                    t["valid_opt"])
    torch.cuda.synchronize()
    out["torch_gpu_rays_ms_per_call"] = (time.perf_counter() - t0) / reps * 1e3
    out["torch_gpu_iterations"] = it_t
    Tw_k, _ = tr.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
    out["kernel_vs_torch_gpu_pose_max_abs_diff"] = float((Tw_k - Tw_t).abs().max())
    if cpu:
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import m3r_oracle as oracle

        c = {k: (v.cpu() if torch.is_tensor(v) else v) for k, v in t.items()}
        t0 = time.perf_counter()
        _, _, it, _ = oracle.track_pose_rays(c["Xf"], c["Xk"], c["T_WCf_init"], c["T_WCk"], c["Qk"], c["valid_opt"])
        out["cpu_port_rays_ms_per_call"] = (time.perf_counter() - t0) * 1e3
        out["cpu_port_iterations"] = it
        out["cpu_port_threads"] = torch.get_num_threads()
    return out


# --------------------------------------------------------------------------------------------
# Gauss-Newton (64-keyframe graph; edges sharded over the ranks at N > 1)
# --------------------------------------------------------------------------------------------
def make_gn_graph(num_kf, device, seed=0):
    import synth

    return synth.make_graph(num_kf, H, W, seed=seed, device=device)


def gn_kwargs(g, D):
    return dict(K=g["K"], height=g["h"], width=g["w"], pixel_border=D["pixel_border"],
                z_eps=D["z_eps"], sigma_ray=D["sigma_ray"], sigma_dist=D["sigma_dist"],
                sigma_pixel=D["sigma_pixel"], sigma_depth=D["sigma_depth"],
                sigma_point=D["sigma_point"], C_thresh=D["C_thresh"], Q_thresh=D["Q_thresh"])


GN_CALLS = {
    "rays": lambda be, g, D, T, iters: be.gauss_newton_rays(
        T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], D["sigma_ray"],
        D["sigma_dist"], D["C_thresh"], D["Q_thresh"], iters, 0.0),
    "calib": lambda be, g, D, T, iters: be.gauss_newton_calib(
        T, g["Xs"], g["Cs"], g["K"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], g["h"], g["w"],
        D["pixel_border"], D["z_eps"], D["sigma_pixel"], D["sigma_depth"], D["C_thresh"], D["Q_thresh"], iters, 0.0),
    "points": lambda be, g, D, T, iters: be.gauss_newton_points(
        T, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"], g["Q"], D["sigma_point"],
        D["C_thresh"], D["Q_thresh"], iters, 0.0),
}


def time_gn(g, D, kind, rank, world, iters=10, reps=3, exchange="peer"):
    """One gauss_newton_<kind> solve of `iters` iterations (delta_thresh = 0), through the
    staged C ABI so that the accumulation/assembly half and the solve half of every iteration
    get their own CUDA events.  At world > 1 each rank linearises its edge shard and the fp64
    systems are summed over the ranks before the replicated solve: `exchange` = "peer" (this
    library's NVLink peer-memory kernels) or "nccl" (torch.distributed all_reduce).
    Returns the repetition with the smallest total (of `reps`, after one warm-up)."""
    from mast3r_slam_backends.staged import PeerExchange, StagedGN, TorchAllReduce, shard_range

    E = g["E"]
    b0, b1 = shard_range(E, rank, world)
    idx = g["idx_ii2jj"][b0:b1].contiguous()
    val = g["valid_match"][b0:b1].contiguous()
    Q = g["Q"][b0:b1].contiguous()
    ex = PeerExchange(g["N"]) if (world > 1 and exchange == "peer") else TorchAllReduce()
    best = None
    for rep in range(reps + 1):  # first repetition is the warm-up
        Twc = g["Twc_init"].clone()
        gn = StagedGN(kind, Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], idx, val, Q, edge_begin=b0,
                      edge_end=b1, **gn_kwargs(g, D))
        torch.cuda.synchronize()
        barrier(world)
        torch.cuda.synchronize()
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(iters)]
        t0, t1, tb = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        t0.record()
        gn.begin()
        tb.record()
        for i in range(iters):
            ev[i][0].record()
            gn.linearize(into=ex.round_buffer(gn))
            ev[i][1].record()
            ex.allreduce(gn)
            ev[i][2].record()
            gn.solve_retract(0.0)
            ev[i][3].record()
        t1.record()
        torch.cuda.synchronize()
        barrier(world)
        if rep == 0:
            continue
        total = max_over_ranks(t0.elapsed_time(t1), world)
        r = dict(total_ms=total, begin_ms=t0.elapsed_time(tb),
                 linearize_ms=statistics.mean(e[0].elapsed_time(e[1]) for e in ev),
                 allreduce_ms=statistics.mean(e[1].elapsed_time(e[2]) for e in ev),
                 solve_ms=statistics.mean(e[2].elapsed_time(e[3]) for e in ev),
                 iters_done=gn.iterations_done(), Twc=Twc)
        if best is None or r["total_ms"] < best["total_ms"]:
            best = r
    del ex
    return best


def time_gn_single_call(be, g, D, kind, iters=10, reps=3):
    """The drop-in export itself (one host call, no events inside); smallest of `reps`."""
    Twc = g["Twc_init"].clone()
    GN_CALLS[kind](be, g, D, Twc, iters)  # warm-up (allocations)
    best = 1e30
    for _ in range(reps):
        Twc = g["Twc_init"].clone()
        torch.cuda.synchronize()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        GN_CALLS[kind](be, g, D, Twc, iters)
        t1.record()
        torch.cuda.synchronize()
        best = min(best, t0.elapsed_time(t1))
    return best, Twc


def pose_groups(a, b):
    """Relative difference of Sim(3) states per component group (t against the largest
    translation, q against 1, s against each pose's scale): the 1e-4 criterion of north_star."""
    a, b = a.double(), b.double()
    return {"t": ((a[:, :3] - b[:, :3]).abs().max() / b[:, :3].abs().max().clamp_min(1e-12)).item(),
            "q": (a[:, 3:7] - b[:, 3:7]).abs().max().item(),
            "s": ((a[:, 7] - b[:, 7]).abs() / b[:, 7].abs()).max().item()}


def gn_parity(be, g, D, kind, Twc_sharded, rank, world, refmod):
    """Correctness fields of a GN line, at every N: the sharded result against the unsharded
    solve of the same library (N > 1), bitwise equality of the replicas (N > 1), the
    reference's own CUDA kernels on the same graph (N = 1, when oracle/_ref is present) and the
    distance to the poses the scene was generated from."""
    out = {"pose_err_vs_truth": (Twc_sharded - g["Twc_true"]).abs().max().item(),
           "tolerance_vs_reference": 1e-4}
    Twc_1 = None
    if rank == 0:  # rank 0 holds the whole graph
        Twc_1 = g["Twc_init"].clone()
        GN_CALLS[kind](be, g, D, Twc_1, 10)
        torch.cuda.synchronize()
    if world > 1:
        gathered = [torch.empty_like(Twc_sharded) for _ in range(world)]
        torch.distributed.all_gather(gathered, Twc_sharded.contiguous())
        if rank == 0:
            out["replicas_bitwise_equal"] = all(torch.equal(gathered[0], t) for t in gathered[1:])
            out["max_abs_diff_vs_1gpu"] = (Twc_sharded - Twc_1).abs().max().item()
            out["rel_diff_vs_1gpu"] = pose_groups(Twc_sharded, Twc_1)
    if rank == 0 and refmod is not None:
        Tr = g["Twc_init"].clone()
        t0 = time.perf_counter()
        GN_CALLS[kind](refmod, g, D, Tr, 10)
        torch.cuda.synchronize()
        out["reference_cuda_ms_per_iter"] = (time.perf_counter() - t0) * 100.0
        out["max_abs_diff_vs_reference_cuda"] = (Twc_1 - Tr).abs().max().item()
        out["rel_diff_vs_reference_cuda"] = pose_groups(Twc_1, Tr)
        out["within_tolerance"] = max(out["rel_diff_vs_reference_cuda"].values()) < 1e-4
    return out


_sp_seq = [0]


def time_gn_single_process(be, g, D, kind, rank, local_rank, world, Twc_spmd, reps=3):
    """The same graph through the UNCHANGED drop-in call from ONE process (rank 0) that drives all
    `world` devices (m3r_gn_multi_*, M3R_GN_DEVICES): what global_opt.py:142-143 would get.  The
    other ranks wait on the host (c10d store: no kernel on their GPUs), keeping their memory.
    `scatter_ms` is the per-call preparation of the non-primary devices: reading the caller's
    Xs / Cs and their shard's idx / valid / Q rows over NVLink + packing + compaction; it is paid on
    every call (the callers' tensors are re-stacked every solve, global_opt.py:114-121) and is
    inside `ms_per_iter`."""
    import datetime

    store = torch.distributed.distributed_c10d._get_default_store()
    _sp_seq[0] += 1
    key = f"m3r_single_process_done_{_sp_seq[0]}"
    torch.cuda.synchronize()
    barrier(world)
    out = None
    if rank == 0:
        try:
            be.set_gn_devices([(local_rank + k) % world for k in range(world)])
            calls = []
            for rep in range(reps + 1):  # the first call also allocates and enables peer access
                Twc = g["Twc_init"].clone()
                torch.cuda.synchronize()
                t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                t0.record()
                GN_CALLS[kind](be, g, D, Twc, 10)
                t1.record()
                torch.cuda.synchronize()
                r = be.last_multi_report(Twc)
                calls.append((t0.elapsed_time(t1), r, Twc))
            ms, r, Twc = min(calls[1:], key=lambda c: c[0])
            out = {"mode": f"one process, {world} devices (M3R_GN_DEVICES), caller's tensors on device {local_rank}",
                   "value": 10 / (ms * 1e-3), "unit": "iters/s", "ms_per_iter": ms / 10,
                   "first_call_ms": calls[0][0], "steady_call_ms": ms,
                   "scatter_ms": {"first_call_max_over_devices": max(calls[0][1]["begin_ms"][1:]),
                                  "steady_max_over_devices": max(r["begin_ms"][1:]),
                                  "steady_per_device": [round(x, 4) for x in r["begin_ms"]],
                                  "primary_begin_ms": r["begin_ms"][0]},
                   "iters_done": r["iterations"], "replicas_bitwise_equal": r["replicas_bitwise_equal"],
                   "max_abs_diff_vs_spmd": (Twc - Twc_spmd).abs().max().item(),
                   "stat": f"smallest of {reps} calls after the first"}
        except Exception as exc:  # report, do not lose the line
            out = {"error": str(exc)[:300]}
        finally:
            be.set_gn_devices(None)
            be.release_workspaces()
            store.set(key, "1")
    else:
        store.wait([key], datetime.timedelta(seconds=600))
    barrier(world)
    return out


def gn_report(be, g, D, kind, rank, world, peak, refmod, reps=3, nccl_too=True, single_call=True,
              single_process=False, local_rank=0):
    """One GN configuration: phase times (smallest-total repetition of `reps`), roofline of the
    linearisation, parity fields, and at N > 1 the NCCL variant beside the peer-memory one."""
    from mast3r_slam_backends import _lib

    l0 = _lib.launch_count()
    gr = time_gn(g, D, kind, rank, world, reps=reps)
    launches = (_lib.launch_count() - l0) // (reps + 1)
    E, n, N = g["E"], g["n"], g["N"]
    gn_bytes = B_GN * E * n + 840 * E + 32 * N
    lin_gbs = gn_bytes / world / (gr["linearize_ms"] * 1e-3) / 1e9
    out = {
        "metric": f"GN iters/s (gauss_newton_{kind}, {N}-keyframe graph, {E} directed edges, {g['w']}x{g['h']})",
        "value": 10 / (gr["total_ms"] * 1e-3), "unit": "iters/s", "scaling": "strong",
        "stat": f"smallest total of {reps} repetitions of one 10-iteration call (after 1 warm-up call); "
                "phase times are means over its 10 iterations",
        "ms_per_iter": gr["total_ms"] / 10, "begin_ms": gr["begin_ms"], "linearize_ms": gr["linearize_ms"],
        "allreduce_ms": gr["allreduce_ms"], "solve_ms": gr["solve_ms"], "iters_done": gr["iters_done"],
        "gpu_launches": int(launches), "system_unknowns": 7 * (N - 1),
        "edges_per_rank": E // world,
        "exchange": "none" if world == 1 else "NVLink peer memory (m3r_gn_comm_*), fp64 system",
        "roofline": {"bound": "hbm", "kernel": f"gn_accumulate2_kernel<{kind}> (+ edge finalize + assemble)",
                     "achieved": lin_gbs, "peak": peak, "unit": "GB/s", "frac": lin_gbs / peak,
                     "traffic": traffic("gn_accumulate2_kernel") if (kind == "rays" and N == 64 and world == 1) else None,
                     "bytes_per_iter": gn_bytes,
                     "note": "algorithmic = 45 B x E x n per iteration (SURVEY 8d) / ranks; the kernel reads "
                             "compacted lists and hits L2 on the pointmaps, so DRAM traffic is below it",
                     "whole_iter_frac": gn_bytes / world / (gr["total_ms"] / 10 * 1e-3) / 1e9 / peak},
    }
    out["parity"] = gn_parity(be, g, D, kind, gr["Twc"], rank, world, refmod)
    if world > 1 and nccl_too:
        gn2 = time_gn(g, D, kind, rank, world, reps=2, exchange="nccl")
        out["nccl_variant"] = {"ms_per_iter": gn2["total_ms"] / 10, "allreduce_ms": gn2["allreduce_ms"],
                               "max_abs_diff_vs_peer": (gn2["Twc"] - gr["Twc"]).abs().max().item()}
    if world == 1 and single_call:
        ms_call, _ = time_gn_single_call(be, g, D, kind)
        out["single_call_ms_per_iter"] = ms_call / 10
    out["mode"] = "one process per GPU (torchrun), edges sharded over the ranks" if world > 1 else "one GPU"
    if world > 1 and single_process:
        out["single_process"] = time_gn_single_process(be, g, D, kind, rank, local_rank, world, gr["Twc"])
    return out


# --------------------------------------------------------------------------------------------
# CPU baseline (oracle port) and the reference's CUDA extension on the same GPU
# --------------------------------------------------------------------------------------------
def cpu_match_pass(oracle, s, M):
    p, conv = oracle.iter_proj(s["rays"], s["pts"], s["p_init"], M["max_iter"], M["lambda_init"],
                               M["convergence_thresh"])
    p1 = p.long()
    (p1n,) = oracle.refine_matches(s["D11"], s["D21"], p1, M["radius"], M["dilation_max"])
    return p1n


def cpu_baseline_matching(M, steps=1, warmup=0, seed=0):
    """The oracle (CPU restatement of the reference kernels, OpenMP over all host threads) on
    whole 512x384 frames of the same workload."""
    import m3r_oracle as oracle

    oracle.build()
    # all host threads, whatever the launcher put in OMP_NUM_THREADS (torchrun sets it to 1)
    oracle.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count())
    sets = make_match_sets("cpu", 1, seed0=seed)
    for _ in range(warmup):
        cpu_match_pass(oracle, sets[0], M)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_match_pass(oracle, sets[0], M)
    dt = time.perf_counter() - t0
    return N_PTS * steps / dt, dt / steps, oracle.num_threads()


def cpu_baseline_gn(D, num_kf=8, iters=1):
    """Bounded GN sample: one iteration on an 8-keyframe 512x384 graph (same per-edge work as
    the 64-keyframe graph; the CPU cost is linear in the number of directed edges)."""
    import m3r_oracle as oracle
    import synth

    oracle.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count())
    g = synth.make_graph(num_kf, H, W, seed=0, device="cpu")
    P = oracle.make_params("rays", sigma_ray=D["sigma_ray"], sigma_dist=D["sigma_dist"],
                           C_thresh=D["C_thresh"], Q_thresh=D["Q_thresh"])
    t0 = time.perf_counter()
    oracle.gauss_newton(P, g["Twc_init"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                        g["valid_match"], g["Q"], iters, 0.0)
    dt = time.perf_counter() - t0
    return dt / iters, g["E"]


def reference_cuda_leg(ref, sets, M, steps):
    """The UNMODIFIED reference extension (oracle/_ref, sm_100 build of the reference's .cu
    files) on the same device and inputs: the number the new kernels have to beat.  (Its GN
    figure comes from gn_parity, which runs it on the same graph.)"""
    out = {}
    try:
        r = time_matching(ref, sets, M, steps, 3, 1)
        out["matched_px_per_s"] = N_PTS * steps / (r["total_ms"] * 1e-3)
        out["iter_proj_ms"], out["refine_ms"] = r["iter_proj_ms"], r["refine_ms"]
    except Exception as exc:  # pragma: no cover
        out["error"] = str(exc)[:200]
    return out


def cpu_baseline_tracking_224():
    """BASELINE configs[0]: the reference's torch tracking GN (tracker.py:173-214, restated in
    oracle.track_pose_rays and pinned by tests/golden/track_*.npz, which the reference's own
    tracker.py produced) on a synthetic 224x224 pair with a known Sim(3), on the host cores."""
    import m3r_oracle as oracle
    import synth

    t = synth.make_track_inputs(224, 224, seed=0, device="cpu", pose_noise=0.03)
    torch.set_num_threads(len(os.sched_getaffinity(0)))
    oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
    t0 = time.perf_counter()
    Tw, _, it, _ = oracle.track_pose_rays(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
    dt = time.perf_counter() - t0
    return {"ms_per_call": dt * 1e3, "iterations": it, "ms_per_iter": dt * 1e3 / max(it, 1),
            "threads": torch.get_num_threads(),
            "pose_err_vs_truth": (Tw - t["T_WCf_true"]).abs().max().item(),
            "pose_err_initial": (t["T_WCf_init"] - t["T_WCf_true"]).abs().max().item()}


WORKLOAD = "configs[1]: 512x384 frame-to-keyframe iter_proj + refine_matches, b=1"
METRIC = "matched px/s (iter_proj + refine_matches, 512x384, b=1)"


def run_reference_arm(args, rank, world):
    """--impl reference: the reference path's CPU restatement on the host cores, rank 0."""
    if rank != 0:
        return
    import synth

    M = synth.MATCH_DEFAULTS
    pxs, s_per_step, cores = cpu_baseline_matching(M, steps=args.steps, warmup=min(args.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC,
        "value": pxs, "unit": "px/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": s_per_step * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32+f16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": "whole 512x384 frames"},
        "cpu_baseline": {"value": pxs, "unit": "px/s", "cores": cores, "kind": "port",
                         "sample": f"{args.steps} whole 512x384 frame(s); the reference has no CPU "
                                   "implementation (setup.py:47-48), this is oracle/m3r_oracle.c"},
        "e2e": {"value": pxs, "unit": "px/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def host_us_per_call(be, s, M, g16, D, calls=40):
    """Host cost of one call through the Python shim + ctypes + launch, per export: wall clock over
    `calls` asynchronous calls issued into an empty stream (the queue never fills), GPU time excluded."""
    out = {}
    p, _ = be.iter_proj(s["rays"], s["pts"], s["p_init"], M["max_iter"], M["lambda_init"], M["convergence_thresh"])
    p1 = p.long()

    def timeit(fn):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(calls):
            fn()
        dt = time.perf_counter() - t0
        torch.cuda.synchronize()
        return dt / calls * 1e6

    out["iter_proj"] = timeit(lambda: be.iter_proj(s["rays"], s["pts"], s["p_init"], M["max_iter"], M["lambda_init"],
                                                   M["convergence_thresh"]))
    out["refine_matches"] = timeit(lambda: be.refine_matches(s["D11"], s["D21"], p1, M["radius"], M["dilation_max"]))
    if g16 is not None:
        for kind in ("rays", "calib", "points"):
            T = g16["Twc_init"].clone()
            calls_gn = 5
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(calls_gn):
                GN_CALLS[kind](be, g16, D, T, 10)
            out[f"gauss_newton_{kind}"] = (time.perf_counter() - t0) / calls_gn * 1e6
            torch.cuda.synchronize()
        out["gauss_newton_note"] = "one 10-iteration call on the 16-keyframe graph: 3 + 10 x 5 launches enqueued"
    return out


def host_topology():
    """NUMA / PCIe facts of the host that the e2e figures depend on (all ranks share them)."""
    try:
        nodes = open("/sys/devices/system/node/online").read().strip()
    except Exception:
        nodes = "?"
    aff = []
    try:
        topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
        lines = [ln for ln in topo.splitlines() if ln.strip()]
        hdr = [h.strip() for h in lines[0].split("\t")]
        for ln in lines[1:]:
            f = [x.strip() for x in ln.split("\t")]
            if f and f[0].startswith("GPU") and len(f) >= len(hdr):
                row = dict(zip(hdr, f[1:])) if hdr and not hdr[0] else dict(zip(hdr[1:], f[1:]))
                aff.append({"gpu": f[0], "cpu_affinity": row.get("CPU Affinity"), "numa": row.get("NUMA Affinity")})
    except Exception:
        pass
    try:
        link = subprocess.run(["nvidia-smi", "--query-gpu=pcie.link.gen.current,pcie.link.width.current",
                               "--format=csv,noheader", "-i", "0"], capture_output=True, text=True, timeout=20).stdout.strip()
    except Exception:
        link = None
    return {"numa_nodes_online": nodes, "host_cores_visible": len(os.sched_getaffinity(0)),
            "gpu_cpu_affinity": aff[:8], "pcie_gen_width_gpu0": link}


def gather_list(x, world):
    if world == 1:
        return [x]
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    out = [torch.empty_like(t) for _ in range(world)]
    torch.distributed.all_gather(out, t)
    return [o.item() for o in out]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=None)
    ap.add_argument("--warmup", type=int, default=None)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--gn-keyframes", type=int, default=64)
    ap.add_argument("--gn-kind", default="rays", choices=["rays", "calib", "points"])
    ap.add_argument("--skip-gn", action="store_true")
    ap.add_argument("--skip-gn-configs", action="store_true",
                    help="only the headline 64-keyframe graph, not BASELINE configs[2..4]")
    ap.add_argument("--skip-cpu", action="store_true")
    ap.add_argument("--skip-refcuda", action="store_true")
    ap.add_argument("--skip-extras", action="store_true", help="no front-end / tracking / frame-pipeline legs")
    ap.add_argument("--profile", action="store_true",
                    help="wrap the timed regions in cudaProfilerStart/Stop (for ncu --profile-from-start off)")
    ap.add_argument("--iter-proj-variant", type=int, default=0)
    args = ap.parse_args()
    rank, local_rank, world = dist_env()
    if args.impl == "reference":
        args.steps = args.steps or 3
        args.warmup = 1 if args.warmup is None else args.warmup
        run_reference_arm(args, rank, world)
        return
    args.steps = args.steps or 200
    args.warmup = 10 if args.warmup is None else max(args.warmup, 3)

    import synth
    import mast3r_slam_backends as be  # fails loudly if libm3r_b200.so is missing
    from mast3r_slam_backends import _lib

    assert torch.cuda.is_available(), "bench.py needs a GPU (no CPU fallback)"
    torch.cuda.set_device(local_rank)
    saved_stdout = None
    if world > 1:
        # NCCL writes its version banner to stdout; stdout must carry the one JSON line only
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    M, D = synth.MATCH_DEFAULTS, synth.GN_DEFAULTS
    peak, peak_src = peaks()
    refmod = None
    if rank == 0 and world == 1 and not args.skip_refcuda:
        import refcuda

        refmod = refcuda.load_ref()

    # ---- matching: K steps, inputs rotated over NUM_SETS frame sets (cold in L2)
    sets = make_match_sets("cuda", NUM_SETS, seed0=100 * rank)
    sampler = ClockSampler(local_rank) if rank == 0 else None  # one sampler per job, not per rank
    time_matching(be, sets, M, 20, 5, world, per_kernel=False)  # clocks up
    if sampler:
        sampler.start()
    l0 = _lib.launch_count()
    if args.profile:
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
    # the headline loop carries no per-kernel events: K x (iter_proj, cast, refine_matches)
    # between one pair of events.  The per-kernel durations come from a second pass.
    r = time_matching(be, sets, M, args.steps, args.warmup, world, per_kernel=False)
    if args.profile:
        torch.cuda.profiler.stop()
    launches = (_lib.launch_count() - l0) * args.steps // (args.steps + args.warmup)  # same launches every step
    clocks = sampler.stop() if sampler else None
    rank_ms = gather_list(r["rank_ms"] / args.steps, world)
    rk = time_matching(be, sets, M, min(args.steps, 100), 3, world, per_kernel=True)
    r.update({k: v for k, v in rk.items() if k.endswith("_ms") and k not in ("total_ms", "rank_ms")})
    ms_step = r["total_ms"] / args.steps
    value = world * N_PTS / (ms_step * 1e-3)
    e2e_steps = max(args.steps, 60)  # long enough that pipeline fill and drain are noise
    e2e_ms, e2e_serial_ms, h2d, d2h = time_matching_e2e(be, sets, M, e2e_steps, 3, world)
    e2e_value = world * N_PTS * e2e_steps / (e2e_ms * 1e-3)

    refine_gbs = B_REFINE * N_PTS / (r["refine_ms"] * 1e-3) / 1e9
    iproj_gbs = B_ITER_PROJ * N_PTS / (r["iter_proj_ms"] * 1e-3) / 1e9
    pair_gbs = (B_REFINE + B_ITER_PROJ) * N_PTS / ((r["refine_ms"] + r["iter_proj_ms"]) * 1e-3) / 1e9
    half_macs = N_PTS * 245 * FDIM  # radius 3, dilation 5: 245 windows x 24 MACs per point

    line = {
        "metric": METRIC, "value": value,
        "unit": "px/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32 (iter_proj) + f16 per-op rounded (refine_matches) + f32/f64 (GN)",
        "data": "synthetic",
        "config": {"workload": WORKLOAD,
                   "replicas": "matching runs as independent replicas per GPU (b = 1 does not shard, SURVEY 8e)",
                   "l2": f"inputs rotated over {NUM_SETS} frame sets ({NUM_SETS}x38 MB > 126 MB L2)",
                   "timing": "CUDA events on the launching stream around all K steps, max over ranks; a ~1 ms "
                             "device-side sleep is enqueued in front of the start event so that the region "
                             "starts with the host ahead of the device (no host launch latency inside it)",
                   "params": M},
        "clocks": clocks,
        "per_rank_ms_per_step": {"min": min(rank_ms), "median": statistics.median(rank_ms), "max": max(rank_ms)},
        "e2e": {"value": e2e_value, "unit": "px/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms / e2e_steps,
                "mode": "streamed: H2D of frame i+1 and D2H of frame i-1 overlap the kernels of frame i "
                        "(double-buffered, three streams); all copies inside the timed region",
                "serial_ms_per_step": e2e_serial_ms / e2e_steps,
                "serial_value": world * N_PTS * e2e_steps / (e2e_serial_ms * 1e-3),
                "host": host_topology() if rank == 0 else None},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": "refine_f16_tiled_kernel", "achieved": refine_gbs,
                     "peak": peak, "unit": "GB/s", "frac": refine_gbs / peak,
                     "traffic": traffic("refine_f16_tiled_kernel"), "peak_source": peak_src,
                     "algorithmic_bytes_per_launch": B_REFINE * N_PTS,
                     "note": "bit-exactness fixes 245 windows x 24 individually rounded half MACs per "
                             "point: the kernel is bound by the fp16x2 pipe (0.49 warp-instr/clk/SMSP "
                             "measured, profiles/r1_ubench_pipes.txt) and shared-memory bandwidth, "
                             "both 62 us per frame at 1965 MHz, not by HBM",
                     "half_op_per_s": 2 * half_macs / (r["refine_ms"] * 1e-3),
                     "fp16_pipe_frac": (2 * half_macs / 2 / 32) / (148 * 4 * 0.486 * 1.965e9) /
                                       (r["refine_ms"] * 1e-3),
                     # 245 windows x 3 LDS.128 per point, 4 wavefronts per warp-wide LDS.128 when
                     # conflict-free, one wavefront per clock per SM
                     "smem_pipe_frac_conflict_free": (N_PTS / 32 * 245 * 3 * 4) / (148 * 1.965e9) /
                                                     (r["refine_ms"] * 1e-3)},
        "kernels": {"iter_proj_ms": r["iter_proj_ms"], "cast_ms": r["cast_ms"], "refine_ms": r["refine_ms"],
                    "iter_proj_gbs": iproj_gbs, "iter_proj_frac": iproj_gbs / peak,
                    "pair_gbs": pair_gbs, "pair_frac": pair_gbs / peak,
                    "note": "per-kernel durations from a second pass with events around each call; the "
                            "headline loop has one pair of events around all K steps"},
    }

    extras = not (args.profile or args.skip_extras)
    if rank == 0 and extras:
        line["front_end"] = time_front_end(be, sets, M, max(args.steps // 4, 10), 3)
    if extras:
        fe = time_front_e2e(sets, M, e2e_steps, 3, world)
        if rank == 0:
            line["front_end"]["e2e"] = fe
    if rank == 0 and extras:
        line["tracking"] = time_tracking(max(args.steps // 4, 10), 3)
        line["frame_pipeline"] = time_frame_pipeline(sets, M, max(args.steps // 2, 20), 3)

    # ---- Gauss-Newton: the headline graph (64 keyframes, rays) and BASELINE configs[2..4]
    g = None
    if not args.skip_gn:
        g = make_gn_graph(args.gn_keyframes, "cuda", seed=0)
        if args.profile:
            torch.cuda.synchronize()
            torch.cuda.profiler.start()
        line["gn"] = gn_report(be, g, D, args.gn_kind, rank, world, peak, refmod, reps=1 if args.profile else 3,
                               single_process=not args.profile, local_rank=local_rank)
        if args.profile:
            torch.cuda.profiler.stop()
        if not (args.profile or args.skip_gn_configs):
            cfgs = {}
            # configs[3]: the same 64-keyframe graph through the calibrated solver, sharded at N > 1
            cfgs["calib64"] = gn_report(be, g, D, "calib", rank, world, peak, refmod, single_call=False)
            if rank == 0 and world == 1 and refmod is not None:
                line["reference_cuda"] = reference_cuda_leg(refmod, sets, M, min(args.steps, 50))
                line["reference_cuda"]["gn_ms_per_iter"] = line["gn"]["parity"].get("reference_cuda_ms_per_iter")
                if line["reference_cuda"]["gn_ms_per_iter"]:
                    line["reference_cuda"]["gn_iters_per_s"] = 1e3 / line["reference_cuda"]["gn_ms_per_iter"]
            del g
            torch.cuda.empty_cache()
            # configs[2]: 16-keyframe rays graph
            g16 = make_gn_graph(16, "cuda", seed=0)
            cfgs["rays16"] = gn_report(be, g16, D, "rays", rank, world, peak, refmod, single_call=False)
            if rank == 0:
                line["host_us_per_call"] = host_us_per_call(be, sets[0], M, g16, D)
            del g16
            torch.cuda.empty_cache()
            # configs[4]: 256-keyframe loop-closure graph, gauss_newton_points, D = 1785
            g256 = make_gn_graph(256, "cuda", seed=0)
            cfgs["points256"] = gn_report(be, g256, D, "points", rank, world, peak, refmod, reps=2,
                                          single_call=False, single_process=True, local_rank=local_rank)
            del g256
            torch.cuda.empty_cache()
            line["gn_configs"] = cfgs
            g = None
    elif rank == 0 and world == 1 and refmod is not None:
        line["reference_cuda"] = reference_cuda_leg(refmod, sets, M, min(args.steps, 50))

    if rank == 0 and world == 1 and not args.skip_cpu:
        pxs, s_per, cores = cpu_baseline_matching(M, steps=2, warmup=1)
        line["cpu_baseline"] = {"value": pxs, "unit": "px/s", "cores": cores, "kind": "port",
                                "sample": "2 whole 512x384 frames (oracle/m3r_oracle.c, OpenMP)"}
        if not args.skip_gn:
            s_it, Ecpu = cpu_baseline_gn(D)
            E64 = line["gn"]["roofline"]["bytes_per_iter"] // (B_GN * N_PTS)
            line["cpu_baseline"]["gn_iters_per_s_64kf_extrapolated"] = 1.0 / (s_it * E64 / Ecpu)
            line["cpu_baseline"]["gn_sample"] = f"1 iteration, 8-keyframe graph ({Ecpu} edges), scaled by edges"
            line["cpu_baseline"]["tracking_224x224"] = cpu_baseline_tracking_224()
    if saved_stdout is not None:
        sys.stdout.flush()
        os.dup2(saved_stdout, 1)
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        os.dup2(2, 1)
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
