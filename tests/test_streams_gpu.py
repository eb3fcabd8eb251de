"""Stream semantics of the drop-in module: every entry point enqueues on torch's CURRENT stream
(the reference uses the legacy default stream), never synchronises the device, and two streams
can run the same entry points concurrently with independent scratch.  Results on a side
stream equal the results on the default stream bit for bit."""
import pytest
import torch

import synth

pytestmark = pytest.mark.gpu
M, D, T = synth.MATCH_DEFAULTS, synth.GN_DEFAULTS, synth.TRACK_DEFAULTS


def _run_all(backends, m, g, t):
    from mast3r_slam_backends import matching as front, pointmaps, tracking

    front.set_config({"matching": dict(M)})
    tracking.set_config({"tracking": dict(T)})
    h, w = m["X11"].shape[1:3]
    rays, pts, p0 = front.prep_for_iter_proj(m["X11"], m["X21"], None)
    p, conv = backends.iter_proj(rays, pts.contiguous(), p0.contiguous(), M["max_iter"], M["lambda_init"],
                                 M["convergence_thresh"])
    (p1n,) = backends.refine_matches(m["D11"].half().contiguous(), m["D21"].view(1, h * w, -1).half().contiguous(),
                                     p.long(), M["radius"], M["dilation_max"])
    idx, valid = front.match(m["X11"], m["X21"], m["D11"], m["D21"])
    Twc = g["Twc_init"].clone()
    (dx,) = backends.gauss_newton_rays(Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"], g["valid_match"],
                                       g["Q"], D["sigma_ray"], D["sigma_dist"], D["C_thresh"], D["Q_thresh"], 3, 0.0)
    Tw, Tk = tracking.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"],
                                             check=False)
    Xc, Cc = t["Xk"].clone(), t["Ck"].clone()
    pointmaps.update_pointmap(Xc, Cc, t["Xf"], t["Cf"], "weighted_pointmap", T=Tk)
    return [p, conv, p1n, idx, valid, Twc, dx, Tw, Tk, Xc, Cc]


def test_side_stream_equals_default_stream(backends):
    m = synth.make_match_inputs(96, 128, seed=2, device="cuda")
    g = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.make_graph(5, 48, 64, seed=2).items()}
    t = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.make_track_inputs(96, 128, seed=2).items()}
    torch.cuda.synchronize()
    ref = _run_all(backends, m, g, t)
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        out = _run_all(backends, m, g, t)
    side.synchronize()
    for a, b in zip(ref, out):
        assert torch.equal(a, b)


def test_two_streams_interleaved(backends):
    """Two streams issue the tracking solve and the fused match alternately; each keeps its own
    scratch (the tracking scratch carries a grid-barrier counter and must not be shared)."""
    from mast3r_slam_backends import matching as front, tracking

    front.set_config({"matching": dict(M)})
    tracking.set_config({"tracking": dict(T)})
    m = synth.make_match_inputs(64, 96, seed=3, device="cuda")
    t = {k: (v.cuda() if torch.is_tensor(v) else v) for k, v in synth.make_track_inputs(120, 160, seed=3).items()}
    torch.cuda.synchronize()
    Tw0, Tk0 = tracking.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"], t["Qk"], t["valid_opt"])
    idx0, valid0 = front.match(m["X11"], m["X21"], m["D11"], m["D21"])
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    results = [[], []]
    for rep in range(6):
        for k, s in enumerate(streams):
            with torch.cuda.stream(s):
                results[k].append(tracking.opt_pose_ray_dist_sim3(t["Xf"], t["Xk"], t["T_WCf_init"], t["T_WCk"],
                                                                  t["Qk"], t["valid_opt"], check=False))
                results[k].append(front.match(m["X11"], m["X21"], m["D11"], m["D21"]))
    torch.cuda.synchronize()
    for k in range(2):
        for j, r in enumerate(results[k]):
            if j % 2 == 0:
                assert torch.equal(r[0], Tw0) and torch.equal(r[1], Tk0)
            else:
                assert torch.equal(r[0], idx0) and torch.equal(r[1], valid0)


def test_matching_under_cuda_graph_capture(backends):
    """iter_proj -> .long() -> refine_matches captured into a CUDA graph and replayed: the refine
    plan scratch is a stream-ordered allocation (a graph memory node under capture), nothing in
    the two calls synchronises or allocates outside the stream.  Replays equal the eager result."""
    m = synth.make_match_inputs(96, 128, seed=4, device="cuda")
    rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"])
    D11 = m["D11"].half().contiguous()
    D21 = m["D21"].view(1, 96 * 128, -1).half().contiguous()

    def run():
        p, conv = backends.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
        (p1n,) = backends.refine_matches(D11, D21, p.long(), M["radius"], M["dilation_max"])
        return p, conv, p1n

    eager = [x.clone() for x in run()]
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):  # warm-up on the capture stream, as torch asks
        run()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out = run()
    for _ in range(3):
        for o in out:
            o.zero_()
        graph.replay()
        torch.cuda.synchronize()
        assert all(torch.equal(a, b) for a, b in zip(eager, out))
