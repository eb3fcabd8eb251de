"""The reference's own callers, unmodified, on this library (BASELINE.json north_star: "so
matching.py, tracker.py and global_opt.py's FactorGraph call it unchanged").

`mast3r_slam/matching.py` (match_iterative_proj :52-88) and `mast3r_slam/global_opt.py`
(FactorGraph.add_factors :31-103, prep_two_way_edges :106-112, get_poses_points :114-121,
solve_GN_rays :123-161, solve_GN_calib :163-226) are imported as they are (tests/refpy.py) and
run twice on the same synthetic keyframes: once with `mast3r_slam_backends` = the drop-in module
over libm3r_b200.so, once = the reference's own CUDA extension (oracle/_ref).  Match indices
must be bit-exact, solved poses within 1e-4 per component group, and the solver must have
written through the `[N,8]` view the caller hands it (global_opt.py:142,161)."""
import pytest
import torch

import refpy
import synth
from test_gn_gpu import TOL, _fmt, _groups_ok, _pose_err_groups

pytestmark = pytest.mark.gpu


def _edges(N, seed):
    si, sj = synth.make_graph_edges(N, seed)
    return si.tolist(), sj.tolist()


def _run(ns, backend, scene, ii, jj, calib):
    """What the backend process does for a batch of new edges (main.py:118-140): add_factors,
    then solve_GN_rays / solve_GN_calib.  Returns the FactorGraph and the keyframe buffer."""
    ns = refpy.load(backend)
    ns.cfg.config["use_calib"] = calib
    ns.mu.mast3r_match_symmetric = refpy.SyntheticDecoder(ns, scene)
    ns.global_opt.mast3r_match_symmetric = ns.mu.mast3r_match_symmetric
    frames = refpy.make_keyframes(ns, scene, "cuda")
    fg = ns.global_opt.FactorGraph(None, frames, scene["K"] if calib else None, "cuda")
    # edges arrive a few at a time, as keyframes are added (global_opt.py:91-98 appends)
    step = 4
    for k in range(0, len(ii), step):
        assert fg.add_factors(ii[k:k + step], jj[k:k + step], ns.cfg.config["local_opt"]["min_match_frac"])
    if calib:
        fg.solve_GN_calib()
    else:
        fg.solve_GN_rays()
    torch.cuda.synchronize()
    return fg, frames


@pytest.mark.parametrize("calib", [False, True], ids=["rays", "calib"])
@pytest.mark.parametrize("N,h,w,seed", [(6, 96, 128, 0), (9, 192, 256, 1)])
def test_factor_graph_unchanged_on_product_vs_reference_cuda(backends, refcuda, N, h, w, seed, calib):
    if refpy.reference_root() is None:
        pytest.skip("baseline/_ref not installed (oracle/install_reference_py.py)")
    if refcuda is None:
        pytest.skip("oracle/_ref/mast3r_slam_backends_ref.so not built")
    ii, jj = _edges(N, seed)
    pairs = {(a, b) for a, b in zip(ii, jj)} | {(b, a) for a, b in zip(ii, jj)}
    scene = refpy.make_scene(N, h, w, seed, "cuda", sorted(pairs))
    ns = refpy.load(backends)
    fg_a, fr_a = _run(ns, backends, scene, ii, jj, calib)
    fg_b, fr_b = _run(ns, refcuda, scene, ii, jj, calib)
    # matching.py on both: bit-exact indices and flags (iter_proj + refine_matches underneath)
    assert torch.equal(fg_a.ii, fg_b.ii) and torch.equal(fg_a.jj, fg_b.jj)
    assert torch.equal(fg_a.idx_ii2jj, fg_b.idx_ii2jj) and torch.equal(fg_a.idx_jj2ii, fg_b.idx_jj2ii)
    assert torch.equal(fg_a.valid_match_j, fg_b.valid_match_j) and torch.equal(fg_a.valid_match_i, fg_b.valid_match_i)
    assert torch.equal(fg_a.Q_ii2jj, fg_b.Q_ii2jj)
    Ta, Tb = fr_a.T_WC[:N, 0], fr_b.T_WC[:N, 0]
    moved = (Ta - scene["T_init"]).abs().max().item()
    gr = _pose_err_groups(Ta, Tb)
    err0 = (scene["T_init"] - scene["T_true"]).abs().max().item()
    err1 = (Ta - scene["T_true"]).abs().max().item()
    print(f"FactorGraph {'calib' if calib else 'rays'} N={N} {h}x{w} E={2 * fg_a.ii.numel()}: product vs reference "
          f"CUDA {_fmt(gr)}; pose error vs truth {err0:.2e} -> {err1:.2e}")
    assert moved > 1e-3, "Twc was not written through the caller's view"
    assert torch.equal(Ta[0], scene["T_init"][0])  # the pinned pose stays
    assert _groups_ok(gr, TOL), gr
    assert err1 < err0


def test_match_unchanged_on_product_vs_reference_cuda(backends, refcuda):
    """mast3r_slam/matching.py::match alone, at the metric's size (512x384, b = 2)."""
    if refpy.reference_root() is None:
        pytest.skip("baseline/_ref not installed (oracle/install_reference_py.py)")
    if refcuda is None:
        pytest.skip("oracle/_ref/mast3r_slam_backends_ref.so not built")
    m = synth.make_match_inputs(384, 512, b=2, seed=3, device="cuda")
    out = []
    for mod in (backends, refcuda):
        ns = refpy.load(mod)
        idx, valid = ns.matching.match(m["X11"], m["X21"], m["D11"], m["D21"])
        out.append((idx, valid))
    assert torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][1], out[1][1])
    assert out[0][1].float().mean() > 0.5
