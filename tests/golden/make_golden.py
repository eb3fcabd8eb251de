#!/usr/bin/env python
"""Generate the golden vectors under tests/golden/ by running the UNMODIFIED reference CUDA
extension (oracle/_ref/mast3r_slam_backends_ref.so, built by oracle/build_ref.py from
/root/reference/mast3r_slam/backend) on a GPU, on seeded synthetic inputs from tests/synth.py.

Run on the B200 box:   python tests/golden/make_golden.py --out gpurun_out/golden
then copy the .npz files into tests/golden/ and commit them.  The reference has no tests or
fixtures of its own for this path; these vectors are what pins the CPU
oracle (tests/test_oracle.py::test_golden_*) and the CUDA kernels (test_golden_*_gpu).
"""
import argparse
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import refcuda  # noqa: E402
import synth  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(HERE))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    ref = refcuda.load_ref()
    assert ref is not None, "oracle/_ref/mast3r_slam_backends_ref.so missing: run oracle/build_ref.py"
    M, D = synth.MATCH_DEFAULTS, synth.GN_DEFAULTS

    # matching: h*w must be a multiple of 16 (the reference kernels have no bounds guard)
    for name, (h, w, b, seed, init) in {"a": (32, 48, 1, 0, "identity"), "b": (40, 64, 2, 1, "true")}.items():
        m = synth.make_match_inputs(h, w, b=b, seed=seed, device="cuda")
        idx_init = None
        if init == "true":
            tu = m["true_uv"].round().long()
            idx_init = tu[..., 1].clamp(0, h - 1) * w + tu[..., 0].clamp(0, w - 1)
        rays, pts, p0 = synth.prep_for_iter_proj(m["X11"], m["X21"], idx_init)
        p, conv = ref.iter_proj(rays, pts, p0, M["max_iter"], M["lambda_init"], M["convergence_thresh"])
        p1 = p.long()
        D11, D21 = m["D11"].half().contiguous(), m["D21"].view(b, h * w, -1).half().contiguous()
        (p1n,) = ref.refine_matches(D11, D21, p1, M["radius"], M["dilation_max"])
        (p1n32,) = ref.refine_matches(D11.float(), D21.float(), p1, M["radius"], M["dilation_max"])
        torch.cuda.synchronize()
        np.savez_compressed(
            os.path.join(args.out, f"match_{name}.npz"), rays=rays.cpu().numpy(), pts=pts.cpu().numpy(),
            p_init=p0.cpu().numpy(), max_iter=M["max_iter"], lambda_init=M["lambda_init"],
            cost_thresh=M["convergence_thresh"], ref_p_new=p.cpu().numpy(),
            ref_converged=conv.cpu().numpy(), D11_f16=D11.cpu().view(torch.int16).numpy(),
            D21_f16=D21.cpu().view(torch.int16).numpy(), ref_p1=p1.cpu().numpy(),
            ref_p1_new=p1n.cpu().numpy(), ref_p1_new_f32=p1n32.cpu().numpy(), radius=M["radius"],
            dilation_max=M["dilation_max"])
        print("wrote match_" + name)

    for kind in ("rays", "calib", "points"):
        g = synth.make_graph(4, 24, 32, seed=11, device="cuda")
        out = {}
        for iters, key in ((1, "ref_1"), (D["max_iter"], "ref_full")):
            Twc = g["Twc_init"].clone()
            if kind == "rays":
                (dx,) = ref.gauss_newton_rays(Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                              g["valid_match"], g["Q"], D["sigma_ray"], D["sigma_dist"],
                                              D["C_thresh"], D["Q_thresh"], iters, D["delta_thresh"])
            elif kind == "calib":
                (dx,) = ref.gauss_newton_calib(Twc, g["Xs"], g["Cs"], g["K"], g["ii"], g["jj"],
                                               g["idx_ii2jj"], g["valid_match"], g["Q"], g["h"], g["w"],
                                               D["pixel_border"], D["z_eps"], D["sigma_pixel"],
                                               D["sigma_depth"], D["C_thresh"], D["Q_thresh"], iters,
                                               D["delta_thresh"])
            else:
                (dx,) = ref.gauss_newton_points(Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx_ii2jj"],
                                                g["valid_match"], g["Q"], D["sigma_point"],
                                                D["C_thresh"], D["Q_thresh"], iters, D["delta_thresh"])
            torch.cuda.synchronize()
            out[key + "_Twc"] = Twc.cpu().numpy()
            out[key + "_dx"] = dx.cpu().numpy()
        np.savez_compressed(
            os.path.join(args.out, f"gn_{kind}.npz"), kind=kind, Twc=g["Twc_init"].cpu().numpy(),
            Xs=g["Xs"].cpu().numpy(), Cs=g["Cs"].cpu().numpy(), K=g["K"].cpu().numpy(),
            ii=g["ii"].cpu().numpy(), jj=g["jj"].cpu().numpy(), idx=g["idx_ii2jj"].cpu().numpy(),
            valid=g["valid_match"].cpu().numpy(), Q=g["Q"].cpu().numpy(), height=g["h"], width=g["w"],
            max_iter=D["max_iter"], delta_thresh=D["delta_thresh"], **out)
        print("wrote gn_" + kind)


if __name__ == "__main__":
    main()
