#!/usr/bin/env python
"""Golden vectors for the matching front end.

Runs the reference's OWN Python code for this path -- mast3r_slam/matching.py (match,
prep_for_iter_proj) and mast3r_slam/image.py (img_gradient), imported unmodified from
/root/reference -- on CPU, with the one thing that cannot run here, its CUDA extension module
`mast3r_slam_backends`, replaced by the CPU oracle's iter_proj / refine_matches (themselves
pinned bit-for-bit against the reference CUDA kernels by match_a/b.npz).  The vectors pin the
oracle's restatement of the Python glue (oracle/m3r_oracle.py::match, prep_for_iter_proj) and
the fused CUDA front end (m3r_prep_rays, m3r_match).

Run in the build container (needs /root/reference):   python tests/golden/make_golden_front.py
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import m3r_oracle as oracle  # noqa: E402
import synth  # noqa: E402


def import_reference_matching():
    stub = types.ModuleType("mast3r_slam_backends")
    stub.iter_proj = oracle.iter_proj
    stub.refine_matches = oracle.refine_matches
    sys.modules["mast3r_slam_backends"] = stub
    # mast3r_slam/__init__ pulls nothing heavy, but keep the import surface minimal anyway
    sys.path.insert(0, "/root/reference")
    import mast3r_slam.config as cfg
    import mast3r_slam.matching as matching

    cfg.config.update({"matching": dict(synth.MATCH_DEFAULTS)})
    return matching


def main():
    matching = import_reference_matching()
    torch.set_num_threads(4)
    cases = {"a": (32, 48, 1, 0, "identity"), "b": (40, 64, 2, 1, "true"), "c": (24, 32, 1, 5, "shifted")}
    for name, (h, w, b, seed, init) in cases.items():
        m = synth.make_match_inputs(h, w, b=b, seed=seed)
        idx_init = None
        if init != "identity":
            tu = m["true_uv"].round().long()
            if init == "shifted":
                tu = tu + torch.tensor([3, -2])
            idx_init = tu[..., 1].clamp(0, h - 1) * w + tu[..., 0].clamp(0, w - 1)
        X11, X21 = m["X11"].float().contiguous(), m["X21"].float().contiguous()
        D11, D21 = m["D11"].half().contiguous(), m["D21"].half().contiguous()
        rays, pts, p_init = matching.prep_for_iter_proj(X11, X21, idx_init)
        idx, valid = matching.match(X11, X21, D11, D21, idx_init)
        out = dict(X11=X11.numpy(), X21=X21.numpy(), D11_f16=D11.view(torch.int16).numpy(),
                   D21_f16=D21.view(torch.int16).numpy(), ref_rays=rays.numpy(), ref_pts=pts.numpy(),
                   ref_p_init=p_init.numpy(), ref_idx=idx.numpy(), ref_valid=valid.numpy(),
                   **{k: v for k, v in synth.MATCH_DEFAULTS.items()})
        if idx_init is not None:
            out["idx_init"] = idx_init.numpy()
        np.savez_compressed(os.path.join(HERE, f"front_{name}.npz"), **out)
        print(f"wrote front_{name}: valid {float(valid.float().mean()):.3f}")


if __name__ == "__main__":
    main()
