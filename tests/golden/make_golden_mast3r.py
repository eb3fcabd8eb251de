#!/usr/bin/env python
"""Golden vectors with the value distribution of the real network heads.

BASELINE.json asks for parity "on the same synthetic pointmaps, descriptors and confidences
(random-init MASt3R weights ...)".  This script builds the production AsymmetricMASt3R
(constructor string: thirdparty/mast3r/README.md:297) from /root/reference with seeded random
weights, runs one asymmetric forward (mast3r_utils.py:61-66 `decoder`) on two seeded random
images on CPU, and feeds the head outputs -- pts3d, conf (>= 1), desc (unit-norm, 24 ch),
desc_conf (>= 0) -- through the reference's OWN Python matching code (matching.py, image.py,
imported unmodified; its CUDA extension replaced by the CPU oracle kernels, which are pinned
bit for bit against the reference CUDA kernels by match_a/b.npz).

A random-init network is geometrically meaningless: the pointmaps are a small blob around the
origin with z of both signs and the descriptors carry no correspondence.  That is the point of
this fixture: it exercises the kernels on values the smooth synthetic scenes never produce
(near-degenerate rays, non-converging LM steps, descriptor scores without a clear maximum).

Two things of the reference that cannot run here are stubbed, neither touching arithmetic:
`torch.cuda.synchronize` (mast3r_slam/profiler.py:18 calls it inside every block) and the
`mast3r_slam_backends` extension (see above).

Run in the build container (needs /root/reference):   python tests/golden/make_golden_mast3r.py
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

H, W = 96, 128


def build_model():
    sys.path[:0] = ["/root/reference", "/root/reference/thirdparty/mast3r",
                    "/root/reference/thirdparty/mast3r/dust3r"]
    torch.cuda.synchronize = lambda *a, **k: None
    from mast3r.model import AsymmetricMASt3R

    inf = float("inf")
    torch.manual_seed(0)
    model = AsymmetricMASt3R(
        pos_embed="RoPE100", patch_embed_cls="ManyAR_PatchEmbed", img_size=(512, 512),
        head_type="catmlp+dpt", output_mode="pts3d+desc24", depth_mode=("exp", -inf, inf),
        conf_mode=("exp", 1, inf), enc_embed_dim=1024, enc_depth=24, enc_num_heads=16,
        dec_embed_dim=768, dec_depth=12, dec_num_heads=12, two_confs=True,
        desc_conf_mode=("exp", 0, inf))
    return model.eval()


def import_reference_matching():
    stub = types.ModuleType("mast3r_slam_backends")
    stub.iter_proj = oracle.iter_proj
    stub.refine_matches = oracle.refine_matches
    sys.modules["mast3r_slam_backends"] = stub
    import mast3r_slam.config as cfg
    import mast3r_slam.matching as matching

    cfg.config.update({"matching": dict(synth.MATCH_DEFAULTS)})
    return matching


def main():
    torch.set_num_threads(os.cpu_count())
    model = build_model()
    matching = import_reference_matching()
    g = torch.Generator().manual_seed(1)
    # a smooth image plus noise, so that the two views differ but are not white noise
    yy, xx = torch.meshgrid(torch.linspace(-1, 1, H), torch.linspace(-1, 1, W), indexing="ij")
    base = torch.stack([torch.sin(3 * xx + yy), torch.cos(2 * yy - xx), xx * yy])[None]
    img1 = (base + 0.3 * torch.randn(1, 3, H, W, generator=g)).clamp(-1, 1)
    img2 = (base.roll(5, dims=3) + 0.3 * torch.randn(1, 3, H, W, generator=g)).clamp(-1, 1)
    shape = torch.tensor([[H, W]])
    with torch.inference_mode():
        f1, p1, _ = model._encode_image(img1, shape)
        f2, p2, _ = model._encode_image(img2, shape)
        dec1, dec2 = model._decoder(f1, p1, f2, p2)  # mast3r_utils.py:61-66
        r11 = model._downstream_head(1, [t.float() for t in dec1], shape)
        r21 = model._downstream_head(2, [t.float() for t in dec2], shape)
    X11, X21 = r11["pts3d"].contiguous(), r21["pts3d"].contiguous()
    D11, D21 = r11["desc"].contiguous(), r21["desc"].contiguous()
    C11, C21 = r11["conf"].contiguous(), r21["conf"].contiguous()
    Q11, Q21 = r11["desc_conf"].contiguous(), r21["desc_conf"].contiguous()
    print("pts3d range", float(X11.min()), float(X11.max()), " conf", float(C11.min()), float(C11.max()),
          " desc_conf", float(Q11.min()), float(Q11.max()), " |desc|", float(D11.norm(dim=-1).mean()))
    rays, pts, p_init = matching.prep_for_iter_proj(X11, X21, None)
    idx, valid = matching.match(X11, X21, D11, D21, None)  # mast3r_utils.py:232-234
    # the intermediate the bit-exactness bar is stated on: iter_proj's output after .long()
    p_proj, conv = oracle.iter_proj(rays, pts, p_init, synth.MATCH_DEFAULTS["max_iter"],
                                    synth.MATCH_DEFAULTS["lambda_init"],
                                    synth.MATCH_DEFAULTS["convergence_thresh"])
    out = dict(X11=X11.numpy(), X21=X21.numpy(), D11_f16=D11.half().view(torch.int16).numpy(),
               D21_f16=D21.half().view(torch.int16).numpy(), C11=C11.numpy(), C21=C21.numpy(),
               Q11=Q11.numpy(), Q21=Q21.numpy(), ref_rays=rays.numpy(), ref_pts=pts.numpy(), ref_idx=idx.numpy(),
               ref_valid=valid.numpy(), ref_p_proj=p_proj.numpy(), ref_converged=conv.numpy(),
               **{k: v for k, v in synth.MATCH_DEFAULTS.items()})
    np.savez_compressed(os.path.join(HERE, "mast3r_rand.npz"), **out)
    print(f"wrote mast3r_rand: converged {float(conv.float().mean()):.3f} valid {float(valid.float().mean()):.3f} "
          f"distinct targets {idx.unique().numel()} of {idx.numel()}")


if __name__ == "__main__":
    main()
