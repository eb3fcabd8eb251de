"""m3r_refine_matches_static: the allocation-free variant of refine_matches (scratch in a
file-scope device array, include/m3r_backends.h).  Called through the C ABI with ctypes, checked
bit for bit against the existing entry point m3r_refine_matches on two calls in a row (the second
with a smaller batch and other inputs, so stale plan/flag bytes of the first call are in the shared
array when it runs), and that problems beyond the array's size are rejected.

Named test_zz_* so that pytest collects it last: a failure here must not hide the other tests
under `pytest -x`.  Passes alone in a fresh process and after the rest of the suite."""
import ctypes

import pytest
import torch

import synth

pytestmark = pytest.mark.gpu
M = synth.MATCH_DEFAULTS


def _inputs(h, w, b, seed, shift):
    m = synth.make_match_inputs(h, w, b=b, seed=seed, device="cuda")
    n = h * w
    D11 = m["D11"].half().contiguous()
    D21 = m["D21"].view(b, n, -1).half().contiguous()
    uv = m["true_uv"].round().long() + torch.tensor(shift, device="cuda")
    p1 = torch.stack((uv[..., 0].clamp(0, w - 1), uv[..., 1].clamp(0, h - 1)), dim=-1).contiguous()
    return D11, D21, p1


def _call(fn, D11, D21, p1, b, h, w):
    from mast3r_slam_backends import _lib

    out = torch.full_like(p1, -7)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = fn(ctypes.c_void_p(D11.data_ptr()), ctypes.c_void_p(D21.data_ptr()), ctypes.c_void_p(p1.data_ptr()),
            ctypes.c_void_p(out.data_ptr()), b, h, w, h * w, 24, _lib.DTYPE_F16, M["radius"], M["dilation_max"],
            stream)
    return rc, out


def test_static_scratch_refine_equals_refine_matches_twice():
    from mast3r_slam_backends import _lib

    assert _lib.refine_matches_static_max_tiles() == 3200
    # call 1: two 384x512 frames (1536 tiles); call 2: one 200x328 frame (ragged tiles, 273 tiles)
    for (h, w, b, seed, shift) in [(384, 512, 2, 0, (2, -3)), (200, 328, 1, 5, (-4, 1))]:
        D11, D21, p1 = _inputs(h, w, b, seed, shift)
        rc0, ref = _call(_lib.refine_matches, D11, D21, p1, b, h, w)
        rc1, got = _call(_lib.refine_matches_static, D11, D21, p1, b, h, w)
        torch.cuda.synchronize()
        assert rc0 == 0 and rc1 == 0, _lib.last_error()
        assert torch.equal(got, ref), f"{(got != ref).any(-1).sum().item()} of {b * h * w} matches differ"
        assert (ref != p1).any(-1).float().mean() > 0.3  # the search did move the matches


def test_static_scratch_limits_are_reported():
    from mast3r_slam_backends import _lib

    h, w, b = 384, 512, 5  # 3840 tiles > 3200
    D11 = torch.zeros((b, h, w, 24), dtype=torch.float16, device="cuda")
    D21 = torch.zeros((b, h * w, 24), dtype=torch.float16, device="cuda")
    p1 = torch.zeros((b, h * w, 2), dtype=torch.int64, device="cuda")
    rc, _ = _call(_lib.refine_matches_static, D11, D21, p1, b, h, w)
    assert rc == 1 and b"exceed the static scratch" in _lib.last_error()
    # a configuration the tiled kernel does not cover (radius 2) is refused, not silently re-routed
    out = torch.empty_like(p1)
    rc = _lib.refine_matches_static(ctypes.c_void_p(D11.data_ptr()), ctypes.c_void_p(D21.data_ptr()),
                                    ctypes.c_void_p(p1.data_ptr()), ctypes.c_void_p(out.data_ptr()), 1, h, w, h * w,
                                    24, _lib.DTYPE_F16, 2, 5, ctypes.c_void_p(0))
    assert rc == 1 and b"only the tiled configuration" in _lib.last_error()
