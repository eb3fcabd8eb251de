"""The C-ABI library loads and exports every symbol include/m3r_backends.h declares.
No compute calls here (no GPU needed)."""
import ctypes
import os
import re

from conftest import PKG, ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "m3r_backends.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(m3r_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    lib = ctypes.CDLL(os.path.join(PKG, "libm3r_b200.so"))
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), f"{n} declared in m3r_backends.h but not exported"


def test_binding_lists_all_exports():
    from mast3r_slam_backends import _lib

    assert sorted(_lib.EXPORTS) == _declared()
    assert _lib.abi_version() == 1
    assert _lib.gn_system_size(64) == 441 * 441 + 441


def test_reference_surface():
    """Same five names as the reference's PYBIND11 module (gn.cpp:116-123)."""
    import mast3r_slam_backends as m

    for name in ["gauss_newton_points", "gauss_newton_rays", "gauss_newton_calib", "iter_proj",
                 "refine_matches"]:
        assert callable(getattr(m, name))


def test_cpu_tensors_rejected():
    """There is no CPU fallback: CPU tensors raise like a CUDA-only extension would."""
    import pytest
    import torch
    import mast3r_slam_backends as m

    with pytest.raises(RuntimeError):
        m.iter_proj(torch.zeros(1, 4, 4, 9), torch.zeros(1, 16, 3), torch.zeros(1, 16, 2), 10, 1e-8, 1e-6)
    with pytest.raises(RuntimeError, match="must be contiguous"):
        m.refine_matches(torch.zeros(1, 4, 4, 48)[..., ::2], torch.zeros(1, 16, 24),
                         torch.zeros(1, 16, 2, dtype=torch.long), 3, 5)
