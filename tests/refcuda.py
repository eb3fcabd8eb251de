"""Loader for the UNMODIFIED reference CUDA extension built by oracle/build_ref.py into
oracle/_ref/ (test infrastructure: the strongest oracle, only runnable on a GPU)."""
import importlib.util
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(_ROOT, "oracle", "_ref", "mast3r_slam_backends_ref.so")
_mod = None


def load_ref():
    global _mod
    if _mod is not None:
        return _mod
    if not os.path.exists(REF_SO):
        return None
    import torch  # noqa: F401  (libtorch symbols)

    spec = importlib.util.spec_from_file_location("mast3r_slam_backends_ref", REF_SO)
    try:
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    except Exception as exc:  # pragma: no cover
        print(f"[refcuda] could not load {REF_SO}: {exc}")
        return None
    _mod = mod
    return mod
