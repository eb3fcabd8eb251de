#!/usr/bin/env python
"""Build the UNMODIFIED reference CUDA extension into oracle/_ref/ (test infrastructure only).

The reference (`/root/reference/mast3r_slam/backend`, module `mast3r_slam_backends`,
setup.py:38-44) has no CPU implementation and stops at sm_86, so it is compiled here,
from the sources where they lie, for sm_100 with nvcc.  Two API-drift edits that do not
touch arithmetic are applied to a scratch copy under /tmp (never inside this repo):

  * `torch::linalg::linalg_norm(dx, ...)`  -> `at::linalg_norm(dx, ...)`
    (gn_kernels.cu:802,1219,1629; torch 2.11 no longer exposes that namespace here)
  * `D11.type()`                           -> `D11.scalar_type()`
    (matching_kernels.cu:103; AT_DISPATCH needs a ScalarType)

Output: oracle/_ref/mast3r_slam_backends_ref.so (git-ignored, NOT gpurun-ignored, so it
travels to the GPU box).  It can only RUN on a GPU; the `-m gpu` parity tests and
`bench.py`'s `reference_cuda` leg load it when present.  Nothing in the product imports it.
"""
import os
import re
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(HERE, "_ref")
NAME = "mast3r_slam_backends_ref"


def main() -> int:
    src_dir = os.path.join(REF, "mast3r_slam", "backend")
    if not os.path.isdir(src_dir):
        print(f"[build_ref] {src_dir} absent: nothing to build (GPU box uses the prebuilt .so)")
        return 0
    os.makedirs(OUT, exist_ok=True)
    so = os.path.join(OUT, NAME + ".so")
    if os.path.exists(so) and "--force" not in sys.argv:
        print(f"[build_ref] {so} already built")
        return 0

    os.environ.setdefault("TORCH_CUDA_ARCH_LIST", "10.0")
    os.environ.setdefault("MAX_JOBS", str(os.cpu_count() or 4))
    from torch.utils.cpp_extension import load

    tmp = tempfile.mkdtemp(prefix="m3r_ref_src_")
    try:
        shutil.copytree(src_dir, os.path.join(tmp, "backend"))
        gn = os.path.join(tmp, "backend", "src", "gn_kernels.cu")
        mk = os.path.join(tmp, "backend", "src", "matching_kernels.cu")
        s = open(gn).read()
        s, n1 = re.subn(
            r"torch::linalg::linalg_norm\(dx, std::optional<c10::Scalar>\(\), \{\}, false, \{\}\)",
            "at::linalg_norm(dx, std::optional<c10::Scalar>(), {}, false, std::nullopt)",
            s,
        )
        open(gn, "w").write(s)
        s = open(mk).read()
        s, n2 = re.subn(r"D11\.type\(\)", "D11.scalar_type()", s)
        open(mk, "w").write(s)
        assert n1 == 3 and n2 == 1, (n1, n2)
        build_dir = os.path.join(tmp, "build")
        os.makedirs(build_dir)
        load(
            name=NAME,
            sources=[
                os.path.join(tmp, "backend", "src", "gn.cpp"),
                gn,
                mk,
            ],
            extra_include_paths=[
                os.path.join(tmp, "backend", "include"),
                os.path.join(REF, "thirdparty", "eigen"),
            ],
            extra_cflags=["-O3"],
            extra_cuda_cflags=["-O3"],  # same flags as the reference's setup.py:28-30
            build_directory=build_dir,
            is_python_module=False,
            verbose=True,
        )
        shutil.copy(os.path.join(build_dir, NAME + ".so"), so)
        print(f"[build_ref] built {so}")
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
