#!/usr/bin/env python
"""Install the reference's pure-Python package `mast3r_slam` into baseline/_ref/ (test
infrastructure only; git-ignored, NOT gpurun-ignored, so it travels to the GPU box).

Why: tests/test_reference_callers_gpu.py runs the reference's OWN callers of the hot path --
mast3r_slam/matching.py and mast3r_slam/global_opt.py (FactorGraph), unmodified -- on a B200
against this library and against the reference's CUDA extension (oracle/_ref).  /root/reference
does not exist on the GPU box, so the package has to be installed somewhere that travels.

How: the contract's install command,
    python -m pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse \
        --target baseline/_ref <copy of /root/reference>
is tried first.  In this container it fails while generating metadata: setup.py:47-48 prints
"CUDA not found, cannot compile backend!" when torch.cuda.is_available() is false and then hits
`NameError: name 'ext_modules' is not defined` (setup.py:51).  The fallback installs exactly what
that build would have put on the path for the Python side: the files setuptools' package discovery
selects (pyproject.toml:19-22: packages.find include = ["mast3r_slam"]), i.e. mast3r_slam/*.py,
plus config/*.yaml (read at run time by mast3r_slam.config.load_config).  The native extension is
NOT taken from here: oracle/build_ref.py builds it for sm_100 into oracle/_ref/.
Nothing in the product imports baseline/_ref.
"""
import glob
import os
import shutil
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.environ.get("M3R_REFERENCE_ROOT", "/root/reference")
OUT = os.path.join(ROOT, "baseline", "_ref")


def main() -> int:
    if not os.path.isdir(os.path.join(REF, "mast3r_slam")):
        print(f"[install_reference_py] {REF}/mast3r_slam absent: nothing to install")
        return 0
    marker = os.path.join(OUT, "mast3r_slam", "global_opt.py")
    if os.path.exists(marker) and "--force" not in sys.argv:
        print(f"[install_reference_py] {OUT} already installed")
        return 0
    os.makedirs(OUT, exist_ok=True)
    how = "pip"
    tmp = tempfile.mkdtemp(prefix="m3r_ref_py_")
    try:
        for name in ("setup.py", "pyproject.toml"):
            shutil.copy(os.path.join(REF, name), tmp)
        shutil.copytree(os.path.join(REF, "mast3r_slam"), os.path.join(tmp, "mast3r_slam"))
        r = subprocess.run([sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation",
                            "--no-deps", "--find-links", "/opt/wheelhouse", "--target", OUT, tmp],
                           capture_output=True, text=True)
        if r.returncode != 0:
            why = [ln.strip() for ln in (r.stdout + r.stderr).splitlines() if "Error" in ln or "CUDA not found" in ln]
            print("[install_reference_py] pip install failed (" + "; ".join(why[:2]) + "); installing the "
                  "pure-Python package by file copy")
            how = "copy"
            dst = os.path.join(OUT, "mast3r_slam")
            os.makedirs(dst, exist_ok=True)
            for f in glob.glob(os.path.join(REF, "mast3r_slam", "*.py")):
                shutil.copy(f, dst)
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    os.makedirs(os.path.join(OUT, "config"), exist_ok=True)
    for f in glob.glob(os.path.join(REF, "config", "*.yaml")):
        shutil.copy(f, os.path.join(OUT, "config"))
    with open(os.path.join(OUT, "INSTALLED_BY"), "w") as fh:
        fh.write(f"oracle/install_reference_py.py ({how}) from {REF}\n")
    print(f"[install_reference_py] installed mast3r_slam ({how}) into {OUT}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
