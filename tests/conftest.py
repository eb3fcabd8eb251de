"""pytest configuration: markers, import paths, shared helpers."""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "lightweight-mast3r-slam_b200")
for p in (ROOT, PKG, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    import m3r_oracle

    m3r_oracle.build()
    return m3r_oracle


@pytest.fixture(scope="session")
def backends():
    """The product: the drop-in `mast3r_slam_backends` module over libm3r_b200.so."""
    import mast3r_slam_backends

    return mast3r_slam_backends


@pytest.fixture(scope="session")
def refcuda():
    """The reference's own CUDA extension compiled for sm_100 (oracle/_ref), or None."""
    import refcuda as rc

    return rc.load_ref()
