"""pytest configuration: the `gpu` marker, import paths and shared fixtures.

`-m "not gpu"` (this container, no GPU): oracle vs golden vectors / closed forms, host logic,
C-ABI export check.  `-m gpu` (B200 box): parity of every C-ABI entry point against the oracle
and, when oracle/_ref holds the prebuilt reference extension, against the reference itself.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "mast3r-slam_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lib_path():
    """libm3r_b200.so, built on demand (nvcc cross-compiles without a GPU)."""
    sys.path.insert(0, PKG)
    import build as m3r_build

    return m3r_build.build()


@pytest.fixture(scope="session")
def backends(lib_path):
    import mast3r_slam_backends

    return mast3r_slam_backends


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle

    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def ref_ext():
    """The UNMODIFIED reference extension prebuilt into oracle/_ref (None if absent)."""
    import torch

    if not torch.cuda.is_available():
        return None
    from oracle import build_ref

    try:
        return build_ref.load_ref()
    except Exception as e:  # pragma: no cover
        print(f"[tests] reference extension not loadable: {e}")
        return None
