"""pytest configuration: `gpu` marker, import paths, shared fixtures.

`-m "not gpu"` : oracle vs golden vectors / analytic anchors, kernel cores in CPU warp-emulation vs
                 the oracle, host-side logic, C-ABI symbol checks.  No CUDA device needed.
`-m gpu`       : parity tests proper -- the CUDA path through the C ABI vs the oracle.  These never
                 read /root/reference (absent on the GPU box).
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "quadruped-reactive-walking_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

# tolerances stated by BASELINE.json north_star: contact forces within 1e-4 N abs / 1e-5 rel
F_ATOL = 1e-4
F_RTOL = 1e-5


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _have_gpu():
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import bindings

    bindings.build()
    return bindings


@pytest.fixture(scope="session")
def emu():
    """ctypes handle of the warp-emulation build of the kernel cores (tests/emu)."""
    import ctypes as C

    d = os.path.join(ROOT, "tests", "emu")
    subprocess.check_call(["make", "-C", d, "-s"], stdout=subprocess.DEVNULL)
    return C.CDLL(os.path.join(d, "libqrw_emu.so"))


@pytest.fixture(scope="session")
def ab_library():
    """Path of the A/B build of the library (tests/ab/libqrw_b200_ab.so: -DQRW_AB_KERNELS adds the superseded MPC kernels and
    the nested-dissection experiment).  Test infrastructure; built by __graft_entry__.build() so that it travels to the GPU box."""
    from qrw_b200 import _lib

    return _lib.build_ab_library()


def assert_forces_close(got, want, what=""):
    import numpy as np

    got, want = np.asarray(got), np.asarray(want)
    err = np.abs(got - want)
    tol = F_ATOL + F_RTOL * np.abs(want)
    assert (err <= tol).all(), f"{what}: max |diff| = {err.max():.3e} (tolerance {F_ATOL} abs + {F_RTOL} rel)"
