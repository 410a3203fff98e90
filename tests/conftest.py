import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return graft.load_package()


@pytest.fixture(scope="session")
def oracle():
    return graft.load_oracle()


@pytest.fixture(scope="session")
def synth(pkg):
    import importlib
    return importlib.import_module("combat-seq_b200.synth")


@pytest.fixture(scope="session")
def hostmath_lib():
    """tests/hostmath/libhostmath.so: the product's device math headers compiled for the host (rebuilt when a header changes)."""
    import ctypes
    import subprocess
    here = os.path.dirname(os.path.abspath(__file__))
    src = os.path.join(here, "hostmath", "hostmath.cu")
    out = os.path.join(here, "hostmath", "libhostmath.so")
    deps = [src] + [os.path.join(ROOT, "combat-seq_b200", "csrc", h) for h in ("nbmath.cuh", "fastmath.cuh", "rng.cuh")]
    if not os.path.exists(out) or os.path.getmtime(out) < max(os.path.getmtime(d) for d in deps):
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc, "-O2", "-std=c++17", "-shared", "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets",
                        "-o", out, src], check=True)
    return ctypes.CDLL(out)


def rel_err(a, b, rtol=1e-6, atol=1e-9):
    """max |a-b| / (rtol*|b| + atol) over the finite entries of b; NaN patterns must agree."""
    a = np.asarray(a); b = np.asarray(b)
    assert np.array_equal(np.isnan(a), np.isnan(b)), "NaN pattern differs"
    m = ~np.isnan(b)
    if not m.any():
        return 0.0
    return float(np.max(np.abs(a[m] - b[m]) / (rtol * np.abs(b[m]) + atol)))
