"""Test configuration.  `-m "not gpu"`: oracle vs golden vectors / reference, host logic, C-ABI symbols (no GPU).
`-m gpu`: parity of the CUDA path (through the C ABI) against the oracle on a B200."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build what is missing (the .so files are git-ignored but travel to the GPU box with the snapshot)."""
    lib = os.path.join(ROOT, "radiative_transfer_b200", "librt_b200.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-C", os.path.join(ROOT, "radiative_transfer_b200", "csrc"), "-j8"], check=True,
                       stdout=subprocess.DEVNULL)
    port = os.path.join(ROOT, "oracle", "_build", "librt_oracle.so")
    if not os.path.exists(port):
        subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "port"], check=True, stdout=subprocess.DEVNULL)
    yield


@pytest.fixture(scope="session")
def port():
    from oracle.loader import Checker
    return Checker("port")


@pytest.fixture(scope="session")
def ref():
    from oracle.loader import Checker, ref_available
    if not ref_available():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference at build time)")
    return Checker("ref")


@pytest.fixture(scope="session")
def best_checker():
    """The compiled reference when it is there, else the C restatement."""
    from oracle.loader import Checker, ref_available
    return Checker("ref" if ref_available() else "port")


def rel_err(a, b, floor=1e-30):
    """max |a-b| / max(|a|,|b|) over entries whose magnitude exceeds `floor`."""
    a = np.asarray(a); b = np.asarray(b)
    den = np.maximum(np.abs(a), np.abs(b))
    mask = den > floor
    if not mask.any():
        return 0.0
    return float(np.max(np.abs(a - b)[mask] / den[mask]))


SMALL = ["generic", "oh", "oh_overlap", "ch3oh_small", "h2o_small"]


@pytest.fixture(scope="session")
def problems():
    from radiative_transfer_b200 import synthetic
    cache = {}

    def get(name, L=4):
        key = (name, L)
        if key not in cache:
            cache[key] = synthetic.small(name, L=L)
        return cache[key]
    return get
