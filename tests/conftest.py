"""tests/conftest.py -- markers and shared fixtures.

`-m "not gpu"`: oracle vs golden vectors / reference, host logic, C-ABI loads.
`-m gpu`: the parity tests proper; every call goes through libnnbvh.so's C ABI.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    from oracle import pyoracle
    if not pyoracle.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference on this machine)")
    return pyoracle.Reference(threads=1)


@pytest.fixture(scope="session")
def golden():
    return {n: np.load(os.path.join(GOLDEN, n + ".npz"))
            for n in ("tri_vectors", "box_vectors", "scene_small", "hlbvh")}


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The CUDA library must exist; tests never fall back to anything else."""
    from nn_bvh_b200 import capi
    if not os.path.exists(capi.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()


def soup(rng, T, s):
    c = rng.random((T, 1, 3), dtype=np.float32)
    return (c + np.float32(s) * (rng.random((T, 3, 3), dtype=np.float32) * 2 - 1)).reshape(T, 9)


def rand_rays(rng, N, lo=-0.25, hi=1.25, tmax=np.inf):
    r = np.zeros((N, 8), dtype=np.float32)
    r[:, :3] = lo + (hi - lo) * rng.random((N, 3))
    d = rng.normal(size=(N, 3))
    r[:, 4:7] = d / np.linalg.norm(d, axis=1, keepdims=True)
    r[:, 3] = tmax
    return r
