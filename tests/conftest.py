"""Shared fixtures.  `-m "not gpu"` runs here (no GPU); `-m gpu` runs on a B200.

oracle/ is test infrastructure: these fixtures are the only way the tests reach it."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (sm_100a) GPU")


@pytest.fixture(scope="session")
def pyoracle():
    return graft.load_oracle()


@pytest.fixture(scope="session")
def orc(pyoracle):
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def ref(pyoracle):
    """The reference's own CPU code, compiled from /root/reference into oracle/_ref (prebuilt on the GPU box)."""
    try:
        pyoracle.build_ref()
        return pyoracle.RefCPU()
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"oracle/_ref/libggml_ref.so unavailable: {e}")


@pytest.fixture(scope="session")
def t4():
    lib = os.path.join(ROOT, "llama.cpp-rrs_b200", "libtcq4_b200.so")
    if not os.path.exists(lib):
        graft.build()
    return graft.load_package()


def nmse(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(((a - b) ** 2).sum() / max((b ** 2).sum(), 1e-300))


def make_weights(rng, N, K, std=0.02):
    return (rng.standard_normal((N, K)) * std).astype(np.float32)


def make_acts(rng, M, K, outliers=True):
    """x ~ N(0,1) with 1% of channels x20 (RRS's motivating case)."""
    x = rng.standard_normal((M, K)).astype(np.float32)
    if outliers:
        idx = rng.choice(K, size=max(1, K // 100), replace=False)
        x[:, idx] *= 20.0
    return x


def block_perm(rng, K):
    """random permutation that maps every 256-block onto itself (quantize.cpp:147-159)."""
    p = np.arange(K, dtype=np.int32).reshape(K // 256, 256)
    for b in range(p.shape[0]):
        p[b] = p[b][rng.permutation(256)]
    return p.reshape(-1)
