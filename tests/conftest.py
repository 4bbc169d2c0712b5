import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        from recommendation_system_pokec_b200.engine import load_library
        return load_library().pk_device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle_libs():
    from oracle import build_oracle
    return build_oracle.build_all()


def make_test_world(n=3000, seed=7, T=None, uid_gap=False, pairs=1500, **kw):
    from recommendation_system_pokec_b200.synth import make_world, sample_normalizers
    w = make_world(n=n, seed=seed, T=T, uid_gap=uid_gap, **kw)
    m, s = sample_normalizers(w, pairs=pairs)
    w.set_normalizers(m, s)
    return w


@pytest.fixture(scope="session")
def small_world():
    return make_test_world(n=3000, seed=7)


@pytest.fixture(scope="session")
def gap_world():
    """non-dense uids, partial normalizers, adjacency rows that point at users without a profile"""
    w = make_test_world(n=1500, seed=11, uid_gap=True)
    has = np.ones(w.n_fields, np.uint8)
    has[[1, 5, 9, 20]] = 0
    sd = w.norm_sd.copy()
    sd[[3, 12]] = 0.0
    w.set_normalizers(w.norm_mean, sd, has)
    # neighbours without a profile and an adjacency row for a uid without a profile
    rng = np.random.default_rng(5)
    adj = w.adj.copy()
    ghost = rng.random(adj.shape[0]) < 0.03
    adj[ghost] = np.int32(10_000_000) + rng.integers(0, 50, int(ghost.sum())).astype(np.int32)
    w.adj = adj
    return w
