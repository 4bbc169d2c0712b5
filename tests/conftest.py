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


def edit_world(w, clubs=None, friends=None, adj=None, cells=None):
    """A copy of world `w` with some per-user lists replaced (tests of the reference's edge cases):
      clubs / friends: {row: array}   UserProfile::clubs / ::friends of that row, verbatim (duplicates allowed)
      adj:             {uid: array}   adjacency row of that uid, verbatim; rows are created when missing
      cells:           {(row, col): (token ids, tfs)}   one token cell (ids unique inside a cell)"""
    from recommendation_system_pokec_b200.world import World

    def csr_replace(ptr, val, repl, n_rows):
        rows = [val[ptr[i]:ptr[i + 1]] for i in range(n_rows)]
        for i, a in repl.items():
            rows[i] = np.asarray(a, val.dtype)
        lens = np.array([len(r) for r in rows], np.int64)
        p = np.zeros(n_rows + 1, np.int64)
        np.cumsum(lens, out=p[1:])
        return p, (np.concatenate(rows) if len(rows) else val[:0]).astype(val.dtype)

    clubs_ptr, clubs_v = (w.clubs_ptr, w.clubs) if not clubs else csr_replace(w.clubs_ptr, w.clubs, clubs, w.n)
    friends_ptr, friends_v = (w.friends_ptr, w.friends) if not friends else csr_replace(w.friends_ptr, w.friends, friends, w.n)
    adj_uid, adj_ptr, adj_v = w.adj_uid, w.adj_ptr, w.adj
    if adj:
        rows = {int(u): w.adj[w.adj_ptr[i]:w.adj_ptr[i + 1]] for i, u in enumerate(w.adj_uid)}
        for u, a in adj.items():
            rows[int(u)] = np.asarray(a, np.int32)
        keys = sorted(rows)
        adj_uid = np.array(keys, np.int32)
        adj_ptr = np.zeros(len(keys) + 1, np.int64)
        np.cumsum([len(rows[k]) for k in keys], out=adj_ptr[1:])
        adj_v = np.concatenate([rows[k] for k in keys]).astype(np.int32) if keys else w.adj[:0]
    tok_ptr, tok_id, tok_tf = w.tok_ptr, w.tok_id, w.tok_tf
    if cells:
        T = w.T
        idx = {r * T + c: v for (r, c), v in cells.items()}
        ids, tfs, lens = [], [], np.diff(w.tok_ptr).copy()
        for c in sorted(idx):
            lens[c] = len(idx[c][0])
        order = sorted(idx)
        prev = 0
        for c in order:
            ids.append(w.tok_id[w.tok_ptr[prev]:w.tok_ptr[c]]); tfs.append(w.tok_tf[w.tok_ptr[prev]:w.tok_ptr[c]])
            ids.append(np.asarray(idx[c][0], np.int32)); tfs.append(np.asarray(idx[c][1], np.int32))
            prev = c + 1
        ids.append(w.tok_id[w.tok_ptr[prev]:]); tfs.append(w.tok_tf[w.tok_ptr[prev]:])
        tok_id, tok_tf = np.concatenate(ids).astype(np.int32), np.concatenate(tfs).astype(np.int32)
        tok_ptr = np.zeros(w.n * T + 1, np.int64)
        np.cumsum(lens, out=tok_ptr[1:])
    out = World(text_columns=list(w.text_columns), uid=w.uid, pub=w.pub, comp=w.comp, gender=w.gender, age=w.age,
                region=w.region, clubs_ptr=clubs_ptr, clubs=clubs_v, friends_ptr=friends_ptr, friends=friends_v,
                tok_ptr=tok_ptr, tok_id=tok_id, tok_tf=tok_tf, adj_uid=adj_uid, adj_ptr=adj_ptr, adj=adj_v,
                norm_mean=w.norm_mean, norm_sd=w.norm_sd, norm_has=w.norm_has)
    out.meta = dict(w.meta, edited=True)
    return out


def best_oracle(world):
    """the compiled unmodified reference where its library travelled with the snapshot, else the CPU restatement"""
    from oracle.oracle import CpuOracle, reference_available
    return CpuOracle(world, "reference" if reference_available() else "port")
