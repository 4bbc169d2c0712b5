"""Host-side helpers of bench.py that the driver's record depends on (no GPU): the parity comparison, the SURVEY 8d
byte formulas of the list modes, the normalizer pair sample."""
import numpy as np

import bench
from recommendation_system_pokec_b200.cabi import MODE_COLLAB, MODE_CLUBS
from recommendation_system_pokec_b200.synth import normalizer_pair_sample, sample_queries
from tests.conftest import make_test_world


def test_compare_lists_counts_every_kind_of_difference():
    ids = np.arange(12, dtype=np.int32).reshape(3, 4)
    sc = np.linspace(0.9, 0.1, 12, dtype=np.float32).reshape(3, 4)
    cnt = np.full(3, 4, np.int32)
    same = bench.compare_lists((ids, sc, cnt), (ids.copy(), sc.copy(), cnt.copy()))
    assert same["id_mismatches"] == 0 and same["one_ulp_flips"] == 0 and same["beyond_one_ulp"] == 0 and same["max_rel_err"] == 0.0
    ids2, sc2, cnt2 = ids.copy(), sc.copy(), cnt.copy()
    ids2[1, 2] = 99
    sc2[0, 0] = np.nextafter(sc2[0, 0], np.float32(2))         # one ulp
    sc2[2, 3] = sc2[2, 3] * np.float32(1.001)                  # far more than one ulp
    cnt2[2] = 3
    d = bench.compare_lists((ids2, sc2, cnt2), (ids, sc, cnt))
    assert d["id_mismatches"] == 1 and d["one_ulp_flips"] == 1 and d["beyond_one_ulp"] == 1 and d["count_mismatches"] == 1
    assert 0.0009 < d["max_rel_err"] < 0.0011 and d["queries"] == 3 and d["entries"] == 12


def test_list_mode_byte_formulas_follow_survey_8d():
    w = make_test_world(n=400, seed=3, T=6, pairs=100)
    q = sample_queries(w, 30, seed=1)
    rec = w.canonical_record_bytes()
    # by hand for one user: unique(friends U cands) records + 4 bytes per adjacency entry read
    look = w.subset_adj_lookup()

    def row(u):
        i = look.get(int(u))
        return w.adj[w.adj_ptr[i]:w.adj_ptr[i + 1]] if i is not None else w.adj[:0]

    u = int(next(x for x in q if len(row(x)) > 0))
    fr = row(u)
    fof = np.concatenate([row(f) for f in fr]) if len(fr) else w.adj[:0]
    fof = fof[fof != u]
    _, first = np.unique(fof, return_index=True)
    cands = fof[np.sort(first)][:5000]
    touched = np.unique(np.concatenate([fr, cands]))
    idx = w.uid_index()
    want = sum(int(rec[idx[int(x)]]) for x in touched if int(x) in idx) + 4 * (len(fr) + sum(len(row(f)) for f in fr))
    got = bench.list_mode_bytes(w, np.array([u], np.int32), MODE_COLLAB, 5000)
    assert got == want
    assert bench.list_mode_bytes(w, q, MODE_CLUBS, 5000) > bench.list_mode_bytes(w, q, MODE_COLLAB, 5000) > 0


def test_normalizer_pair_sample_is_distinct_and_reproducible():
    w = make_test_world(n=300, seed=5, T=4, pairs=50)
    a, b = normalizer_pair_sample(w, 2000, 5)
    a2, b2 = normalizer_pair_sample(w, 2000, 5)
    assert np.array_equal(a, a2) and np.array_equal(b, b2) and len(a) == 10000 and (a != b).all()
    key = np.minimum(a, b).astype(np.int64) * 10**7 + np.maximum(a, b)
    assert len(np.unique(key)) == len(key)
