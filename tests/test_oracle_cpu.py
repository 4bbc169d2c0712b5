"""Pins the CPU restatement (oracle/fas_oracle.cpp) before anything trusts it:

  * against the golden fixtures of tests/golden/ -- outputs of the UNMODIFIED compiled reference, generated
    by tests/golden/make_golden.py (runs everywhere, including the GPU box where /root/reference is absent);
  * against the compiled reference itself on larger seeded samples, where oracle/_ref/libpokec_ref.so exists.

Contract: ids / counts / idf bits / per-field floats bit-exact; final scores bit-exact up to the
documented 1-ulp flips of order-dependent double sums (counted and bounded)."""
import os

import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS
from recommendation_system_pokec_b200.synth import sample_queries
from recommendation_system_pokec_b200.world import World
from tests.conftest import make_test_world

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.int32)


def ulp_diff(a, b):
    return np.abs(bits(a).astype(np.int64) - bits(b).astype(np.int64))


def check_scores(got, want, what, max_flips=1):
    d = ulp_diff(got, want)
    assert int(d.max(initial=0)) <= 1, f"{what}: {int(d.max())} ulps"
    assert int((d != 0).sum()) <= max_flips, f"{what}: {int((d != 0).sum())} one-ulp flips"


def check_topk(got, want, what):
    assert np.array_equal(got[2], want[2]), f"{what}: counts"
    assert np.array_equal(got[0], want[0]), f"{what}: ids"
    check_scores(got[1], want[1], what)


@pytest.fixture(scope="module")
def golden(oracle_libs):
    from oracle.oracle import CpuOracle
    w = World.load(os.path.join(GOLDEN, "world_small.npz"))
    g = np.load(os.path.join(GOLDEN, "golden_small.npz"))
    orc = CpuOracle(w, "port")
    yield w, g, orc
    orc.close()


def test_golden_pairs(golden):
    w, g, orc = golden
    a, b = g["pair_a"], g["pair_b"]
    assert np.array_equal(bits(orc.pair_fields(a, b)), bits(g["pair_fields"]))
    check_scores(orc.pair_similarity(a, b), g["pair_fas"], "pair FAS vs golden")
    assert (g["pair_fields"][:, 1:3] > 0).sum() > 20      # the fixture holds real club / friend overlaps


def test_golden_idf(golden):
    w, g, orc = golden
    for col in (0, 2, 22, 32, 48):
        tok = g[f"idf_tok_{col}"]
        assert orc.idf_size(col) == len(tok)
        val, has = orc.idf_lookup(col, tok)
        assert has.all() and np.array_equal(bits(val), bits(g[f"idf_val_{col}"]))
    orc.compute_idf(double_log=True)            # TFIDFIndex::build variant
    try:
        for col in (0, 2, 22, 32, 48):
            val, has = orc.idf_lookup(col, g[f"idf_tok_{col}"])
            assert has.all() and np.array_equal(bits(val), bits(g[f"idf_dbl_{col}"]))
    finally:
        orc.compute_idf()


@pytest.mark.parametrize("name,mode,limit", [("graph", MODE_GRAPH, 5000), ("graph_l25", MODE_GRAPH, 25),
                                             ("collab", MODE_COLLAB, 5000), ("collab_l40", MODE_COLLAB, 40),
                                             ("clubs", MODE_CLUBS, 5000)])
def test_golden_recommend(golden, name, mode, limit):
    w, g, orc = golden
    got = orc.recommend(mode, g["queries"], 20, limit, nthreads=4)
    check_topk(got, (g[f"{name}_id"], g[f"{name}_score"], g[f"{name}_cnt"]), name)
    assert g[f"{name}_cnt"].max() > 0


@pytest.mark.parametrize("ex", [0, 1, 2])
def test_golden_fullscan(golden, ex):
    w, g, orc = golden
    got = orc.fullscan(g["queries"], 30, ex, nthreads=4)
    check_topk(got, (g[f"fullscan{ex}_id"], g[f"fullscan{ex}_score"], g[f"fullscan{ex}_cnt"]), f"fullscan exclude={ex}")


def test_port_vs_compiled_reference(oracle_libs):
    """larger differential run against the reference itself (this container only)"""
    from oracle.oracle import CpuOracle, reference_available
    if not reference_available():
        pytest.skip("oracle/_ref/libpokec_ref.so not present")
    w = make_test_world(n=2500, seed=13)
    port, ref = CpuOracle(w, "port"), CpuOracle(w, "reference")
    rng = np.random.default_rng(1)
    a = w.uid[rng.integers(0, w.n, 20000)].astype(np.int32)
    b = w.uid[rng.integers(0, w.n, 20000)].astype(np.int32)
    assert np.array_equal(bits(port.pair_fields(a[:4000], b[:4000], nthreads=8)), bits(ref.pair_fields(a[:4000], b[:4000], nthreads=8)))
    check_scores(port.pair_similarity(a, b, nthreads=8), ref.pair_similarity(a, b, nthreads=8), "pair FAS vs reference", max_flips=2)
    q = np.concatenate([sample_queries(w, 40, seed=5), np.array([-1, 99_999_999], np.int32)])
    for mode, limit in ((MODE_GRAPH, 5000), (MODE_GRAPH, 17), (MODE_COLLAB, 5000), (MODE_COLLAB, 33), (MODE_CLUBS, 5000)):
        check_topk(port.recommend(mode, q, 50, limit, nthreads=8), ref.recommend(mode, q, 50, limit, nthreads=8), f"mode {mode} limit {limit}")
    check_topk(port.fullscan(q[:12], 50, 2, nthreads=8), ref.fullscan(q[:12], 50, 2, nthreads=8), "fullscan vs reference")
    port.close()
    ref.close()


def test_normalizer_pair_sample_matches_the_reference(oracle_libs):
    """oracle/ref_driver.cpp restates the pair sampling of compute_column_normalizers (src/utils.cpp:163-181); the
    scalar fields recomputed from those pairs in numpy (double sums in pair order) must reproduce the reference's
    mean / sd floats bit for bit -- which pins the sample the GPU test feeds to pk_normalizers_from_pairs"""
    from oracle.oracle import CpuOracle, reference_available
    if not reference_available():
        pytest.skip("oracle/_ref was not built (no /root/reference here)")
    from tests.conftest import make_test_world
    w = make_test_world(n=2000, seed=13, T=8, pairs=500)
    orc = CpuOracle(w, "reference")
    mean, sd = orc.column_normalizers(300, 5)
    a, b = orc.normalizer_pairs(300, 5)
    assert len(a) == 1500 and (a != b).all()
    row = {int(u): i for i, u in enumerate(w.uid)}
    ra = np.array([row[int(x)] for x in a])
    rb = np.array([row[int(x)] for x in b])

    def stats(v):
        mu = 0.0
        for x in v:
            mu += float(x)
        mu /= len(v)
        acc = 0.0
        for x in v:
            acc += (float(x) - mu) ** 2
        s = np.sqrt(acc / (len(v) - 1))
        return np.float32(mu), np.float32(1.0 if s == 0 else s)

    pub = ((w.pub[ra] >= 0) & (w.pub[rb] >= 0) & (w.pub[ra] == w.pub[rb])).astype(np.float64)
    x, y = w.age[ra].astype(np.float64), w.age[rb].astype(np.float64)
    age = np.where((x > 0) & (y > 0), np.minimum(x, y) / np.maximum(np.maximum(x, y), 1.0), 0.0)
    for fld, v in ((0, pub), (3, age)):
        m, s = stats(v)
        assert m.tobytes() == mean[fld].tobytes() and s.tobytes() == sd[fld].tobytes(), f"field {fld}"
    orc.close()


def test_world_file_round_trip_and_rejects_damage(tmp_path):
    """SURVEY 8 f2 (pk_world_save / pk_world_open, host only): every array of the world survives the file bit for
    bit; truncated, corrupted and foreign files are refused."""
    from recommendation_system_pokec_b200 import engine
    from tests.conftest import make_test_world
    w = make_test_world(n=300, seed=41, T=5, pairs=100, uid_gap=True)
    path = str(tmp_path / "world.pkw")
    engine.world_save(w, path)
    assert os.path.getsize(path) % 64 == 0 and not os.path.exists(path + ".tmp")
    with engine.WorldFile(path) as f:
        assert (f.view.n_users, f.view.n_text_cols, f.view.n_adj_rows) == (w.n, w.T, w.n_adj_rows)
        back = f.to_world(w.text_columns)
    for name in w._DTYPES:
        a, b = getattr(w, name), getattr(back, name)
        assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), name
    blob = open(path, "rb").read()
    cases = {"short": blob[:100], "truncated": blob[:-64], "magic": b"X" + blob[1:],
             "header": blob[:40] + bytes([blob[40] ^ 1]) + blob[41:], "empty": b"",
             # one flipped bit deep inside the payload (the last section with content), header untouched
             "payload": blob[:len(blob) - 100] + bytes([blob[len(blob) - 100] ^ 4]) + blob[len(blob) - 99:],
             "payload_mid": blob[:len(blob) // 2] + bytes([blob[len(blob) // 2] ^ 0x80]) + blob[len(blob) // 2 + 1:]}
    for tag, data in cases.items():
        bad = str(tmp_path / f"{tag}.pkw")
        open(bad, "wb").write(data)
        with pytest.raises(engine.PokecError):
            engine.WorldFile(bad)
    with pytest.raises(engine.PokecError):
        engine.WorldFile(str(tmp_path / "missing.pkw"))
    # an empty world is a valid file
    e = make_test_world(n=300, seed=41, T=5, pairs=100)
    e2 = type(e)(text_columns=e.text_columns, uid=e.uid[:0], pub=e.pub[:0], comp=e.comp[:0], gender=e.gender[:0], age=e.age[:0],
                 region=e.region[:0], clubs_ptr=np.zeros(1, np.int64), clubs=e.clubs[:0], friends_ptr=np.zeros(1, np.int64),
                 friends=e.friends[:0], tok_ptr=np.zeros(1, np.int64), tok_id=e.tok_id[:0], tok_tf=e.tok_tf[:0],
                 adj_uid=e.adj_uid[:0], adj_ptr=np.zeros(1, np.int64), adj=e.adj[:0])
    p0 = str(tmp_path / "empty_world.pkw")
    engine.world_save(e2, p0)
    with engine.WorldFile(p0) as f:
        assert f.view.n_users == 0 and f.to_world(e.text_columns).n == 0


def test_world_file_round_trip_through_reference_containers(tmp_path):
    """SURVEY 8 f2 on the reference side: pokec_b200::save_world_file writes the drop-in Recommender's profile map,
    adjacency map and normalizer maps; load_world_file rebuilds fresh unordered_maps of UserProfile from the file;
    the C++ check in oracle/ref_driver.cpp (drop-in build) compares every member.  Host only: no GPU call is made."""
    import ctypes as C
    from oracle.oracle import CpuOracle, dropin_available
    if not dropin_available():
        pytest.skip("dropin/_build/libpokec_dropin.so not built (needs the reference headers)")
    w = make_test_world(n=400, seed=5, T=6, pairs=200, uid_gap=True)
    o = CpuOracle(w, kind="dropin", compute_idf=False)
    o.lib.ref_world_file_round_trip.argtypes = [C.c_void_p, C.c_char_p]
    path = str(tmp_path / "w.pkw")
    assert o.lib.ref_world_file_round_trip(o.h, path.encode()) == 0
    # the file the drop-in wrote is the same format the Python side reads
    from recommendation_system_pokec_b200 import engine
    with engine.WorldFile(path) as f:
        back = f.to_world(w.text_columns)
    order = np.argsort(w.uid)
    assert np.array_equal(back.uid, w.uid[order]) and np.array_equal(back.age, w.age[order])
    assert np.array_equal(back.norm_mean, w.norm_mean) and np.array_equal(back.norm_has, w.norm_has)
    o.close()


def test_restatement_vs_reference_on_edge_case_worlds(oracle_libs):
    """the worlds of tests/test_edge_cases_gpu.py (friends without profiles, duplicated ids inside clubs / friends /
    adjacency rows): the CPU restatement must agree with the compiled reference there too, because the GPU box may
    only have the restatement to check the CUDA path against"""
    from oracle.oracle import CpuOracle, reference_available
    if not reference_available():
        pytest.skip("oracle/_ref/libpokec_ref.so not present")
    from tests.test_edge_cases_gpu import ghost_friend_world, duplicate_world
    from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS
    from recommendation_system_pokec_b200.synth import sample_queries
    w, a, b = ghost_friend_world()
    port, ref = CpuOracle(w, "port"), CpuOracle(w, "reference")
    q = np.concatenate([[a, b], sample_queries(w, 20, seed=2)]).astype(np.int32)
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        for x, y in zip(port.recommend(mode, q, 20, 5000, 4), ref.recommend(mode, q, 20, 5000, 4)):
            assert np.array_equal(x, y), f"ghost world, mode {mode}"
    one = ref.recommend(MODE_COLLAB, np.array([a], np.int32), 20, 5000)
    assert one[2][0] == 20 and (one[1][0] == 0).all()
    port.close(); ref.close()
    w, rows = duplicate_world()
    port, ref = CpuOracle(w, "port"), CpuOracle(w, "reference")
    rng = np.random.default_rng(5)
    a = np.concatenate([w.uid[rows], w.uid[rng.integers(0, w.n, len(rows))]]).astype(np.int32)
    b = np.concatenate([w.uid[rng.integers(0, w.n, len(rows))], w.uid[rows]]).astype(np.int32)
    assert np.array_equal(port.pair_fields(a, b, 4).view(np.int32), ref.pair_fields(a, b, 4).view(np.int32))
    q = np.concatenate([w.uid[rows[:20]], sample_queries(w, 12, seed=6)]).astype(np.int32)
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        for x, y in zip(port.recommend(mode, q, 30, 5000, 4), ref.recommend(mode, q, 30, 5000, 4)):
            assert np.array_equal(x, y), f"duplicate world, mode {mode}"
    for x, y in zip(port.fullscan(q, 30, 2, 4), ref.fullscan(q, 30, 2, 4)):
        assert np.array_equal(x, y)
    port.close(); ref.close()
