"""The drop-in boundary end to end: the reference's UNCHANGED C++ `Recommender` / `TFIDFIndex` headers,
implemented by dropin/recommender_b200.cpp + dropin/tfidf_index_b200.cpp on top of the CUDA C ABI, driven
through the same C++ driver (oracle/ref_driver.cpp) as the compiled reference -- so both sides are called
exactly the way api_cli.cpp / test.cpp call them (include/recommender.h:24-44), one user per call.

Compared against the compiled reference when oracle/_ref travelled with the snapshot, else the CPU port."""
import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS
from recommendation_system_pokec_b200.synth import sample_queries
from tests.conftest import make_test_world
from tests.test_gpu_parity import assert_topk_equal, assert_scores_close, random_pairs, bits

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sides(oracle_libs):
    from oracle.oracle import CpuOracle, reference_available, dropin_available
    if not dropin_available():
        pytest.skip("dropin/_build/libpokec_dropin.so not present (built where the reference headers are)")
    w = make_test_world(n=2500, seed=31)
    drop = CpuOracle(w, "dropin")                  # Recommender::compute_idf_from_profiles runs on the device
    ref = CpuOracle(w, "reference" if reference_available() else "port")
    yield w, drop, ref
    drop.close()
    ref.close()


def test_idf_public_member_bit_exact(sides):
    """compute_idf_from_profiles fills the public idf_per_col map (callers copy it: src/test.cpp:40)"""
    w, drop, ref = sides
    for col in (0, 3, 22, 48):
        cell = np.arange(w.n) * w.T + col
        toks = np.unique(np.concatenate([w.tok_id[w.tok_ptr[c]:w.tok_ptr[c + 1]] for c in cell]))
        assert drop.idf_size(col) == ref.idf_size(col) == len(toks)
        g, gh = drop.idf_lookup(col, toks)
        r, rh = ref.idf_lookup(col, toks)
        assert gh.all() and rh.all()
        assert np.array_equal(bits(g), bits(r))


@pytest.mark.parametrize("mode,limit", [(MODE_GRAPH, 5000), (MODE_GRAPH, 30), (MODE_COLLAB, 5000), (MODE_CLUBS, 5000)])
def test_recommend_calls(sides, mode, limit):
    w, drop, ref = sides
    q = np.concatenate([sample_queries(w, 24, seed=40 + mode), np.array([-7, 77_000_000], np.int32)])
    assert_topk_equal(drop.recommend(mode, q, 20, limit), ref.recommend(mode, q, 20, limit, nthreads=8), f"drop-in mode {mode}")


def test_profile_similarity_members(sides):
    w, drop, ref = sides
    a, b = random_pairs(w, 300, 17)
    assert_scores_close(drop.pair_similarity(a, b), ref.pair_similarity(a, b, nthreads=8), "drop-in profile_similarity")


def test_private_field_helpers(sides):
    """vec_set_similarity / region_similarity_local / tfidf_cosine_for_column through the friend hook"""
    w, drop, ref = sides
    a, b = random_pairs(w, 12, 19)
    assert np.array_equal(bits(drop.pair_fields(a, b)), bits(ref.pair_fields(a, b, nthreads=4)))


def test_public_members_mutated_directly(sides):
    """callers write idf_per_col directly (src/test.cpp:37-40); the next call must see it"""
    w, drop, ref = sides
    a, b = random_pairs(w, 200, 23)
    tok = np.unique(w.tok_id[:400])
    try:
        for o in (drop, ref):
            o.erase_idf_column(2)
            o.erase_idf_tokens(22, tok)
        assert_scores_close(drop.pair_similarity(a, b), ref.pair_similarity(a, b, nthreads=8), "after idf edits")
        q = sample_queries(w, 8, seed=3)
        assert_topk_equal(drop.recommend(MODE_GRAPH, q, 20, 5000), ref.recommend(MODE_GRAPH, q, 20, 5000, nthreads=8), "graph after idf edits")
    finally:
        for o in (drop, ref):
            o.compute_idf()


def test_adjacency_rewritten_behind_the_recommender(sides):
    """src/test.cpp:59-73 rewrites adj rows while a Recommender holds the pointer"""
    from oracle.oracle import reference_available
    if not reference_available():
        pytest.skip("needs the compiled reference driver on both sides")
    w, drop, ref = sides
    q = sample_queries(w, 6, seed=9)
    look = w.subset_adj_lookup()
    for u in q:
        i = look.get(int(u))
        if i is None:
            continue
        row = w.adj[w.adj_ptr[i]:w.adj_ptr[i + 1]]
        if len(row) < 3:
            continue
        drop.set_adj_row(int(u), row[: len(row) // 2])
        ref.set_adj_row(int(u), row[: len(row) // 2])
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(drop.recommend(mode, q, 20, 1000), ref.recommend(mode, q, 20, 1000, nthreads=4), f"mode {mode} after adjacency edits")


def test_tfidf_index_build(sides):
    """TFIDFIndex::build (src/tfidf_index.cpp:7-30): double-log idf, rounded to float"""
    from oracle.oracle import reference_available
    if not reference_available():
        pytest.skip("needs the compiled reference")
    w, drop, ref = sides
    toks = np.unique(w.tok_id[w.tok_ptr[7]:w.tok_ptr[7 + 49 * 50]])
    for col in (1, 7):
        g, gh = drop.tfidf_index_lookup(col, toks)
        r, rh = ref.tfidf_index_lookup(col, toks)
        assert np.array_equal(gh, rh) and np.array_equal(bits(g), bits(r))
