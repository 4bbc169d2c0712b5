"""Reference edge cases the first round left untested, CUDA path through the C
ABI against the compiled unmodified reference (oracle/_ref, when it travelled) or the CPU restatement:

  * a user none of whose friends has a profile still gets zero-score collaborative candidates
    (src/recommender_graph.cpp:167-213)
  * duplicate ids inside UserProfile::clubs / ::friends: vec_set_similarity de-duplicates A and counts B with
    multiplicity over the raw sizes (src/recommender.cpp:119-128)
  * a friend listed twice in adj[u] is visited twice by the collaborative sum (src/recommender_graph.cpp:173)
  * candidate_limit = 10000 (the header default, include/recommender.h:24-28) with a hub user that truncates
  * topk beyond 256 for the list modes (src/test.cpp:75 passes |friends| / 5)
  * token cells of hundreds of tokens: the pre-filter's error bound scales with the longest cell, and gives the
    pre-filter up when the bound is useless
  * a prepared full scan is refused after any setter changed what it was planned for
"""
import os

import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS, EXCLUDE_NONE, EXCLUDE_SELF_AND_ADJ)
from recommendation_system_pokec_b200.synth import sample_queries
from tests.conftest import make_test_world, edit_world, best_oracle
from tests.test_gpu_parity import assert_topk_equal, assert_scores_close, bits, random_pairs

pytestmark = pytest.mark.gpu
NT = min(32, os.cpu_count() or 1)


def _dev(w):
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    return DeviceRecommender(w)


def ghost_friend_world():
    """user A's adjacency row holds only uids without a profile, whose own rows point at real users"""
    w = make_test_world(n=800, seed=31, T=8, pairs=300)
    a, b = int(w.uid[10]), int(w.uid[11])
    g1, g2, g3 = 9_000_001, 9_000_002, 9_000_003
    adj = {a: [g1, g2], g1: [int(u) for u in w.uid[20:40]] + [a], g2: [int(u) for u in w.uid[35:60]] + [g3],
           b: [g3], g3: []}
    return edit_world(w, adj=adj), a, b


def test_collab_user_without_profiled_friends(oracle_libs):
    w, a, b = ghost_friend_world()
    dev, orc = _dev(w), best_oracle(w)
    alone = dev.recommend(MODE_COLLAB, np.array([a], np.int32), 20, 5000)
    want = orc.recommend(MODE_COLLAB, np.array([a], np.int32), 20, 5000)
    assert want[2][0] == 20 and (want[1][0] == 0).all()            # the reference emits the candidates with score 0
    assert_topk_equal(alone, want, "collab, friends without profiles, single-user call")
    # the same user inside a batch, next to users with ordinary rows and one whose friends lead nowhere
    q = np.concatenate([[a, b], sample_queries(w, 30, seed=2)]).astype(np.int32)
    assert_topk_equal(dev.recommend(MODE_COLLAB, q, 20, 5000), orc.recommend(MODE_COLLAB, q, 20, 5000, NT), "collab batch")
    for mode in (MODE_GRAPH, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q, 20, 5000), orc.recommend(mode, q, 20, 5000, NT), f"mode {mode}")
    dev.close()
    orc.close()


def duplicate_world():
    w = make_test_world(n=1500, seed=17, T=10, pairs=500)
    rng = np.random.default_rng(3)
    clubs, friends, adj = {}, {}, {}
    adj_idx = {int(u): i for i, u in enumerate(w.adj_uid)}
    for r in rng.choice(w.n, 200, replace=False):
        c = w.clubs[w.clubs_ptr[r]:w.clubs_ptr[r + 1]]
        if len(c):
            clubs[int(r)] = np.concatenate([c, c[: 1 + len(c) // 2], c[:1]])           # repeated club ids
        f = w.friends[w.friends_ptr[r]:w.friends_ptr[r + 1]]
        if len(f):
            friends[int(r)] = np.concatenate([f[:2], f, f[-1:]])                       # repeated friend ids
        u = int(w.uid[r])
        if u in adj_idx:
            row = w.adj[w.adj_ptr[adj_idx[u]]:w.adj_ptr[adj_idx[u] + 1]]
            if len(row) > 1:
                adj[u] = np.concatenate([row, row[:2], [u]])                           # friend twice + a self loop
    return edit_world(w, clubs=clubs, friends=friends, adj=adj), np.array(sorted(clubs), np.int64)


def test_duplicate_ids_in_clubs_friends_and_adjacency(oracle_libs):
    w, rows = duplicate_world()
    dev, orc = _dev(w), best_oracle(w)
    rng = np.random.default_rng(5)
    # pairs with a duplicated list on the A side, on the B side, on both (the similarity is asymmetric then)
    a = np.concatenate([w.uid[rows], w.uid[rng.integers(0, w.n, len(rows))], w.uid[rows]])
    b = np.concatenate([w.uid[rng.integers(0, w.n, len(rows))], w.uid[rows], w.uid[rng.permutation(rows)]])
    # plus (user, friend) pairs: real overlaps
    a2, b2 = random_pairs(w, 4000, 9)
    a, b = np.concatenate([a, a2]).astype(np.int32), np.concatenate([b, b2]).astype(np.int32)
    assert np.array_equal(bits(dev.pair_fields(a, b)), bits(orc.pair_fields(a, b, NT)))
    assert_scores_close(dev.pair_similarity(a, b), orc.pair_similarity(a, b, NT), "FAS with duplicated list ids")
    asym = np.abs(dev.pair_fields(a, b)[:, 1:3] - dev.pair_fields(b, a)[:, 1:3]).max()
    assert asym > 0, "the sample does not exercise the asymmetric multiplicity"
    q = np.concatenate([w.uid[rows[:40]], sample_queries(w, 24, seed=6)]).astype(np.int32)
    assert_topk_equal(dev.fullscan(q, 30, EXCLUDE_SELF_AND_ADJ), orc.fullscan(q, 30, EXCLUDE_SELF_AND_ADJ, NT), "fullscan, duplicates")
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q, 30, 5000), orc.recommend(mode, q, 30, 5000, NT), f"mode {mode}, duplicates")
    dev.close()
    orc.close()


def hub_world():
    """a hub whose friend-of-friend expansion exceeds 10000 candidates: 40 friends with ~400 neighbours each"""
    w = make_test_world(n=30000, seed=23, T=12, pairs=500)
    rng = np.random.default_rng(8)
    hub = int(w.uid[77])
    fr = rng.choice(w.uid, 40, replace=False).astype(np.int32)
    adj = {hub: fr}
    for f in fr:
        adj[int(f)] = rng.choice(w.uid, 400, replace=False).astype(np.int32)
    return edit_world(w, adj=adj), hub


def test_hub_user_truncates_at_candidate_limit(oracle_libs):
    from oracle.oracle import CpuOracle
    w, hub = hub_world()
    dev, orc, port = _dev(w), best_oracle(w), CpuOracle(w, "port", compute_idf=False)
    for mode in (MODE_GRAPH, MODE_COLLAB):
        for limit in (10000, 5000, 401, 17, 1):
            got = dev.candidates(mode, hub, limit)
            assert np.array_equal(got, port.candidates(mode, hub, limit)), (mode, limit)
            assert len(got) == limit                                  # the hub really truncates
    q = np.concatenate([[hub], sample_queries(w, 6, seed=3)]).astype(np.int32)
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q, 50, 10000), orc.recommend(mode, q, 50, 10000, NT), f"hub, mode {mode}, limit 10000")
    assert_topk_equal(dev.recommend(MODE_COLLAB, q, 20, 5000), orc.recommend(MODE_COLLAB, q, 20, 5000, NT), "hub, collab, limit 5000")
    dev.close()
    orc.close()
    port.close()


@pytest.mark.parametrize("mode,k,limit", [(MODE_COLLAB, 1000, 1000), (MODE_GRAPH, 1000, 5000), (MODE_GRAPH, 300, 5000),
                                          (MODE_CLUBS, 700, 5000), (MODE_COLLAB, 2500, 10000)])
def test_topk_beyond_256_for_the_list_modes(small_world, oracle_libs, mode, k, limit):
    w = small_world
    dev, orc = _dev(w), best_oracle(w)
    q = sample_queries(w, 24, seed=40 + mode)
    got, want = dev.recommend(mode, q, k, limit), orc.recommend(mode, q, k, limit, NT)
    assert want[2].max() > 256 or mode == MODE_CLUBS
    assert_topk_equal(got, want, f"mode {mode} k={k}")
    dev.close()
    orc.close()


def long_cell_world(n_tok):
    """some users carry one cell of n_tok tokens in a free-text column, drawn from a shared pool so that the
    cosines between them have hundreds of matching products"""
    w = make_test_world(n=2500, seed=29, pairs=800)
    rng = np.random.default_rng(4)
    col = w.text_columns.index("hobbies")
    pool = np.arange(1, 2 * n_tok + 1)
    cells = {}
    for r in rng.choice(w.n, 60, replace=False):
        ids = np.sort(rng.choice(pool, n_tok, replace=False))
        cells[(int(r), col)] = (ids, rng.integers(1, 4, n_tok))
    return edit_world(w, cells=cells), np.array(sorted(r for r, _ in cells), np.int64)


def test_long_token_cells_scale_the_prefilter_margin(oracle_libs, monkeypatch):
    from recommendation_system_pokec_b200.engine import STAT_MARGIN_E9, STAT_PREFILTER
    base = make_test_world(n=2500, seed=29, pairs=800)
    d0 = _dev(base)
    d0.fullscan(base.uid[:4], 10, EXCLUDE_NONE)
    m0 = d0.stat(STAT_MARGIN_E9)
    d0.close()
    w, rows = long_cell_world(300)
    dev, orc = _dev(w), best_oracle(w)
    q = np.concatenate([w.uid[rows[:20]], sample_queries(w, 44, seed=7)]).astype(np.int32)
    got = dev.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ)
    m1 = dev.stat(STAT_MARGIN_E9)
    assert dev.stat(STAT_PREFILTER) == 1
    assert m1 > 2.5 * m0, f"margin did not follow the 300-token cells: {m0} -> {m1}"
    # the measured fp32 error against the exact kernel, on queries that own the long cells
    # (a few of them per group: 32 such queries would not fit one group's shared-memory tables)
    qd = np.concatenate([w.uid[rows[:5]], sample_queries(w, 27, seed=9)]).astype(np.int32)
    approx, exact = dev.debug_prefilter(qd, 0, w.n)
    worst = float(np.abs(approx.astype(np.float64) - exact).max())
    assert worst <= m1 / 1e9 / 2, f"pre-filter error {worst:.3e} vs bound {m1 / 1e9:.3e}"
    assert_topk_equal(got, orc.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ, NT), "fullscan with 300-token cells")
    monkeypatch.setenv("POKEC_B200_NO_PREFILTER", "1")
    exact_scan = dev.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ)
    monkeypatch.delenv("POKEC_B200_NO_PREFILTER")
    for x, y in zip(got, exact_scan):
        assert np.array_equal(x, y)
    dev.close()
    orc.close()
    # cells so long that the fp32 bound is useless: the pre-filter is given up, the result is still exact
    w2, rows2 = long_cell_world(4000)
    dev2, orc2 = _dev(w2), best_oracle(w2)
    q2 = np.concatenate([w2.uid[rows2[:6]], sample_queries(w2, 10, seed=8)]).astype(np.int32)
    got2 = dev2.fullscan(q2, 20, EXCLUDE_SELF_AND_ADJ)
    assert dev2.stat(STAT_PREFILTER) == 0 and dev2.stat(STAT_MARGIN_E9) > 1e6
    assert_topk_equal(got2, orc2.fullscan(q2, 20, EXCLUDE_SELF_AND_ADJ, NT), "fullscan with 4000-token cells")
    dev2.close()
    orc2.close()


def test_prepared_scan_is_refused_after_a_setter(small_world):
    from recommendation_system_pokec_b200.engine import PokecError
    w = small_world
    dev = _dev(w)
    q = sample_queries(w, 40, seed=1)

    def fresh():
        dev.fullscan_prepare(q, 10, EXCLUDE_SELF_AND_ADJ)
        dev.fullscan_execute()
        return dev.fullscan_fetch()

    ref = fresh()
    tok, idf, _ = dev.get_idf(3)
    u = int(q[0])
    changes = [lambda: dev.set_normalizers(w.norm_mean, w.norm_sd, w.norm_has),
               lambda: dev.set_shard(0, w.n // 2),
               lambda: dev.set_idf(3, tok, idf),
               lambda: dev.compute_idf(),
               lambda: dev.set_adj_row(u, np.array([int(w.uid[5])], np.int32)),
               lambda: dev.lib.pk_store_load_world(dev.h, __import__("ctypes").byref(dev._pw))]
    for i, change in enumerate(changes):
        dev.fullscan_prepare(q, 10, EXCLUDE_SELF_AND_ADJ)
        change()
        for call in (dev.fullscan_execute, dev.fullscan_execute_shortlists, dev.fullscan_fetch,
                     lambda: dev.fullscan_execute_finish(0, 1, 0, 1)):
            with pytest.raises(PokecError, match="stale|prepare"):
                call()
    # back to the original state: the same answer as before
    dev.compute_idf()
    dev.set_shard(0, w.n)
    dev.set_adj_row(u, w.adj[w.adj_ptr[np.searchsorted(w.adj_uid, u)]:w.adj_ptr[np.searchsorted(w.adj_uid, u) + 1]]
                    if u in set(w.adj_uid.tolist()) else np.zeros(0, np.int32))
    again = fresh()
    for x, y in zip(ref, again):
        assert np.array_equal(x, y)
    dev.close()


def test_device_expansion_equals_host_lists(small_world, gap_world, oracle_libs):
    """the friend-of-friend expansion on the device (pk_candidates_batch: adjacency CSR in HBM, one
    warp per user, ordered de-duplication + truncation) against the host walk (pk_candidates) and the restatement --
    every list bit for bit, at limits that truncate (1, 17), that do not (5000, 10000), on dense and on gapped uids,
    with neighbours that have no profile, self loops and duplicated neighbours."""
    from oracle.oracle import CpuOracle
    worlds = [small_world, gap_world, duplicate_world()[0]]
    hub_w, hub = hub_world()
    for wi, w in enumerate(worlds + [hub_w]):
        dev, port = _dev(w), CpuOracle(w, "port", compute_idf=False)
        users = sample_queries(w, 300, seed=60 + wi)
        if w is hub_w:
            users = np.concatenate([[hub], users[:40]]).astype(np.int32)
        users = np.concatenate([users, np.array([-7, 2_000_000_001], np.int32)])          # unknown users: empty lists
        for mode in (MODE_GRAPH, MODE_COLLAB):
            for limit in (1, 17, 401, 5000, 10000):
                ids, cnt = dev.candidates_batch(mode, users, limit)
                for qi, u in enumerate(users):
                    want = dev.candidates(mode, int(u), limit)
                    assert cnt[qi] == len(want) and np.array_equal(ids[qi, :cnt[qi]], want), (wi, mode, limit, int(u))
                for qi in (0, 5, 11):
                    assert np.array_equal(ids[qi, :cnt[qi]], port.candidates(mode, int(users[qi]), limit))
        dev.close()
        port.close()


def test_batched_list_modes_device_pipeline_equals_host_pipeline(small_world, oracle_libs, monkeypatch):
    """pk_recommend on a batch (device expansion -> lane-per-pair FAS -> reduce -> selection) against the host
    enumeration path (POKEC_B200_HOST_LISTS=1, the path single-user drop-in calls take) and the oracle"""
    w = small_world
    dev, orc = _dev(w), best_oracle(w)
    q = np.concatenate([sample_queries(w, 200, seed=77), np.array([-1, 123456789], np.int32)])
    for mode in (MODE_GRAPH, MODE_COLLAB):
        for k, limit in ((50, 5000), (20, 40), (300, 10000)):
            monkeypatch.delenv("POKEC_B200_HOST_LISTS", raising=False)
            a = dev.recommend(mode, q, k, limit)
            monkeypatch.setenv("POKEC_B200_HOST_LISTS", "1")
            b = dev.recommend(mode, q, k, limit)
            monkeypatch.delenv("POKEC_B200_HOST_LISTS")
            for x, y in zip(a, b):
                assert np.array_equal(x, y), (mode, k, limit)
            assert_topk_equal(a, orc.recommend(mode, q, k, limit, NT), f"device pipeline mode {mode} k={k} limit={limit}")
    # clubs: the edge stream built on the device, FAS values from the pair kernel or from the per-edge cache
    for cached in (False, True):
        if cached:
            dev.edge_cache_build()
        for k in (50, 700):
            monkeypatch.delenv("POKEC_B200_HOST_LISTS", raising=False)
            a = dev.recommend(MODE_CLUBS, q, k, 5000)
            monkeypatch.setenv("POKEC_B200_HOST_LISTS", "1")
            b = dev.recommend(MODE_CLUBS, q, k, 5000)
            monkeypatch.delenv("POKEC_B200_HOST_LISTS")
            for x, y in zip(a, b):
                assert np.array_equal(x, y), ("clubs", cached, k)
            assert_topk_equal(a, orc.recommend(MODE_CLUBS, q, k, 5000, NT), f"device clubs pipeline cached={cached} k={k}")
    dev.close()
    orc.close()
