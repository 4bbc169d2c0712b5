"""Parity of the CUDA path (through the C ABI) with the CPU oracle on the same seeded inputs.

Contract:
  bit-exact   df counts, idf table, set-overlap counts, per-field float similarities, candidate id lists,
              top-k id sequences including the uid tie-break
  tolerance   final FAS / collaborative / club scores: <= 1e-5 relative (north_star); in practice they are
              bit-identical except where the device exp() and the host exp() round differently, which moves
              a double by 1 ulp and flips the final float with probability ~1e-8 per value.  Such flips are
              counted and bounded here, never hidden.
"""
import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS, EXCLUDE_NONE, EXCLUDE_SELF,
                                                   EXCLUDE_SELF_AND_ADJ)
from recommendation_system_pokec_b200.synth import sample_queries
from tests.conftest import make_test_world

pytestmark = pytest.mark.gpu

REL_TOL = 1e-5          # north_star tolerance for floating-point scores
MAX_ULP_FLIP_FRAC = 1e-5  # fraction of scores allowed to differ by one float ulp


def bits(x):
    return np.ascontiguousarray(x, np.float32).view(np.int32)


def assert_scores_close(got, want, what):
    got = np.asarray(got, np.float32)
    want = np.asarray(want, np.float32)
    assert got.shape == want.shape
    denom = np.maximum(np.abs(want), 1e-30)
    rel = np.abs(got.astype(np.float64) - want) / denom
    assert float(rel.max(initial=0.0)) <= REL_TOL, f"{what}: max rel err {rel.max()}"
    ulp = np.abs(bits(got).astype(np.int64) - bits(want).astype(np.int64))
    assert int(ulp.max(initial=0)) <= 1, f"{what}: a score differs by {ulp.max()} ulps"
    nflip = int((ulp != 0).sum())
    assert nflip <= max(1, int(MAX_ULP_FLIP_FRAC * got.size)), f"{what}: {nflip} of {got.size} scores differ by 1 ulp"
    return nflip


def assert_topk_equal(got, want, what):
    gi, gs, gc = got
    wi, ws, wc = want
    assert np.array_equal(gc, wc), f"{what}: result counts differ"
    if not np.array_equal(gi, wi):
        # an id swap is only acceptable between neighbours whose reference scores are within one ulp
        bad = np.argwhere(gi != wi)
        for q, j in bad:
            near = np.abs(bits(ws[q]).astype(np.int64) - int(bits(ws[q, j:j + 1])[0])) <= 1
            assert gi[q, j] in wi[q][near], f"{what}: query {q} rank {j}: id {gi[q, j]} vs {wi[q, j]} not explained by a 1-ulp tie"
        assert len(bad) <= 4, f"{what}: {len(bad)} id mismatches"
    assert_scores_close(gs, ws, what)


@pytest.fixture(scope="module")
def env(small_world, oracle_libs):
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    dev = DeviceRecommender(small_world)
    orc = CpuOracle(small_world, "port")
    yield small_world, dev, orc
    dev.close()
    orc.close()


def random_pairs(w, n, seed):
    rng = np.random.default_rng(seed)
    a = w.uid[rng.integers(0, w.n, n)]
    b = w.uid[rng.integers(0, w.n, n)]
    # half of the pairs are (user, one of the user's friends): non-trivial overlaps
    rows = rng.integers(0, w.n, n // 2)
    deg = np.diff(w.friends_ptr)[rows]
    ok = deg > 0
    pick = w.friends_ptr[rows[ok]] + (rng.random(int(ok.sum())) * deg[ok]).astype(np.int64)
    a[:int(ok.sum())] = w.uid[rows[ok]]
    b[:int(ok.sum())] = w.friends[pick].astype(np.int32)
    return a.astype(np.int32), b.astype(np.int32)


def test_idf_table_bit_exact(env):
    w, dev, orc = env
    for col in (0, 2, 7, 22, 48):
        tok, idf, df = dev.get_idf(col)
        assert len(tok) == orc.idf_size(col)
        want, has = orc.idf_lookup(col, tok)
        assert has.all()
        assert np.array_equal(bits(idf), bits(want)), f"idf of column {col}"
        cell = np.arange(w.n) * w.T + col
        cnt = np.zeros(int(tok.max()) + 1 if len(tok) else 1, np.int64)
        for c in cell:
            ids = w.tok_id[w.tok_ptr[c]:w.tok_ptr[c + 1]]
            cnt[ids] += 1
        assert np.array_equal(df, cnt[tok]), f"df of column {col}"


def test_pair_fields_and_overlaps_bit_exact(env):
    w, dev, orc = env
    a, b = random_pairs(w, 6000, 1)
    got = dev.pair_fields(a, b)
    want = orc.pair_fields(a, b, nthreads=8)
    assert np.array_equal(bits(got), bits(want))
    for which in (0, 1):
        assert np.array_equal(dev.pair_overlap(a, b, which), orc.pair_overlap(a, b, which))
    assert (orc.pair_overlap(a, b, 1) > 0).sum() > 100   # the sample exercises real overlaps


def test_pair_similarity(env):
    w, dev, orc = env
    a, b = random_pairs(w, 40000, 2)
    a[:5] = [-5, 0, 2_000_000_000, a[5], a[6]]          # unknown uids -> -1
    got = dev.pair_similarity(a, b)
    want = orc.pair_similarity(a, b, nthreads=8)
    assert (got[:3] == -1).all() and (want[:3] == -1).all()
    assert_scores_close(got, want, "pair FAS")


@pytest.mark.parametrize("exclude", [EXCLUDE_NONE, EXCLUDE_SELF, EXCLUDE_SELF_AND_ADJ])
def test_fullscan_topk(env, exclude):
    w, dev, orc = env
    q = sample_queries(w, 70, seed=3)
    got = dev.fullscan(q, 50, exclude)
    want = orc.fullscan(q, 50, exclude, nthreads=8)
    assert_topk_equal(got, want, f"fullscan exclude={exclude}")
    assert dev.stat(1) == len(q) * w.n


def test_fullscan_edge_cases(env):
    w, dev, orc = env
    q = np.array([w.uid[0], -3, w.uid[5], 999_999_999, w.uid[0]], np.int32)   # unknown + repeated users
    got = dev.fullscan(q, 7, EXCLUDE_SELF)
    want = orc.fullscan(q, 7, EXCLUDE_SELF, nthreads=2)
    assert_topk_equal(got, want, "fullscan with unknown users")
    assert got[2][1] == 0 and got[2][3] == 0
    # k larger than the candidate set, k = 1, empty batch, k = 0
    w2 = make_test_world(n=40, seed=3, pairs=200)
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    d2, o2 = DeviceRecommender(w2), CpuOracle(w2, "port")
    for k in (1, 64, 256):
        assert_topk_equal(d2.fullscan(w2.uid[:9], k, EXCLUDE_SELF_AND_ADJ), o2.fullscan(w2.uid[:9], k, EXCLUDE_SELF_AND_ADJ), f"tiny world k={k}")
    ids, sc, cnt = d2.fullscan(np.zeros(0, np.int32), 5)
    assert ids.shape == (0, 5)
    ids, sc, cnt = d2.fullscan(w2.uid[:3], 0)
    assert (cnt == 0).all()
    from recommendation_system_pokec_b200.engine import PokecError
    with pytest.raises(PokecError):
        d2.fullscan(w2.uid[:3], 257)
    d2.close()


@pytest.mark.parametrize("mode,limit", [(MODE_GRAPH, 5000), (MODE_GRAPH, 40), (MODE_COLLAB, 5000), (MODE_COLLAB, 25),
                                        (MODE_CLUBS, 5000)])
def test_recommend_modes(env, mode, limit):
    w, dev, orc = env
    q = sample_queries(w, 48, seed=4 + mode)
    q = np.concatenate([q, np.array([-1, 123456789], np.int32)])
    got = dev.recommend(mode, q, 50, limit)
    want = orc.recommend(mode, q, 50, limit, nthreads=8)
    assert_topk_equal(got, want, f"recommend mode={mode} limit={limit}")
    assert want[2].max() > 0


def test_candidate_lists_bit_exact(env):
    w, dev, orc = env
    for u in sample_queries(w, 25, seed=9):
        for mode in (MODE_GRAPH, MODE_COLLAB):
            for limit in (5000, 17, 1):
                assert np.array_equal(dev.candidates(mode, u, limit), orc.candidates(mode, int(u), limit))


def test_gap_world_all_modes(gap_world, oracle_libs):
    """non-dense uids, neighbours / adjacency rows without profiles, missing normalizer entries, sd == 0"""
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    w = gap_world
    dev, orc = DeviceRecommender(w), CpuOracle(w, "port")
    q = sample_queries(w, 40, seed=1)
    assert_topk_equal(dev.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ), orc.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ, nthreads=8), "gap fullscan")
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q, 20, 300), orc.recommend(mode, q, 20, 300, nthreads=8), f"gap mode {mode}")
    a, b = random_pairs(w, 3000, 5)
    assert np.array_equal(bits(dev.pair_fields(a, b)), bits(orc.pair_fields(a, b, nthreads=8)))
    dev.close()


def test_idf_variants(oracle_libs):
    """set_tfidf_index semantics: a column without an idf entry uses the raw-count cosine
    (src/recommender_similarity.cpp:99-104); a token missing from a column's table scores with idf 1.0f
    (src/recommender.cpp:78); TFIDFIndex::build's double-log idf (src/tfidf_index.cpp:25)."""
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    w = make_test_world(n=1200, seed=21, pairs=600)
    dev, orc = DeviceRecommender(w), CpuOracle(w, "port")
    a, b = random_pairs(w, 4000, 3)
    for col in (0, 2, 24):
        dev.erase_idf_column(col)
        orc.erase_idf_column(col)
    tok, _, _ = dev.get_idf(4)
    drop = tok[::3]
    dev.erase_idf_tokens(4, drop)
    orc.erase_idf_tokens(4, drop)
    assert dev.idf_size(0) == -1 and dev.idf_size(4) == orc.idf_size(4)
    assert np.array_equal(bits(dev.pair_fields(a, b)), bits(orc.pair_fields(a, b, nthreads=8)))
    assert_scores_close(dev.pair_similarity(a, b), orc.pair_similarity(a, b, nthreads=8), "FAS with partial idf")
    dev.compute_idf(double_log=True)
    orc.compute_idf(double_log=True)
    t2, i2, _ = dev.get_idf(2)
    want, has = orc.idf_lookup(2, t2)
    assert has.all() and np.array_equal(bits(i2), bits(want))
    assert np.array_equal(bits(dev.pair_fields(a, b)), bits(orc.pair_fields(a, b, nthreads=8)))
    dev.close()


def test_normalizer_update(env):
    w, dev, orc = env
    from oracle.oracle import CpuOracle
    w2 = make_test_world(n=3000, seed=7)
    rng = np.random.default_rng(0)
    mean = rng.random(w2.n_fields).astype(np.float32) * 0.4
    sd = (rng.random(w2.n_fields).astype(np.float32) * 0.3 + 0.05)
    has = (rng.random(w2.n_fields) < 0.8).astype(np.uint8)
    w2.set_normalizers(mean, sd, has)
    o2 = CpuOracle(w2, "port")
    dev.set_normalizers(mean, sd, has)
    try:
        a, b = random_pairs(w2, 5000, 8)
        assert_scores_close(dev.pair_similarity(a, b), o2.pair_similarity(a, b, nthreads=8), "FAS after set_normalizers")
    finally:
        dev.set_normalizers(w.norm_mean, w.norm_sd, w.norm_has)


def test_sharded_scan_and_key_merge(env):
    """two candidate shards scanned separately, their key lists merged (the multi-GPU data path on one GPU)"""
    import torch
    w, dev, orc = env
    q = sample_queries(w, 33, seed=12)
    k = 50
    full = dev.fullscan(q, k, EXCLUDE_SELF_AND_ADJ)
    keys = torch.zeros((2, len(q), k), dtype=torch.int64, device="cuda")
    cut = w.n // 2 + 17
    for r, (b, e) in enumerate(((0, cut), (cut, w.n))):
        dev.set_shard(b, e)
        dev.fullscan_local_keys(q, k, EXCLUDE_SELF_AND_ADJ, keys[r].data_ptr())
        dev.sync()
    dev.set_shard(0, w.n)
    merged = dev.merge_keys(keys.data_ptr(), 2, len(q), k)
    for x, y in zip(merged, full):
        assert np.array_equal(x, y)


def sharded_pipeline(dev, w, q, k, exclude, cuts):
    """the N-rank full scan on one GPU (rank r = candidate rows [cuts[r], cuts[r+1])): pre-filter per shard, gather the
    approximate shortlists, prove + re-score the owned groups per rank, merge the per-rank keys"""
    import torch
    from recommendation_system_pokec_b200.engine import STAT_SHORTLIST
    n_ranks = len(cuts) - 1
    shorts = None
    for r in range(n_ranks):
        dev.set_shard(cuts[r], cuts[r + 1])
        dev.fullscan_prepare(q, k, exclude)
        ks = dev.stat(STAT_SHORTLIST)
        assert ks > 0, "the pre-filter was expected to run"
        if shorts is None:
            shorts = torch.zeros((n_ranks, len(q), ks), dtype=torch.int64, device="cuda")
        dev.fullscan_execute_shortlists()
        dev.fullscan_shortlist_keys(shorts[r].data_ptr())
        dev.sync()
    keys = torch.zeros((n_ranks, len(q), k), dtype=torch.int64, device="cuda")
    for r in range(n_ranks):
        dev.set_shard(cuts[r], cuts[r + 1])
        dev.fullscan_prepare(q, k, exclude)
        dev.fullscan_execute_shortlists()        # rebuilds the group tables of this `rank`
        dev.fullscan_execute_finish(shorts.data_ptr(), n_ranks, r, n_ranks)
        dev.fullscan_keys(keys[r].data_ptr())
        dev.sync()
    dev.set_shard(0, w.n)
    return dev.merge_keys(keys.data_ptr(), n_ranks, len(q), k), keys


def test_sharded_scan_with_shared_rescore(env):
    """every rank proves the merged shortlist and re-scores only the query groups it owns; the merged result is
    identical to the unsharded scan (pk_fullscan_execute_shortlists / _finish)"""
    w, dev, orc = env
    q = sample_queries(w, 150, seed=14)          # 5 groups over 3 ranks: 1 / 2 / 2
    full = dev.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ)
    merged, keys = sharded_pipeline(dev, w, q, 50, EXCLUDE_SELF_AND_ADJ, [0, w.n // 3 + 5, 2 * w.n // 3 - 11, w.n])
    owned = (keys != 0).any(dim=2).sum(dim=0).cpu().numpy()
    assert (owned == 1).all(), "every query must be re-scored by exactly one rank"
    for x, y in zip(merged, full):
        assert np.array_equal(x, y)


def test_sharded_scan_thresholds_shared_over_peer_memory(small_world):
    """three `ranks` = three stores of this process on three candidate shards, each reading the other two's pruning
    thresholds while it scans (pk_fullscan_theta_peers with device pointers; across processes the same arrays travel as
    cudaIpc handles).  A rank's own shortlist may now lack candidates that cannot reach the GLOBAL top-k -- the merged,
    proven, re-scored result must still equal the unsharded scan.  A store that prepared other queries, or is one
    execution ahead, must be ignored (tag mismatch), not believed."""
    import torch
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_SHORTLIST
    w = small_world
    cuts = [0, w.n // 3 + 5, 2 * w.n // 3 - 11, w.n]
    devs = [DeviceRecommender(w) for _ in range(3)]
    ptrs = [d.theta_ptr() for d in devs]
    assert all(ptrs)
    q = sample_queries(w, 150, seed=14)
    k = 50
    ref = devs[0].fullscan(q, k, EXCLUDE_SELF_AND_ADJ)                      # unsharded, no peers yet
    for r, d in enumerate(devs):
        d.set_shard(cuts[r], cuts[r + 1])
        d.theta_peers(None, 3, r, ptrs=[0 if i == r else ptrs[i] for i in range(3)])

    def run(queries_of):
        shorts = None
        for r, d in enumerate(devs):
            d.fullscan_prepare(queries_of(r), k, EXCLUDE_SELF_AND_ADJ)
            ks = d.stat(STAT_SHORTLIST)
            if shorts is None:
                shorts = torch.zeros((3, len(q), ks), dtype=torch.int64, device="cuda")
        for r, d in enumerate(devs):                                        # lock-step: execution n of every rank
            d.fullscan_execute_shortlists()
            d.fullscan_shortlist_keys(shorts[r].data_ptr())
            d.sync()
        keys = torch.zeros((3, len(q), k), dtype=torch.int64, device="cuda")
        for r, d in enumerate(devs):
            d.fullscan_execute_finish(shorts.data_ptr(), 3, r, 3)
            d.fullscan_keys(keys[r].data_ptr())
            d.sync()
        return devs[0].merge_keys(keys.data_ptr(), 3, len(q), k)

    for _ in range(2):                                                      # two executions: the tags move together
        merged = run(lambda r: q)
        for x, y in zip(merged, ref):
            assert np.array_equal(x, y)
    # rank 1 prepares a permutation of the batch: same size, other fingerprint -> its thresholds must not be used
    q2 = q[::-1].copy()
    d1_alone = devs[1].fullscan(q2, k, EXCLUDE_SELF_AND_ADJ)               # (also puts rank 1 one execution ahead)
    devs[0].fullscan_prepare(q, k, EXCLUDE_SELF_AND_ADJ)
    devs[0].fullscan_execute()
    got0 = devs[0].fullscan_fetch()
    solo = DeviceRecommender(w)
    solo.set_shard(cuts[0], cuts[1])
    want0 = solo.fullscan(q, k, EXCLUDE_SELF_AND_ADJ)
    # rank 0 saw only stale / foreign thresholds from its peers: its shard-local result is the plain shard scan
    for x, y in zip(got0, want0):
        assert np.array_equal(x, y)
    assert d1_alone[2].min() > 0
    for d in devs + [solo]:
        d.close()


def test_sharded_scan_tie_heavy_fallback(oracle_libs):
    """unproven queries in the N-rank pipeline: every rank runs their exact scan over its own shard, the key merge
    puts the shards together -- identical to the unsharded result"""
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_FALLBACK_QUERIES
    w = make_test_world(n=4000, seed=5, T=2, pairs=800)
    w.age[:] = 20 + (w.age % 3)
    w.comp[:] = 50
    dev = DeviceRecommender(w)
    q = sample_queries(w, 40, seed=2)            # 2 groups, one per rank
    full = dev.fullscan(q, 50, EXCLUDE_SELF)
    assert dev.stat(STAT_FALLBACK_QUERIES) > 0, "the world was meant to defeat the shortlist proof"
    merged, _ = sharded_pipeline(dev, w, q, 50, EXCLUDE_SELF, [0, 1700, w.n])
    for x, y in zip(merged, full):
        assert np.array_equal(x, y)
    dev.close()


def test_edge_cache_clubs_identical(small_world, oracle_libs):
    """the per-edge FAS cache holds exactly the floats pk_pair_similarity returns, the clubs call
    served from it is bit-identical to the scan-based one and to the oracle, and every setter drops it"""
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_EDGE_CACHE, STAT_PAIRS_SCORED
    w = small_world
    dev, orc = DeviceRecommender(w), CpuOracle(w, "port")
    q = sample_queries(w, 120, seed=31)
    q = np.concatenate([q, np.array([10 ** 8], np.int32)])            # an unknown user rides along
    plain = dev.recommend(MODE_CLUBS, q, 30, 5000)
    assert dev.stat(STAT_EDGE_CACHE) == 0 and dev.stat(STAT_PAIRS_SCORED) > 0
    dev.edge_cache_build()
    assert dev.stat(STAT_EDGE_CACHE) == len(w.adj)
    # cached floats == the pair kernel's floats, in adjacency order (-1 for neighbours without a profile)
    rng = np.random.default_rng(5)
    for r in rng.integers(0, len(w.adj_uid), 25):
        u = int(w.adj_uid[r])
        nb = w.adj[w.adj_ptr[r]:w.adj_ptr[r + 1]].astype(np.int32)
        got = dev.edge_cache_row(u)
        want = dev.pair_similarity(np.full(len(nb), u, np.int32), nb)
        assert got.tobytes() == want.tobytes()
    cached = dev.recommend(MODE_CLUBS, q, 30, 5000)
    assert dev.stat(STAT_PAIRS_SCORED) == 0, "the cached call must not score pairs"
    for x, y in zip(cached, plain):
        assert np.array_equal(x, y)
    assert_topk_equal(cached, orc.recommend(MODE_CLUBS, q, 30, 5000, 4), "clubs from the edge cache")
    # a row replaced behind the store bypasses the cache (and still matches the oracle) ...
    u = int(w.adj_uid[3])
    new_row = w.uid[:7].astype(np.int32)
    dev.set_adj_row(u, new_row)
    got = dev.recommend(MODE_CLUBS, np.array([u], np.int32), 30, 5000)
    assert dev.stat(STAT_PAIRS_SCORED) > 0
    dev.set_adj_row(u, w.adj[w.adj_ptr[3]:w.adj_ptr[4]].astype(np.int32))     # restored: served from the cache again
    dev.recommend(MODE_CLUBS, np.array([u], np.int32), 30, 5000)
    assert dev.stat(STAT_PAIRS_SCORED) == 0
    # ... and the setters drop it
    dev.set_normalizers(w.norm_mean, w.norm_sd)
    assert dev.stat(STAT_EDGE_CACHE) == 0
    dev.edge_cache_build()
    dev.compute_idf()
    assert dev.stat(STAT_EDGE_CACHE) == 0
    dev.close()
    orc.close()


def test_column_normalizers_bit_exact(small_world, oracle_libs):
    """compute_column_normalizers (src/utils.cpp:155-240) -- per-pair raw similarities from the
    device (raw-count cosines whatever idf is set), sums in the reference's order: every mean / sd float identical;
    the scoring path afterwards still sees the tf-idf weights"""
    from oracle.oracle import CpuOracle, reference_available
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    if not reference_available():
        pytest.skip("oracle/_ref did not travel")
    w = small_world
    ref = CpuOracle(w, "reference")
    dev = DeviceRecommender(w)
    q = sample_queries(w, 40, seed=3)
    before = dev.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ)
    for sample, comps in ((400, 5), (1, 1), (50, 3)):
        mean, sd = ref.column_normalizers(sample, comps)
        a, b = ref.normalizer_pairs(sample, comps)
        gm, gs = dev.normalizers_from_pairs(a, b)
        assert gm.tobytes() == mean.tobytes(), f"means differ ({sample} x {comps}): {np.flatnonzero(gm != mean)}"
        assert gs.tobytes() == sd.tobytes(), f"sds differ ({sample} x {comps}): {np.flatnonzero(gs != sd)}"
    after = dev.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ)
    for x, y in zip(before, after):
        assert np.array_equal(x, y), "the idf state must be restored after the raw-count pass"
    gm, gs = dev.normalizers_from_pairs(np.zeros(0, np.int32), np.zeros(0, np.int32))
    assert (gm == 0).all() and (gs == 1).all()
    with pytest.raises(Exception):
        dev.normalizers_from_pairs(np.array([int(w.uid[0])], np.int32), np.array([-5], np.int32))
    dev.close()
    ref.close()


def test_vs_compiled_reference(small_world, oracle_libs):
    """the same comparison against the UNMODIFIED reference built from /root/reference (when the prebuilt
    library travelled with the snapshot)"""
    from oracle.oracle import CpuOracle, reference_available
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    if not reference_available():
        pytest.skip("oracle/_ref/libpokec_ref.so not present")
    w = small_world
    dev, ref = DeviceRecommender(w), CpuOracle(w, "reference")
    a, b = random_pairs(w, 20000, 13)
    assert_scores_close(dev.pair_similarity(a, b), ref.pair_similarity(a, b, nthreads=8), "pair FAS vs reference")
    q = sample_queries(w, 32, seed=14)
    assert_topk_equal(dev.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ), ref.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ, nthreads=8), "fullscan vs reference")
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q, 20, 5000), ref.recommend(mode, q, 20, 5000, nthreads=8), f"mode {mode} vs reference")
    # adjacency mutated behind the recommender (src/test.cpp:59-73)
    u = int(q[0])
    ai = int(np.searchsorted(w.adj_uid, u))
    row = w.adj[w.adj_ptr[ai]:w.adj_ptr[ai + 1]] if ai < w.n_adj_rows and w.adj_uid[ai] == u else np.zeros(0, np.int32)
    if len(row) > 2:
        dev.set_adj_row(u, row[: len(row) // 2])
        ref.set_adj_row(u, row[: len(row) // 2])
        assert_topk_equal(dev.recommend(MODE_COLLAB, q[:4], 20, 1000), ref.recommend(MODE_COLLAB, q[:4], 20, 1000), "collab after set_adj_row")
    dev.close()


# ---- two-stage full scan (pk_prefilter.cuh): fp32 pre-filter -> exact re-score -> exact fallback -------------

def _no_norm_world():
    w = make_test_world(n=2000, seed=41)
    w.set_normalizers(w.norm_mean, w.norm_sd, np.zeros(w.n_fields, np.uint8))   # every field uses z = 6(s - 0.5)
    return w


@pytest.mark.parametrize("which", ["normalized", "default_z", "gaps"])
def test_prefilter_error_within_its_bound(which, small_world, gap_world, oracle_libs):
    """|fp32 pre-filter score - exact score| must stay below the bound the shortlist proof relies on"""
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_MARGIN_E9
    w = {"normalized": small_world, "default_z": _no_norm_world(), "gaps": gap_world}[which]
    dev = DeviceRecommender(w)
    worst = 0.0
    for seed in (1, 2, 3):
        q = sample_queries(w, 32, seed=seed)
        approx, exact = dev.debug_prefilter(q, 0, w.n)
        worst = max(worst, float(np.abs(approx.astype(np.float64) - exact).max()))
    margin = dev.stat(STAT_MARGIN_E9) / 1e9
    assert 0 < margin < 1e-3
    assert worst <= margin / 2, f"pre-filter error {worst:.3e} vs bound {margin:.3e}"
    dev.close()


def test_fullscan_with_and_without_prefilter_identical(env, monkeypatch):
    from recommendation_system_pokec_b200.engine import STAT_PREFILTER, STAT_SHORTLIST
    w, dev, orc = env
    q = sample_queries(w, 150, seed=21)
    for k in (50, 5, 100):
        monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
        two_stage = dev.fullscan(q, k, EXCLUDE_SELF_AND_ADJ)
        assert dev.stat(STAT_PREFILTER) == 1 and dev.stat(STAT_SHORTLIST) >= k + 16
        monkeypatch.setenv("POKEC_B200_NO_PREFILTER", "1")
        exact = dev.fullscan(q, k, EXCLUDE_SELF_AND_ADJ)
        assert dev.stat(STAT_PREFILTER) == 0
        for x, y in zip(two_stage, exact):
            assert np.array_equal(x, y), f"k={k}"
    monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
    # k too large for a shortlist with slack: the exact scan is used directly
    dev.fullscan(q[:4], 250, EXCLUDE_SELF)
    assert dev.stat(STAT_PREFILTER) == 0


def test_fullscan_more_groups_than_sms(monkeypatch):
    """dynamic schedule of the pre-filter: 6000 queries = 188 groups > 148 persistent CTAs and a shard of fewer rows
    than the `worth joining` threshold -- every group must still be scanned completely (groups picked in later rounds,
    several instances per group); result identical to the exact scan on every pair, also on a sub-shard"""
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_PREFILTER, STAT_FALLBACK_QUERIES
    w = make_test_world(n=6000, seed=3, T=12, pairs=800)
    dev = DeviceRecommender(w)
    q = w.uid.astype(np.int32)
    for lo, hi in ((0, w.n), (1000, 2500)):
        dev.set_shard(lo, hi)
        monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
        two_stage = dev.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ)
        assert dev.stat(STAT_PREFILTER) == 1
        monkeypatch.setenv("POKEC_B200_NO_PREFILTER", "1")
        exact = dev.fullscan(q, 20, EXCLUDE_SELF_AND_ADJ)
        assert dev.stat(STAT_PREFILTER) == 0
        for x, y in zip(two_stage, exact):
            assert np.array_equal(x, y), f"shard [{lo},{hi})"
        assert int(two_stage[2].min()) == 20
    monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
    dev.close()


@pytest.mark.parametrize("case", range(10))
def test_fullscan_fuzz_two_stage_equals_exact(case, monkeypatch):
    """random small worlds / batch shapes / k / exclusion modes / shards: the two-stage scan (cuckoo tables, dynamic
    schedule, shortlist proof, partial re-score) must return exactly what the exact kernel returns on every pair"""
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    rng = np.random.default_rng(1000 + case)
    n = int(rng.choice([40, 150, 700, 2500]))
    T = int(rng.choice([1, 3, 10, 49]))
    w = make_test_world(n=n, seed=200 + case, T=T, pairs=300, uid_gap=bool(rng.integers(0, 2)),
                        mean_deg=float(rng.choice([2.0, 12.0])))
    dev = DeviceRecommender(w)
    nq = int(rng.choice([1, 31, 33, 200]))
    q = w.uid[rng.integers(0, w.n, nq)].astype(np.int32)         # repeated users allowed
    q[rng.integers(0, nq)] = -7 if case % 3 == 0 else q[0]        # sometimes an unknown user
    k = int(rng.choice([1, 7, 50, 100]))
    mode = int(rng.choice([EXCLUDE_NONE, EXCLUDE_SELF, EXCLUDE_SELF_AND_ADJ]))
    lo = int(rng.integers(0, w.n // 2))
    hi = int(rng.integers(lo + 1, w.n + 1))
    for shard in ((0, w.n), (lo, hi)):
        dev.set_shard(*shard)
        monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
        two_stage = dev.fullscan(q, k, mode)
        monkeypatch.setenv("POKEC_B200_NO_PREFILTER", "1")
        exact = dev.fullscan(q, k, mode)
        for x, y in zip(two_stage, exact):
            assert np.array_equal(x, y), f"case {case}: n={n} T={T} nq={nq} k={k} mode={mode} shard={shard}"
    monkeypatch.delenv("POKEC_B200_NO_PREFILTER", raising=False)
    dev.close()


def test_fullscan_tie_heavy_world_falls_back_to_exact(oracle_libs):
    """few fields and tiny vocabularies: thousands of candidates tie at the k-th score, the shortlist cannot be
    proven complete, and the flagged queries must come out of the exact full scan -- identical to the oracle"""
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.engine import DeviceRecommender, STAT_FALLBACK_QUERIES, STAT_PREFILTER
    w = make_test_world(n=4000, seed=5, T=2, pairs=800)
    # strip the distinguishing fields: no clubs / friends similarity, coarse ages
    w.age[:] = 20 + (w.age % 3)
    w.comp[:] = 50
    dev, orc = DeviceRecommender(w), CpuOracle(w, "port")
    q = sample_queries(w, 40, seed=2)
    got = dev.fullscan(q, 50, EXCLUDE_SELF)
    assert dev.stat(STAT_PREFILTER) == 1
    assert dev.stat(STAT_FALLBACK_QUERIES) > 0, "the world was meant to defeat the shortlist proof"
    assert_topk_equal(got, orc.fullscan(q, 50, EXCLUDE_SELF, nthreads=8), "tie-heavy fullscan")
    dev.close()
    orc.close()


def test_cuda_path_vs_golden_fixtures(oracle_libs):
    """the CUDA path against the committed outputs of the unmodified reference (tests/golden/make_golden.py)"""
    import os
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    from recommendation_system_pokec_b200.world import World
    gdir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    w = World.load(os.path.join(gdir, "world_small.npz"))
    g = np.load(os.path.join(gdir, "golden_small.npz"))
    dev = DeviceRecommender(w)
    a, b, q = g["pair_a"], g["pair_b"], g["queries"]
    assert np.array_equal(bits(dev.pair_fields(a, b)), bits(g["pair_fields"]))
    assert_scores_close(dev.pair_similarity(a, b), g["pair_fas"], "pair FAS vs golden")
    for col in (0, 2, 22, 32, 48):
        tok, idf, _ = dev.get_idf(col)
        assert np.array_equal(tok, g[f"idf_tok_{col}"]) and np.array_equal(bits(idf), bits(g[f"idf_val_{col}"]))
    for name, mode, limit in (("graph", MODE_GRAPH, 5000), ("graph_l25", MODE_GRAPH, 25), ("collab", MODE_COLLAB, 5000),
                              ("collab_l40", MODE_COLLAB, 40), ("clubs", MODE_CLUBS, 5000)):
        assert_topk_equal(dev.recommend(mode, q, 20, limit), (g[f"{name}_id"], g[f"{name}_score"], g[f"{name}_cnt"]), f"golden {name}")
    for ex in (0, 1, 2):
        assert_topk_equal(dev.fullscan(q, 30, ex), (g[f"fullscan{ex}_id"], g[f"fullscan{ex}_score"], g[f"fullscan{ex}_cnt"]), f"golden fullscan {ex}")
    dev.compute_idf(double_log=True)
    for col in (0, 22):
        tok, idf, _ = dev.get_idf(col)
        assert np.array_equal(bits(idf), bits(g[f"idf_dbl_{col}"]))
    dev.close()


def test_store_from_world_file_identical(tmp_path):
    """a store built straight from the mapped world file (pk_world_open -> pk_store_load_world) answers
    every call exactly like the store built from the in-memory world"""
    from recommendation_system_pokec_b200 import engine
    w = make_test_world(n=3000, seed=61, pairs=500, uid_gap=True)
    path = str(tmp_path / "world.pkw")
    engine.world_save(w, path)
    a = engine.DeviceRecommender(w)
    b = engine.DeviceRecommender.from_world_file(path, text_columns=w.text_columns)
    q = sample_queries(w, 48, seed=4)
    for x, y in zip(a.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ), b.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ)):
        assert np.array_equal(x, y)
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        for x, y in zip(a.recommend(mode, q, 20, 5000), b.recommend(mode, q, 20, 5000)):
            assert np.array_equal(x, y), f"mode {mode}"
    for t in range(0, w.T, 7):
        ia, ib = a.get_idf(t), b.get_idf(t)
        assert (ia is None) == (ib is None)
        if ia is not None:
            for x, y in zip(ia, ib):
                assert np.array_equal(x, y)
    a.close()
    b.close()
