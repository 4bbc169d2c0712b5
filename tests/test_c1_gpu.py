"""BASELINE.json configs[0] ("C1"): a self-contained 100 000-user synthetic world -- the size the
reference's loader caps at (src/user_loader.cpp:34) -- 1 000 query users, recommend_by_interest(uid, 50, 5000)
(src/recommender_graph.cpp:224-227), every list compared with the compiled unmodified reference; plus the other two
list modes and the north-star full scan on samples of the same queries.  Through the C ABI, on the GPU."""
import os
import time

import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS, EXCLUDE_SELF_AND_ADJ
from recommendation_system_pokec_b200.synth import make_world, sample_normalizers, sample_queries
from tests.conftest import best_oracle
from tests.test_gpu_parity import assert_topk_equal

pytestmark = pytest.mark.gpu
NT = min(32, os.cpu_count() or 1)


@pytest.fixture(scope="module")
def c1(oracle_libs):
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    w = make_world(n=100_000)
    m, s = sample_normalizers(w, pairs=3000)
    w.set_normalizers(m, s)
    dev, orc = DeviceRecommender(w), best_oracle(w)
    q = sample_queries(w, 1000, seed=1)
    yield w, dev, orc, q
    dev.close()
    orc.close()


def test_c1_interest_all_1000_queries(c1):
    w, dev, orc, q = c1
    t = time.time()
    got = dev.recommend(MODE_GRAPH, q, 50, 5000)
    t_gpu = time.time() - t
    t = time.time()
    want = orc.recommend(MODE_GRAPH, q, 50, 5000, NT)
    t_cpu = time.time() - t
    assert want[2].max() == 50 and (want[2] > 0).mean() > 0.8
    assert_topk_equal(got, want, "C1 recommend_by_interest(uid, 50, 5000)")
    print(f"\nC1 interest: {len(q)} queries, GPU call {t_gpu * 1e3:.1f} ms, {orc.kind} on {NT} threads {t_cpu:.2f} s")


@pytest.mark.parametrize("mode,nq", [(MODE_COLLAB, 1000), (MODE_CLUBS, 1000)])
def test_c1_collaborative_and_clubs(c1, mode, nq):
    w, dev, orc, q = c1
    got = dev.recommend(mode, q[:nq], 50, 5000)
    want = orc.recommend(mode, q[:nq], 50, 5000, NT)
    assert_topk_equal(got, want, f"C1 mode {mode} (50, 5000)")


def test_c1_api_cli_shape_topk20(c1):
    """the server's own call shape: topk 20, candidate_limit 5000 (src/api_cli.cpp:213-234)"""
    w, dev, orc, q = c1
    for mode in (MODE_GRAPH, MODE_COLLAB, MODE_CLUBS):
        assert_topk_equal(dev.recommend(mode, q[:200], 20, 5000), orc.recommend(mode, q[:200], 20, 5000, NT), f"C1 mode {mode} (20, 5000)")


def test_c1_fullscan_sample(c1):
    w, dev, orc, q = c1
    s = q[::16][:64]
    got = dev.fullscan(s, 50, EXCLUDE_SELF_AND_ADJ)
    want = orc.fullscan(s, 50, EXCLUDE_SELF_AND_ADJ, NT)
    assert_topk_equal(got, want, "C1 full scan, 64 queries x 100 000 candidates")
