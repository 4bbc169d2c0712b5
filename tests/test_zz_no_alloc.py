"""pk_candidates_batch_noalloc: the device expansion on DEVICE buffers with its scratch in a static device array (no
allocation inside the call; include/pokec_b200.h).  Checked against the existing entry point pk_candidates_batch --
integer lists, so bit for bit -- on two calls in a row (the second continues the epochs the first left in the shared
scratch), plus the stated limits.  Collected last on purpose (file name)."""
import numpy as np
import pytest

from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB
from recommendation_system_pokec_b200.synth import sample_queries
from tests.conftest import make_test_world

pytestmark = pytest.mark.gpu

NA_MAX_LIMIT = 992
NA_MAX_Q = 8192


def _noalloc(dev, torch, mode, q, limit):
    """one call through the C ABI; the device buffers are the CALLER's (torch tensors here), filled with a sentinel"""
    d_q = torch.from_numpy(np.ascontiguousarray(q, np.int32)).cuda()
    d_ids = torch.full((len(q), limit), -7, dtype=torch.int32, device="cuda")
    d_cnt = torch.full((len(q),), -7, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()                                   # inputs complete before the store's stream reads them
    dev.candidates_batch_noalloc(mode, d_q.data_ptr(), len(q), limit, d_ids.data_ptr(), d_cnt.data_ptr())
    dev.sync()
    return d_ids.cpu().numpy(), d_cnt.cpu().numpy()


def _assert_same_lists(got, want, what):
    ids_g, cnt_g = got
    ids_w, cnt_w = want
    assert np.array_equal(cnt_g, cnt_w), f"{what}: list lengths differ"
    for i in range(len(cnt_w)):
        assert np.array_equal(ids_g[i, :cnt_w[i]], ids_w[i, :cnt_w[i]]), f"{what}: list {i} differs"


def test_no_alloc_expansion_equals_the_allocating_entry_point():
    import torch
    from recommendation_system_pokec_b200.engine import DeviceRecommender, PokecError
    w = make_test_world(n=3000, seed=91, mean_deg=24.0)
    dev = DeviceRecommender(w)
    q1 = np.concatenate([sample_queries(w, 200, seed=5), np.array([-3, 88_000_000], np.int32)])   # + users without a row
    q2 = sample_queries(w, 333, seed=6)

    # before the adjacency is resident the call must refuse (it may not allocate) ...
    with pytest.raises(PokecError, match="not resident"):
        _noalloc(dev, torch, MODE_GRAPH, q1, 100)
    # ... the existing entry point makes it resident and gives the lists to compare with
    want1 = dev.candidates_batch(MODE_GRAPH, q1, 100)
    want2 = dev.candidates_batch(MODE_COLLAB, q2, NA_MAX_LIMIT)
    want3 = dev.candidates_batch(MODE_GRAPH, q2, 17)
    assert want1[1].max() == 100 and want2[1].max() > 100, "the world must exercise the truncation and long lists"

    # two calls in a row on the shared static scratch (+ a third with another shape), then the first again
    _assert_same_lists(_noalloc(dev, torch, MODE_GRAPH, q1, 100), want1, "call 1 (graph, limit 100)")
    _assert_same_lists(_noalloc(dev, torch, MODE_COLLAB, q2, NA_MAX_LIMIT), want2, "call 2 (collaborative, limit 992)")
    _assert_same_lists(_noalloc(dev, torch, MODE_GRAPH, q2, 17), want3, "call 3 (graph, limit 17)")
    _assert_same_lists(_noalloc(dev, torch, MODE_GRAPH, q1, 100), want1, "call 4 (= call 1)")

    # a second store on the same device shares the array (serialised calls): still the same lists
    dev2 = DeviceRecommender(w)
    dev2.candidates_batch(MODE_GRAPH, q1[:16], 10)
    _assert_same_lists(_noalloc(dev2, torch, MODE_GRAPH, q1, 100), want1, "second store")
    dev2.close()

    # sizes beyond the static scratch are rejected with a message, nothing is launched
    with pytest.raises(PokecError, match="limits in"):
        _noalloc(dev, torch, MODE_GRAPH, q1, NA_MAX_LIMIT + 1)
    many = np.resize(q2, NA_MAX_Q + 1).astype(np.int32)
    with pytest.raises(PokecError, match="at most"):
        _noalloc(dev, torch, MODE_GRAPH, many, 4)
    # the maximum batch itself is taken
    full = np.resize(q2, NA_MAX_Q).astype(np.int32)
    _assert_same_lists(_noalloc(dev, torch, MODE_GRAPH, full, 8), dev.candidates_batch(MODE_GRAPH, full, 8), "8192 users")
    dev.close()
