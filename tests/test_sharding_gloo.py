"""The N > 1 data path of the full scan on CPU: two processes over gloo, candidate rows split by row range, one
all-gather of packed top-k keys, merge with the reference comparator (recommendation_system_pokec_b200/sharding.py).
The per-shard scan itself is played by the CPU oracle here (the CUDA scan of a shard is covered on the GPU by
tests/test_gpu_parity.py::test_sharded_scan_and_key_merge); what this pins is the host logic around it: shard
bounds, key packing, the collective, the merge, and that the merged result equals the unsharded one."""
import os
import socket

import numpy as np
import pytest

from recommendation_system_pokec_b200.sharding import shard_rows, owned_groups, pack_keys, unpack_keys, merge_key_lists


def test_shard_rows_cover_and_balance():
    for n in (0, 1, 7, 1000, 1_632_803):
        for ws in (1, 2, 3, 8):
            b = [shard_rows(n, r, ws) for r in range(ws)]
            assert b[0][0] == 0 and b[-1][1] == n
            assert all(b[i][1] == b[i + 1][0] for i in range(ws - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_key_order_is_the_reference_comparator():
    rng = np.random.default_rng(0)
    ids = rng.integers(-5, 2_000_000_000, 5000).astype(np.int32)
    sc = rng.choice(np.array([0.0, 0.25, 0.5, 0.5000001, 0.9], np.float32), 5000)
    keys = pack_keys(ids, sc)
    order = np.argsort(keys, kind="stable")[::-1]
    want = sorted(range(5000), key=lambda i: (-float(sc[i]), int(ids[i])))
    got_pairs = [(float(sc[i]), int(ids[i])) for i in order]
    want_pairs = [(float(sc[i]), int(ids[i])) for i in want]
    assert got_pairs == want_pairs
    i2, s2, ok = unpack_keys(keys)
    assert ok.all() and np.array_equal(i2, ids) and np.array_equal(s2, sc)


def _worker(rank, world_size, port, k, out_dir):
    import torch
    import torch.distributed as dist
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.sharding import all_gather_keys
    from recommendation_system_pokec_b200.synth import sample_queries
    from tests.conftest import make_test_world
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        w = make_test_world(n=1200, seed=17, pairs=400)
        orc = CpuOracle(w, "port")
        q = sample_queries(w, 12, seed=3)
        lo, hi = shard_rows(w.n, rank, world_size)
        look = w.subset_adj_lookup()
        local = np.zeros((len(q), k), np.uint64)
        for qi, u in enumerate(q):
            sc = orc.scores_of_query(int(u))                      # exact FAS of u against every row
            excl = {int(u)}
            ai = look.get(int(u))
            if ai is not None:
                excl.update(int(x) for x in w.adj[w.adj_ptr[ai]:w.adj_ptr[ai + 1]])
            rows = np.array([r for r in range(lo, hi) if int(w.uid[r]) not in excl], np.int64)
            keys = np.sort(pack_keys(w.uid[rows], sc[rows]))[::-1][:k]
            local[qi, :len(keys)] = keys
        gathered = all_gather_keys(torch.from_numpy(local.view(np.int64)), world_size).numpy().view(np.uint64)
        merged = merge_key_lists(gathered, k)
        want = orc.fullscan(q, k, 2)
        ids, scores, valid = unpack_keys(merged)
        assert np.array_equal(ids, want[0]) and np.array_equal(scores, want[1])
        assert np.array_equal(valid.sum(1), want[2])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_sharded_fullscan_over_gloo(tmp_path, oracle_libs):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, 20, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()


def test_owned_groups_partition():
    for ng in (0, 1, 5, 128):
        for ws in (1, 2, 3, 8):
            b = [owned_groups(ng, r, ws) for r in range(ws)]
            assert b[0][0] == 0 and b[-1][1] == ng and all(b[i][1] == b[i + 1][0] for i in range(ws - 1))


def _worker_shared_rescore(rank, world_size, port, k, ks, out_dir):
    """the split pipeline of the two-stage scan (bench.py N > 1): shortlists of length ks per shard -> all-gather ->
    merged global shortlist -> the owner rank of each query group turns it into the top-k -> all-gather -> merge.
    The shard scan and the exact re-score are played by the CPU oracle's exact scores."""
    import torch
    import torch.distributed as dist
    from oracle.oracle import CpuOracle
    from recommendation_system_pokec_b200.sharding import all_gather_keys
    from recommendation_system_pokec_b200.synth import sample_queries
    from tests.conftest import make_test_world
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world_size)
    try:
        w = make_test_world(n=900, seed=19, pairs=300)
        orc = CpuOracle(w, "port")
        q = sample_queries(w, 70, seed=4)                          # 3 groups of 32 / 32 / 6 queries
        lo, hi = shard_rows(w.n, rank, world_size)
        short = np.zeros((len(q), ks), np.uint64)                  # stage 1: the shard's shortlist, keyed by ROW
        scores = []
        for qi, u in enumerate(q):
            sc = orc.scores_of_query(int(u))
            scores.append(sc)
            rows = np.array([r for r in range(lo, hi) if int(w.uid[r]) != int(u)], np.int64)
            keys = np.sort(pack_keys(rows.astype(np.int32), sc[rows]))[::-1][:ks]
            short[qi, :len(keys)] = keys
        g_short = all_gather_keys(torch.from_numpy(short.view(np.int64)), world_size).numpy().view(np.uint64)
        merged_short = merge_key_lists(g_short, ks)                # identical on every rank
        g_lo, g_hi = owned_groups((len(q) + 31) // 32, rank, world_size)
        local = np.zeros((len(q), k), np.uint64)                   # stage 2: exact top-k of the owned groups only
        for qi in range(len(q)):
            if not (g_lo <= qi // 32 < g_hi):
                continue
            rows, _, valid = unpack_keys(merged_short[qi])
            rows = rows[valid]
            keys = np.sort(pack_keys(w.uid[rows], scores[qi][rows]))[::-1][:k]
            local[qi, :len(keys)] = keys
        gathered = all_gather_keys(torch.from_numpy(local.view(np.int64)), world_size).numpy().view(np.uint64)
        assert ((gathered != 0).any(axis=2).sum(axis=0) == 1).all(), "every query is owned by exactly one rank"
        ids, sc2, valid = unpack_keys(merge_key_lists(gathered, k))
        want = orc.fullscan(q, k, 1)
        assert np.array_equal(ids, want[0]) and np.array_equal(sc2, want[1]) and np.array_equal(valid.sum(1), want[2])
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_shared_rescore_over_gloo(tmp_path, oracle_libs):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker_shared_rescore, args=(2, port, 20, 64, str(tmp_path)), nprocs=2, join=True)
    assert (tmp_path / "ok0").exists() and (tmp_path / "ok1").exists()

