"""Multi-GPU plumbing of the full scan: one process per GPU, candidate rows partitioned by row
(= uid) range, every rank scans its shard for the whole query batch, the per-rank top-k lists travel as packed
64-bit sort keys through an all-gather, and every rank merges them with the reference comparator
(score desc, uid asc; reference src/recommender_graph.cpp:97-100).  With the two-stage scan a second all-gather
sits in the middle: the shards' approximate shortlists are gathered and merged, and every rank re-scores exactly
only the query groups it owns (`owned_groups`, the host mirror of pk_fullscan_execute_finish's split).

The key layout is csrc/pk_layout.h:make_key; the numpy forms below are the host mirror used for decoding and by
the CPU (gloo) tests of this plumbing.
"""
from __future__ import annotations

import numpy as np


def shard_rows(n: int, rank: int, world_size: int):
    """rows [lo, hi) of rank `rank`: contiguous, balanced to within one row, covering [0, n)"""
    return n * rank // world_size, n * (rank + 1) // world_size


def owned_groups(n_groups: int, rank: int, world_size: int):
    """query groups [lo, hi) whose merged shortlist rank `rank` re-scores (csrc/pk_host.cu:fullscan_finish); groups
    are consecutive blocks of 32 queries that have a profile, in caller order"""
    return n_groups * rank // world_size, n_groups * (rank + 1) // world_size


def pack_keys(ids, scores):
    """(float32 score >= 0, int32 id) -> uint64 key; larger key = better under the reference comparator"""
    ids = np.asarray(ids, np.int32)
    sb = np.ascontiguousarray(scores, np.float32).view(np.uint32).astype(np.uint64)
    low = (~(ids.view(np.uint32) ^ np.uint32(0x80000000))).astype(np.uint64)
    return (sb << np.uint64(33)) | (low << np.uint64(1)) | np.uint64(1)


def unpack_keys(keys):
    """uint64 keys (0 = empty slot) -> (ids, scores, valid)"""
    keys = np.asarray(keys, np.uint64)
    valid = keys != 0
    sb = (keys >> np.uint64(33)).astype(np.uint32)
    low = ((keys >> np.uint64(1)) & np.uint64(0xFFFFFFFF)).astype(np.uint32)
    ids = ((~low) ^ np.uint32(0x80000000)).view(np.int32)
    return np.where(valid, ids, -1).astype(np.int32), np.where(valid, sb.view(np.float32), 0).astype(np.float32), valid


def merge_key_lists(lists, k):
    """[n_lists, nq, k] keys -> [nq, k] best keys (descending, 0-padded): the host statement of pk_merge_keys"""
    lists = np.asarray(lists, np.uint64)
    n_lists, nq, kk = lists.shape
    flat = np.transpose(lists, (1, 0, 2)).reshape(nq, n_lists * kk)
    return np.sort(flat, axis=1)[:, ::-1][:, :k].copy()


def all_gather_keys(local_keys, world_size):
    """torch.distributed all-gather of a [nq, k] int64/uint64 key tensor -> [world_size, nq, k]"""
    import torch
    import torch.distributed as dist
    out = torch.zeros((world_size,) + tuple(local_keys.shape), dtype=local_keys.dtype, device=local_keys.device)
    dist.all_gather_into_tensor(out.view(-1), local_keys.reshape(-1).contiguous())
    return out



def share_thetas(dev, rank, world_size, device):
    """Gives every rank's store the cudaIpc handles of the other ranks' pruning-threshold arrays
    (pk_fullscan_theta_export -> all-gather of 64 bytes per rank -> pk_fullscan_theta_peers): the pre-filter then reads
    its peers' thresholds over NVLink while it scans.  Returns True when the peers were mapped; a box without peer
    access keeps scanning without sharing (the results do not depend on it)."""
    import torch
    import torch.distributed as dist
    if world_size <= 1:
        return False
    mine = torch.tensor(list(dev.theta_export()), dtype=torch.uint8, device=device)
    allh = torch.zeros(world_size * 64, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allh, mine)
    ok = True
    try:
        dev.theta_peers(bytes(allh.cpu().numpy().tobytes()), world_size, rank)
    except Exception as e:                 # PokecError: no peer access
        import sys
        print(f"[sharding] rank {rank}: thresholds not shared ({e})", file=sys.stderr, flush=True)
        ok = False
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if int(flag.item()) == 0 and ok:       # all or nothing: a half-connected group would still be correct, but keep it simple
        dev.theta_peers(None, 1, 0, ptrs=[0])
        ok = False
    return ok
