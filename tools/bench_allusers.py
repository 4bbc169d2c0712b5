#!/usr/bin/env python
"""bench_allusers.py -- BASELINE configs[4]: all-users batch scoring (every user is a query, FAS top-50 against all
1.63 M candidates), candidates sharded by row range over the GPUs, NCCL all-gather of the per-rank keys + merge.

    python tools/bench_allusers.py [--batches B] [--batch 4096]                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
        --master-port 29533 tools/bench_allusers.py --batches B                                   # N GPUs

The full sweep is ceil(1 632 803 / 4096) = 399 batches; the sustained rate is measured on a window of B consecutive
batches (distinct query users, walked in uid order) through the public end-to-end call -- host uids in, host top-k
out, every batch -- and extrapolated to the sweep.  Prints one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.synth import POKEC_N  # noqa: E402
from recommendation_system_pokec_b200.cabi import EXCLUDE_SELF_AND_ADJ  # noqa: E402
from recommendation_system_pokec_b200.sharding import shard_rows, share_thetas  # noqa: E402


def sweep(w, ws, rank, lr, B, k, n_batches, depth=2, keep_batches=(), batch_fn=None, stats=None, reserve_sms=None):
    """`n_batches` consecutive batches of B distinct query users (walked in uid order) through the end-to-end calls --
    host uids in, host top-k out, host planning + H2D, both all-gathers and the merge every batch.  `depth` stores on the
    same GPU take the batches in turn, each on its own stream: the host planning and H2D of batch i + 1 overlap the
    kernels of batch i, and its all-gather overlaps the next scan.  Returns (query-users/s, seconds,
    [(uids, lists) of the batches named in keep_batches]); timed with a barrier + synchronize on both sides, max over
    ranks.  batch_fn(b) overrides the uid walk (bench.py's end-to-end leg repeats its query batch); `stats` (a dict)
    receives the host<->device bytes per batch counted by the stores."""
    import torch
    import torch.distributed as dist
    from recommendation_system_pokec_b200 import engine
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    lo, hi = shard_rows(w.n, rank, ws)
    if reserve_sms is None:
        # SMs the persistent pre-filter leaves to the tail of the previous batch (merges, re-score, all-gathers).  Measured
        # at 8 GPUs (profiles/r02v_reserve_n8.jsonl: 0 -> 286 / 272 k, 4 -> 280 k, 8 -> 278 k, 16 -> 274 k query-users/s): the
        # tail needs throughput, not a few SMs, so nothing is gained -- the default stays 0.
        reserve_sms = int(os.environ.get("POKEC_B200_RESERVE_SMS", "0"))

    def batch_uids(b):
        if batch_fn is not None:
            return np.ascontiguousarray(batch_fn(b), np.int32)
        q = w.uid[(b * B) % w.n:(b * B) % w.n + B].astype(np.int32)
        if len(q) < B:
            q = np.concatenate([q, w.uid[:B - len(q)].astype(np.int32)])
        return q

    class Slot:
        def __init__(self):
            self.dev = DeviceRecommender(w, device=lr)
            self.dev.set_shard(lo, hi)
            if reserve_sms > 0:
                self.dev.set_option(engine.OPT_PREFILTER_RESERVE_SMS, reserve_sms)
            share_thetas(self.dev, rank, ws, f"cuda:{lr}")      # slot i of every rank pairs with slot i of the others
            self.stream = torch.cuda.ExternalStream(self.dev.stream(), device=lr)
            self.local = torch.zeros((B, k), dtype=torch.int64, device=f"cuda:{lr}")
            self.gathered = torch.zeros((ws, B, k), dtype=torch.int64, device=f"cuda:{lr}")
            self.short = None
            self.pending = None

        def launch(self, b):
            """everything of batch b up to (not including) the fetch: asynchronous on this slot's stream"""
            dev = self.dev
            q = batch_uids(b)
            dev.fullscan_prepare(q, k, EXCLUDE_SELF_AND_ADJ)
            ks = dev.stat(engine.STAT_SHORTLIST)
            if ws > 1 and ks > 0:      # shared exact re-score
                if self.short is None:
                    self.short = (torch.zeros((B, ks), dtype=torch.int64, device=f"cuda:{lr}"),
                                  torch.zeros((ws, B, ks), dtype=torch.int64, device=f"cuda:{lr}"))
                ls, gs = self.short
                dev.fullscan_execute_shortlists()
                dev.fullscan_shortlist_keys(ls.data_ptr())
                with torch.cuda.stream(self.stream):
                    dist.all_gather_into_tensor(gs.view(-1), ls.view(-1))
                dev.fullscan_execute_finish(gs.data_ptr(), ws, rank, ws)
            else:
                dev.fullscan_execute()
            if ws > 1:
                dev.fullscan_keys(self.local.data_ptr())
                with torch.cuda.stream(self.stream):
                    dist.all_gather_into_tensor(self.gathered.view(-1), self.local.view(-1))
            self.pending = b

        def fetch(self):
            """host top-k of the batch launched last on this slot (synchronises this slot's stream only)"""
            b, self.pending = self.pending, None
            if ws == 1:
                return b, self.dev.fullscan_fetch()
            return b, self.dev.merge_keys(self.gathered.data_ptr(), ws, B, k)

    slots = [Slot() for _ in range(max(1, depth))]
    dev = slots[0].dev
    kept = []
    state = {"complete": True}

    def take(b, res):
        state["complete"] = state["complete"] and int(res[2].min()) == k
        if b in keep_batches:
            kept.append((batch_uids(b), res))

    def run_window(b0, nb, record):
        for b in range(b0, b0 + nb):
            sl = slots[b % len(slots)]
            if sl.pending is not None:
                bb, res = sl.fetch()
                if record:
                    take(bb, res)
            sl.launch(b)
        for i in range(len(slots)):                # drain in launch order
            sl = slots[(b0 + nb + i) % len(slots)]
            if sl.pending is not None:
                bb, res = sl.fetch()
                if record:
                    take(bb, res)

    def barrier():
        for sl in slots:
            sl.dev.sync()
        torch.cuda.synchronize()
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run_window(n_batches, len(slots), False)       # warm-up on batches past the window (wrapping around)
    barrier()
    h2d0 = sum(sl.dev.stat(engine.STAT_H2D_BYTES) for sl in slots)
    d2h0 = sum(sl.dev.stat(engine.STAT_D2H_BYTES) for sl in slots)
    t = time.perf_counter()
    run_window(0, n_batches, True)
    barrier()
    dt = time.perf_counter() - t
    if ws > 1:
        tt = torch.tensor([dt], dtype=torch.float64, device=f"cuda:{lr}")
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dt = float(tt.item())
    if stats is not None:
        stats["h2d_bytes_per_batch"] = (sum(sl.dev.stat(engine.STAT_H2D_BYTES) for sl in slots) - h2d0) // max(n_batches, 1) + 4 * B
        stats["d2h_bytes_per_batch"] = (sum(sl.dev.stat(engine.STAT_D2H_BYTES) for sl in slots) - d2h0) // max(n_batches, 1)
        stats["fallback_queries_last"] = slots[(n_batches - 1) % len(slots)].dev.stat(engine.STAT_FALLBACK_QUERIES)
    for sl in slots:
        sl.dev.close()
    assert state["complete"], "a batch of the sweep returned incomplete top-k lists"
    return n_batches * B / dt, dt, kept


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=POKEC_N)
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--batches", type=int, default=16)
    ap.add_argument("--topk", type=int, default=50)
    ap.add_argument("--depth", type=int, default=2, help="batches in flight (stores on the GPU); 1 = no overlap")
    ap.add_argument("--reserve", default="", help="comma list: time the window once per value of PK_OPT_PREFILTER_RESERVE_SMS")
    args = ap.parse_args()
    fd = os.dup(1)
    os.dup2(2, 1)                      # NCCL prints its banner on stdout
    rank = int(os.environ.get("RANK", "0"))
    ws = int(os.environ.get("WORLD_SIZE", "1"))
    lr = int(os.environ.get("LOCAL_RANK", "0"))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(lr)
    if ws > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    from bench import build_world          # the bench world (cached on disk between invocations and ranks)
    w = build_world(args.users)
    if args.reserve:                       # A/B of the SMs the persistent pre-filter leaves to the tail of the previous batch
        sweep(w, ws, rank, lr, args.batch, args.topk, 4, args.depth)          # warm-up (allocations, NCCL channels)
        for r in [int(x) for x in args.reserve.split(",")]:
            rate, dt, _ = sweep(w, ws, rank, lr, args.batch, args.topk, args.batches, args.depth, reserve_sms=r)
            if rank == 0:
                os.write(fd, (json.dumps({"n_gpus": ws, "reserve_sms": r, "query_users_per_s": rate,
                                          "ms_per_batch": dt / args.batches * 1e3}) + "\n").encode())
        if ws > 1:
            dist.barrier()
            dist.destroy_process_group()
        return 0
    rate, dt, _ = sweep(w, ws, rank, lr, args.batch, args.topk, args.batches, args.depth)
    if rank == 0:
        out = {"workload": "all-users FAS top-50 (BASELINE configs[4])", "n_gpus": ws, "users": w.n, "batch": args.batch,
               "window_batches": args.batches, "batches_in_flight": max(1, args.depth), "query_users_per_s": rate,
               "ms_per_batch": dt / args.batches * 1e3, "full_sweep_batches": -(-w.n // args.batch),
               "full_sweep_seconds_extrapolated": w.n / rate, "pairs_per_s": rate * w.n, "complete_lists": True}
        os.write(fd, (json.dumps(out) + "\n").encode())
    if ws > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
