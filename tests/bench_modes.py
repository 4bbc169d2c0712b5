#!/usr/bin/env python
"""bench_modes.py -- BASELINE configs[2] / configs[3] (and the list-mode configs[0] shape) on one GPU.

    python tests/bench_modes.py [--users N] [--queries Q] [--steps K] [--cpu-queries C]

For each of the reference's list modes (include/recommender.h:24-28)
    graph   recommend_graph_registration == recommend_by_interest(uid, 50, 5000)
    collab  recommend_collaborative(uid, 50, 5000)
    clubs   recommend_clubs_collab(uid, 50, .)
it times pk_recommend over a batch of Q query users END TO END (host uids in, host top-k out: candidate
enumeration on the host, H2D of the plan, kernels, D2H) and the reference's own CPU code (oracle/_ref, all
host threads) on the first C of the same queries, checks the ids of those C queries against the reference,
and prints one JSON line per mode.  These are parity-test configurations, not the headline bench line
(bench.py is configs[1]); the numbers go to profiles/.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))   # repo root (this file lives in tests/: it times the oracle)
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.synth import make_world, sample_normalizers, sample_queries, POKEC_N  # noqa: E402
from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS  # noqa: E402
from recommendation_system_pokec_b200 import engine  # noqa: E402
from recommendation_system_pokec_b200.engine import DeviceRecommender  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=POKEC_N)
    ap.add_argument("--queries", type=int, default=4096)
    ap.add_argument("--topk", type=int, default=50)
    ap.add_argument("--limit", type=int, default=5000)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--cpu-queries", type=int, default=64)
    ap.add_argument("--modes", default="graph,collab,clubs")
    args = ap.parse_args()
    threads = len(os.sched_getaffinity(0))

    w = make_world(n=args.users)
    m, s = sample_normalizers(w, pairs=4000)
    w.set_normalizers(m, s)
    dev = DeviceRecommender(w, device=0)
    q = sample_queries(w, args.queries, seed=2)
    from oracle.oracle import CpuOracle, reference_available
    kind = "reference" if reference_available() else "port"
    orc = CpuOracle(w, kind) if args.cpu_queries > 0 else None
    names = {"graph": MODE_GRAPH, "collab": MODE_COLLAB, "clubs": MODE_CLUBS, "clubs_cached": MODE_CLUBS}
    for name in args.modes.split(","):
        if name == "normalizers":
            # compute_column_normalizers(profiles, cols, 100000, 5) (src/utils.cpp:155-240, src/main.cpp:124): the reference on one host core (as shipped) against pk_normalizers_from_pairs on the
            # reference's own pair sample; every mean / sd float must be identical
            if orc is None or kind != "reference":
                continue
            t = time.perf_counter()
            rm, rs = orc.column_normalizers(100000, 5)
            ct = time.perf_counter() - t
            a, b = orc.normalizer_pairs(100000, 5)
            dev.normalizers_from_pairs(a[:1000], b[:1000])
            t = time.perf_counter()
            gm, gs = dev.normalizers_from_pairs(a, b)
            gt = time.perf_counter() - t
            print(json.dumps({"mode": name, "users": w.n, "pairs": int(len(a)), "fields": int(len(rm)),
                              "reference_cpu_s": ct, "cpu_cores": 1, "gpu_e2e_s": gt, "speedup": ct / gt,
                              "bit_identical": bool(gm.tobytes() == rm.tobytes() and gs.tobytes() == rs.tobytes())}),
                  flush=True)
            continue
        cache_s = None
        if name == "clubs_cached":
            # the clubs call served from the per-edge FAS cache (built once per world / normalizer set)
            t = time.perf_counter()
            dev.edge_cache_build()
            cache_s = time.perf_counter() - t
            cache_pairs, cache_scan_ms = dev.stat(engine.STAT_PAIRS_SCORED), dev.stat(engine.STAT_SCAN_KERNEL_NS) / 1e6
        mode = names[name]
        dev.recommend(mode, q, args.topk, args.limit)          # warm-up (also sizes the arenas)
        dev.sync()
        l0 = dev.stat(engine.STAT_KERNEL_LAUNCHES)
        t = time.perf_counter()
        for _ in range(args.steps):
            got = dev.recommend(mode, q, args.topk, args.limit)
        dt = (time.perf_counter() - t) / args.steps
        out = {"mode": name, "users": w.n, "queries": len(q), "topk": args.topk, "limit": args.limit,
               "e2e_query_users_per_s": len(q) / dt, "ms_per_batch": dt * 1e3,
               "pairs_scored_per_batch": dev.stat(engine.STAT_PAIRS_SCORED),
               "scan_kernel_ms": dev.stat(engine.STAT_SCAN_KERNEL_NS) / 1e6,
               "gpu_launches_per_batch": (dev.stat(engine.STAT_KERNEL_LAUNCHES) - l0) // args.steps}
        if cache_s is not None:
            out["edge_cache"] = {"edges": dev.stat(engine.STAT_EDGE_CACHE), "build_s": cache_s,
                                 "pairs_scored": cache_pairs, "scan_kernel_ms": cache_scan_ms}
        if orc is not None:
            qc = q[:args.cpu_queries]
            t = time.perf_counter()
            want = orc.recommend(mode, qc, args.topk, args.limit, threads)
            ct = time.perf_counter() - t
            sub = dev.recommend(mode, qc, args.topk, args.limit)
            out["cpu"] = {"kind": kind, "cores": threads, "queries": len(qc), "query_users_per_s": len(qc) / ct}
            out["ids_equal"] = bool(np.array_equal(sub[0], want[0]) and np.array_equal(sub[2], want[2]))
            rel = np.abs(sub[1].astype(np.float64) - want[1]) / np.maximum(np.abs(want[1]), 1e-30)
            out["max_rel_score_err"] = float(rel[want[0] >= 0].max()) if (want[0] >= 0).any() else 0.0
            out["speedup_vs_cpu"] = out["e2e_query_users_per_s"] / out["cpu"]["query_users_per_s"]
        print(json.dumps(out), flush=True)
    dev.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
