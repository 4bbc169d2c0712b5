#!/usr/bin/env python
"""variants.py -- A/B timing of experimental builds of libpokec_b200 (build.build_variant) on the bench workload.

    python tools/variants.py [--users N] [--queries Q] [--steps K] name1 name2 ...

`base` = the product library; any other name = recommendation_system_pokec_b200/libpokec_b200.<name>.so.
The world is generated once; every variant runs the same prepared full scan, its top-k ids are compared with the
first variant's (they must be identical: the output is always the exact kernel's), and one JSON line is printed
per variant.  Experiment tooling only -- bench.py is the measurement of record.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.synth import sample_queries, POKEC_N  # noqa: E402
from recommendation_system_pokec_b200.cabi import EXCLUDE_SELF_AND_ADJ  # noqa: E402
from recommendation_system_pokec_b200 import engine  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=POKEC_N)
    ap.add_argument("--queries", type=int, default=4096)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--shard", type=int, default=1, help="scan only the first 1/shard of the rows (multi-GPU shard size)")
    ap.add_argument("names", nargs="+")
    args = ap.parse_args()
    from bench import build_world          # the bench world (cached on disk between invocations)
    w = build_world(args.users)
    q = sample_queries(w, args.queries, seed=2)
    ref_ids = None
    for spec in args.names:
        # "name" or "name:ENV=VALUE[,ENV=VALUE]" (environment switches read by the library at prepare / execute time)
        name, _, envs = spec.partition(":")
        for kv in filter(None, envs.split(",")):
            k_, _, v_ = kv.partition("=")
            os.environ[k_] = v_
        path = engine.LIB_PATH if name == "base" else os.path.join(os.path.dirname(engine.LIB_PATH), f"libpokec_b200.{name}.so")
        os.environ["POKEC_B200_LIB"] = path
        engine._lib = None
        dev = engine.DeviceRecommender(w, device=0)
        dev.set_shard(0, w.n // args.shard)
        dev.fullscan_prepare(q, 50, EXCLUDE_SELF_AND_ADJ)
        for _ in range(2):
            dev.fullscan_execute()
        dev.sync()
        t = time.perf_counter()
        for _ in range(args.steps):
            dev.fullscan_execute()
        dev.sync()
        dt = (time.perf_counter() - t) / args.steps
        ids, sc, cnt = dev.fullscan_fetch()
        same = True if ref_ids is None else bool(np.array_equal(ids, ref_ids))
        if ref_ids is None:
            ref_ids = ids
        for kv in filter(None, envs.split(",")):
            os.environ.pop(kv.partition("=")[0], None)
        print(json.dumps({"variant": spec, "shard": args.shard, "ms_per_step": dt * 1e3, "query_users_per_s": len(q) / dt,
                          "prefilter_ms": dev.stat(engine.STAT_SCAN_KERNEL_NS) / 1e6,
                          "rescore_ms": dev.stat(engine.STAT_RESCORE_NS) / 1e6,
                          "fallback_queries": dev.stat(engine.STAT_FALLBACK_QUERIES), "ids_equal_first": same}), flush=True)
        dev.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
