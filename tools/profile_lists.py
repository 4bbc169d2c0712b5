#!/usr/bin/env python
"""profile_lists.py -- one batched list-mode call on the bench world, for ncu / timing of its kernels.

    python tools/profile_lists.py --mode collab [--queries 4096] [--reps 1]
    ncu --set full -k regex:fas_pairs -c 1 ... python tools/profile_lists.py --mode collab

The normalizers are the numpy placeholders of the world cache (no device pair launch before the profiled call)."""
import argparse
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="collab", choices=["graph", "collab", "clubs"])
    ap.add_argument("--queries", type=int, default=4096)
    ap.add_argument("--reps", type=int, default=1)
    ap.add_argument("--users", type=int, default=0)
    args = ap.parse_args()
    from bench import build_world
    from recommendation_system_pokec_b200.synth import sample_queries, POKEC_N
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    from recommendation_system_pokec_b200 import engine
    from recommendation_system_pokec_b200.cabi import MODE_GRAPH, MODE_COLLAB, MODE_CLUBS
    w = build_world(args.users or POKEC_N)
    dev = DeviceRecommender(w)
    q = sample_queries(w, args.queries, seed=2)
    mode = {"graph": MODE_GRAPH, "collab": MODE_COLLAB, "clubs": MODE_CLUBS}[args.mode]
    for r in range(args.reps):
        t = time.perf_counter()
        res = dev.recommend(mode, q, 50, 5000)
        dt = time.perf_counter() - t
        print(f"{args.mode}: {len(q)} queries in {dt * 1e3:.1f} ms, scan kernels {dev.stat(engine.STAT_SCAN_KERNEL_NS) / 1e6:.2f} ms, "
              f"{dev.stat(engine.STAT_PAIRS_SCORED)} pairs, lists with entries: {(res[2] > 0).sum()}", file=sys.stderr, flush=True)
    dev.close()


if __name__ == "__main__":
    main()
