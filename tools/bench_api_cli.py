#!/usr/bin/env python
"""bench_api_cli.py -- latency of the reference's serving loop, one `USER <id>` line at a time (src/api_cli.cpp:175-255:
four recommend_* calls with topk 20, candidate_limit 5000 per line), on BASELINE configs[0]'s world (100 000 synthetic
users -- the size the reference's loader caps at, src/user_loader.cpp:34):

    oracle/_ref/api_cli_ref          the UNMODIFIED reference binary
    dropin/_build/api_cli_b200       the same api_cli.cpp linked against dropin/*.cpp + libpokec_b200.so

Both read the same data directory; every reply is compared byte for byte.  Prints one JSON object: start-up time to
READY, p50 / p90 / p99 / mean milliseconds per USER line for each binary, and the number of differing replies.

    python tools/bench_api_cli.py [--users 100000] [--lines 300] [--out gpurun_out/api_cli_latency.json]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.refdata import write_reference_data_dir  # noqa: E402
from recommendation_system_pokec_b200.synth import make_world, sample_normalizers, sample_queries  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "api_cli_ref")
OUR_BIN = os.path.join(ROOT, "dropin", "_build", "api_cli_b200")


def serve(binary, cwd, users, warm):
    t0 = time.perf_counter()
    p = subprocess.Popen([binary], stdin=subprocess.PIPE, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, cwd=cwd,
                         text=True, bufsize=1)
    while True:
        line = p.stdout.readline()
        if not line:
            raise RuntimeError(f"{binary} ended before READY")
        if line.strip() == "READY":
            break
    ready_s = time.perf_counter() - t0
    replies, lat = [], []
    for i, u in enumerate(list(users[:warm]) + list(users)):
        t = time.perf_counter()
        p.stdin.write(f"USER {u}\n")
        p.stdin.flush()
        rep = p.stdout.readline()
        dt = time.perf_counter() - t
        if i >= warm:
            replies.append(rep)
            lat.append(dt * 1e3)
    p.stdin.write("EXIT\n")
    p.stdin.flush()
    p.wait(timeout=60)
    lat = np.sort(np.array(lat))
    return replies, {"ready_s": ready_s, "lines": len(lat), "p50_ms": float(lat[len(lat) // 2]), "p90_ms": float(lat[int(len(lat) * 0.9)]),
                     "p99_ms": float(lat[min(len(lat) - 1, int(len(lat) * 0.99))]), "mean_ms": float(lat.mean()), "max_ms": float(lat[-1])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--users", type=int, default=100_000)
    ap.add_argument("--lines", type=int, default=300)
    ap.add_argument("--ref-lines", type=int, default=100, help="USER lines sent to the reference binary (it takes ~0.1 s per line)")
    ap.add_argument("--out", default="")
    args = ap.parse_args()
    if not (os.path.exists(REF_BIN) and os.path.exists(OUR_BIN)):
        raise SystemExit("api_cli binaries not built (dropin/build_api_cli.py needs the reference sources)")
    w = make_world(n=args.users)
    m, s = sample_normalizers(w, pairs=3000)
    w.set_normalizers(m, s)
    users = [int(u) for u in sample_queries(w, args.lines, seed=1)]
    with tempfile.TemporaryDirectory() as d:
        t = time.time()
        write_reference_data_dir(w, d, club_names=False)
        print(f"[api_cli bench] data dir for {w.n} users written in {time.time() - t:.1f}s", file=sys.stderr, flush=True)
        ours, so = serve(OUR_BIN, d, users, warm=5)
        ref, sr = serve(REF_BIN, d, users[:args.ref_lines], warm=2)
    differ = sum(a != b for a, b in zip(ours, ref))
    out = {"workload": f"api_cli USER lines (4 recommend calls each, topk 20, limit 5000), {w.n} users (BASELINE configs[0] world)",
           "drop_in": so, "reference": sr, "replies_compared": len(ref), "replies_differing": differ,
           "p50_speedup": sr["p50_ms"] / so["p50_ms"], "p99_speedup": sr["p99_ms"] / so["p99_ms"]}
    txt = json.dumps(out)
    print(txt)
    if args.out:
        with open(args.out, "w") as f:
            f.write(txt + "\n")
    return 0 if differ == 0 else 1


if __name__ == "__main__":
    sys.exit(main())
