#!/usr/bin/env python
"""bench.py -- query-users/sec of the FAS top-50 full scan (BASELINE.json configs[1]).

One "step" = one pass of the hot path over one batch: Q = 4096 query users, each scored with
Fill-Aware Similarity against EVERY one of the 1 632 803 synthetic Pokec-shaped profiles, top-50 per
query by the reference comparator.

    python bench.py [--gpus N] [--steps K] [--warmup W]            # our CUDA path
    python bench.py --impl reference [...]                         # the reference's CPU code, host cores

N > 1 is launched by the driver with torch.distributed.run: one process per GPU, candidates partitioned by
row range (user-id range), per-rank top-k keys all-gathered over NCCL and merged (north_star).

Prints ONE JSON line on rank 0.  Besides the contract keys it carries
  parity         the lists the timed end-to-end step returned, compared id by id and bit by bit with the compiled
                 UNMODIFIED reference run on the full 1.63 M-user world for a sample of the timed queries
  cpu_baseline   that same reference run, timed (not extrapolated: every sampled query scans all candidates)
  roofline       the dominant kernel against the bound that actually limits it (issue slots), with the SURVEY 8d
                 algorithmic-bytes figure, the physical DRAM fraction and the shared-memory wavefront fraction beside it
  breakdown      device time of one step by phase (table build, pre-filter, merges, re-score, all-gathers, host plan)
  extra.configs  the other BASELINE configs: C1 (100 k users, 1 k interest queries, CPU timed in full), C3
                 (collaborative) and C4 (clubs) at 1.63 M users with their SURVEY 8d byte-formula rooflines, C5 (the
                 all-users sweep, N > 1) -- each with a parity sample against the reference
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from recommendation_system_pokec_b200.synth import (make_world, sample_normalizers, sample_queries, normalizer_pair_sample,  # noqa: E402
                                                    POKEC_N, SEED)
from recommendation_system_pokec_b200.cabi import EXCLUDE_SELF_AND_ADJ, MODE_GRAPH, MODE_COLLAB, MODE_CLUBS  # noqa: E402
from recommendation_system_pokec_b200.sharding import shard_rows, share_thetas  # noqa: E402

METRIC = "query-users/sec, FAS top-50 full scan vs 1.63M users"
UNIT = "query-users/s"
SM_COUNT = 148

_RESULT_FD = 1


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def emit(obj):
    os.write(_RESULT_FD, (json.dumps(obj) + "\n").encode())


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons while the timed region runs"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
                for nme, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nme)
            except Exception:
                pass
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------- inputs
def build_world(n_users):
    """the seeded synthetic world of the bench (synth.make_world).  Generating it takes ~27 s per process; it is
    deterministic, so the first process leaves a copy under POKEC_B200_WORLD_CACHE (default /tmp/pokec_b200_worlds)
    and later invocations / ranks load that (outside every timed region either way).  The normalizers it carries are
    placeholders (synth.sample_normalizers): our arm replaces them with bench_normalizers() below."""
    from recommendation_system_pokec_b200.world import World
    t = time.time()
    cache_dir = os.environ.get("POKEC_B200_WORLD_CACHE", "/tmp/pokec_b200_worlds")
    path = os.path.join(cache_dir, f"world_v1_{n_users}_{SEED:x}.npz")
    if cache_dir != "off" and os.path.exists(path):
        try:
            w = World.load(path)
            with open(path + ".meta.json") as f:
                w.meta = json.load(f)
            w.validate()
            log(f"[bench] world (cached {path}): {w.meta} ({time.time() - t:.1f}s)")
            return w
        except Exception as e:      # a torn or stale file: fall back to the generator
            log(f"[bench] world cache unusable ({e}); regenerating")
    w = make_world(n=n_users)
    mean, sd = sample_normalizers(w, pairs=4000)
    w.set_normalizers(mean, sd)
    log(f"[bench] world: {w.meta} ({time.time() - t:.1f}s)")
    if cache_dir != "off" and int(os.environ.get("LOCAL_RANK", "0")) == 0:
        try:
            os.makedirs(cache_dir, exist_ok=True)
            tmp = os.path.join(cache_dir, f".tmp_{os.getpid()}.npz")
            w.save(tmp)
            with open(path + ".meta.json", "w") as f:
                json.dump(w.meta, f)
            os.replace(tmp, path)
        except Exception as e:
            log(f"[bench] world cache not written ({e})")
    return w


def bench_normalizers(world, dev):
    """the column normalizers are those of the reference's compute_column_normalizers(profiles, cols,
    100000, 5) (src/utils.cpp:155-240, seed 12345): 500 000 distinct random pairs, mean / sample sd of the RAW per-field
    similarities, raw-count cosine for text.  Computed by the device path of SURVEY 8 f3 (pk_normalizers_from_pairs,
    bit-identical to the reference on the same pairs -- tests/test_gpu_parity.py) on the pair sample of
    synth.normalizer_pair_sample, then set on the world (the CPU reference gets the same floats) and on the store."""
    t = time.time()
    a, b = normalizer_pair_sample(world, 100_000, 5)
    mean, sd = dev.normalizers_from_pairs(a, b)
    world.set_normalizers(mean, sd)
    dev.set_normalizers(mean, sd)
    log(f"[bench] normalizers from {len(a)} pairs on the device (SURVEY 8 f3) in {time.time() - t:.2f}s")
    return len(a)


def cpu_reference_run(world, n_full, sample_users, q_per_thread, steps, warmup, threads):
    """`--impl reference`: the reference's own CPU implementation (oracle/_ref, the unmodified reference compiled from
    /root/reference) -- or the CPU restatement when that library did not travel -- on a bounded sample of the
    workload: `threads * q_per_thread` query users x the first `sample_users` candidates per step."""
    from oracle.oracle import CpuOracle, reference_available
    kind = "reference" if reference_available() else "port"
    sub = world.head(sample_users)
    t = time.time()
    orc = CpuOracle(sub, kind)
    if kind == "reference":     # the reference's own normalizers on its own store (values do not change the timing)
        m, s = orc.column_normalizers(100_000, 5)
        sub.set_normalizers(m, s)
        orc.close()
        orc = CpuOracle(sub, kind)
    log(f"[bench] cpu baseline ({kind}) store over {sub.n} users built in {time.time() - t:.1f}s")
    nq = threads * q_per_thread
    q = sample_queries(sub, nq, seed=2)
    times = []
    for i in range(warmup + steps):
        t = time.perf_counter()
        orc.fullscan(q, 50, EXCLUDE_SELF_AND_ADJ, nthreads=threads)
        dt = time.perf_counter() - t
        if i >= warmup:
            times.append(dt)
    orc.close()
    total = sum(times)
    pairs_per_s = len(q) * sub.n * len(times) / total
    value = pairs_per_s / n_full            # query-users/s against the full candidate set
    sample = (f"{len(q)} query users x first {sub.n} of {n_full} candidates per step, {len(times)} timed step(s), "
              f"scaled by candidates ({pairs_per_s / 1e6:.2f} M pairs/s)")
    return dict(value=value, unit=UNIT, cores=threads, kind=kind, sample=sample, extrapolated=True), total / len(times)


# ---------------------------------------------------------------------------------------------------- parity
def compare_lists(got, want):
    """got / want = (ids [q][k], scores [q][k], counts [q]).  Every position compared: ids exactly, scores by bits."""
    gi, gs, gc = got
    wi, ws, wc = want
    gb, wb = np.ascontiguousarray(gs, np.float32).view(np.int32).astype(np.int64), np.ascontiguousarray(ws, np.float32).view(np.int32).astype(np.int64)
    ulp = np.abs(gb - wb)
    rel = np.abs(gs.astype(np.float64) - ws) / np.maximum(np.abs(ws.astype(np.float64)), 1e-30)
    return {"queries": int(len(gc)), "entries": int(gi.size), "id_mismatches": int((gi != wi).sum()),
            "count_mismatches": int((gc != wc).sum()), "one_ulp_flips": int((ulp == 1).sum()),
            "beyond_one_ulp": int((ulp > 1).sum()), "max_rel_err": float(rel.max(initial=0.0))}


# ---------------------------------------------------------------------------------------------------- SURVEY 8d bytes
def list_mode_bytes(world, q, mode, limit, sample=256):
    """SURVEY 8d algorithmic bytes per query of the list modes, evaluated on the generator side (numpy, never by the
    kernel under test) for a sample of the queries:
      C3  sum over unique(friends U cands) of rec_bytes + 4 (deg(u) + sum_f deg(f))
      C4  sum over unique({u} U friends U fofs) of rec_bytes + 4 (deg(u) + sum_f deg(f)) + 8 (#club increments)
    (club increments counted for every edge whose target has a profile; the FAS > 0 gate is not evaluated here)."""
    rec = world.canonical_record_bytes()
    row_of = {int(u): i for i, u in enumerate(world.uid)} if world.n <= 200_000 else None
    uid_sorted = world.uid if np.all(np.diff(world.uid) > 0) else None

    def rows_of(u):
        u = np.asarray(u, np.int64)
        if uid_sorted is not None:
            p = np.searchsorted(uid_sorted, u)
            p = np.minimum(p, world.n - 1)
            return np.where(uid_sorted[p] == u, p, -1)
        return np.array([row_of.get(int(x), -1) for x in u], np.int64)

    def adj_row(u):
        i = int(np.searchsorted(world.adj_uid, u))
        if i < world.n_adj_rows and world.adj_uid[i] == u:
            return world.adj[world.adj_ptr[i]:world.adj_ptr[i + 1]]
        return world.adj[:0]

    ncl = np.diff(world.clubs_ptr)
    total = []
    for u in q[:sample]:
        fr = adj_row(int(u))
        adj_bytes = 4 * len(fr)
        fof_all = []
        for f in fr:
            r = adj_row(int(f))
            adj_bytes += 4 * len(r)
            fof_all.append(r)
        fof = np.concatenate(fof_all) if fof_all else world.adj[:0]
        fof = fof[fof != u]
        if mode == MODE_COLLAB:
            _, first = np.unique(fof, return_index=True)
            cands = fof[np.sort(first)][:limit]
            touched = np.unique(np.concatenate([fr, cands]))
            tr = rows_of(touched)
            total.append(int(rec[tr[tr >= 0]].sum()) + adj_bytes)
        else:
            touched = np.unique(np.concatenate([[u], fr, fof]))
            tr = rows_of(touched)
            inc = 0
            for lst in (fr, fof):
                r = rows_of(lst)
                inc += int(ncl[r[r >= 0]].sum())
            total.append(int(rec[tr[tr >= 0]].sum()) + adj_bytes + 8 * inc)
    return float(np.mean(total)) if total else 0.0


def run_list_config(name, dev, orc, world, q, mode, k, limit, threads, peak, reps=2, parity_q=64):
    """one BASELINE list-mode config on the resident store: end-to-end query-users/s (host uids in, host lists out),
    parity of a query sample with the reference, the reference timed on that sample, the 8d byte-formula roofline"""
    from recommendation_system_pokec_b200 import engine
    dev.recommend(mode, q, k, limit)                             # warm-up at the full batch size (grow-only device buffers)
    t = time.perf_counter()
    for _ in range(reps):
        got = dev.recommend(mode, q, k, limit)
    dt = (time.perf_counter() - t) / reps
    scan_ms = dev.stat(engine.STAT_SCAN_KERNEL_NS) / 1e6
    pairs = dev.stat(engine.STAT_PAIRS_SCORED)
    out = {"workload": name, "queries": int(len(q)), "topk": k, "candidate_limit": limit,
           "e2e": {"value": len(q) / dt, "unit": UNIT, "ms_per_batch": dt * 1e3}, "scan_kernel_ms": scan_ms,
           "pairs_scored": int(pairs)}
    sel = np.linspace(0, len(q) - 1, min(parity_q, len(q))).astype(np.int64)
    if orc is not None:
        t = time.perf_counter()
        want = orc.recommend(mode, q[sel], k, limit, threads)
        cdt = time.perf_counter() - t
        out["parity"] = compare_lists(tuple(x[sel] for x in got), want)
        out["cpu_baseline"] = {"value": len(sel) / cdt, "unit": UNIT, "cores": threads, "kind": orc.kind,
                               "sample": f"{len(sel)} of the {len(q)} queries, every one in full"}
    bpq = list_mode_bytes(world, q, mode, limit)
    ach = len(q) / dt * bpq / 1e9
    out["roofline"] = {"bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                       "algorithmic_bytes_per_query": bpq,
                       "note": "SURVEY 8d byte formula on a 256-query sample, times the END-TO-END rate (host enumeration included)"}
    return out


def run_c1(threads, peak):
    """BASELINE configs[0]: 100 000-user world, 1 000 query users, recommend_by_interest(uid, 50, 5000); the reference
    timed in full on all host cores and on one core (it is single-threaded as shipped); parity on all 1 000 lists"""
    from oracle.oracle import CpuOracle, reference_available
    from recommendation_system_pokec_b200.engine import DeviceRecommender
    w = make_world(n=100_000)
    dev = DeviceRecommender(w)
    a, b = normalizer_pair_sample(w, 100_000, 5)
    mean, sd = dev.normalizers_from_pairs(a, b)
    w.set_normalizers(mean, sd)
    dev.set_normalizers(mean, sd)
    q = sample_queries(w, 1000, seed=1)
    kind = "reference" if reference_available() else "port"
    orc = CpuOracle(w, kind)
    out = run_list_config("C1: reference api_cli interest recommendation, 100000 users, topk=50, 1k query users "
                          "(recommend_by_interest(uid, 50, 5000))", dev, None, w, q, MODE_GRAPH, 50, 5000, threads, peak, reps=3)
    got = dev.recommend(MODE_GRAPH, q, 50, 5000)
    t = time.perf_counter()
    want = orc.recommend(MODE_GRAPH, q, 50, 5000, threads)
    t_all = time.perf_counter() - t
    t = time.perf_counter()
    orc.recommend(MODE_GRAPH, q[:200], 50, 5000, 1)
    t_one = (time.perf_counter() - t) / 200
    out["parity"] = compare_lists(got, want)
    out["cpu_baseline"] = {"value": len(q) / t_all, "unit": UNIT, "cores": threads, "kind": kind,
                           "sample": "all 1000 queries, timed in full", "one_core_ms_per_query": t_one * 1e3}
    # the per-call latency of the drop-in's shape (one user per call, as src/api_cli.cpp:213 issues them)
    lat = []
    for u in q[:200]:
        t = time.perf_counter()
        dev.recommend(MODE_GRAPH, np.array([u], np.int32), 20, 5000)
        lat.append(time.perf_counter() - t)
    lat = np.sort(np.array(lat)) * 1e3
    out["single_call_ms"] = {"p50": float(lat[len(lat) // 2]), "p99": float(lat[int(len(lat) * 0.99)]),
                             "shape": "pk_recommend(GRAPH, 1 user, topk 20, limit 5000) through the C ABI"}
    dev.close()
    orc.close()
    return out


def main():
    # stdout carries exactly ONE JSON line: libraries that print to fd 1 (NCCL's version banner under torchrun)
    # are sent to stderr, the result line goes to a private copy of the original stdout
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--users", type=int, default=POKEC_N)
    ap.add_argument("--queries", type=int, default=4096)
    ap.add_argument("--topk", type=int, default=50)
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the reference run (no parity block, no cpu_baseline)")
    ap.add_argument("--parity-queries", type=int, default=-1, help="queries of the timed batch checked against the reference "
                    "on the full world (default 64 at 1 GPU, 32 at N > 1)")
    ap.add_argument("--skip-extra", action="store_true", help="skip extra.configs (C1, C3, C4, C5)")
    ap.add_argument("--allusers-batches", type=int, default=-1, help="C5 window in batches (default: the full sweep at N > 1, none at N = 1)")
    ap.add_argument("--cpu-sample-users", type=int, default=200_000)
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    config = {"workload": f"FAS interest top-{args.topk} full scan, {args.queries} query users x {args.users} candidates "
                          f"(BASELINE configs[1])",
              "users": args.users, "queries_per_step": args.queries, "topk": args.topk,
              "exclude": "self + existing friends (recommender_graph.cpp:46-52)",
              "l2": "inputs larger than L2 (packed store >> 126 MB)",
              "normalizers": "compute_column_normalizers(.., 100000, 5) (src/utils.cpp:155-240)",
              "parallelism": f"candidate rows sharded over {world_size} GPU(s); NCCL all-gather of the approximate shortlists, exact re-score split by query group, NCCL all-gather of the top-k keys + merge" if world_size > 1
              else "1 GPU"}

    # ------------------------------------------------------------------ reference arm (CPU)
    if args.impl == "reference":
        if rank != 0:
            return 0
        w = build_world(args.users)      # the same world as our arm; the sample is its first rows
        base, step_s = cpu_reference_run(w, args.users, min(args.cpu_sample_users, w.n), 2, max(args.steps, 1), args.warmup, threads)
        out = {"impl": "reference", "metric": METRIC, "value": base["value"], "unit": UNIT, "n_gpus": args.gpus,
               "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
               "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
               "cpu_baseline": base, "normalizer_source": "the reference's own compute_column_normalizers on the sampled store",
               "e2e": {"value": base["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        emit(out)
        return 0

    # ------------------------------------------------------------------ our arm (CUDA)
    import torch
    import torch.distributed as dist
    from recommendation_system_pokec_b200 import engine
    from recommendation_system_pokec_b200.engine import DeviceRecommender

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the scoring path has no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), timeout=datetime.timedelta(minutes=30))

    world = build_world(args.users)
    n = world.n
    t0 = time.time()
    dev = DeviceRecommender(world, device=local_rank)
    n_norm_pairs = bench_normalizers(world, dev)
    lo, hi = shard_rows(n, rank, world_size)
    dev.set_shard(lo, hi)
    q = sample_queries(world, args.queries, seed=2)
    nq, k = len(q), args.topk
    thetas_shared = share_thetas(dev, rank, world_size, f"cuda:{local_rank}")     # pruning thresholds over NVLink peer memory
    dev.fullscan_prepare(q, k, EXCLUDE_SELF_AND_ADJ)       # builds idf, packed store, plan
    dev.sync()
    log(f"[bench] rank {rank}: store {dev.stat(engine.STAT_STORE_BYTES) / 1e9:.2f} GB packed, shard rows [{lo},{hi}) "
        f"ready in {time.time() - t0:.1f}s")

    stream = torch.cuda.ExternalStream(dev.stream(), device=local_rank)
    cuda = f"cuda:{local_rank}"
    gathered = torch.zeros((world_size, nq, k), dtype=torch.int64, device=cuda)
    local = torch.zeros((nq, k), dtype=torch.int64, device=cuda)

    # N > 1: the two-stage scan is split at its exchange point -- every rank pre-filters its candidate shard, the
    # approximate shortlists are all-gathered, every rank proves the merged shortlists and re-scores (fp64) only the
    # query groups it owns, then the per-rank [Q][k] keys are all-gathered and merged
    ks = dev.stat(engine.STAT_SHORTLIST)
    split = world_size > 1 and ks > 0
    if split:
        local_s = torch.zeros((nq, ks), dtype=torch.int64, device=cuda)
        gathered_s = torch.zeros((world_size, nq, ks), dtype=torch.int64, device=cuda)
    ag_ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]

    def scan_and_exchange(fetch, mark=False):
        if split:
            dev.fullscan_execute_shortlists()
            dev.fullscan_shortlist_keys(local_s.data_ptr())
            with torch.cuda.stream(stream):
                if mark:
                    ag_ev[0].record(stream)
                dist.all_gather_into_tensor(gathered_s.view(-1), local_s.view(-1))
                if mark:
                    ag_ev[1].record(stream)
            dev.fullscan_execute_finish(gathered_s.data_ptr(), world_size, rank, world_size)
        else:
            dev.fullscan_execute()
        if world_size > 1:
            dev.fullscan_keys(local.data_ptr())
            with torch.cuda.stream(stream):
                if mark:
                    ag_ev[2].record(stream)
                dist.all_gather_into_tensor(gathered.view(-1), local.view(-1))
                if mark:
                    ag_ev[3].record(stream)
            return dev.merge_keys(gathered.data_ptr(), world_size, nq, k, fetch=fetch)   # fetch=False: keys stay on the device
        return None

    def device_step():
        return scan_and_exchange(False)

    def barrier():
        dev.sync()
        torch.cuda.synchronize()
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world_size == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=cuda)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident measurement: prepared plan, kernels only --------------------------------
    for _ in range(args.warmup):
        device_step()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = dev.stat(engine.STAT_KERNEL_LAUNCHES)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        device_step()
    ev1.record(stream)
    barrier()
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
    launches = dev.stat(engine.STAT_KERNEL_LAUNCHES) - launches0
    # scan-kernel time of the LAST timed step (CUDA events on the launching stream, inside the timed region)
    scan_ns = dev.stat(engine.STAT_SCAN_KERNEL_NS)
    scan_launches = dev.stat(engine.STAT_SCAN_LAUNCHES)
    clocks = sampler.stop() if rank == 0 else None
    value = nq * args.steps / (dev_ms / 1e3)
    breakdown = {"prefilter_ms": max_over_ranks(scan_ns / 1e6), "table_build_ms": max_over_ranks(dev.stat(engine.STAT_PREP_NS) / 1e6),
                 "merge_ms": max_over_ranks(dev.stat(engine.STAT_MERGE_NS) / 1e6),
                 "rescore_ms": max_over_ranks(dev.stat(engine.STAT_RESCORE_NS) / 1e6)}

    # ---- end to end: host uids in, host top-k out, every step -------------------------------------
    # The public calls with HOST buffers, every step: host planning, H2D of the plan + exclusion ids, the kernels, both
    # all-gathers (N > 1), D2H of the merged keys, decode.  Two batches are kept in flight on two stores / streams (as the
    # all-users sweep does): the host planning of step i + 1 overlaps the scan of step i.
    from tools.bench_allusers import sweep
    e2e_stats = {}
    e2e_value, e2e_s, kept = sweep(world, world_size, rank, local_rank, nq, k, args.steps, depth=2, keep_batches=(args.steps - 1,),
                                   batch_fn=lambda b: q, stats=e2e_stats)
    res = kept[-1][1]
    h2d, d2h = e2e_stats["h2d_bytes_per_batch"], e2e_stats["d2h_bytes_per_batch"]
    assert res is not None and int(res[2].min()) == k, "e2e step returned incomplete top-k lists"
    fallback_q = e2e_stats["fallback_queries_last"]

    # one more (untimed) step with marks around the host planning and the two all-gathers
    t = time.perf_counter()
    dev.fullscan_prepare(q, k, EXCLUDE_SELF_AND_ADJ)
    breakdown["host_plan_ms"] = max_over_ranks((time.perf_counter() - t) * 1e3)
    scan_and_exchange(False, mark=True)
    barrier()
    if world_size > 1:
        ag = (ag_ev[0].elapsed_time(ag_ev[1]) if split else 0.0) + ag_ev[2].elapsed_time(ag_ev[3])
        breakdown["allgather_ms"] = max_over_ranks(ag)
    breakdown["step_ms"] = dev_ms / args.steps
    breakdown["note"] = ("device time of one step by phase, max over ranks; allgather_ms includes waiting for the slowest rank; "
                         "host_plan_ms is paid by the end-to-end step only")

    # ---- roofline of the dominant kernel -----------------------------------------------------------
    rec_bytes = world.canonical_record_bytes()
    b_shard = int(rec_bytes[lo:hi].sum())             # canonical bytes of this rank's candidate shard
    alg_bytes_per_launch = nq * b_shard / max(scan_launches, 1)
    scan_s = scan_ns / 1e9 / max(scan_launches, 1)
    peak, peak_src = measured_peak_hbm()
    alg_achieved = alg_bytes_per_launch / scan_s / 1e9 if scan_s > 0 else 0.0
    two_stage = dev.stat(engine.STAT_PREFILTER) == 1
    # counters of ONE launch of the dominant kernel on this workload from the committed `ncu --set full` capture of the
    # shipped build (profiles/ncu_traffic.json, written by tools/ncu_counters.py): warp instructions, DRAM bytes,
    # shared-memory wavefronts.  They are properties of (build, workload), not of a run; the launch time they are
    # divided by is measured live (CUDA events inside the timed region).
    tr = None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tr = json.load(f)
        if not (two_stage and tr.get("users") == args.users and tr.get("queries") == args.queries
                and tr.get("warp_instructions_per_launch") and tr.get("smem_wavefronts_per_launch")):
            tr = None
    except Exception:
        tr = None
    clk_hz = (clocks or {}).get("sm_mhz") or 1965.0
    clk_hz *= 1e6
    roofline = {"kernel": "fas_prefilter_kernel<false>" if two_stage else "fas_scan_kernel<true,0,false>",
                "launch_ms": scan_s * 1e3, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes_per_launch, "algorithmic_gbps": alg_achieved,
                "algorithmic_frac": alg_achieved / peak,
                "note": "the kernel is ISSUE-bound: 32 queries share every streamed record, so DRAM moves ~0.4 % of the "
                        "algorithmic bytes. frac = warp instructions per launch / (SMs x 4 schedulers x SM clock x launch time). "
                        "algorithmic_frac is the SURVEY 8d figure (queries x canonical record bytes / time / HBM peak; "
                        "cross-query reuse not credited, can exceed 1); dram_frac is the physical DRAM fraction."}
    if tr is not None and scan_s > 0:
        # N > 1: a rank scans 1/N of the candidates, so its share of the unsharded launch's counters is the ALGORITHMIC
        # work of its launch; what a small shard executes beyond that (weaker pruning, fixed costs) counts as waste, i.e.
        # lowers frac rather than raising it
        inst = float(tr["warp_instructions_per_launch"]) / world_size
        tr = dict(tr)
        for key in ("dram_bytes_per_launch", "smem_wavefronts_per_launch", "smem_bank_conflicts_per_launch"):
            if tr.get(key) is not None:
                tr[key] = float(tr[key]) / world_size
        if world_size > 1:
            tr["source"] = f"1/{world_size} of: " + str(tr.get("source"))
        issue_peak = SM_COUNT * 4 * clk_hz
        roofline.update({"bound": "issue", "achieved": inst / scan_s / 1e9, "peak": issue_peak / 1e9, "unit": "G warp-instr/s",
                         "frac": inst / scan_s / issue_peak, "issue_frac": inst / scan_s / issue_peak,
                         "traffic": tr["dram_bytes_per_launch"],
                         "dram_frac": tr["dram_bytes_per_launch"] / scan_s / 1e9 / peak,
                         "smem_wavefront_frac": float(tr["smem_wavefronts_per_launch"]) / scan_s / (SM_COUNT * clk_hz),
                         "smem_bank_conflict_share": float(tr.get("smem_bank_conflicts_per_launch", 0)) / float(tr["smem_wavefronts_per_launch"]),
                         "counters_from": tr.get("source"), "hbm_peak_gbps": peak})
    else:       # no capture of this (build, workload): only the algorithmic figure can be stated
        roofline.update({"bound": "hbm", "achieved": alg_achieved, "peak": peak, "unit": "GB/s", "frac": alg_achieved / peak,
                         "traffic": None})

    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f32 pre-filter + f64 exact re-score (every returned score is the f64 path's)", "data": "synthetic",
           "config": config, "clocks": clocks,
           "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                   "ms_per_step": e2e_s / args.steps * 1e3, "batches_in_flight": 2},
           "gpu_launches": int(launches), "roofline": roofline, "breakdown": breakdown,
           "pipeline": {"two_stage": two_stage, "shortlist": dev.stat(engine.STAT_SHORTLIST),
                        "prefilter_margin": dev.stat(engine.STAT_MARGIN_E9) / 1e9,
                        "fallback_queries_last_step": fallback_q,
                        "rescore_and_fallback_ms": dev.stat(engine.STAT_RESCORE_NS) / 1e6,
                        "lean_record_bytes_shard": dev.stat(engine.STAT_SHARD_LEAN_BYTES),
                        "exact_record_bytes_shard": dev.stat(engine.STAT_SHARD_RECORD_BYTES),
                        "normalizer_pairs": n_norm_pairs, "normalizer_source": "pk_normalizers_from_pairs on the device (SURVEY 8 f3)",
                        "thresholds_shared_over_nvlink": bool(thetas_shared)}}

    # ---- C5 (N > 1): the all-users sweep through the end-to-end calls, two batches in flight ----------------
    extra = {}
    sweep_keep = None
    n_sweep = args.allusers_batches if args.allusers_batches >= 0 else (-(-n // nq) if world_size > 1 else 0)
    if n_sweep > 0 and not args.skip_extra:
        from tools.bench_allusers import sweep
        rate, dt, sweep_keep = sweep(world, world_size, rank, local_rank, nq, k, n_sweep, depth=2, keep_batches=(0, n_sweep // 2, n_sweep - 1))
        full = -(-n // nq)
        extra["C5"] = {"workload": "C5: all-users batch scoring (every user a query, FAS top-50 vs all candidates), candidate-sharded, "
                                   "NCCL merges, two batches in flight (BASELINE configs[4])",
                       "n_gpus": world_size, "batches": n_sweep, "full_sweep_batches": full, "complete_sweep": bool(n_sweep >= full),
                       "e2e": {"value": rate, "unit": UNIT, "ms_per_batch": dt / n_sweep * 1e3}, "seconds": dt,
                       "pairs_per_s": rate * n}

    # ---- the reference on the FULL world (rank 0): parity of the timed lists, cpu_baseline, C3 / C4 / C1 -------
    if rank == 0 and not args.no_cpu_baseline:
        from oracle.oracle import CpuOracle, reference_available
        kind = "reference" if reference_available() else "port"
        t = time.time()
        orc = CpuOracle(world, kind)
        log(f"[bench] {kind} store over all {n} users built in {time.time() - t:.1f}s")
        npar = args.parity_queries if args.parity_queries >= 0 else (64 if world_size == 1 else 32)
        npar = min(npar, nq)
        sel = np.linspace(0, nq - 1, npar).astype(np.int64)
        t = time.perf_counter()
        want = orc.fullscan(q[sel], k, EXCLUDE_SELF_AND_ADJ, nthreads=threads)
        cdt = time.perf_counter() - t
        par = compare_lists(tuple(x[sel] for x in res), want)
        par.update({"against": f"{kind} (oracle/_ref = the unmodified reference compiled from its sources)" if kind == "reference" else kind,
                    "candidates_per_query": n, "fallback_queries": int(fallback_q),
                    "what": "lists returned by the LAST timed end-to-end step (merged over all ranks)"})
        out["parity"] = par
        out["cpu_baseline"] = {"value": npar / cdt, "unit": UNIT, "cores": threads, "kind": kind, "extrapolated": False,
                               "sample": f"{npar} of the {nq} timed query users x all {n} candidates, one pass "
                                         f"({npar * n / cdt / 1e6:.2f} M pairs/s)"}
        log(f"[bench] parity {par}")
        if sweep_keep is not None:           # C5: a sample of the sweep's lists against the reference
            qs = np.concatenate([uq[:: max(1, len(uq) // 8)][:8] for uq, _ in sweep_keep])
            gs = tuple(np.concatenate([r[i][:: max(1, len(uq) // 8)][:8] for uq, r in sweep_keep]) for i in range(3))
            extra["C5"]["parity"] = compare_lists(gs, orc.fullscan(qs, k, EXCLUDE_SELF_AND_ADJ, nthreads=threads))
        if world_size == 1 and not args.skip_extra:
            dev.set_shard(0, n)
            extra["C3"] = run_list_config("C3: collaborative friend-of-friend propagation weighted by FAS, 1.63M users / 30.5M edges "
                                          "(recommend_collaborative(uid, 50, 5000))", dev, orc, world, q, MODE_COLLAB, k, 5000, threads, peak)
            extra["C4"] = run_list_config("C4: collaborative club recommendation (recommend_clubs_collab(uid, 50, .)), scan-based",
                                          dev, orc, world, q, MODE_CLUBS, k, 5000, threads, peak)
            extra["C2_interest_lists"] = run_list_config("recommend_by_interest(uid, 50, 5000) on the 1.63M world (the reference's own "
                                                         "candidate sets)", dev, orc, world, q, MODE_GRAPH, k, 5000, threads, peak)
            t = time.perf_counter()
            dev.edge_cache_build()
            t_cache = time.perf_counter() - t
            c4c = run_list_config("C4 served from the per-edge FAS cache (SURVEY 8 f4)", dev, orc, world, q, MODE_CLUBS, k, 5000, threads, peak)
            c4c["edge_cache_build_s"] = t_cache
            extra["C4_cached"] = c4c
            dev.edge_cache_drop()
        orc.close()
        if world_size == 1 and not args.skip_extra:
            dev.close()
            extra["C1"] = run_c1(threads, peak)
    if extra:
        out["extra"] = {"configs": extra}
    if rank == 0:
        emit(out)
    if world_size > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
