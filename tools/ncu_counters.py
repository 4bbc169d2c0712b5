#!/usr/bin/env python
"""ncu_counters.py -- the per-launch counters bench.py's roofline block needs, from one `ncu --set full` capture.

    python tools/ncu_counters.py gpurun_out/x.ncu-rep --users 1632803 --queries 4096 --note "..." \
        [--raw profiles/x_ncu_full.txt] > profiles/ncu_traffic.json

Reads `ncu -i REP --page raw` (first kernel of the report), optionally saves that raw page, and prints a JSON object:
warp instructions, DRAM bytes, shared-memory wavefronts and bank conflicts of ONE launch, plus the launch shape.
These are properties of (build, workload); bench.py divides them by the launch time it measures live.
"""
import json
import re
import subprocess
import sys


def main():
    rep = sys.argv[1]
    arg = lambda name, d=None: sys.argv[sys.argv.index(name) + 1] if name in sys.argv else d
    if rep.endswith(".txt"):          # an already exported raw page
        txt = open(rep).read()
    else:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw"], capture_output=True, text=True).stdout
    if arg("--raw"):
        open(arg("--raw"), "w").write(txt)
    UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3,
            "inst": 1, "": 1, "cycle": 1, "register/thread": 1, "block": 1, "thread": 1}
    val = {}
    kernel = None
    for line in txt.splitlines():
        if kernel is None and "_kernel" in line and "(" in line and not line.lstrip().startswith("#"):
            kernel = line.strip().split("(")[0].strip()
        m = re.match(r"^\s*(\S+)\s+(?:(\S+)\s+)?([0-9][0-9,\.]*)\s*$", line)
        if not m:
            continue
        name, unit, v = m.group(1), m.group(2) or "", float(m.group(3).replace(",", ""))
        if name not in val:
            val[name] = v * UNIT.get(unit, 1)
    g = lambda k: val.get(k)
    out = {"kernel": kernel, "users": int(arg("--users", 0)), "queries": int(arg("--queries", 0)), "n_gpus": 1,
           "warp_instructions_per_launch": g("smsp__inst_executed.sum"),
           "dram_bytes_per_launch": (g("dram__bytes_read.sum") or 0) + (g("dram__bytes_write.sum") or 0),
           "dram_bytes_read": g("dram__bytes_read.sum"), "dram_bytes_write": g("dram__bytes_write.sum"),
           "smem_wavefronts_per_launch": g("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"),
           "smem_bank_conflicts_per_launch": g("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
           "gpu_time_duration_ms_under_ncu": g("gpu__time_duration.sum"),
           "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active") or g("sm__inst_issued.avg.pct_of_peak_sustained_active"),
           "block_size": g("launch__block_size"), "registers_per_thread": g("launch__registers_per_thread"),
           "grid_size": g("launch__grid_size"),
           "source": arg("--note", rep)}
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
