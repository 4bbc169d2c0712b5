#!/usr/bin/env python
"""ncu_lines.py -- per-source-line instruction / stall summary of one kernel from an .ncu-rep.

    python tools/ncu_lines.py gpurun_out/x.ncu-rep [--top 40] > profiles/x_lines.txt

Reads `ncu -i REP --page source --csv --print-source cuda,sass` (the kernel must have been compiled with
-lineinfo and captured with --import-source on) and sums, per CUDA source line, the warp instructions
executed, the stall samples and the shared-memory wavefronts of the SASS instructions attributed to it.
"""
import csv
import subprocess
import sys
from collections import defaultdict


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    agg = defaultdict(lambda: [0, 0, 0, 0, ""])     # (file, line) -> [inst, samples, smem wavefronts, thread inst, source]
    fname, hdr, cur = None, None, None
    total = 0
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Function Name":
            continue
        if r[0] == "Line No":
            hdr = r
            ci = {n: i for i, n in enumerate(hdr) if n not in ("Source",)}
            continue
        if hdr is None:
            continue
        if r[0] != "":
            cur = (fname, int(r[0]))
            agg[cur][4] = r[1].strip()
            continue
        if cur is None:
            continue
        try:
            inst = int(r[ci["Instructions Executed"]])
            smp = int(r[ci["# Samples"]])
            wf = int(r[ci["L1 Wavefronts Shared"]] or 0) if "L1 Wavefronts Shared" in ci else 0
            ti = int(r[ci["Thread Instructions Executed"]])
        except (ValueError, KeyError):
            continue
        a = agg[cur]
        a[0] += inst; a[1] += smp; a[2] += wf; a[3] += ti
        total += inst
    tot_s = sum(a[1] for a in agg.values()) or 1
    print(f"# {rep}: {total} warp instructions, {tot_s} stall samples")
    print(f"# {'file:line':28s} {'inst%':>6s} {'smp%':>6s} {'thr/inst':>8s} {'smem_wf/inst':>12s}  source")
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        if a[0] == 0:
            continue
        print(f"{f + ':' + str(ln):30s} {100 * a[0] / total:6.2f} {100 * a[1] / tot_s:6.2f} {a[3] / a[0]:8.1f} {a[2] / a[0]:12.2f}  {a[4][:110]}")


if __name__ == "__main__":
    main()
