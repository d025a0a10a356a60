#!/usr/bin/env python
"""Developer tool: warp-stall samples of one kernel in an .ncu-rep attributed to CUDA source lines
(`ncu --page source --print-source cuda,sass`; needs -lineinfo builds).  usage: ncu_lines.py rep kernel-regex [top]"""
import csv
import io
import os
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "-k", "regex:" + kre],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
recs, fname, hdr, seen_fn = [], None, None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        fname = os.path.basename(r[1]); continue
    if r[0] == "Function Name":
        if seen_fn is None:
            seen_fn = r[1]
        cur_fn = r[1]; continue
    if r[0] == "Line No":
        hdr = r; continue
    if hdr is None or r[0] == "" or cur_fn != seen_fn:      # first captured launch of the kernel only
        continue
    ci = {}
    for k, x in enumerate(hdr):
        ci.setdefault(x, k)
    try:
        s = int(r[ci["# Samples"]] or 0)
    except ValueError:
        continue
    g = lambda n: int(r[ci[n]] or 0) if n in ci else 0
    recs.append((s, fname, r[0], r[1].strip()[:110], g("stall_long_sb"), g("stall_barrier"), g("stall_short_sb"),
                 g("stall_wait"), g("stall_mio"), g("stall_lg"), g("stall_math"), g("stall_selected") + g("stall_not_selected"),
                 g("Instructions Executed")))
tot = sum(r[0] for r in recs)
print(f"{seen_fn}: {tot} samples over {len(recs)} source lines")
print("  %   samples  long_sb barrier short_sb wait  mio   lg  math  issue   inst   file:line  source")
for r in sorted(recs, key=lambda r: -r[0])[:top]:
    print(f"{100.0 * r[0] / max(tot, 1):5.1f} {r[0]:7d} {r[4]:7d} {r[5]:7d} {r[6]:7d} {r[7]:5d} {r[8]:5d} {r[9]:4d} {r[10]:5d} {r[11]:6d} {r[12]:7d}  {r[1]}:{r[2]}  {r[3]}")
