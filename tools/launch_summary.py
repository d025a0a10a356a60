"""Developer tool: per-kernel summary of an `ncu --metrics gpu__time_duration.sum --csv` launch list.
usage: python tools/launch_summary.py gpurun_out/launches.csv "<command that was profiled>" > profiles/rNN_launches.md"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[h]
ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
d = collections.defaultdict(list)
for r in rows[h + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    v = v / 1000 if r[ui] == "ns" else (v * 1000 if r[ui] == "ms" else v)
    d[r[ki].split("(")[0][:60]].append(v)
tot = sum(sum(v) for v in d.values())
print(f"# ncu launch list: `{sys.argv[2] if len(sys.argv) > 2 else ''}`\n")
print("`ncu --metrics gpu__time_duration.sum --clock-control none` (cold-cache, serialised launches: "
      "shares are comparable with the bench, absolute times are not).\n")
print(f"{sum(len(v) for v in d.values())} launches, {tot / 1000:.2f} ms total\n")
print("| kernel | launches | mean us | share |\n|---|---|---|---|")
for n, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
    print(f"| `{n}` | {len(v)} | {sum(v) / len(v):.1f} | {100 * sum(v) / tot:.1f} % |")
