"""Developer tool: compact summary of an .ncu-rep (raw page + top stalled source lines)."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
keys = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "smsp__inst_executed.sum",
        "smsp__cycles_active.avg", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
        "launch__grid_size", "launch__block_size"]
idx = {k: hdr.index(k) for k in keys if k in hdr}
for r in rows[2:]:
    print("=" * 100)
    for k, i in idx.items():
        print(f"  {k.split('.')[0] if k.startswith('smsp__average') else k:70s} {units[i]:10s} {r[i][:70]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
seen = set()
for a, b in zip(secs[:-1], secs[1:]):
    name = rows[a][1][:60]
    if name in seen:
        continue
    seen.add(name)
    h = rows[a + 1]
    ci = {x: k for k, x in enumerate(h)}
    data = rows[a + 2:b]
    tot = sum(int(r[ci["# Samples"]] or 0) for r in data)
    print("=" * 100)
    print(name, "total samples", tot, "instructions", len(data))
    for r in sorted(data, key=lambda r: -int(r[ci["# Samples"]] or 0))[:topn]:
        s = int(r[ci["# Samples"]] or 0)
        print(f"{100.0 * s / max(tot, 1):5.1f}% lsb={r[ci['stall_long_sb']]:>5} ssb={r[ci['stall_short_sb']]:>4} bar={r[ci['stall_barrier']]:>4} "
              f"wait={r[ci['stall_wait']]:>4} exec={r[ci['Instructions Executed']]:>8}  {r[ci['Source']][:80]}")
