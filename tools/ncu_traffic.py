#!/usr/bin/env python
"""Per-kernel table of one or more `ncu --set full` captures of bench.py, and profiles/traffic.json -- the measured
DRAM bytes per launch that bench.py reports as `roofline.traffic` (only while the kernel sources still hash to the
build the capture was made on: bench.kernel_source_hash).

    python tools/ncu_traffic.py --config 2 --out-md profiles/r02_ncu_kernels.md gpurun_out/fwd.ncu-rep gpurun_out/bwd.ncu-rep

Runs here (no GPU): `ncu -i <rep> --page raw --csv` only reads the report.  Per kernel name the launches in the
reports are averaged; every figure is per launch."""
import argparse
import collections
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

M = {
    "t_us": "gpu__time_duration.sum",
    "rd": "dram__bytes_read.sum",
    "wr": "dram__bytes_write.sum",
    "dram_pct": "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "tensor_pct": "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "warps_pct": "sm__warps_active.avg.pct_of_peak_sustained_active",
    "regs": "launch__registers_per_thread",
    "inst": "smsp__inst_executed.sum",
    "issue_pct": "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "l2_hit": "lts__t_sector_hit_rate.pct",
    "lsb": "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "bar": "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "grid": "launch__grid_size",
}
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6,
         "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3, "second": 1e6}


def read(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        raise SystemExit(f"{rep}: no launches in the report")
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        rec = {"name": r[hdr.index("Kernel Name")].split("(")[0]}
        for k, col in M.items():
            if col not in hdr:
                rec[k] = None
                continue
            i = hdr.index(col)
            try:
                v = float(r[i].replace(",", ""))
            except ValueError:
                rec[k] = None
                continue
            rec[k] = v * SCALE.get(units[i], 1.0)
        out.append(rec)
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("reps", nargs="+")
    ap.add_argument("--config", type=int, default=2)
    ap.add_argument("--out-md", default=None)
    ap.add_argument("--traffic-json", default=os.path.join(ROOT, "profiles", "traffic.json"))
    ap.add_argument("--command", default="python bench.py --steps 1 --warmup 3 --cpu-seconds 0 --no-partitioned --no-cuda-graph")
    args = ap.parse_args()
    import bench
    recs = [r for rep in args.reps for r in read(rep)]
    by = collections.OrderedDict()
    for r in recs:
        by.setdefault(r["name"], []).append(r)

    def mean(rs, k):
        v = [r[k] for r in rs if r[k] is not None]
        return sum(v) / len(v) if v else float("nan")
    lines = [f"# ncu --set full, per kernel (config {args.config}; `{args.command}`)\n",
             f"kernel source hash `{bench.kernel_source_hash()}`; reports: " + ", ".join(os.path.basename(r) for r in args.reps),
             "\nPer launch, averaged over the captured launches.  ncu serialises launches and replays each one ~40 times "
             "(cold L2 on every pass): durations are longer than in the bench; DRAM bytes, tensor-pipe and issue "
             "utilisation are what this table is for.\n",
             "| kernel | launches | us | DRAM read MB | DRAM write MB | DRAM % | tensor pipe % | issue slots % | warps active % | "
             "L2 hit % | long-scoreboard / issue | barrier / issue | regs | grid |",
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    traffic = {}
    for name, rs in by.items():
        lines.append(f"| `{name}` | {len(rs)} | {mean(rs, 't_us'):.1f} | {mean(rs, 'rd') / 1e6:.1f} | {mean(rs, 'wr') / 1e6:.1f} | "
                     f"{mean(rs, 'dram_pct'):.1f} | {mean(rs, 'tensor_pct'):.1f} | {mean(rs, 'issue_pct'):.1f} | "
                     f"{mean(rs, 'warps_pct'):.1f} | {mean(rs, 'l2_hit'):.1f} | {mean(rs, 'lsb'):.2f} | {mean(rs, 'bar'):.2f} | "
                     f"{mean(rs, 'regs'):.0f} | {mean(rs, 'grid'):.0f} |")
        key = name.replace("void ", "").split("<")[0]
        rec = {"dram_bytes_per_launch": int(mean(rs, "rd") + mean(rs, "wr")), "launches": len(rs),
               "us_under_ncu": round(mean(rs, "t_us"), 1), "dram_pct": round(mean(rs, "dram_pct"), 1),
               "tensor_pct": round(mean(rs, "tensor_pct"), 1),
               "source": f"ncu --set full, {', '.join(os.path.basename(r) for r in args.reps)} "
                         f"({os.path.basename(args.out_md) if args.out_md else 'profiles/'})"}
        # several template instances share a key: keep the one with more launches (the processor-loop instance)
        if key not in traffic or traffic[key][f"config{args.config}"]["launches"] < len(rs):
            traffic[key] = {f"config{args.config}": rec}
    text = "\n".join(lines) + "\n"
    if args.out_md:
        with open(args.out_md, "w") as f:
            f.write(text)
    sys.stdout.write(text)
    tj = {"source_hash": bench.kernel_source_hash(), "kernels": {}}
    if os.path.exists(args.traffic_json):
        try:
            old = json.load(open(args.traffic_json))
            if old.get("source_hash") == tj["source_hash"]:
                tj = old
        except Exception:
            pass
    for k, v in traffic.items():
        tj["kernels"].setdefault(k, {}).update(v)
    with open(args.traffic_json, "w") as f:
        json.dump(tj, f, indent=1)
    print(f"wrote {args.traffic_json} ({len(tj['kernels'])} kernels, build {tj['source_hash']})", file=sys.stderr)


if __name__ == "__main__":
    main()
