#!/usr/bin/env python3
"""Summarise ncu outputs brought back in gpurun_out/ into profiles/ (tracked).

  python tools/prof_summary.py launches gpurun_out/launches_X.csv  > profiles/X_launches.txt
  python tools/prof_summary.py raw      gpurun_out/prof_X.ncu-rep   > profiles/X_kernels.txt
"""
import collections
import csv
import io
import subprocess
import sys

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors.sum",
    "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__cycles_active.avg", "sm__cycles_elapsed.avg.per_second",
]


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = r[ki].split("(")[0]
        if len(name) > 70:
            name = name[:67] + "..."
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui] in ("us", "usecond") else v
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {len(rows) - 1} launches, {tot:.3f} ms total (ncu-serialised, cold cache: compare SHARES)")
    print(f"{'kernel':72s} {'n':>5s} {'ms':>10s} {'ms/launch':>10s} {'share':>7s}")
    for n, (c, t) in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{n:72s} {c:5d} {t:10.3f} {t / c:10.4f} {100 * t / tot:6.1f}%")


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}: ncu --set full --clock-control none, per launch")
    for r in rows[2:]:
        print(f"\n== {r[hdr.index('Kernel Name')]}  (launch id {r[0]})")
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                print(f"  {m:68s} {r[i]:>22s} {units[i]}")


if __name__ == "__main__":
    {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
