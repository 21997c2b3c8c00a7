#!/usr/bin/env python
"""Turns the raw ncu outputs that gpurun brings back (gpurun_out/, scratch) into the small text summaries committed here.

    python profiles/summarize.py launches gpurun_out/launches_X.csv  > profiles/rNN_launches_X.txt
    python profiles/summarize.py kernel   gpurun_out/prof_X.ncu-rep  > profiles/rNN_kernel_X.txt
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_static",
    "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "smsp__inst_executed.sum",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
]


def launches(path):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr = rows[h]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    unit = "ns"
    for r in rows[h + 1:]:
        if len(r) <= vi:
            continue
        agg.setdefault(r[ki].split("(")[0], []).append(float(r[vi].replace(",", "")))
        unit = r[ui]
    total = sum(sum(v) for v in agg.values())
    print(f"# {path}: gpu__time_duration.sum per launch ({unit}); cold-cache, serialised -- compare shares")
    print(f"{'kernel':90s} {'n':>5s} {'avg':>12s} {'min':>12s} {'max':>12s} {'total':>14s} {'share':>7s}")
    for k, v in agg.items():
        print(f"{k[:90]:90s} {len(v):5d} {sum(v)/len(v):12.1f} {min(v):12.1f} {max(v):12.1f} {sum(v):14.1f} {100*sum(v)/total:6.2f}%")


def kernel(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {path}: ncu --set full, selected metrics per captured launch")
    for r in rows[2:]:
        print("== " + r[hdr.index("Kernel Name")][:140])
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"  {k:90s} {r[i]:>18s} {units[i]}")
        if "dram__bytes_read.sum" in hdr:
            def b(name):
                i = hdr.index(name)
                mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[units[i]]
                return float(r[i]) * mult
            print(f"  {'traffic = dram read + write (bytes)':90s} {b('dram__bytes_read.sum') + b('dram__bytes_write.sum'):18.0f}")


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2])
