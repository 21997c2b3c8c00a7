#!/usr/bin/env python
"""Turns the raw ncu outputs that gpurun brings back (gpurun_out/, scratch) into the small text summaries committed here.

    python profiles/summarize.py launches gpurun_out/launches_X.csv  > profiles/rNN_launches_X.txt
    python profiles/summarize.py kernel   gpurun_out/prof_X.ncu-rep  > profiles/rNN_kernel_X.txt
    python profiles/summarize.py sass     [quadrants_b200/lib/libqdsp.so] > profiles/sass_census.txt
    python profiles/summarize.py stalls   gpurun_out/prof_X.ncu-rep  > profiles/rNN_stalls_X.txt   (needs --import-source on)
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


MNEMONICS = ["UBLKCP", "SYNCS", "LDGSTS", "LDS", "STS", "LDG", "STG", "LD.E", "ST.E", "ATOM", "RED", "VOTE", "SHFL", "MATCH",
             "BAR", "MEMBAR", "CCTL", "FFMA", "DFMA", "DADD", "DMUL", "NANOSLEEP", "HMMA", "UTCMMA", "LDL", "STL"]


def sass(path="quadrants_b200/lib/libqdsp.so"):
    """Per-kernel SASS census (cuobjdump -sass): which instruction classes each kernel of the shipped library contains.
    UBLKCP = cp.async.bulk (the TMA bulk-copy engine), SYNCS = mbarrier ops; no HMMA / UTCMMA anywhere: the path is
    bandwidth-bound and uses no tensor cores.  LDL / STL > 0 would mean register spills."""
    out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    demangled = {}
    fn, counts, order = None, {}, []
    for line in out.splitlines():
        line = line.strip()
        if line.startswith("Function :"):
            fn = line.split(":", 1)[1].strip()
            counts[fn] = collections.Counter()
            order.append(fn)
        elif fn and line.startswith("/*") and "*/" in line:
            body = line.split("*/", 1)[1].strip()
            if not body or body.startswith("/*"):
                continue
            toks = body.replace("@P", " @P").split()
            op = next((t for t in toks if not t.startswith("@") and not t.startswith("{")), "")
            counts[fn]["_all"] += 1
            for m in MNEMONICS:
                if op == m or op.startswith(m + "."):
                    counts[fn][m] += 1
    names = subprocess.run(["cu++filt"] + order, capture_output=True, text=True).stdout.splitlines() if order else []
    for a, b in zip(order, names):
        demangled[a] = b
    arch = subprocess.run(["cuobjdump", "-lelf", path], capture_output=True, text=True).stdout
    print(f"# {path}: SASS census per kernel ({len(order)} kernels; ELF images: "
          f"{sorted(set(w for w in arch.replace('.', ' ').split() if w.startswith('sm_')))})")
    print(f"{'kernel':100s} {'insts':>7s} " + " ".join(f"{m:>6s}" for m in MNEMONICS))
    tot = collections.Counter()
    for f in order:
        c = counts[f]
        tot.update(c)
        name = demangled.get(f, f).replace("qdsp::", "").replace("(anonymous namespace)::", "")
        name = name.replace("void ", "")[:100]
        print(f"{name:100s} {c['_all']:7d} " + " ".join(f"{c[m]:6d}" for m in MNEMONICS))
    print(f"{'TOTAL':100s} {tot['_all']:7d} " + " ".join(f"{tot[m]:6d}" for m in MNEMONICS))


def stalls(path):
    """Warp-stall samples per SASS line of the first kernel in the report, top lines and totals by stall reason."""
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    print("# " + path + ": " + rows[0][1][:150])
    hdr = rows[1]
    wi, si = hdr.index("Warp Stall Sampling (All Samples)"), hdr.index("Source")
    reasons = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot, by = 0, collections.Counter()
    lines = []
    for n, r in enumerate(rows[2:]):
        try:
            w = int(r[wi])
        except (ValueError, IndexError):
            continue
        tot += w
        top = max(reasons, key=lambda k: int(r[hdr.index(k)] or 0))
        for k in reasons:
            by[k] += int(r[hdr.index(k)] or 0)
        lines.append((w, n, r[si][:70], top))
    print(f"total samples {tot}; by reason: " + ", ".join(f"{k[6:]} {100 * v / max(tot, 1):.1f}%" for k, v in by.most_common(8)))
    for w, n, src, top in sorted(lines, reverse=True)[:40]:
        print(f"  {100 * w / max(tot, 1):5.1f}%  line {n:5d}  {top[6:]:14s} {src}")


if __name__ == "__main__":
    fn = {"launches": launches, "kernel": kernel, "sass": sass, "stalls": stalls}[sys.argv[1]]
    fn(*sys.argv[2:3])
