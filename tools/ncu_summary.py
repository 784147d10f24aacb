"""Summarises ncu output for profiles/: (a) a launch list CSV (gpu__time_duration.sum per
launch) -> per-kernel count / total / share; (b) a --set full .ncu-rep -> one row per
captured launch with duration, DRAM bytes, DRAM %, fp64 pipe %, occupancy, registers.
Usage: python tools/ncu_summary.py launches <csv> | full <ncu-rep>"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict


def short(name):
    m = re.search(r"fm_kernel<fm::(\w+)", name)
    if m:
        return m.group(1)
    return re.sub(r"\(.*", "", name).split("::")[-1][:60]


def launches(path):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1.0)
        a = agg[short(r[ki])]
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{'kernel':40s} {'launches':>8s} {'total_us':>12s} {'avg_us':>10s} {'share':>7s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:40s} {a[0]:8d} {a[1]:12.1f} {a[1] / a[0]:10.2f} {a[1] / tot:7.3f}")
    print(f"{'TOTAL':40s} {sum(a[0] for a in agg.values()):8d} {tot:12.1f}")


WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum"]


def full(path):
    out = subprocess.check_output(["ncu", "-i", path, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    ki = hdr.index("Kernel Name")
    cols = [(w, hdr.index(w)) for w in WANT if w in hdr]
    print("kernel | " + " | ".join(f"{w} [{units[i]}]" for w, i in cols))
    for r in rows[2:]:
        print(short(r[ki]) + " | " + " | ".join(r[i] for _, i in cols))


if __name__ == "__main__":
    {"launches": launches, "full": full}[sys.argv[1]](sys.argv[2])
