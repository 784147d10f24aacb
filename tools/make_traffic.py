#!/usr/bin/env python
"""profiles/traffic.json from ncu --set full captures (.ncu-rep): per kernel the DRAM bytes per launch
(dram__bytes_read.sum + dram__bytes_write.sum), DRAM / fp64-pipe utilisation and the duration under
ncu, stamped with the hash of the kernel sources so that bench.py attaches it only to the code it was
taken on.   python tools/make_traffic.py <tag> <n> <density> <rep> [<rep> ...]"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

NAMES = {"KSweepT<0>": "sweep", "KSweepT<(bool)0>": "sweep", "fm_centroid_update": "centroid_update", "fm_flip_coop": "flip_loop"}


def main():
    tag, n, density, reps = sys.argv[1], int(sys.argv[2]), sys.argv[3], sys.argv[4:]
    kernels = {}
    for rep in reps:
        out = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
        rows = list(csv.reader(io.StringIO(out)))
        hdr = rows[0]
        col = {k: hdr.index(k) for k in ("Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
                                         "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                                         "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
                                         "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread")}
        units = rows[1]
        for r in rows[2:]:
            name = next((v for k, v in NAMES.items() if k in r[col["Kernel Name"]]), None)
            if not name:
                continue
            f = lambda k: float(r[col[k]].replace(",", ""))  # noqa: E731
            scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}
            rd = f("dram__bytes_read.sum") * scale[units[col["dram__bytes_read.sum"]]]
            wr = f("dram__bytes_write.sum") * scale[units[col["dram__bytes_write.sum"]]]
            kernels[name] = {"dram_bytes_per_launch": int(rd + wr), "dram_read_bytes": int(rd), "dram_write_bytes": int(wr),
                             "duration_under_ncu_us": f("gpu__time_duration.sum"),
                             "dram_pct_of_peak": round(f("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), 1),
                             "fp64_pipe_pct": round(f("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"), 1),
                             "warps_active_pct": round(f("sm__warps_active.avg.pct_of_peak_sustained_active"), 1),
                             "registers": int(f("launch__registers_per_thread"))}
    res = {"csrc_sha16": bench.csrc_sha16(), "generators": n, "density": density, "n_gpus": 1,
           "source": f"profiles/{tag}_full_*.txt (ncu --set full --clock-control none --import-source on, FM_NO_GRAPH=1 python "
                     "tools/profile_step.py: a warm Lloyd iteration after 3+ others; per launch: dram__bytes_read.sum + dram__bytes_write.sum)",
           "kernels": kernels}
    with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
        json.dump(res, f, indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
