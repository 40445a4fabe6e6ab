#!/usr/bin/env python
"""scripts/ncu_summary.py -- turns the `ncu --set full` captures of scripts/gpu_profile_r2.sh (their raw pages, gpurun_out/p2_<workload>_raw.csv)
into what the repository keeps: profiles/traffic.json (dram__bytes_read.sum + dram__bytes_write.sum per FRAME and per launch, by
workload -- bench.py's roofline.traffic reads it), profiles/r2_<workload>_raw.csv (the raw metric page) and a one-line digest
per workload on stdout for profiles/r2_kernels.md."""
import csv
import glob
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench


def unit_scale(unit):
    return {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)


def main():
    traffic = {}
    for exported in sorted(glob.glob(os.path.join(ROOT, "gpurun_out", "p2_*_raw.csv"))):
        workload = os.path.basename(exported)[3:-8]
        if workload not in bench.WORKLOADS:
            continue
        raw = open(exported).read()
        if not raw.strip():
            continue
        with open(os.path.join(ROOT, "profiles", "r2_%s_raw.csv" % workload), "w") as f:
            f.write(raw)
        digest = os.path.join(ROOT, "gpurun_out", "p2_%s_digest.txt" % workload)
        if os.path.exists(digest):
            with open(os.path.join(ROOT, "profiles", "r2_%s_digest.txt" % workload), "w") as f:
                f.write(open(digest).read())
        rows = list(csv.reader(raw.splitlines()))
        m = {h: (v, u) for h, u, v in zip(rows[0], rows[1], rows[-1])}
        num = lambda k: float(m[k][0].replace(",", "")) * unit_scale(m[k][1])
        dram = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
        ms = float(m["gpu__time_duration.sum"][0].replace(",", "")) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[m["gpu__time_duration.sum"][1]]
        kernel = re.sub(r"\(.*", "", m["Kernel Name"][0])
        algorithmic = bench.workload_bytes(workload)
        frames = max(1, round(dram / algorithmic))              # a batch launch renders many frames: traffic ~ algorithmic bytes
        traffic[workload] = {"dram_bytes_per_frame": dram / frames, "dram_bytes_per_launch": dram, "frames_per_launch": frames,
                             "algorithmic_bytes_per_frame": algorithmic, "kernel": kernel, "launch_ms": ms,
                             "registers_per_thread": int(float(m["launch__registers_per_thread"][0])),
                             "warps_active_per_sm": float(m["sm__warps_active.avg.per_cycle_active"][0]),
                             "issue_active_pct": float(m["smsp__issue_active.avg.pct_of_peak_sustained_active"][0]),
                             "source": "profiles/r2_%s_raw.csv (ncu --set full, scripts/gpu_profile_r2.sh)" % workload}
        print("%-22s %-28s %8.3f ms  dram %.1f MB/frame vs algorithmic %.1f MB (x%.3f)  regs %d  warps/SM %.1f  issue %.0f%%" % (
            workload, kernel[:28], ms, dram / frames / 1e6, algorithmic / 1e6, dram / frames / algorithmic,
            traffic[workload]["registers_per_thread"], traffic[workload]["warps_active_per_sm"], traffic[workload]["issue_active_pct"]))
    if traffic:
        with open(os.path.join(ROOT, "profiles", "traffic.json"), "w") as f:
            json.dump(traffic, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
