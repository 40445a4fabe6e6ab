#!/usr/bin/env python
"""scripts/ncu_digest.py REPORT.ncu-rep [top] -- the numbers of one `ncu --set full` capture that the profiles/*.md
notes quote: duration, DRAM bytes and throughput, issue rate, pipes, stall reasons per issue, shared-memory conflicts,
and the `top` SASS instructions by stall samples (with -lineinfo, their source lines)."""
import csv
import subprocess
import sys


def page(report, name, extra=()):
    out = subprocess.run(["ncu", "-i", report, "--page", name, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(out.splitlines()))


def main():
    report = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    rows = page(report, "raw")
    hdr, units, vals = rows[0], rows[1], rows[-1]
    m = {h: (v, u) for h, u, v in zip(hdr, units, vals)}

    def num(k):
        try:
            return float(m[k][0].replace(",", ""))
        except Exception:
            return float("nan")

    print("kernel:", m.get("Kernel Name", ("?",))[0][:110])
    print("grid %s x block %s, regs/thread %s, warps active/SM %.1f" % (m["launch__grid_size"][0], m["launch__block_size"][0],
          m["launch__registers_per_thread"][0], num("sm__warps_active.avg.per_cycle_active")))
    dur = num("gpu__time_duration.sum")
    dur_unit = m["gpu__time_duration.sum"][1]
    rd, wr = num("dram__bytes_read.sum"), num("dram__bytes_write.sum")
    print("duration %.3f %s; dram read %.3f %s write %.3f %s; dram throughput %.1f %% of peak" % (dur, dur_unit, rd, m["dram__bytes_read.sum"][1],
          wr, m["dram__bytes_write.sum"][1], num("dram__throughput.avg.pct_of_peak_sustained_elapsed")))
    print("instructions %.4g; issue active %.1f %%; fma pipe %.1f %%; lsu pipe %.1f %%; eligible warps/scheduler %.2f" % (
          num("smsp__inst_executed.sum"), num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
          num("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active"), num("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
          num("smsp__warps_eligible.avg.per_cycle_active")))
    print("shared wavefronts %.4g of which bank conflicts %.4g; local loads %.4g stores %.4g" % (
          num("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"), num("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
          num("smsp__inst_executed_op_local_ld.sum"), num("smsp__inst_executed_op_local_st.sum")))
    stalls = []
    for h in hdr:
        if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
            stalls.append((num(h), h.split("issue_stalled_")[1].split("_per_issue")[0]))
    print("stalls per issue:", ", ".join("%s %.2f" % (n, v) for v, n in sorted(stalls, reverse=True)[:7]))

    src = page(report, "source", ("--print-source", "cuda,sass")) if False else page(report, "source")
    hdr = src[1]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in src[2:] if len(r) == len(hdr)]
    total = sum(int(r[ix["# Samples"]]) for r in data) or 1
    order = sorted(range(len(data)), key=lambda i: -int(data[i][ix["# Samples"]]))[:top]
    print("top %d of %d SASS instructions by stall samples (%d samples):" % (top, len(data), total))
    for i in sorted(order):
        r = data[i]
        reasons = sorted(((int(r[ix[k]]), k[6:]) for k in ix if k.startswith("stall_") and "Not Issued" not in k and r[ix[k]].isdigit()), reverse=True)[:2]
        print("  #%-5d %5.2f%%  exec %-10s %-14s %s" % (i, 100.0 * int(r[ix["# Samples"]]) / total, r[ix["Instructions Executed"]],
              ",".join("%s" % n for v, n in reasons if v), r[ix["Source"]].strip()[:80]))


if __name__ == "__main__":
    main()
