#!/usr/bin/env python
"""Key figures of an .ncu-rep (first kernel): speed-of-light, occupancy, stall samples, DRAM bytes,
hottest SASS lines.  Usage: ncu_brief.py report.ncu-rep [--sass]"""
import csv, io, subprocess, sys


def page(rep, *args):
    out = subprocess.run(["ncu", "-i", rep, "--csv", *args], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    r = page(rep, "--page", "raw")
    h, v = r[0], r[2]
    d = {}
    for k, x in zip(h, v):
        try:
            d[k] = float(x.replace(",", ""))
        except ValueError:
            pass
    keys = ["gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
            "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed.avg.per_cycle_active",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
            "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__cycles_active.avg",
            "smsp__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_fp64.sum",
            "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
            "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers"]
    for k in keys:
        if k in d:
            print(f"{k:70s} {d[k]:.6g}")
    stalls = sorted(((x, k.split("issue_stalled_")[1]) for k, x in d.items()
                     if "pcsamp_warps_issue_stalled" in k and "not_issued" not in k and x > 0), reverse=True)
    tot = sum(x for x, _ in stalls)
    print("stall samples:", ", ".join(f"{k} {100 * x / tot:.0f}%" for x, k in stalls[:8]))
    for k in sorted(d):
        if "pipe" in k and "pct_of_peak_sustained_active" in k and d[k] > 15:
            print(f"{k:70s} {d[k]:.4g}")
    if "--sass" in sys.argv:
        s = page(rep, "--page", "source", "--print-source", "sass")
        hh = s[1]
        rows = [dict(zip(hh, x)) for x in s[2:]]
        tot_s = sum(int(x["Warp Stall Sampling (All Samples)"] or 0) for x in rows)
        for i, x in enumerate(rows):
            st = int(x["Warp Stall Sampling (All Samples)"] or 0)
            if st > 0.012 * tot_s:
                print(i, f"{100 * st / tot_s:5.1f}%", x["Source"][:100])


if __name__ == "__main__":
    main()
