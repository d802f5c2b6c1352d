#!/usr/bin/env python3
"""Turns an .ncu-rep (brought back in gpurun_out/) into the small text summaries committed here.

usage: python profiles/summarize.py gpurun_out/prof.ncu-rep profiles/r1_<name>
writes <out>_raw.csv (every metric of every captured launch, one metric per line) and
<out>_summary.md (the handful of numbers DESIGN.md / bench.py quote).
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__waves_per_multiprocessor", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fp64.sum",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    txt = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    with open(out + "_raw.csv", "w") as fh:
        w = csv.writer(fh)
        w.writerow(["launch", "metric", "unit", "value"])
        for li, r in enumerate(data):
            for h, u, v in zip(hdr, units, r):
                w.writerow([li, h, u, v])
    with open(out + "_summary.md", "w") as fh:
        fh.write("# ncu summary: %s\n\n`ncu --set full --clock-control none --import-source on`, "
                 "%d launch(es) captured. Cold-cache, serialised: compare shares, not absolutes.\n\n"
                 % (rep.split("/")[-1], len(data)))
        for li, r in enumerate(data):
            name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
            fh.write("## launch %d: `%s`\n\n| metric | value | unit |\n|---|---|---|\n" % (li, name))
            for k in KEYS:
                if k in hdr:
                    i = hdr.index(k)
                    fh.write("| %s | %s | %s |\n" % (k, r[i], units[i]))
            fh.write("\n")
    print("wrote", out + "_raw.csv", out + "_summary.md")


if __name__ == "__main__":
    main()
