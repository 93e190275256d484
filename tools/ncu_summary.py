#!/usr/bin/env python
"""Condense `ncu --set full` reports (.ncu-rep) into the per-launch CSV kept under profiles/.

    python tools/ncu_summary.py "title line" out.csv report1.ncu-rep [report2.ncu-rep ...]

Reads each report with `ncu -i <rep> --page raw --csv` (works without a GPU) and keeps the columns
the roofline discussion in DESIGN.md uses."""
import csv
import io
import subprocess
import sys

KEEP = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__block_size", "launch__grid_size",
        "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]


def main():
    title, out, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
    lines = []
    header_done = False
    for rep in reps:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        hdr, units = rows[0], rows[1]
        idx = [hdr.index(k) for k in KEEP if k in hdr]
        if not header_done:
            lines.append([hdr[i] for i in idx] + ["report"])
            lines.append([units[i] for i in idx] + [""])
            header_done = True
        for r in rows[2:]:
            lines.append([r[i][:60] if hdr[i] == "Kernel Name" else r[i] for i in idx] + [rep.split("/")[-1]])
    with open(out, "w", newline="") as f:
        f.write("# " + title + "\n")
        csv.writer(f).writerows(lines)
    print("wrote", out, len(lines) - 2, "launches")


if __name__ == "__main__":
    main()
