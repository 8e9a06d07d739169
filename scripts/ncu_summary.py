#!/usr/bin/env python
"""Condenses the ncu --set full captures under gpurun_out/ into profiles/<round>_ncu_full_summary.csv
(capture, kernel, metric, unit, value): the metrics profiles/README.md quotes, per captured launch."""
import csv
import io
import os
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
    "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__cycles_elapsed.avg.per_second",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
]


def main():
    src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
    out = sys.argv[2] if len(sys.argv) > 2 else "profiles/r01_ncu_full_summary.csv"
    caps = sys.argv[3].split(",") if len(sys.argv) > 3 else ["tc2_search_full", "helpers_full", "kmodes_full", "smooth_full"]
    rows = [("capture", "kernel", "metric", "unit", "value")]
    for cap in caps:
        rep = os.path.join(src, cap + ".ncu-rep")
        if not os.path.exists(rep):
            continue
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        r = list(csv.reader(io.StringIO(txt)))
        hdr, units = r[0], r[1]
        kcol = hdr.index("Kernel Name")
        for launch in r[2:]:
            kname = launch[kcol].split("(")[0].replace("void <unnamed>::", "")
            for h, u, v in zip(hdr, units, launch):
                if any(h == w or h.endswith("." + w) for w in WANT):
                    rows.append((cap, kname, h, u, v))
    with open(out, "w", newline="") as f:
        csv.writer(f).writerows(rows)
    print(out, len(rows) - 1, "rows")


if __name__ == "__main__":
    main()
