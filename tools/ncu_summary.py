#!/usr/bin/env python
"""Condense an .ncu-rep into the handful of numbers the roofline discussion needs.

  python tools/ncu_summary.py gpurun_out/prof.ncu-rep [> profiles/<name>.txt]
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__shared_mem_per_block_static", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__inst_executed.avg.per_cycle_elapsed",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
]


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    print(f"# {rep}: {len(rows) - 2} profiled launch(es); ncu --set full --clock-control none")
    for r in rows[2:]:
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                print(f"{k:95s} {r[i]} {units[i]}")
        try:
            t = float(r[hdr.index("gpu__time_duration.sum")].replace(",", ""))
            rd = float(r[hdr.index("dram__bytes_read.sum")].replace(",", ""))
            wr = float(r[hdr.index("dram__bytes_write.sum")].replace(",", ""))
            ut, ub = units[hdr.index("gpu__time_duration.sum")], units[hdr.index("dram__bytes_read.sum")]
            scale_t = {"ns": 1e-9, "us": 1e-6, "ms": 1e-3, "s": 1.0}.get(ut.replace("second", "s").replace("n", "n"), None)
            scale_b = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(ub, None)
            if scale_t and scale_b:
                print(f"{'derived: DRAM traffic / duration':95s} {(rd + wr) * scale_b / (t * scale_t) / 1e9:.1f} GB/s")
        except Exception:
            pass
        print("-" * 60)


if __name__ == "__main__":
    main()
