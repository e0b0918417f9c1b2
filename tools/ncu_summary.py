"""Summarise an .ncu-rep into the `metric,unit,value` csv kept under profiles/ (one block per captured launch).

usage: python tools/ncu_summary.py report.ncu-rep > profiles/rNN_<kernel>_ncu.csv"""
import csv
import io
import subprocess
import sys

KEEP = [
    "dram__bytes_read.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "gpu__time_duration.sum",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__registers_per_thread",
    "lts__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
    "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "l1tex__t_sector_hit_rate.pct", "sm__cycles_elapsed.max", "launch__waves_per_multiprocessor",
    "sm__maximum_warps_per_active_cycle_pct", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units, launches = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(head)}
    for r in launches:
        print(f"Kernel Name,,{r[col['Kernel Name']]}")
        print(f"Block Size,,{r[col['Block Size']]}")
        print(f"Grid Size,,{r[col['Grid Size']]}")
        for k in KEEP:
            if k in col:
                print(f"{k},{units[col[k]]},{r[col[k]]}")
        print()


if __name__ == "__main__":
    main()
