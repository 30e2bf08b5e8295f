"""Summarise an ncu report of the fused stage kernel: python tools/ncu_summary.py <file.ncu-rep> [kernel regex].

Reads `ncu -i <rep> --page raw --csv` and prints the handful of metrics DESIGN.md argues with, per captured launch."""
import csv
import io
import re
import subprocess
import sys

KEEP = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum",
    "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_sector_hit_rate.pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread",
    "launch__cluster_size",
    "launch__grid_size",
    "launch__block_size",
    "launch__shared_mem_per_block_dynamic",
    "launch__shared_mem_per_block_static",
    "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "launch__occupancy_limit_shared_mem",
    "launch__waves_per_multiprocessor",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "sm__icc_request_hit_rate.pct",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_membar_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_active_per_inst_executed.ratio",
]


def main():
    rep = sys.argv[1]
    pat = re.compile(sys.argv[2] if len(sys.argv) > 2 else "fusion_stage")
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, body = rows[0], rows[1], rows[2:]
    col = {n: i for i, n in enumerate(head)}
    n = 0
    for r in body:
        name = r[col["Kernel Name"]]
        if not pat.search(name):
            continue
        print("launch %d: %s" % (n, name))
        n += 1
        for k in KEEP:
            if k in col:
                print("   %-70s %s %s" % (k, r[col[k]], units[col[k]]))


if __name__ == "__main__":
    main()
