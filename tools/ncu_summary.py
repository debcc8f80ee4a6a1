"""Condense an `ncu -i <rep> --page raw --csv` dump into the per-launch summary committed under profiles/.

    python tools/ncu_summary.py gpurun_out/r02_full_raw.csv profiles/r02_ncu_full_summary.csv
"""
import csv
import sys

COLS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "sm__inst_issued.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.per_cycle_active", "sm__cycles_elapsed.avg", "sm__cycles_active.avg",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
]


def main(src, dst):
    rows = list(csv.reader(open(src)))
    h = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    hdr, units = rows[h], rows[h + 1]
    ki = hdr.index("Kernel Name")
    have = [c for c in COLS if c in hdr]
    idx = [hdr.index(c) for c in have]
    with open(dst, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["ID", "Kernel Name"] + have)
        w.writerow(["", ""] + [units[i] for i in idx])
        for r in rows[h + 2:]:
            if len(r) <= max(idx):
                continue
            w.writerow([r[0], r[ki].split("(")[0]] + [r[i] for i in idx])
    missing = [c for c in COLS if c not in hdr]
    if missing:
        print("not in the dump:", missing, file=sys.stderr)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
