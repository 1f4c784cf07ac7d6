"""Turns `ncu -i X.ncu-rep --page raw --csv` output into the short per-kernel summary kept
under profiles/ (FP64 and XU pipes, occupancy and what limits it, branch uniformity, DRAM
bytes, shared-memory bank conflicts).  Usage: python tools/ncu_summary.py raw.csv [header]"""
import csv
import sys

METRICS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps",
    "launch__occupancy_limit_blocks", "sm__maximum_warps_per_active_cycle_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__cycles_elapsed.max", "smsp__cycles_active.avg",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed",
    "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_fp64.sum",
    "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__sass_average_branch_targets_threads_uniform.pct",
    "smsp__sass_branch_targets_threads_divergent.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    col = {name: i for i, name in enumerate(hdr)}
    if len(sys.argv) > 2:
        print(sys.argv[2])
        print()
    for r in rows[2:]:
        print("Kernel Name [] = %s" % r[col["Kernel Name"]])
        for m in METRICS:
            if m in col:
                print("%s [%s] = %s" % (m, units[col[m]], r[col[m]]))
        print()


if __name__ == "__main__":
    main()
