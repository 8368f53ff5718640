"""Text summary of one kernel of an .ncu-rep (raw page), for profiles/.  usage: ncu_summary.py rep "header line" [extra metric substrings]"""
import csv, subprocess, sys
KEEP = ["dram__bytes_read.sum ", "dram__bytes_write.sum ", "gpu__time_duration.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "launch__block_size", "launch__grid_size", "launch__registers_per_thread ", "sm__cycles_elapsed.max",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum ", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active", "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active", "smsp__average_warps_issue_stalled_not_selected_per_issue_active",
        "smsp__average_warps_issue_stalled_wait_per_issue_active", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active",
        "smsp__average_warps_issue_stalled_membar_per_issue_active", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active"]
rep, header = sys.argv[1], sys.argv[2]
keep = KEEP + sys.argv[3:]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units, vals = rows[0], rows[1], rows[2]
print(header)
for h, u, v in zip(hdr, units, vals):
    if h == "Kernel Name" or any((h + " ").startswith(k) or h == k.strip() for k in keep):
        print(f"{h} [{u}] = {v}")
