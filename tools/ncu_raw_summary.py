"""`ncu -i x.ncu-rep --page raw --csv` -> small CSV with the metrics the roofline discussion uses, the last (warm) launch of every
(kernel, grid):   python tools/ncu_raw_summary.py gpurun_out/x_raw.csv profiles/x_ncu_full.csv "header comment" """
import csv, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio"]
rows = list(csv.reader(open(sys.argv[1])))
h, u = rows[0], rows[1]
ki, gi = h.index("Kernel Name"), h.index("launch__grid_size")
last = {}
for i, r in enumerate(rows[2:]):
    last[(r[ki], r[gi])] = i
keep = [r for i, r in enumerate(rows[2:]) if last[(r[ki], r[gi])] == i]
with open(sys.argv[2], "w") as f:
    f.write("# %s\n" % (sys.argv[3] if len(sys.argv) > 3 else "ncu --set full --clock-control none"))
    f.write("metric,unit," + ",".join("launch%d" % i for i in range(len(keep))) + "\n")
    f.write("Kernel Name,," + ",".join('"%s"' % r[ki][:70] for r in keep) + "\n")
    for m in WANT:
        if m in h:
            i = h.index(m)
            f.write("%s,%s,%s\n" % (m, u[i], ",".join('"%s"' % r[i] for r in keep)))
