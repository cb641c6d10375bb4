"""Print the key metrics of an `ncu --page raw --csv` export."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warps_issue_stalled", "launch__occupancy_limit", "sm__throughput.avg.pct", "gpu__dram_throughput.avg.pct",
        "sass__inst_executed_local", "launch__grid_size", "launch__block_size", "smsp__inst_executed_op_shfl", "sm__inst_executed_pipe_fmaheavy", "sm__inst_executed_pipe_fmalite"]
for vals in rows[2:]:
    for h, u, v in zip(hdr, units, vals):
        if any(h.startswith(w) or w in h for w in want) and not any(x in h for x in (".max", ".min", ".sum.pct", "per_second", "realtime")):
            print(f"{h:95s} {u:16s} {v}")
    print("-" * 60)
