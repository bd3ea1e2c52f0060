"""Turn `ncu -i X.ncu-rep --page raw --csv` (stdin) into the small summary CSV committed under profiles/: header row,
units row, one row per captured launch, a fixed set of columns + `launch__steps` (given on the command line)."""
import csv, sys
steps = [int(x) for x in sys.argv[1].split(",")]
rows = list(csv.reader(sys.stdin))
hdr = rows[0]
keep = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_bytes.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "sm__inst_executed_pipe_fma.sum",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]
cols = [hdr.index(k) for k in keep if k in hdr]
w = csv.writer(sys.stdout)
w.writerow([hdr[c] for c in cols] + ["launch__steps"])
w.writerow([rows[1][c] for c in cols] + ["steps"])
for i, r in enumerate(rows[2:]):
    w.writerow([r[c] for c in cols] + [steps[i] if i < len(steps) else ""])
