"""Print selected metrics from `ncu -i X.ncu-rep --page raw --csv` (stdin)."""
import csv, sys
rows = list(csv.reader(sys.stdin))
h = rows[0]
pats = sys.argv[1:] or ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "registers_per_thread", "pipe_fp64", "issue_stalled",
                        "sm__throughput.avg.pct", "warps_active.avg.pct", "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts.sum", "grid_size"]
for v in rows[2:]:
    for k, u, x in zip(h, rows[1], v):
        if any(p in k for p in pats):
            print(f"{k} [{u}] = {x}")
