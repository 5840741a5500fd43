"""Headline counters of the first kernel in an ncu report (development aid)."""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
h, u, v = rows[0], rows[1], rows[2]
want = ["gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct", "smsp__inst_executed.sum ",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__warps_eligible.avg.per_cycle_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum ", "dram__bytes_write.sum ",
        "lts__t_bytes.sum ", "l1tex__t_bytes.sum ", "sm__inst_executed_pipe_fma", "sm__inst_executed_pipe_alu",
        "sm__pipe_fma_cycles_active.avg.pct", "sm__pipe_alu_cycles_active.avg.pct", "lts__t_sectors.sum ", "sm__cycles_elapsed.avg ",
        "lts__throughput.avg.pct", "l1tex__throughput.avg.pct", "sm__throughput.avg.pct", "smsp__inst_executed.sum"]
for i, k in enumerate(h):
    kk = k + " "
    if any(kk.startswith(w) or k == w.strip() for w in want) or ("issue_stalled" in k and k.endswith("per_issue_active.ratio")):
        try:
            if "issue_stalled" in k and float(v[i]) < 0.15: continue
        except ValueError: pass
        print(f"{k:80s} {u[i]:12s} {v[i]}")
