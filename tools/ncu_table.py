"""Print a compact per-launch table from an .ncu-rep (raw page CSV on stdin)."""
import csv
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_xu.sum", "sm__inst_executed_pipe_alu.sum",
        "sm__inst_executed_pipe_lsu.sum", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.sum"]
rows = list(csv.reader(sys.stdin))
h = rows[0]
idx = {n: i for i, n in enumerate(h)}
cols = [k for k in KEEP if k in idx]
short = [c.split(".")[0].replace("sm__inst_executed_pipe_", "pipe_").replace("pct_of_peak", "%") for c in cols]
print("kernel | " + " | ".join(short))
print("units  | " + " | ".join(rows[1][idx[c]] for c in cols))
for r in rows[2:]:
    print(r[idx["Kernel Name"]][:36] + " | " + " | ".join(r[idx[c]] for c in cols))
if len(sys.argv) > 1:
    pat = sys.argv[1]
    print([n for n in h if pat in n])
