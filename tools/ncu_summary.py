"""Summarise an .ncu-rep (raw page) into the handful of numbers we track per kernel."""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
    "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "sm__cycles_elapsed.avg",
    "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("====", d.get("Kernel Name", "")[:70], "id", d.get("ID"))
    for k in KEYS:
        if k in d:
            print(f"  {k:70s} {d[k]}")
    st = sorted(((float(d[k]), k) for k in hdr if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and d[k]), reverse=True)
    print("  stalls (warps per issue):", ", ".join(f"{k.split('issue_stalled_')[1].split('_per_')[0]}={v:.2f}" for v, k in st[:7]))
