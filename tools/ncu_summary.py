"""Summarise an `ncu --set full` report into the few numbers the design argues with (python tools/ncu_summary.py rep.ncu-rep)."""
import csv
import io
import subprocess
import sys
from collections import Counter

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__cluster", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor", "sm__pipe_tensor",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warps_issue_stalled"]
for i, h in enumerate(hdr):
    if any(h == w or (w.endswith("stalled") and h.startswith(w)) or (w in ("sm__inst_executed_pipe_tensor", "sm__pipe_tensor", "launch__cluster") and h.startswith(w)) for w in want):
        print(f"{h} [{units[i]}]: " + " | ".join(r[i] for r in data))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = rows[1]
i_src, i_ex, i_sm = h.index("Source"), h.index("Instructions Executed"), h.index("# Samples")
c, cs = Counter(), Counter()
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) < 10 or r[0] == "Address":
        continue
    t = r[i_src].split()
    op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
    c[op] += int(r[i_ex]); cs[op] += int(r[i_sm])
tot = sum(c.values())
print(f"warp instructions executed (first launch): {tot}; by opcode: " + ", ".join(f"{k} {v}" for k, v in c.most_common(16)))
print("stall samples by opcode: " + ", ".join(f"{k} {v}" for k, v in cs.most_common(10)))
