#!/usr/bin/env python
"""Summarise .ncu-rep files (raw page) into a markdown table: python benchmarks/ncu_summary.py gpurun_out/ncu_*.ncu-rep"""
import csv
import io
import subprocess
import sys

KEYS = [("gpu__time_duration.sum", "time us"), ("dram__bytes_read.sum", "DRAM rd MB"), ("dram__bytes_write.sum", "DRAM wr MB"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps %"), ("launch__registers_per_thread", "regs"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem conflicts")]
STALLS = ["barrier", "branch_resolving", "long_scoreboard", "short_scoreboard", "math_pipe_throttle", "mio_throttle",
          "no_instruction", "wait", "sleeping", "membar", "lg_throttle", "dispatch_stall", "not_selected"]
print("| report | kernel | " + " | ".join(k[1] for k in KEYS) + " | top stalls (per issue) |")
print("|---|---|" + "---|" * (len(KEYS) + 1))
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    if len(rows) < 3:
        continue
    hdr = rows[0]
    seen = set()
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")].split("(")[0][-40:]
        if name in seen:
            continue
        seen.add(name)
        vals = []
        for k, _ in KEYS:
            vals.append(next((r[i] for i, h in enumerate(hdr) if h == k), "-"))
        st = []
        for s in STALLS:
            h = f"smsp__average_warps_issue_stalled_{s}_per_issue_active.ratio"
            if h in hdr:
                try:
                    st.append((float(r[hdr.index(h)]), s))
                except ValueError:
                    pass
        st.sort(reverse=True)
        print(f"| {path.split('/')[-1]} | {name} | " + " | ".join(vals) + " | " + ", ".join(f"{s} {v:.2f}" for v, s in st[:4]) + " |")
