#!/usr/bin/env python
"""Short per-kernel digest of an ncu report: time, instructions, issue rate, pipes above 15 %, stall reasons above 0.1.
  python tools/ncu_brief.py gpurun_out/x.ncu-rep [kernel substring]"""
import csv
import io
import subprocess
import sys

rep, pat = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else "")
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, kn = rows[0], rows[0].index("Kernel Name")
base = ["gpu__time_duration.sum", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__average_warp_latency_per_inst_issued.ratio"]
seen = set()
for r in rows[2:]:
    name = r[kn].split("(")[0]
    if pat not in name or name in seen:
        continue
    seen.add(name)
    print("==", name)
    for w in base:
        if w in hdr:
            print("   %-78s %s" % (w, r[hdr.index(w)]))
    for h, v in zip(hdr, r):
        try:
            f = float(v.replace(",", ""))
        except ValueError:
            continue
        if "average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio") and f >= 0.1:
            print("   stall %-72s %.2f" % (h.split("stalled_")[1].split("_per_issue")[0], f))
        if h.startswith("sm__inst_executed_pipe_") and h.endswith(".avg.pct_of_peak_sustained_active") and f >= 15:
            print("   pipe  %-72s %.1f" % (h.split("pipe_")[1].split(".")[0], f))
