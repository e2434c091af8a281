#!/usr/bin/env python
"""profiles/r02_traffic.json from an ncu --set full report (run here, no GPU):
  python tools/ncu_traffic.py gpurun_out/r02_prof.ncu-rep profiles/r02_traffic.json
Per kernel: DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum), duration, issue-slot use."""
import csv
import io
import json
import re
import subprocess
import sys


def main():
    rep, out = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}

    def val(r, name):
        v, u = float(r[ix[name]].replace(",", "")), units[ix[name]]
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0,
                 "msecond": 1e-3, "usecond": 1e-6, "nsecond": 1e-9, "second": 1.0}.get(u, 1.0)
        return v * scale
    kernels = {}
    for r in rows[2:]:
        m = re.search(r"(k_[a-z0-9_]+)(<[^>]*>)?", r[ix["Kernel Name"]])
        if not m:
            continue
        name = m.group(1)
        if m.group(2) and m.group(2) not in ("<0>", "<(bool)0>"):
            name += m.group(2).replace("(bool)", "")
        rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
        kernels[name] = {"dram_bytes_per_launch": int(rd + wr), "dram_read": int(rd), "dram_write": int(wr),
                         "ncu_seconds": val(r, "gpu__time_duration.sum"),
                         "issue_active_pct": float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]]),
                         "warp_instructions": int(float(r[ix["smsp__inst_executed.sum"]])),
                         "registers": int(float(r[ix["launch__registers_per_thread"]]))}
    json.dump({"source": rep.split("/")[-1], "note": "one launch each, bench.py --steps 2 --warmup 3 (cfg3, 2 M pairs)",
               "kernels": kernels}, open(out, "w"), indent=1)
    for k, v in kernels.items():
        print(k, v)


if __name__ == "__main__":
    main()
