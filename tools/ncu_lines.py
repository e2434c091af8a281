#!/usr/bin/env python
"""Aggregate an ncu report's per-SASS-instruction counters by CUDA source line.

  python tools/ncu_lines.py gpurun_out/prof.ncu-rep k_compact [--top 30] [--lib quade_b200/libquade_b200.so]

ncu's source page is per SASS instruction; the line table comes from `nvdisasm -g` of the cubin inside the
in-tree library (same build as the profiled one), matched by instruction order inside the kernel."""
import argparse
import csv
import glob
import io
import os
import re
import subprocess
import tempfile


def sass_lines(lib, kernel):
    tmp = tempfile.mkdtemp(prefix="cubin_")
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, check=True, stdout=subprocess.DEVNULL)
    out = []
    for cubin in glob.glob(os.path.join(tmp, "*sm_100a*.cubin")):
        txt = subprocess.run(["nvdisasm", "-g", "-c", cubin], capture_output=True, text=True).stdout
        cur_fn, line = None, ("?", 0)
        for l in txt.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", l)
            if m:
                cur_fn = m.group(1)
                continue
            m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
            if m:
                line = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            if cur_fn and kernel in cur_fn and re.match(r"\s*/\*[0-9a-f]{4,}\*/", l):
                out.append((line, l.split("*/", 1)[1].strip()))
        if out:
            break
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("rep")
    ap.add_argument("kernel")
    ap.add_argument("--top", type=int, default=30)
    ap.add_argument("--lib", default="quade_b200/libquade_b200.so")
    ap.add_argument("--launch", type=int, default=0, help="which matching launch in the report")
    ap.add_argument("--by", default="inst", choices=["inst", "stall"])
    ap.add_argument("--section", default=None, help="substring of the mangled name in the cubin (template instances)")
    a = ap.parse_args()
    txt = subprocess.run(["ncu", "-i", a.rep, "--page", "source", "--csv", "--kernel-name", "regex:" + a.kernel],
                         capture_output=True, text=True).stdout
    blocks, cur = [], None
    for r in csv.reader(io.StringIO(txt)):
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "hdr": None, "rows": []}
            blocks.append(cur)
        elif cur is not None and cur["hdr"] is None and r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and cur["hdr"] and len(r) >= len(cur["hdr"]) - 2 and r[0].startswith("0x"):
            cur["rows"].append(r)
    b = blocks[a.launch]
    h = b["hdr"]
    ie, ns, src = h.index("Instructions Executed"), h.index("# Samples"), h.index("Source")
    lines = sass_lines(a.lib, a.section or a.kernel)
    if len(lines) != len(b["rows"]):
        print("warning: %d SASS instructions in the library vs %d in the report" % (len(lines), len(b["rows"])))
    agg, tot_i, tot_s = {}, 0.0, 0.0
    for i, r in enumerate(b["rows"]):
        key = lines[i][0] if i < len(lines) else ("?", 0)
        x = agg.setdefault(key, [0.0, 0.0])
        x[0] += float(r[ie] or 0)
        x[1] += float(r[ns] or 0)
        tot_i += float(r[ie] or 0)
        tot_s += float(r[ns] or 0)
    print("%s: %d SASS instructions, %.0f warp-instructions executed, %.0f samples" % (b["name"], len(b["rows"]), tot_i, tot_s))
    srcs = {}
    for (f, n), (i_, s_) in sorted(agg.items(), key=lambda kv: -kv[1][0 if a.by == "inst" else 1])[:a.top]:
        if f not in srcs:
            p = os.path.join("quade_b200", "csrc", f)
            srcs[f] = open(p).read().splitlines() if os.path.exists(p) else []
        text = srcs[f][n - 1].strip()[:100] if 0 < n <= len(srcs[f]) else ""
        print("%5.1f%% inst %5.1f%% stall  %s:%d  %s" % (100 * i_ / tot_i, 100 * s_ / max(tot_s, 1), f, n, text))


if __name__ == "__main__":
    main()
