#!/usr/bin/env python
"""Per-kernel SASS evidence of the in-tree library (run here, no GPU): architecture, instruction count and the
Blackwell / Hopper-class mnemonics the design relies on -- UBLKCP (cp.async.bulk, the TMA bulk copy), UBLKPF (bulk L2
prefetch), SYNCS (mbarrier), IDP (dp4a mask gather), SHF/PRMT (byte funnel), REDUX, MATCH, SHFL, BAR.
usage: python tools/sass_summary.py [lib] > profiles/r02_sass_summary.txt"""
import collections
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else "quade_b200/libquade_b200.so"
txt = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
WATCH = ["UBLKCP", "UBLKPF", "SYNCS", "UTMALDG", "IDP", "SHF", "PRMT", "LOP3", "POPC", "FLO", "BREV", "SHFL", "MATCH", "REDUX", "VOTE",
         "BAR", "LDS", "STS", "LDG", "STG", "ATOMS", "ATOMG", "RED", "HMMA", "UTCHMMA"]
arch, fn, counts, total = None, None, {}, {}
for line in txt.splitlines():
    m = re.search(r"arch = (sm_\w+)", line)
    if m:
        arch = m.group(1)
    m = re.search(r"Function : (\S+)", line)
    if m:
        fn = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        fn = re.sub(r"quade::\(anonymous namespace\)::", "", fn).split("(")[0].replace("void ", "")
        counts[fn], total[fn] = collections.Counter(), 0
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
    if m and fn:
        total[fn] += 1
        op = m.group(1)
        for w in WATCH:
            if op == w or op.startswith(w + ".") or (w in ("LDS", "STS", "LDG", "STG") and op.startswith(w)):
                counts[fn][w] += 1
print("library %s, arch %s, %d kernels" % (lib, arch, len(counts)))
print("%-28s %6s  %s" % ("kernel", "SASS", "  ".join("%s" % w for w in WATCH if any(counts[f][w] for f in counts))))
for f in sorted(counts):
    print("%-28s %6d  %s" % (f[:28], total[f], "  ".join("%*d" % (len(w), counts[f][w]) for w in WATCH if any(counts[g][w] for g in counts))))
print("no HMMA / UTC*MMA / UTMALDG expected: byte-streaming integer work, 1-D bulk copies need no tensor map")
