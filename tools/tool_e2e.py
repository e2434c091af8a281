"""Whole-tool timing: synthetic cfg3 chunks as FASTQ files -> `Quade(conf)()` (the reference's entry point,
src/Quade.py:169-193) -> per-bin fastq.gz files + Quade_report.csv.  Reports pairs/s from the tool's wall time and
checks the report totals.  usage: python tools/tool_e2e.py [pairs_per_chunk] [chunks] [gz|plain] [zlib level] [dir]
This is SURVEY.md section 8(f) territory (host inflate / deflate around the kernels), not the bench metric."""
import gzip
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from quade_b200.synth import Synth  # noqa: E402

per = int(sys.argv[1]) if len(sys.argv) > 1 else 250_000
chunks = int(sys.argv[2]) if len(sys.argv) > 2 else 4
gz = (sys.argv[3] if len(sys.argv) > 3 else "plain") == "gz"
level = int(sys.argv[4]) if len(sys.argv) > 4 else 1
s = Synth.from_config("cfg3")
wd = tempfile.mkdtemp(prefix="quade_tool_", dir=sys.argv[5] if len(sys.argv) > 5 else None)
os.chdir(wd)
names = [[] for _ in range(4)]
t0 = time.time()
in_bytes = 0
for c in range(chunks):
    arrays = s.generate(c * per, per)
    for k, a in enumerate(arrays):
        fn = "C%d_%d.fastq%s" % (c, k, ".gz" if gz else "")
        if gz:
            with gzip.open(fn, "wb", compresslevel=1) as fp:
                fp.write(memoryview(a))
        else:
            a.tofile(fn)
        names[k].append(fn)
        in_bytes += a.size
print("inputs: %d chunks x %d pairs, %.2f GB of FASTQ text (%s) written to %s in %.1f s" % (
    chunks, per, in_bytes / 1e9, "gz" if gz else "plain", wd, time.time() - t0))
with open("conf.txt", "w") as fp:
    fp.write(s.conf_text(names))

from quade_b200.fastq_writer import FastqWriter  # noqa: E402
from quade_b200.quade import Quade  # noqa: E402

FastqWriter.COMPRESSLEVEL = level
t0 = time.time()
rc = Quade(conf_file="conf.txt")()
dt = time.time() - t0
total = None
for line in open("Quade_report.csv"):
    f = line.rstrip("\n").split("\t")
    if len(f) == 2 and f[1].isdigit() and total is None and "otal" in f[0]:
        total = int(f[1])
out_files = [f for f in os.listdir(".") if f.endswith(".fastq.gz") and not f.startswith("C")]
print("tool (%s in, zlib level %d out): rc %d, %.2f s, %.0f pairs/s (%d pairs, report total %s), %d output files, %.2f GB written" % (
    "gz" if gz else "plain", level, rc, dt, chunks * per / dt, chunks * per, total, len(out_files),
    sum(os.path.getsize(f) for f in out_files) / 1e9))
assert rc == 0 and total == chunks * per
import shutil  # noqa: E402
os.chdir("/")
shutil.rmtree(wd, ignore_errors=True)
