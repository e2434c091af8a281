"""Whole-tool timing: synthetic cfg3 chunks as FASTQ files -> `Quade(conf)()` (the reference's entry point,
src/Quade.py:169-193) -> per-bin fastq.gz files + Quade_report.csv.  Reports pairs/s from the tool's wall time and
checks the report totals.  usage: python tools/tool_e2e.py [pairs_per_chunk] [chunks] [gz|plain]
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
s = Synth.from_config("cfg3")
wd = tempfile.mkdtemp(prefix="quade_tool_")
os.chdir(wd)
names = {"seq_R1": [], "seq_R2": [], "index_R1": [], "index_R2": []}
order = ["seq_R1", "seq_R2", "index_R1", "index_R2"]
t0 = time.time()
in_bytes = 0
for c in range(chunks):
    arrays = s.generate(c * per, per)
    for key, a in zip(order, arrays):
        fn = "C%d_%s.fastq%s" % (c, key, ".gz" if gz else "")
        if gz:
            with gzip.open(fn, "wb", compresslevel=1) as fp:
                fp.write(memoryview(a))
        else:
            a.tofile(fn)
        names[key].append(fn)
        in_bytes += a.size
print("inputs: %d chunks x %d pairs, %.2f GB of FASTQ text (%s) written in %.1f s" % (chunks, per, in_bytes / 1e9, "gz" if gz else "plain", time.time() - t0))
il, ml = s.index_len, s.mol_len
conf = ["[quality]", "minimal_qual : %d" % s.min_qual, "", "[fastq]"]
conf += ["%s : %s" % (k, " ".join(names[k])) for k in order]
conf += ["", "[index]", "index2 : True", "molecular1 : True", "molecular2 : True",
         "index1_start : 1", "index1_end : %d" % il, "index2_start : 1", "index2_end : %d" % il,
         "molecular1_start : %d" % (il + 1), "molecular1_end : %d" % (il + ml),
         "molecular2_start : %d" % (il + 1), "molecular2_end : %d" % (il + ml),
         "", "[output]", "write_pass : True", "write_fail : True", "write_undetermined : True", ""]
for i, (n, k) in enumerate(zip(s.sample_names(), s.sample_keys())):
    conf += ["[sample%d]" % (i + 1), "name : %s" % n, "index1_seq : %s" % k[:il], "index2_seq : %s" % k[il:], ""]
with open("conf.txt", "w") as fp:
    fp.write("\n".join(conf))

from quade_b200.quade import Quade  # noqa: E402

t0 = time.time()
rc = Quade(conf_file="conf.txt")()
dt = time.time() - t0
total = None
for line in open("Quade_report.csv"):
    f = line.rstrip("\n").split("\t")
    if len(f) == 2 and f[1].isdigit() and total is None and "otal" in f[0]:
        total = int(f[1])
out_files = [f for f in os.listdir(".") if f.endswith(".fastq.gz") and not f.startswith("C")]
print("tool: rc %d, %.2f s, %.0f pairs/s (%d pairs, report total %s), %d output files, %.2f GB gz written" % (
    rc, dt, chunks * per / dt, chunks * per, total, len(out_files), sum(os.path.getsize(f) for f in out_files) / 1e9))
assert rc == 0 and total == chunks * per
