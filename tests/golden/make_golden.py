#!/usr/bin/env python
"""Generate the golden fixtures of tests/golden/ FROM THE REFERENCE ITSELF.

Run in the build container (needs /root/reference):   python tests/golden/make_golden.py

1. quade_dataset/ : a copy of the reference's only golden vector -- the DATA files test/dataset/*.fastq.gz,
   test/result/*.fastq.gz, Quade_report.csv and the conf (no source code is copied).
2. adversarial.json : purpose-built inputs (lower-case / N bases, short index reads, CRLF, trailing blanks,
   decorated '+' lines, skipped records, threshold edges, single index, disabled writers, several chunks,
   seeded random fuzz) together with the outputs the reference's own source produces for them when run
   through oracle/ref_shim.py.  Strings are latin-1; outputs are the decompressed bytes per created file
   and the report from line 2 on (line 1 carries the date).
"""
import gzip
import json
import os
import random
import shutil
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import ref_shim  # noqa: E402

REF = ref_shim.REFERENCE_ROOT


def copy_dataset():
    dst = os.path.join(HERE, "quade_dataset")
    os.makedirs(os.path.join(dst, "dataset"), exist_ok=True)
    os.makedirs(os.path.join(dst, "result"), exist_ok=True)
    for f in sorted(os.listdir(os.path.join(REF, "test", "dataset"))):
        shutil.copyfile(os.path.join(REF, "test", "dataset", f), os.path.join(dst, "dataset", f))
    for f in sorted(os.listdir(os.path.join(REF, "test", "result"))):
        shutil.copyfile(os.path.join(REF, "test", "result", f), os.path.join(dst, "result", f))
    for d in ("dataset", "result"):
        for f in os.listdir(os.path.join(dst, d)):
            os.chmod(os.path.join(dst, d, f), 0o644)


def write_conf(path, case, files):
    """Emit a config in the grammar of src/Conf_file.py:18-104."""
    c = case["conf"]
    L = ["[quality]", "minimal_qual : %d" % c["minimal_qual"], "[fastq]"]
    L.append("seq_R1 : " + " ".join(f[0] for f in files))
    L.append("seq_R2 : " + " ".join(f[1] for f in files))
    L.append("index_R1 : " + " ".join(f[2] for f in files))
    if c["index2"]:
        L.append("index_R2 : " + " ".join(f[3] for f in files))
    L += ["[index]", "index2 : %s" % c["index2"], "molecular1 : %s" % c["molecular1"], "molecular2 : %s" % c["molecular2"]]
    for k in ("index1", "index2", "molecular1", "molecular2"):
        if k in c["pos"]:
            L.append("%s_start : %d" % (k, c["pos"][k][0]))
            L.append("%s_end : %d" % (k, c["pos"][k][1]))
    L += ["[output]", "write_pass : %s" % c["write_pass"], "write_fail : %s" % c["write_fail"],
          "write_undetermined : %s" % c["write_undetermined"]]
    for i, s in enumerate(case["samples"]):
        L += ["[sample%d]" % (i + 1), "name : " + s[0], "index1_seq : " + s[1]]
        if c["index2"]:
            L.append("index2_seq : " + s[2])
    with open(path, "w") as fh:
        fh.write("\n".join(L) + "\n")


def run_case(case):
    tmp = tempfile.mkdtemp(prefix="quade_golden_")
    try:
        files = []
        for n, chunk in enumerate(case["chunks"]):
            row = []
            for s, data in enumerate(chunk):
                p = os.path.join(tmp, "in_c%d_s%d.fastq" % (n, s))
                with open(p, "wb") as fh:
                    fh.write(data.encode("latin-1"))
                row.append(p)
            files.append(row)
        conf = os.path.join(tmp, "conf.txt")
        write_conf(conf, case, files)
        out = os.path.join(tmp, "out")
        os.mkdir(out)
        ref_shim.run_reference(conf, out)
        exp = {"files": {}, "report": None}
        for f in sorted(os.listdir(out)):
            if f.endswith(".fastq.gz"):
                with gzip.open(os.path.join(out, f), "rb") as fh:
                    exp["files"][f] = fh.read().decode("latin-1")
            elif f == "Quade_report.csv":
                with open(os.path.join(out, f), "r", newline="") as fh:
                    exp["report"] = fh.readlines()[1:]
        case["expect"] = exp
    finally:
        shutil.rmtree(tmp)
    return case


def conf(minimal_qual=25, index2=True, molecular1=True, molecular2=True, pos=None, write_pass=True, write_fail=True,
         write_undetermined=True):
    if pos is None:
        pos = {"index1": [1, 4], "index2": [1, 4], "molecular1": [4, 6], "molecular2": [4, 6]}
    return {"minimal_qual": minimal_qual, "index2": index2, "molecular1": molecular1, "molecular2": molecular2,
            "pos": pos, "write_pass": write_pass, "write_fail": write_fail, "write_undetermined": write_undetermined}


def rec(name, seq, qual, descr="", plus="+", eol="\n"):
    return "@" + name + ((" " + descr) if descr else "") + eol + seq + eol + plus + eol + qual + eol


def cases():
    S2 = [["S1", "ACAG", "ACAG"], ["S2", "CTTG", "CTTG"]]
    out = []
    ins = "ACGTTGCATTGACCAGTNAC"
    q20 = "IIIIIHHHHHGGGGGFFFF#"

    # 1. lower-case and N handling, raw-case suffix
    out.append({"name": "lowercase_and_N", "conf": conf(), "samples": S2 + [["S3", "ACNG", "TTNA"]], "chunks": [[
        rec("r1", ins, q20, "1:N:0:") + rec("r2", ins, q20, "1:N:0:") + rec("r3", ins, q20) + rec("r4", ins, q20) + rec("r5", ins, q20),
        rec("r1", ins[::-1], q20, "4:N:0:") + rec("r2", ins[::-1], q20) + rec("r3", ins[::-1], q20) + rec("r4", ins[::-1], q20) + rec("r5", ins[::-1], q20),
        rec("r1", "acagTT", "IIIIII") + rec("r2", "ACNGAA", "IIIIII") + rec("r3", "AcnGaa", "IIIIII") + rec("r4", "ACXGAA", "IIIIII") + rec("r5", "NNNNNN", "IIIIII"),
        rec("r1", "AcAgGG", "IIIIII") + rec("r2", "TTNACC", "IIIIII") + rec("r3", "ttnaCC", "IIIIII") + rec("r4", "TTNACC", "IIIIII") + rec("r5", "NNNNNN", "IIIIII"),
    ]]})
    # 2. short index reads: slice truncation, vanishing molecular part (A.4, A.6)
    out.append({"name": "short_index_reads", "conf": conf(pos={"index1": [1, 4], "index2": [1, 4], "molecular1": [5, 6], "molecular2": [5, 6]}),
                "samples": S2 + [["S3", "ACA", "CTTG"], ["S4", "ACAG", "CT"]], "chunks": [[
        "".join(rec("r%d" % i, ins, q20) for i in range(7)),
        "".join(rec("m%d" % i, ins, q20) for i in range(7)),
        rec("i0", "ACAGTT", "IIIIII") + rec("i1", "ACAG", "IIII") + rec("i2", "ACA", "III") + rec("i3", "", "") + rec("i4", "ACAGT", "IIIII") + rec("i5", "ACAG", "IIII") + rec("i6", "ACAGTT", "IIIIII"),
        rec("i0", "ACAGGG", "IIIIII") + rec("i1", "ACAG", "IIII") + rec("i2", "CTTGAA", "IIIIII") + rec("i3", "ACAGTT", "IIIIII") + rec("i4", "ACAG", "IIII") + rec("i5", "CT", "II") + rec("i6", "", ""),
    ]]})
    # 3. CRLF files, trailing blanks / tabs, tab in header, decorated '+', leading blanks after '@'
    out.append({"name": "crlf_and_whitespace", "conf": conf(), "samples": S2, "chunks": [[
        rec("a1", ins, q20, "1:N:0:", eol="\r\n") + "@a2\tdescr with tab  \n" + ins + "  \t\n+a2 decorated\n" + q20 + " \n" + "@  a3 x y\n" + ins + "\n+\n" + q20 + "\n" + "@a4:x:y \x0b\x0c\n" + ins + "\x0b\n+\n" + q20 + "\x0c\n",
        rec("b1", ins, q20, "4:N:0:", eol="\r\n") + rec("b2", ins, q20) + rec("b3", ins, q20, plus="+b3") + rec("b4", ins, q20),
        rec("i1", "ACAGTC", "IIIIII", eol="\r\n") + rec("i2", "CTTGAA", "IIIIII") + "@i3\nACAGTT \n+\nIIIIII\t\n" + rec("i4", "CTTGCA", "IIIIII"),
        rec("j1", "ACAGGA", "IIIIII", eol="\r\n") + rec("j2", "CTTGCC", "IIIIII") + rec("j3", "ACAGAA", "IIIIII") + "@j4\nCTTGAC\r\n+\nIIIIII\r\n",
    ]]})
    # 4. quality threshold edges: ':' = 25 passes, '9' = 24 fails; low quality at a molecular-only position is ignored;
    #    overlapping position 4 (index 1-4, molecular 4-6) is checked (A.5)
    out.append({"name": "threshold_edges", "conf": conf(), "samples": S2, "chunks": [[
        "".join(rec("t%d" % i, ins, q20) for i in range(6)),
        "".join(rec("u%d" % i, ins, q20) for i in range(6)),
        rec("i0", "ACAGTT", "::::II") + rec("i1", "ACAGTT", ":::9II") + rec("i2", "ACAGTT", "IIII!!") + rec("i3", "CTTGAA", "9IIIII") + rec("i4", "CTTGAA", "IIIIII") + rec("i5", "GGGGGG", "!!!!!!"),
        rec("j0", "ACAGTT", "::::II") + rec("j1", "ACAGTT", "IIIIII") + rec("j2", "ACAGTT", "IIII##") + rec("j3", "CTTGAA", "IIIIII") + rec("j4", "CTTGAA", "III:!!") + rec("j5", "GGGGGG", "!!!!!!"),
    ]]})
    # 5. single index, no molecular, 8-base barcode (BASELINE config 2 layout)
    S8 = [["A1", "ACGTACGT"], ["A2", "TTGGCCAA"], ["A3", "NACGTACG"]]
    out.append({"name": "single_index", "conf": conf(index2=False, molecular1=False, molecular2=False, pos={"index1": [1, 8]}),
                "samples": S8, "chunks": [[
        "".join(rec("s%d" % i, ins, q20, "1:N:0:0") for i in range(6)),
        "".join(rec("s%d" % i, ins[::-1], q20, "3:N:0:0") for i in range(6)),
        rec("s0", "ACGTACGT", "IIIIIIII") + rec("s1", "TTGGCCAA", "IIII5III") + rec("s2", "TTGGCCAT", "IIIIIIII") + rec("s3", "nacgtacg", "IIIIIIII") + rec("s4", "ACGTACG", "IIIIIII") + rec("s5", "ACGTACGTA", "IIIIIIII#"),
    ]]})
    # 6. disabled writers still count (A.5-4), min_qual 0
    out.append({"name": "writers_off", "conf": conf(minimal_qual=30, write_fail=False, write_undetermined=False), "samples": S2, "chunks": [[
        "".join(rec("w%d" % i, ins, q20) for i in range(4)), "".join(rec("x%d" % i, ins, q20) for i in range(4)),
        rec("i0", "ACAGTT", "IIIIII") + rec("i1", "ACAGTT", "I5IIII") + rec("i2", "GGGGGG", "IIIIII") + rec("i3", "CTTGAA", "IIIIII"),
        rec("j0", "ACAGTT", "IIIIII") + rec("j1", "ACAGTT", "IIIIII") + rec("j2", "GGGGGG", "IIIIII") + rec("j3", "CTTGAA", "IIIIII"),
    ]]})
    out.append({"name": "pass_off_minqual0", "conf": conf(minimal_qual=0, write_pass=False), "samples": S2, "chunks": [[
        "".join(rec("w%d" % i, ins, q20) for i in range(3)), "".join(rec("x%d" % i, ins, q20) for i in range(3)),
        rec("i0", "ACAGTT", "!!!!!!") + rec("i1", "GGGGGG", "IIIIII") + rec("i2", "CTTGAA", "IIIIII"),
        rec("j0", "ACAGTT", "!!!!!!") + rec("j1", "GGGGGG", "IIIIII") + rec("j2", "CTTGAA", "IIIIII"),
    ]]})
    # 7. skipped (invalid) records in several streams, unequal stream lengths, partial last record,
    #    final line without newline, three chunks incl. an empty one (A.1, A.3)
    bad = "@bad\nACGT\n+\nIII\n"
    out.append({"name": "desync_and_tails", "conf": conf(), "samples": S2, "chunks": [
        [rec("a0", ins, q20) + bad + rec("a1", ins, q20) + rec("a2", ins, q20) + rec("a3", ins, q20) + "@partial\nACGT\n",
         rec("b0", ins, q20) + rec("b1", ins, q20) + bad + bad + rec("b2", ins, q20) + rec("b3", ins, q20) + rec("b4", ins, q20),
         rec("i0", "ACAGTT", "IIIIII") + rec("i1", "CTTGAA", "IIIIII") + rec("i2", "ACAGCC", "IIIIII") + "@ibad\nACAGTT\n+\nIIIII\n" + rec("i3", "CTTGGG", "IIIIII"),
         rec("j0", "ACAGTT", "IIIIII") + rec("j1", "CTTGAA", "IIIIII") + rec("j2", "ACAGCC", "IIIIII") + rec("j3", "CTTGGG", "IIIIII") + "@jlast\nACAGTT\n+\nIIIIII"],
        ["", "", "", ""],
        [rec("c0", ins, q20) + rec("c1", ins, q20)[:-1],
         rec("d0", ins, q20) + rec("d1", ins, q20),
         rec("k0", "CTTGTT", "IIIIII") + rec("k1", "ACAGAA", "IIIII9")[:-1],
         rec("l0", "CTTGTT", "IIIIII") + rec("l1", "ACAGAA", "IIIIII") + rec("l2", "ACAGAA", "IIIIII")],
    ]})
    # 8. dual index without molecular, longer barcodes, spans reaching past the read end
    S10 = [["P1", "ACGTACGTAC", "TTGGCCAATT"], ["P2", "GGGGGCCCCC", "AAAAATTTTT"]]
    out.append({"name": "dual_no_molecular", "conf": conf(molecular1=False, molecular2=False, pos={"index1": [1, 10], "index2": [1, 10]}),
                "samples": S10, "chunks": [[
        "".join(rec("n%d:%d" % (i, i * 7919), ins * 3, q20 * 3, "1:N:0:0") for i in range(5)),
        "".join(rec("n%d:%d" % (i, i * 7919), ins * 2, q20 * 2, "4:N:0:0") for i in range(5)),
        rec("n0", "ACGTACGTAC", "IIIIIIIIII") + rec("n1", "GGGGGCCCCC", "IIIIIIIII5") + rec("n2", "ACGTACGTACGG", "IIIIIIIIIIII") + rec("n3", "ACGTACGTA", "IIIIIIIII") + rec("n4", "GGGGGCCCCC", "IIIIIIIIII"),
        rec("n0", "TTGGCCAATT", "IIIIIIIIII") + rec("n1", "AAAAATTTTT", "IIIIIIIIII") + rec("n2", "TTGGCCAATTAA", "IIIIIIIIIIII") + rec("n3", "TTGGCCAATT", "IIIIIIIIII") + rec("n4", "AAAAATTTTA", "IIIIIIIIII"),
    ]]})
    # 9. molecular on index read 1 only, index 2 without molecular, non-adjacent spans
    out.append({"name": "mol1_only", "conf": conf(molecular2=False, pos={"index1": [3, 6], "index2": [2, 5], "molecular1": [7, 10]}),
                "samples": [["Q1", "ACAG", "CTTG"], ["Q2", "CTTG", "ACAG"]], "chunks": [[
        "".join(rec("q%d" % i, ins, q20) for i in range(4)), "".join(rec("q%d" % i, ins, q20) for i in range(4)),
        rec("q0", "TTACAGGATCGG", "IIIIIIIIIIII") + rec("q1", "TTCTTGAATCGG", "II9IIIIIIIII") + rec("q2", "TTACAGG", "IIIIIII") + rec("q3", "TTCTTG", "IIIIII"),
        rec("q0", "ACTTGA", "IIIIII") + rec("q1", "AACAGA", "IIIIII") + rec("q2", "ACTTGA", "IIIIII") + rec("q3", "AACAGA", "I!IIII"),
    ]]})
    # 10. seeded random fuzz in the style of the synthetic benchmark data
    for seed in range(6):
        out.append(fuzz_case(seed))
    return out


def fuzz_case(seed):
    rng = random.Random(0xC0FFEE + seed)
    dual = rng.random() < 0.7
    ilen = rng.choice([4, 6, 8])
    mlen = rng.choice([0, 0, 3, 6])
    nsamp = rng.choice([2, 5, 17])
    alphabet = "ACGT"

    def rb(n, extra=""):
        return "".join(rng.choice(alphabet + extra) for _ in range(n))
    keys = set()
    samples = []
    while len(samples) < nsamp:
        k1, k2 = rb(ilen, "N" if rng.random() < 0.1 else ""), rb(ilen)
        if (k1, k2 if dual else "") in keys:
            continue
        keys.add((k1, k2 if dual else ""))
        samples.append(["F%02d" % len(samples), k1, k2])
    pos = {"index1": [1, ilen]}
    if dual:
        pos["index2"] = [1, ilen]
    if mlen:
        pos["molecular1"] = [ilen + 1, ilen + mlen]
        if dual:
            pos["molecular2"] = [ilen + 1, ilen + mlen]
    cf = conf(minimal_qual=rng.choice([0, 20, 25, 30]), index2=dual, molecular1=bool(mlen) and dual, molecular2=bool(mlen) and dual,
              pos=pos, write_fail=rng.random() < 0.8, write_undetermined=rng.random() < 0.8)
    if not cf["molecular1"]:
        pos.pop("molecular1", None)
        pos.pop("molecular2", None)
    chunks = []
    for _ in range(rng.choice([1, 2, 3])):
        n = rng.randrange(0, 60)
        st = [[], [], [], []]
        for i in range(n):
            name = "NB:%d:%d:%d" % (rng.randrange(1, 9), rng.randrange(1101, 2679), rng.randrange(1000, 100000))
            rl = rng.choice([20, 36, 75])
            for s, d in ((0, "1:N:0:0"), (1, "4:N:0:0")):
                seq = rb(rl, "N" if rng.random() < 0.2 else "")
                qual = "".join(chr(33 + rng.choice([2, 12, 23, 37])) for _ in range(rl))
                if rng.random() < 0.03:
                    qual = qual[:-1]                      # invalid record -> skipped in this stream only
                eol = "\r\n" if rng.random() < 0.05 else "\n"
                st[s].append(rec(name, seq, qual, d if rng.random() < 0.8 else "", plus="+" if rng.random() < 0.9 else "+" + name, eol=eol))
            smp = rng.choice(samples)
            for s, key in ((2, smp[1]), (3, smp[2])):
                r = rng.random()
                b = key if r < 0.7 else rb(ilen)
                if r > 0.95:
                    b = b.lower()
                full = b + rb(mlen + rng.choice([0, 0, 1]))
                if rng.random() < 0.05:
                    full = full[:rng.randrange(0, len(full) + 1)]
                floor = 30 if rng.random() < 0.7 else 2
                qual = "".join(chr(33 + rng.randrange(floor, 41)) for _ in range(len(full)))
                if rng.random() < 0.02:
                    qual += "I"
                st[s].append(rec(name, full, qual, "2:N:0:0"))
        if rng.random() < 0.3 and st[0]:
            st[rng.randrange(4)].pop()                    # unequal stream lengths
        data = ["".join(x) for x in st]
        if rng.random() < 0.3 and data[1]:
            data[1] = data[1][:-1]                        # no trailing newline
        chunks.append(data if dual else data[:3])
    return {"name": "fuzz_%d" % seed, "conf": cf, "samples": samples, "chunks": chunks}


def main():
    if not ref_shim.available():
        sys.exit("reference not present under %s" % REF)
    copy_dataset()
    done = [run_case(c) for c in cases()]
    with open(os.path.join(HERE, "adversarial.json"), "w") as fh:
        json.dump({"generator": "tests/golden/make_golden.py", "source": "a-slide/Quade reference via oracle/ref_shim.py",
                   "cases": done}, fh, indent=1, ensure_ascii=True)
    for c in done:
        print("%-22s files=%d report_lines=%d" % (c["name"], len(c["expect"]["files"]), len(c["expect"]["report"] or [])))


if __name__ == "__main__":
    main()
