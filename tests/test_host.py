"""Host side of the drop-in: config grammar, Sample / FastqWriter object model, batching with carry-over, report,
chunk sharding over ranks (gloo, world_size 2).  The device is replaced by tests.util.OracleEngine here (the
GPU twin of the golden run is tests/test_gpu_host.py)."""
import gzip
import os
import shutil
import subprocess
import sys

import numpy as np
import pytest

from tests import util

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATASET = os.path.join(util.GOLDEN, "quade_dataset")


def _fresh():
    from quade_b200.sample import Sample
    from quade_b200.fastq_writer import FastqWriter
    Sample.RESET()
    FastqWriter.PART = None


def _stage_dataset(tmp_path):
    """tmp/dataset + tmp/result laid out like the reference's test/ directory (the template's relative paths)."""
    shutil.copytree(os.path.join(DATASET, "dataset"), tmp_path / "dataset")
    (tmp_path / "result").mkdir()
    shutil.copy(os.path.join(DATASET, "result", "Quade_conf_file.txt"), tmp_path / "result" / "Quade_conf_file.txt")
    return tmp_path / "result"


def _check_golden_outputs(workdir):
    _, _, exp, report = util.quade_dataset()
    got = sorted(f for f in os.listdir(workdir) if f.endswith(".fastq.gz"))
    assert got == sorted(exp)                      # same file set: no *_fail_* files, no leftover parts
    for f in exp:
        with gzip.open(os.path.join(workdir, f), "rb") as fh:
            assert fh.read() == exp[f], f
    with open(os.path.join(workdir, "Quade_report.csv"), newline="") as fh:
        assert fh.readlines()[1:] == report


def test_template_has_the_reference_values(tmp_path, monkeypatch):
    from quade_b200.quade import Quade
    from quade_b200.sample import Sample
    _fresh()
    wd = _stage_dataset(tmp_path)
    monkeypatch.chdir(wd)
    with pytest.raises(SystemExit) as e:
        Quade(init_conf=True)
    assert e.value.code == 0
    q = Quade(conf_file="Quade_conf_file.txt")
    ref = Quade.__new__(Quade)
    _fresh()
    shutil.copy(os.path.join(DATASET, "result", "Quade_conf_file.txt"), "golden_conf.txt")
    ref.__init__(conf_file="golden_conf.txt")
    for k in ("minimal_qual", "idx1", "idx2", "mol1", "mol2", "idx1_pos", "idx2_pos", "mol1_pos", "mol2_pos",
              "seq_R1", "seq_R2", "index_R1", "index_R2"):
        assert getattr(q, k) == getattr(ref, k), k
    assert q.spans() == [(0, 4), (0, 4), (3, 6), (3, 6)]
    assert [(s.name, s.index) for s in Sample.SAMPLE_LIST] == [("S1", "ACAGACAG"), ("S2", "CTTGCTTG")]
    assert (Sample.WRITE_PASS, Sample.WRITE_FAIL, Sample.WRITE_UNDETERMINED, Sample.MIN_QUAL) == (True, True, True, 25)


@pytest.mark.parametrize("batch_bytes", [1 << 20, 2500, 997])
def test_golden_run_through_the_host_loop(tmp_path, monkeypatch, batch_bytes):
    """The documented smoke test (README.md:67-72) through Quade.__call__: staging, ragged batches with
    carry-over, per-stream end-of-file flags, lazy file creation, report."""
    from quade_b200.quade import Quade
    _fresh()
    wd = _stage_dataset(tmp_path)
    monkeypatch.chdir(wd)
    util.patch_host_engine(Quade, monkeypatch)
    monkeypatch.setattr(Quade, "BATCH_BYTES", batch_bytes)
    q = Quade(conf_file="Quade_conf_file.txt")
    assert q() == 0
    _check_golden_outputs(wd)


@pytest.mark.parametrize("batch_bytes", [1 << 20, 1499])
def test_golden_run_with_plain_text_inputs(tmp_path, monkeypatch, batch_bytes):
    """Same run with the inputs decompressed to `.fastq` files: a path that does not end in "gz" is opened as a plain
    file (FastqReader.py:30) and read straight into the staging buffers."""
    from quade_b200.quade import Quade
    _fresh()
    wd = _stage_dataset(tmp_path)
    for f in sorted(os.listdir(tmp_path / "dataset")):
        if f.endswith(".fastq.gz"):
            with gzip.open(tmp_path / "dataset" / f, "rb") as src, open(tmp_path / "dataset" / f[:-3], "wb") as dst:
                dst.write(src.read())
            os.remove(tmp_path / "dataset" / f)
    conf = (wd / "Quade_conf_file.txt").read_text().replace(".fastq.gz", ".fastq")
    (wd / "Quade_conf_file.txt").write_text(conf)
    monkeypatch.chdir(wd)
    util.patch_host_engine(Quade, monkeypatch)
    monkeypatch.setattr(Quade, "BATCH_BYTES", batch_bytes)
    q = Quade(conf_file="Quade_conf_file.txt")
    assert all(f.endswith(".fastq") for f in q.seq_R1 + q.seq_R2 + q.index_R1 + q.index_R2)
    assert q() == 0
    _check_golden_outputs(wd)


def test_config_errors_exit_1(tmp_path, monkeypatch, capsys):
    from quade_b200.quade import Quade
    wd = _stage_dataset(tmp_path)
    monkeypatch.chdir(wd)
    good = open("Quade_conf_file.txt").read()
    cases = {
        "missing option": (good.replace("minimal_qual : 25", ""), "Option or section missing"),
        "bad value": (good.replace("minimal_qual : 25", "minimal_qual : 41"), "One of the value in the configuration file is not correct"),
        "not an int": (good.replace("index1_end : 4", "index1_end : x"), "One of the value in the configuration file is not correct"),
        "lower-case index": (good.replace("index1_seq : ACAG", "index1_seq : acag"), "Non canonical DNA base in index"),
        "duplicate index": (good.replace("CTTG", "ACAG"), "Index is not unique"),
        "unreadable file": (good.replace("../dataset/C2_R1.fastq.gz", "../dataset/nope.fastq.gz"), "One of the file is incorrect or unreadable"),
        "ragged lists": (good.replace("../dataset/C3_R4.fastq.gz", ""), "have to contain the same number of files"),
        "end before start": (good.replace("molecular2_end : 6", "molecular2_end : 2"), "One of the value in the configuration file is not correct"),
    }
    for name, (text, msg) in cases.items():
        _fresh()
        with open("bad.txt", "w") as fh:
            fh.write(text)
        with pytest.raises(SystemExit) as e:
            Quade(conf_file="bad.txt")
        assert e.value.code == 1, name
        assert msg in capsys.readouterr().out, name
    _fresh()
    with pytest.raises(SystemExit) as e:
        Quade(conf_file=None)
    assert e.value.code == 1
    assert "A path to the configuration file is mandatory" in capsys.readouterr().out


def test_writer_object_model(tmp_path, monkeypatch):
    from quade_b200.fastq_writer import FastqWriter
    monkeypatch.chdir(tmp_path)
    _fresh()
    w = FastqWriter(name="X_pass")
    assert (w.R1_fastq_name, w.R2_fastq_name, w.counter, w.buffer_size) == ("X_pass_R1.fastq.gz", "X_pass_R2.fastq.gz", -1, 20)
    assert not os.path.exists(w.R1_fastq_name)                  # lazy creation (FastqWriter.py:55-57)
    w(b"@a:AC\nAC\n+\nII\n", b"@a:AC\nGT\n+\nII\n")
    assert os.path.exists(w.R1_fastq_name) and w.counter == 1
    w(memoryview(b"@b:AC\nAC\n+\nII\n"), memoryview(b"@b:AC\nGT\n+\nII\n"))
    w.flush_buffers()
    assert w.counter == 0
    assert gzip.open(w.R1_fastq_name).read() == b"@a:AC\nAC\n+\nII\n@b:AC\nAC\n+\nII\n"
    assert gzip.open(w.R2_fastq_name).read() == b"@a:AC\nGT\n+\nII\n@b:AC\nGT\n+\nII\n"
    with pytest.raises(TypeError):
        w(object(), object())                                   # no per-record CPU serialisation


def test_report_matches_oracle_formatter():
    from quade_b200.sample import Sample
    from oracle import oracle as O
    _fresh()
    Sample.CLASS_INIT(True, True, True, 25)
    for n, i in (("A", "ACGT"), ("B", "TTTT"), ("C", "GGGG")):
        Sample(n, i)
    for counters in ([0, 0, 0, 0, 0, 0, 0], [5, 0, 0, 0, 0, 0, 0], [3, 7, 2, 0, 0, 1, 0], [247, 25, 0, 27, 0, 0, 0]):
        Sample.SET_COUNTS(counters)
        lines = ["\n"] + ["{}\t{}\n".format(d, v) for d, v in Sample.REPORT()]
        assert lines == O.report_lines(["A", "B", "C"], counters)
        assert Sample.COUNTS() == counters
    with pytest.raises(TypeError):
        Sample.FINDER("read1", "read2", "index")                # the per-record CPU form is not provided


def test_sharded_run_world_size_2_gloo(tmp_path):
    """Two ranks (gloo), chunks round-robin, counters all-reduced, parts merged in chunk order: same files and
    report as the single-process run."""
    wd = _stage_dataset(tmp_path)
    code = (
        "import sys; sys.path.insert(0, %r)\n"
        "from tests import util\n"
        "from quade_b200.quade import Quade\n"
        "util.patch_host_engine(Quade)\n"
        "Quade.BATCH_BYTES = 4096\n"
        "sys.exit(Quade(conf_file='Quade_conf_file.txt')())\n" % ROOT)
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), WORLD_SIZE="2", LOCAL_RANK=str(rank), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT="29641")
        procs.append(subprocess.Popen([sys.executable, "-c", code], cwd=str(wd), env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=240)[0] for p in procs]
    assert [p.returncode for p in procs] == [0, 0], outs
    _check_golden_outputs(wd)
    assert "Start parsing chunk 2" in outs[1] and "Start parsing chunk 2" not in outs[0]


def test_native_gzip_sink_keeps_member_order_and_handles_stored_blocks(tmp_path):
    """libquade_io.so: members of one file are appended in submission order whatever thread compressed them; level 0
    writes stored blocks (valid gzip); empty bins produce nothing; the arenas are released by wait()."""
    import numpy as np
    from quade_b200._io import GzipSink
    rng = np.random.default_rng(5)
    nb = 7
    r1 = [str(tmp_path / ("b%d_R1.fastq.gz" % b)) for b in range(nb)]
    r2 = [str(tmp_path / ("b%d_R2.fastq.gz" % b)) for b in range(nb)]
    for level in (0, 1):
        for p in r1 + r2:
            gzip.open(p, "wb").close()                       # lazy creation is the caller's job
        sink = GzipSink(n_threads=5, level=level)
        sink.set_paths(r1, r2)
        expect = [[bytearray(), bytearray()] for _ in range(nb)]
        tickets = []
        for batch in range(12):
            arenas, offs = [], []
            for k in range(2):
                sizes = rng.integers(0, 3000, nb) * (rng.random(nb) < 0.7)
                off = np.zeros(nb + 1, dtype=np.uint64)
                np.cumsum(sizes, out=off[1:])
                a = rng.integers(65, 91, int(off[-1]) + 1, dtype=np.uint8)
                for b in range(nb):
                    expect[b][k] += a[int(off[b]):int(off[b + 1])].tobytes()
                arenas.append(a)
                offs.append(off)
            tickets.append(sink.append(arenas[0].ctypes.data, offs[0], arenas[1].ctypes.data, offs[1], keep=arenas))
        sink.wait(tickets[-1])
        n_in, n_out = sink.stats()
        assert n_in == sum(len(e[0]) + len(e[1]) for e in expect) and n_out > 0
        sink.close()
        for b in range(nb):
            for k, paths in enumerate((r1, r2)):
                with gzip.open(paths[b], "rb") as fh:
                    assert fh.read() == bytes(expect[b][k]), (level, b, k)


def test_a_buffer_of_invalid_records_only_makes_progress(tmp_path, monkeypatch):
    """One file starts with more records of len(seq) != len(qual) than a staging buffer holds: the reference skips them
    in that file only (FastqReader.py:48-50); the batched host loop must pass them (consumed, counted, printed) and
    pair what follows, not stop with 'record larger than the staging buffer'."""
    from quade_b200.quade import Quade
    from quade_b200.sample import Sample
    from oracle import oracle as O
    _fresh()
    good = [b"@r%d\nACGTAC\n+\nIIIIII\n" % i for i in range(50)]
    bad = [b"@b%d\nACGTAC\n+\nIIIII\n" % i for i in range(300)]
    streams = [b"".join(bad + good), b"".join(good), b"".join(good)]
    wd = tmp_path / "run"
    wd.mkdir()
    monkeypatch.chdir(wd)
    for f, s in zip(("R1.fastq", "R2.fastq", "I1.fastq"), streams):
        (wd / f).write_bytes(s)
    (wd / "conf.txt").write_text("\n".join([
        "[quality]", "minimal_qual : 20", "[fastq]", "seq_R1 : R1.fastq", "seq_R2 : R2.fastq", "index_R1 : I1.fastq",
        "[index]", "index2 : False", "molecular1 : False", "molecular2 : False", "index1_start : 1", "index1_end : 4",
        "[output]", "write_pass : True", "write_fail : True", "write_undetermined : True",
        "[sample1]", "name : S1", "index1_seq : ACGT"]))
    util.patch_host_engine(Quade, monkeypatch)
    monkeypatch.setattr(Quade, "BATCH_BYTES", 1024)              # ~50 of the 300 invalid records per buffer
    assert Quade(conf_file="conf.txt")() == 0
    p = O.OracleParams(False, [(0, 4), (0, 0), (0, 0), (0, 0)], 20, ["ACGT"])
    ref = O.demux_batch(p, streams, last=True)
    assert ref.n_pairs == 50 and Sample.TOTAL == 50 and Sample.COUNTS() == [int(x) for x in ref.counters]
    with gzip.open(wd / "S1_pass_R1.fastq.gz", "rb") as fh:
        assert fh.read() == ref.bin_bytes(0, 1)


# ---------------------------------------------------------------------------------------------------------------
# native gzip reader (libquade_io.so) against Python's gzip module
# ---------------------------------------------------------------------------------------------------------------
def _drain(reader, sizes):
    out, i = bytearray(), 0
    while not reader.eof:
        buf = np.empty(sizes[i % len(sizes)], dtype=np.uint8)
        n = reader.readinto(buf)
        out += buf[:n].tobytes()
        i += 1
    return bytes(out)


def test_native_gzip_reader_matches_pythons_gzip(tmp_path, monkeypatch):
    import gzip
    from quade_b200.pyfastq import FastqReader
    rng = np.random.default_rng(5)
    text = b"".join(b"@r%d\n%s\n+\n%s\n" % (i, bytes(rng.choice(list(b"ACGT"), 40)), bytes(rng.integers(35, 70, 40, dtype=np.uint8)))
                    for i in range(6000))
    cases = {}
    cases["one_member"] = gzip.compress(text, 1)
    cut = [0, 1000, 1001, 70000, len(text) // 2, len(text)]
    cases["members"] = b"".join(gzip.compress(text[a:b], 6) for a, b in zip(cut, cut[1:]))      # an empty member too
    cases["members_20"] = b"".join(gzip.compress(text[a:a + 1800], 1) for a in range(0, len(text), 1800))
    cases["padded"] = gzip.compress(text, 9) + b"\0" * 777
    cases["empty_file"] = b""
    cases["empty_member"] = gzip.compress(b"")
    monkeypatch.setattr(FastqReader, "BLOCK", 4096)                     # many blocks: the ring turns over
    for name, blob in cases.items():
        p = tmp_path / (name + ".fastq.gz")
        p.write_bytes(blob)
        want = gzip.decompress(blob) if blob else b""
        for depth, sizes in ((1, [100000]), (3, [1, 4095, 4096, 4097, 50000]), (64, [1 << 20])):
            r = FastqReader(str(p), depth=depth)
            assert r._native is not None
            assert _drain(r, sizes) == want, (name, depth)
            assert r.readinto(np.empty(10, dtype=np.uint8)) == 0 and r.eof
            r.close()
        r = FastqReader(str(p), prefetch=False)                         # Python's gzip in the calling thread: same bytes
        assert _drain(r, [65536]) == want, name
        r.close()


def test_native_gzip_reader_reports_damage_after_the_good_bytes(tmp_path, monkeypatch):
    import gzip
    from quade_b200.pyfastq import FastqReader
    text = bytes(np.random.default_rng(6).integers(33, 126, 300000, dtype=np.uint8))
    blob = gzip.compress(text, 1)
    monkeypatch.setattr(FastqReader, "BLOCK", 8192)
    for name, bad in (("truncated", blob[:len(blob) // 2]), ("garbage_after", blob + b"not a member"),
                      ("bad_crc", blob[:-8] + b"\x00\x00\x00\x00" + blob[-4:]), ("not_gzip", b"@r1\nACGT\n+\nIIII\n")):
        p = tmp_path / (name + ".gz")
        p.write_bytes(bad)
        r = FastqReader(str(p))
        got = bytearray()
        with pytest.raises(IOError):
            while not r.eof:
                buf = np.empty(50000, dtype=np.uint8)
                n = r.readinto(buf)
                got += buf[:n].tobytes()
        assert text.startswith(bytes(got)), name                        # nothing wrong was handed out before the error
        if name == "truncated":
            assert len(got) > 0
        r.close()
    r = FastqReader(str(tmp_path / "truncated.gz"))                     # closing with blocks still in flight
    r.readinto(np.empty(10, dtype=np.uint8))
    r.close()
    with pytest.raises(IOError):
        FastqReader(str(tmp_path / "missing.gz"))


def test_feeder_windows_reassemble_the_file(tmp_path):
    """quade_b200/feeder.py: whatever the window limits and the bytes a batch consumes, the consumed bytes are the
    file, in order, through any number of wraps of the buffer."""
    from quade_b200.feeder import Feeder
    from quade_b200.pyfastq import FastqReader
    rng = np.random.default_rng(11)
    data = bytes(rng.integers(0, 256, 200_000, dtype=np.uint8))
    p = tmp_path / "blob.fastq"
    p.write_bytes(data)
    for head, ring, chunk in ((1024, 3, 512), (4096, 3, 2048), (64, 2, 37), (1 << 16, 3, 1 << 15)):
        buf = np.zeros(2 * head * (1 + ring), dtype=np.uint8)          # two halves of head + ring * head bytes
        f = Feeder(FastqReader(str(p)), buf, head, chunk)
        out = bytearray()
        stuck = 0
        while True:
            limit = int(rng.integers(1, 2 * head + 1))
            cap = int(rng.integers(0, 3 * head))
            f.wait_for(limit)
            view, last = f.window(limit, cap)
            assert len(view) <= max(limit, cap)
            take = int(rng.integers(0, len(view) + 1)) if len(view) else 0
            if rng.random() < 0.3:
                take = len(view)
            out += view[:take].tobytes()
            rest = f.consume(take)
            if last and take == len(view):
                assert rest == 0
                break
            stuck = stuck + 1 if take == 0 else 0
            assert stuck < 200
        f.close()
        assert bytes(out) == data, (head, ring, chunk)


def test_feeder_surfaces_reader_errors(tmp_path):
    from quade_b200.feeder import Feeder

    class Broken:
        eof = False

        def readinto(self, arr):
            raise IOError("disk gone")

    f = Feeder(Broken(), np.zeros(8192, dtype=np.uint8), 1024, 512)
    with pytest.raises(IOError):
        f.wait_for(100)
    f.close()
