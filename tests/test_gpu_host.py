"""GPU twin of tests/test_host.py: the reference's documented smoke test (README.md:67-72) through
Quade.__call__ with the real CUDA engine, compared with the reference's golden outputs."""
import os
import subprocess
import sys

import pytest

from tests.test_host import ROOT, _check_golden_outputs, _fresh, _stage_dataset

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("batch_bytes", [8 << 20, 4096])
def test_golden_run_on_the_gpu(tmp_path, monkeypatch, batch_bytes):
    from quade_b200.quade import Quade
    _fresh()
    wd = _stage_dataset(tmp_path)
    monkeypatch.chdir(wd)
    monkeypatch.setattr(Quade, "BATCH_BYTES", batch_bytes)
    q = Quade(conf_file="Quade_conf_file.txt")
    assert q() == 0
    _check_golden_outputs(wd)


def test_command_line_entry(tmp_path):
    wd = _stage_dataset(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, QUADE_BATCH_BYTES=str(1 << 20))
    out = subprocess.run([sys.executable, os.path.join(ROOT, "Quade.py"), "-c", "Quade_conf_file.txt"], cwd=str(wd),
                         env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "Start parsing files: 3 chunks to be parsed" in out.stdout and "Done in" in out.stdout
    assert out.stdout.count("Skipping the sequence") == 1          # C1_R1 record 0
    _check_golden_outputs(wd)


def test_sharded_run_two_gpus_nccl(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (the gloo twin runs in tests/test_host.py)")
    wd = _stage_dataset(tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT, QUADE_BATCH_BYTES=str(1 << 20))
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", "29642", "-m", "quade_b200", "-c",
                          "Quade_conf_file.txt"], cwd=str(wd), env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stdout + out.stderr
    _check_golden_outputs(wd)


@pytest.mark.parametrize("gz_in,level", [(False, 1), (True, 0)])
def test_overlapped_pipeline_three_chunks_many_batches_matches_oracle(tmp_path, monkeypatch, gz_in, level):
    """The whole tool through the overlapped path: 3 chunks of 60 k synthetic cfg3 pairs as files (plain or gzip),
    staged 4 MiB at a time (>= 5 batches per chunk, three in flight: reader threads, H2D + kernels, D2H, native gzip
    sink), 769 bins.  EVERY output file is compared, decompressed, with the oracle's per-bin streams over the chunks in
    order; so are the file set (lazy creation) and the report counters."""
    import gzip
    import numpy as np
    from oracle import oracle as O
    from quade_b200.fastq_writer import FastqWriter
    from quade_b200.quade import Quade
    from quade_b200.sample import Sample
    from quade_b200.synth import Synth
    _fresh()
    s = Synth.from_config("cfg3")
    per, chunks = 60_000, 3
    wd = tmp_path / "run"
    wd.mkdir()
    monkeypatch.chdir(wd)
    names = [[] for _ in range(4)]
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    acc = [[bytearray(), bytearray()] for _ in range(p.nbins)]
    counters = np.zeros(p.nbins, dtype=np.int64)
    for c in range(chunks):
        arrays = s.generate(c * per, per)
        for k, a in enumerate(arrays):
            fn = "C%d_%d.fastq%s" % (c, k, ".gz" if gz_in else "")
            if gz_in:
                with gzip.open(fn, "wb", compresslevel=1) as fh:
                    fh.write(memoryview(a))
            else:
                a.tofile(fn)
            names[k].append(fn)
        ref = O.demux_batch(p, arrays, last=True)
        counters += ref.counters
        for b in range(p.nbins):
            acc[b][0] += ref.bin_bytes(0, b)
            acc[b][1] += ref.bin_bytes(1, b)
    with open("conf.txt", "w") as fh:
        fh.write(s.conf_text(names))
    monkeypatch.setattr(Quade, "BATCH_BYTES", 4 << 20)
    monkeypatch.setattr(FastqWriter, "COMPRESSLEVEL", level)
    assert Quade(conf_file="conf.txt")() == 0
    assert Sample.COUNTS() == [int(x) for x in counters]
    bins = O.bin_names(s.sample_names())
    expect = {}
    for n, (a, b) in zip(bins, acc):
        if len(a) or len(b):
            expect[n + "_R1.fastq.gz"], expect[n + "_R2.fastq.gz"] = bytes(a), bytes(b)
    got = sorted(f for f in os.listdir(".") if f.endswith(".fastq.gz") and not f.startswith("C"))
    assert got == sorted(expect)
    for f in got:
        with gzip.open(f, "rb") as fh:
            assert fh.read() == expect[f], f


def _short_index_chunk(n, seed=3):
    """n pairs whose index reads are 30-byte records (SRA-style headers, 4-base index): lines of 7.5 bytes on
    average, five times the records per byte of the insert reads."""
    import numpy as np
    rng = np.random.default_rng(seed)
    keys = ["ACGT", "TTGA", "CCAT"]
    r1, r2, i1 = [], [], []
    for j in range(n):
        name = b"@S.%d" % j
        seq = bytes(rng.choice(list(b"ACGT"), 60).tolist())
        r1.append(name + b" 1\n" + seq + b"\n+\n" + b"I" * 60 + b"\n")
        r2.append(name + b" 2\n" + seq[::-1] + b"\n+\n" + b"F" * 60 + b"\n")
        idx = keys[j % 3].encode() if j % 7 else b"GGGG"
        i1.append(b"@%d\n" % (j % 1000) + idx + b"\n+\n" + (b"IIII" if j % 5 else b"I#II") + b"\n")
    return [b"".join(r1), b"".join(r2), b"".join(i1)], keys


def test_short_index_records_fill_batches_without_capacity_errors(tmp_path, monkeypatch):
    """An index file of ~17-byte records beside 140-byte insert records, staged in small batches: the files are read
    in proportion to their bytes per record, the index tiles hold more line ends than the fast pass lists (redo
    path), and nothing overflows.  Outputs equal the oracle's."""
    import gzip
    from oracle import oracle as O
    from quade_b200.quade import Quade
    from quade_b200.sample import Sample
    _fresh()
    n = 40_000
    streams, keys = _short_index_chunk(n)
    wd = tmp_path / "run"
    wd.mkdir()
    monkeypatch.chdir(wd)
    names = ["R1.fastq", "R2.fastq", "I1.fastq"]
    for f, s in zip(names, streams):
        (wd / f).write_bytes(s)
    conf = ["[quality]", "minimal_qual : 20", "[fastq]", "seq_R1 : R1.fastq", "seq_R2 : R2.fastq", "index_R1 : I1.fastq",
            "[index]", "index2 : False", "molecular1 : False", "molecular2 : False", "index1_start : 1", "index1_end : 4",
            "[output]", "write_pass : True", "write_fail : True", "write_undetermined : True"]
    for i, k in enumerate(keys):
        conf += ["[sample%d]" % (i + 1), "name : S%d" % (i + 1), "index1_seq : " + k]
    (wd / "conf.txt").write_text("\n".join(conf))
    monkeypatch.setattr(Quade, "BATCH_BYTES", 1 << 20)
    assert Quade(conf_file="conf.txt")() == 0
    p = O.OracleParams(False, [(0, 4), (0, 0), (0, 0), (0, 0)], 20, keys)
    ref = O.demux_batch(p, streams, last=True)
    assert Sample.COUNTS() == [int(x) for x in ref.counters] and Sample.TOTAL == n
    for b, name in enumerate(O.bin_names(["S1", "S2", "S3"])):
        for k in range(2):
            exp = ref.bin_bytes(k, b)
            path = wd / ("%s_R%d.fastq.gz" % (name, k + 1))
            if len(exp):
                with gzip.open(path, "rb") as fh:
                    assert fh.read() == exp, path
            else:
                assert not path.exists()
