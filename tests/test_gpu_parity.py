"""GPU parity: the CUDA path (through the C-ABI) against the pinned C oracle and the committed golden vectors.
Bit-exact: byte-identical per-bin streams, identical counters / consumed / skipped."""
import numpy as np
import pytest

from oracle import oracle as O
from tests import util

pytestmark = pytest.mark.gpu


def _ctx_for(case_or_params, **kw):
    from quade_b200 import DemuxContext
    p = case_or_params if isinstance(case_or_params, O.OracleParams) else util.oracle_params(case_or_params)
    return DemuxContext(p.dual, p.spans, p.min_qual, p.sample_keys, p.write_undetermined, p.write_pass, p.write_fail,
                        **kw), p


def _assert_same(res, ref, nbins, tag=""):
    assert res.n_pairs == ref.n_pairs, tag
    assert (res.counters == ref.counters).all(), tag
    assert res.consumed[:len(ref.consumed)] == ref.consumed, tag
    assert res.n_skipped[:len(ref.n_skipped)] == ref.n_skipped, tag
    for k in range(2):
        assert (res.bin_off[k] == ref.bin_off[k]).all(), tag
        assert res.arena(k).tobytes() == ref.out[k], tag


def test_golden_dataset_byte_identical():
    """The reference's own golden vector (test/dataset -> test/result), incl. the skipped-record desync (299 pairs)."""
    case, chunks, exp, report = util.quade_dataset()
    ctx, p = _ctx_for(case, max_stream_bytes=1 << 20, max_pairs=4096)
    acc = [[bytearray(), bytearray()] for _ in range(p.nbins)]
    for ch in chunks:
        r = ctx.run(ch, last=True)
        for b in range(p.nbins):
            acc[b][0] += r.bin_bytes(0, b)
            acc[b][1] += r.bin_bytes(1, b)
    names = O.bin_names([s[0] for s in case["samples"]])
    got = util.files_from_bins(names, acc)
    assert sorted(got) == sorted(exp)
    for f in exp:
        assert got[f] == exp[f], f
    counters = ctx.counters()
    assert int(counters.sum()) == 299
    assert O.report_lines([s[0] for s in case["samples"]], counters) == report
    ctx.close()


@pytest.mark.parametrize("case", util.load_cases(), ids=lambda c: c["name"])
def test_adversarial_cases_match_reference_and_oracle(case):
    ctx, p = _ctx_for(case, max_stream_bytes=1 << 20, max_pairs=4096)
    acc = [[bytearray(), bytearray()] for _ in range(p.nbins)]
    for ch in util.case_chunks(case):
        ref = O.demux_batch(p, ch, last=True)
        r = ctx.run(ch, last=True)
        _assert_same(r, ref, p.nbins, case["name"])
        for b in range(p.nbins):
            acc[b][0] += r.bin_bytes(0, b)
            acc[b][1] += r.bin_bytes(1, b)
    names = O.bin_names([s[0] for s in case["samples"]])
    got = util.files_from_bins(names, acc)
    exp = util.expected_files(case)
    assert sorted(got) == sorted(exp)
    for f in exp:
        assert got[f] == exp[f], f
    assert O.report_lines([s[0] for s in case["samples"]], ctx.counters()) == case["expect"]["report"]
    ctx.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_synthetic_configs_match_oracle(name):
    from quade_b200.synth import Synth
    s = Synth.from_config(name)
    n = 50000
    streams = s.generate(0, n)
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ref = O.demux_batch(p, streams, last=True)
    ctx, _ = _ctx_for(p, max_stream_bytes=32 << 20, max_pairs=n + 16)
    r = ctx.run(streams, last=True)
    _assert_same(r, ref, p.nbins, name)
    assert r.n_pairs == n
    ctx.close()


def test_batched_with_carry_over_matches_one_shot():
    """A chunk fed in ragged byte batches (double-buffered submit/collect) gives the same bins."""
    from quade_b200.synth import Synth
    s = Synth.from_config("cfg3")
    n = 30000
    streams = [bytes(a) for a in s.generate(7, n)]
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ref = O.demux_batch(p, streams, last=True)
    ctx, _ = _ctx_for(p, max_stream_bytes=4 << 20, max_pairs=16384)
    rng = np.random.default_rng(3)
    pos = [0] * 4
    acc = [[bytearray(), bytearray()] for _ in range(p.nbins)]
    total = 0
    while True:
        take = [min(len(streams[k]) - pos[k], int(rng.integers(100_000, 3_000_000)) // (1 if k < 2 else 4)) for k in range(4)]
        last = all(pos[k] + take[k] == len(streams[k]) for k in range(4))
        r = ctx.run([streams[k][pos[k]:pos[k] + take[k]] for k in range(4)], last=last)
        total += r.n_pairs
        for b in range(p.nbins):
            acc[b][0] += r.bin_bytes(0, b)
            acc[b][1] += r.bin_bytes(1, b)
        for k in range(4):
            pos[k] += r.consumed[k]
        if last:
            break
    assert total == ref.n_pairs == n
    assert (ctx.counters() == ref.counters).all()
    for b in range(p.nbins):
        assert bytes(acc[b][0]) == ref.bin_bytes(0, b) and bytes(acc[b][1]) == ref.bin_bytes(1, b)
    ctx.close()


def test_blank_header_is_fatal():
    from quade_b200 import DemuxContext, QuadeError
    ctx = DemuxContext(True, [(0, 4), (0, 4), (0, 0), (0, 0)], 0, ["ACAGACAG"], max_stream_bytes=1 << 16, max_pairs=64)
    good = b"@a\nACGT\n+\nIIII\n"
    with pytest.raises(QuadeError) as e:
        ctx.run([b"@\nACGT\n+\nIIII\n", good, good, good])
    assert e.value.code == -1
    r = ctx.run([good, good, good, good])        # the context stays usable
    assert r.n_pairs == 1
    ctx.close()


def test_error_codes_of_the_c_abi():
    """include/quade_b200.h error enum through the C-ABI: every error leaves the context usable."""
    from quade_b200 import DemuxContext, QuadeError
    good = b"@a\nACGT\n+\nIIII\n"
    spans = [(0, 4), (0, 0), (0, 0), (0, 0)]
    for keys, code in ((["ACGX"], -3), (["ACGT", "ACGT"], -3), ([""], -3), (["A" * 33], -7)):   # Sample.py:140-141,171-175
        with pytest.raises(QuadeError) as e:
            DemuxContext(False, spans, 0, keys, max_stream_bytes=1 << 16, max_pairs=64)
        assert e.value.code == code, keys
    ctx = DemuxContext(False, spans, 0, ["ACGT"], max_stream_bytes=1 << 12, max_pairs=32, n_slots=1)
    with pytest.raises(QuadeError) as e:
        ctx.collect()                                           # nothing submitted
    assert e.value.code == -8
    with pytest.raises(QuadeError) as e:
        ctx.run([good * 300, good, good])                       # 4500 bytes > max_stream_bytes
    assert e.value.code == -6
    r = ctx.run([good * 200, good * 200, good * 200], last=False)   # 200 pairs > max_pairs: the rest is carried over
    assert r.n_pairs == 32 and r.consumed[:3] == [32 * len(good)] * 3
    ctx.submit([good, good, good])
    with pytest.raises(QuadeError) as e:
        ctx.submit([good, good, good])                          # the only slot is in flight
    assert e.value.code == -8
    assert ctx.collect().n_pairs == 1
    r = ctx.run([good * 32, good * 32, good * 32])              # exactly max_pairs
    assert r.n_pairs == 32 and int(r.counters.sum()) == 32
    ctx.close()


def test_empty_and_tiny_batches():
    from quade_b200 import DemuxContext
    ctx = DemuxContext(False, [(0, 4), (0, 0), (0, 0), (0, 0)], 10, ["ACAG"], max_stream_bytes=1 << 16, max_pairs=64)
    p = O.OracleParams(False, [(0, 4), (0, 0), (0, 0), (0, 0)], 10, ["ACAG"])
    for streams in ([b"", b"", b""], [b"@a\nAC\n+\nII", b"@a\nAC\n+\nII\n", b"@a\nACAG\n+\nIII5"],
                    [b"@a\nAC\n+\nII\n", b"", b"@a\nACAG\n+\nIIII\n"]):
        ref = O.demux_batch(p, streams, last=True)
        _assert_same(ctx.run(streams, last=True), ref, p.nbins, repr(streams))
        ref = O.demux_batch(p, streams, last=False)
        _assert_same(ctx.run(streams, last=False), ref, p.nbins, repr(streams))
    ctx.close()


def test_per_stream_eof_flags_match_oracle_batch_by_batch():
    """Files of a chunk reach end-of-file in different batches (QUADE_BATCH_EOF(s)); the final record of every
    file has no trailing newline.  Every batch result equals the oracle's."""
    from quade_b200.synth import Synth
    s = Synth.from_config("cfg3")
    n = 6000
    streams = [bytes(a)[:-1] for a in s.generate(11, n)]       # drop the last '\n' of every file
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ctx, _ = _ctx_for(p, max_stream_bytes=1 << 20, max_pairs=8192)
    rng = np.random.default_rng(5)
    pos, avail, total = [0] * 4, [0] * 4, 0
    for it in range(10000):
        for k in range(4):
            avail[k] = min(len(streams[k]), max(avail[k], pos[k]) + int(rng.integers(1000, 400_000)))
        eof = [avail[k] == len(streams[k]) for k in range(4)]
        batch = [streams[k][pos[k]:avail[k]] for k in range(4)]
        ref = O.demux_batch(p, batch, last=eof)
        r = ctx.run(batch, last=eof)
        _assert_same(r, ref, p.nbins, "batch %d" % it)
        total += r.n_pairs
        for k in range(4):
            pos[k] += r.consumed[k]
        if all(eof):
            break
    assert total == n
    ctx.close()


def _mutate_records(stream, rng, frac, fn):
    """Apply fn(lines[4]) -> lines[4] to a random fraction of the 4-line records of one FASTQ byte stream."""
    lines = stream.split(b"\n")
    tail = lines[-1]
    lines = lines[:-1]
    for j in range(0, len(lines) - 3, 4):
        if rng.random() < frac:
            lines[j:j + 4] = fn(lines[j:j + 4])
    return b"\n".join(lines) + b"\n" + tail


@pytest.mark.parametrize("variant", ["headers_without_at", "plus_line_text", "qual_starts_at", "invalid_and_crlf", "tiny_lines"])
def test_multi_tile_malformed_text_matches_oracle(variant):
    """Many 16 KiB tiles per file whose text defeats any guess about where a record starts: headers that do not
    start with '@' (the reference drops the first byte whatever it is, FastqReader.py:42), third lines that are
    not '+' (ignored by the reference, :43), quality strings starting with '@' or '+', records skipped for
    len(seq) != len(qual) in one file only (:48-50), CRLF, and a stretch of very short lines (line-end list
    overflow).  The GPU framing must not depend on what the text looks like."""
    from quade_b200.synth import Synth
    s = Synth.from_config("cfg3")
    n = 12000
    streams = [bytes(a) for a in s.generate(3, n)]
    rng = np.random.default_rng(17)

    def no_at(r):
        return [b"X" + r[0][1:], r[1], r[2], r[3]]

    def plus_text(r):
        return [r[0], r[1], b"@" + r[0][1:8], r[3]]             # third line looks like a header

    def qual_at(r):
        return [r[0], r[1], r[2], (b"@" if rng.random() < 0.5 else b"+") + r[3][1:]]

    def invalid(r):
        return [r[0], r[1], r[2], r[3][:-1]]                    # len(seq) != len(qual): skipped in this file only

    def crlf(r):
        return [x + b"\r" for x in r]

    if variant == "headers_without_at":
        streams = [_mutate_records(x, rng, 0.5, no_at) for x in streams]
    elif variant == "plus_line_text":
        streams = [_mutate_records(x, rng, 0.5, plus_text) for x in streams]
    elif variant == "qual_starts_at":
        streams = [_mutate_records(x, rng, 0.5, qual_at) for x in streams]
    elif variant == "invalid_and_crlf":
        streams = [_mutate_records(x, rng, 0.3, crlf) for x in streams]
        streams[1] = _mutate_records(streams[1], rng, 0.001, invalid)
        streams[2] = _mutate_records(streams[2], rng, 0.002, invalid)
    elif variant == "tiny_lines":
        tiny = b"".join(b"@t%d\nA\n+\nI\n" % i for i in range(3000))      # ~11 bytes per record, 2 bytes per line
        streams = [tiny + _mutate_records(x[x.index(b"\n@") + 1:], rng, 0.2, no_at) for x in streams]
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ref = O.demux_batch(p, streams, last=True)
    ctx, _ = _ctx_for(p, max_stream_bytes=8 << 20, max_pairs=n + 4096)
    r = ctx.run(streams, last=True)
    _assert_same(r, ref, p.nbins, variant)
    assert ref.n_pairs > 1000
    import os
    if os.environ.get("QUADE_FRAME", "10") == "10" and variant in ("headers_without_at", "plus_line_text", "tiny_lines"):
        assert r.reframed_tiles > 0, "the wrong-guess path was not exercised"
    ctx.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_full_size_batch_matches_oracle_and_is_split_invariant(name):
    """BASELINE.json's step size: ONE batch of 2 M pairs (1.5-1.7 GB of FASTQ text; cfg3 is the bench workload).
    (1) byte-identical to the C oracle on the same input; (2) conservation: every pair lands in exactly one bin and
    the output size is the size of the written records plus their name suffixes; (3) split invariance: the same
    chunk fed as three ragged batches with carry-over gives the same per-bin streams (compared by digest)."""
    import hashlib
    from quade_b200.synth import Synth
    s = Synth.from_config(name)
    n = 2_000_000
    streams = s.generate(0, n)
    sizes = [len(a) for a in streams]
    ns = len(streams)
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ref = O.demux_batch(p, streams, last=True)
    ctx, _ = _ctx_for(p, max_stream_bytes=max(sizes) + 4096, max_pairs=n + 1024, n_slots=1)
    r = ctx.run(streams, last=True)
    _assert_same(r, ref, p.nbins, "2 M pairs")
    assert r.n_pairs == n and int(r.counters.sum()) == n
    assert r.reframed_tiles <= 2 * ns, "only the tiles at the end of each buffer take the generic pass on well-formed input"
    one_shot = [[hashlib.blake2b(r.bin_bytes(k, b), digest_size=16).digest() for b in range(p.nbins)] for k in range(2)]
    counters = r.counters.copy()
    # three ragged batches: cut points are arbitrary byte positions, different in every file
    views = [memoryview(a) for a in streams]
    cuts = [[0, int(sz * 0.31) + 17 * k, int(sz * 0.64) + 5 * k, sz] for k, sz in enumerate(sizes)]
    h = [[hashlib.blake2b(digest_size=16) for _ in range(p.nbins)] for _ in range(2)]
    pos, total, csum = [0] * ns, 0, np.zeros_like(counters)
    for i in range(1, 4):
        last = i == 3
        rb = ctx.run([views[k][pos[k]:cuts[k][i]] for k in range(ns)], last=last)
        total += rb.n_pairs
        csum += rb.counters
        for k in range(2):
            for b in range(p.nbins):
                h[k][b].update(rb.bin_bytes(k, b))
        for k in range(ns):
            pos[k] += rb.consumed[k]
    assert total == n and (csum == counters).all()
    assert [[x.digest() for x in row] for row in h] == one_shot
    ctx.close()


@pytest.mark.parametrize("max_len", [3000, 40000])
def test_long_reads_spanning_tiles_match_oracle(max_len):
    """Insert reads far longer than Illumina's: records larger than the 2 KiB halo of a framing tile and (40000)
    larger than a whole 16 KiB tile, so that a record's four lines lie in different tiles; long names and
    descriptions; index reads stay short.  Same bytes as the oracle, whatever path the tiles take."""
    rng = np.random.default_rng(max_len)
    n = 400
    keys = ["ACGTACGT", "TTGCAAGC", "GGATCCAA"]
    p = O.OracleParams(False, [(0, 8), (0, 0), (0, 0), (0, 0)], 20, keys)
    bases = np.frombuffer(b"ACGTN", dtype=np.uint8)
    r1, r2, i1 = [], [], []
    for j in range(n):
        name = b"read%d/%s" % (j, b"x" * int(rng.integers(0, 120)))
        for out in (r1, r2):
            L = int(rng.integers(1, max_len))
            seq = bases[rng.integers(0, 5, L)].tobytes()
            qual = (rng.integers(33, 74, L).astype(np.uint8)).tobytes()
            out.append(b"@" + name + b" desc " + b"d" * int(rng.integers(0, 60)) + b"\n" + seq + b"\n+" +
                       (name if j % 7 == 0 else b"") + b"\n" + qual + b"\n")
        idx = keys[j % 3].encode() if j % 4 else bases[rng.integers(0, 4, 8)].tobytes()
        iq = (rng.integers(33 + 15, 74, 8).astype(np.uint8)).tobytes()
        i1.append(b"@" + name + b"\n" + idx + b"\n+\n" + iq + b"\n")
    streams = [b"".join(r1), b"".join(r2), b"".join(i1)]
    ref = O.demux_batch(p, streams, last=True)
    ctx, _ = _ctx_for(p, max_stream_bytes=max(len(x) for x in streams) + 4096, max_pairs=n + 16)
    r = ctx.run(streams, last=True)
    _assert_same(r, ref, p.nbins, "long reads %d" % max_len)
    assert r.n_pairs == n
    ctx.close()


@pytest.mark.parametrize("name", ["cfg2", "cfg3", "cfg4", "cfg5"])
def test_device_generator_writes_the_bytes_of_the_host_generator(name):
    """quade_synth_fill_device (CUDA kernels, HBM-resident buffers: what bench.py streams) == quade_synth_fill (host
    threads: what the oracle is fed) for 1 M pairs starting in the middle of the data set."""
    import torch
    from quade_b200.synth import Synth
    s = Synth.from_config(name)
    first, n = 3_000_017, 1_000_000
    host = s.generate(first, n)
    sizes = [len(a) for a in host]
    dev = [torch.zeros(x + 64, dtype=torch.uint8, device="cuda") for x in sizes]
    assert s.fill_device(first, n, [t.data_ptr() for t in dev]) == sizes
    for k in range(s.n_streams):
        assert torch.equal(dev[k][:sizes[k]].cpu(), torch.from_numpy(host[k])), "stream %d" % k
        assert int(dev[k][sizes[k]:].sum()) == 0, "the generator wrote past the end of stream %d" % k


def _bin_checksums(arena, bin_off):
    """per bin: (bytes, 64-bit running checksum of the byte stream: two CRC-32 with different seeds)"""
    import zlib
    out = []
    mv = memoryview(arena)
    for b in range(len(bin_off) - 1):
        seg = mv[int(bin_off[b]):int(bin_off[b + 1])]
        out.append((len(seg), zlib.crc32(seg) | zlib.crc32(seg, 0x9E3779B9) << 32))
    return out


def test_sampled_chunks_of_the_scale_config_match_the_oracle_by_checksum():
    """BASELINE config 4 (dual 10+10, sharded over 8 GPUs, 2 B reads) cannot be compared byte for byte as a whole:
    chunks sampled from the ranks' shares are generated ON THE DEVICE (as bench.py does), demultiplexed from HBM, and
    compared with the oracle fed by the host generator through per-bin byte counts, a 64-bit running checksum per
    bin and read, and the counter vector (SURVEY.md section 8(d) "parity at scale")."""
    import torch
    from quade_b200.synth import Synth
    s = Synth.from_config("cfg4")
    P = 500_000
    p = O.OracleParams(s.dual, s.spans(), s.min_qual, s.sample_keys())
    ctx = None
    for chunk in (0, 777, 3999):                                  # 4000 chunks x 500 k pairs = 2 B pairs
        first = chunk * P
        sizes = s.sizes(first, P)
        if ctx is None:
            ctx, _ = _ctx_for(p, max_stream_bytes=max(sizes) + (1 << 20), max_pairs=P + 1024, n_slots=1)
        dev = [torch.empty(x + 64, dtype=torch.uint8, device="cuda") for x in sizes]
        assert s.fill_device(first, P, [t.data_ptr() for t in dev]) == sizes
        host_in = [t[:n].cpu().numpy() for t, n in zip(dev, sizes)]          # D2H of the device-made inputs
        r = ctx.run(host_in, last=True)
        ref = O.demux_batch(p, s.generate(first, P), last=True)
        assert r.n_pairs == ref.n_pairs == P and (r.counters == ref.counters).all()
        for k in range(2):
            assert (r.bin_off[k] == ref.bin_off[k]).all()
            assert _bin_checksums(r.arena(k), r.bin_off[k]) == _bin_checksums(np.frombuffer(ref.out[k], dtype=np.uint8), ref.bin_off[k])
    ctx.close()


def test_allreduce_counters_c_abi():
    """quade_allreduce_counters: the single-process form of the path's only collective (one context per GPU, NCCL).
    With one GPU it is the identity and refuses two contexts on one device; with two it sums."""
    import torch
    from quade_b200 import DemuxContext, QuadeError
    from quade_b200.demux import allreduce_counters
    good = b"@a\nACGT\n+\nIIII\n"
    spans = [(0, 4), (0, 0), (0, 0), (0, 0)]
    n_dev = torch.cuda.device_count()
    ctxs = [DemuxContext(False, spans, 0, ["ACGT", "TTTT"], device=d % n_dev, max_stream_bytes=1 << 16, max_pairs=64) for d in range(2)]
    ctxs[0].run([good * 3, good * 3, good * 3])
    ctxs[1].run([good * 5, good * 5, good.replace(b"ACGT", b"GGGG") * 5])
    allreduce_counters(ctxs[:1])
    assert list(ctxs[0].counters()) == [0, 3, 0, 0, 0]
    if n_dev < 2:
        with pytest.raises(QuadeError) as e:
            allreduce_counters(ctxs)
        assert e.value.code == -5                     # two contexts on one device
    else:
        allreduce_counters(ctxs)
        assert list(ctxs[0].counters()) == list(ctxs[1].counters()) == [5, 3, 0, 0, 0]
    for c in ctxs:
        c.close()
