"""Generated FASTQ text through the CUDA path against the oracle (last bullet).

hypothesis builds small chunks out of the things real files do and the reference tolerates or trips over: CRLF line
ends, tabs and blanks in and after headers, blank lines, '@' / '+' at the start of quality strings, decorated '+'
lines, lower-case and N bases, index reads shorter than the configured spans, records with len(seq) != len(qual)
(skipped in their own file only, FastqReader.py:48-50), a missing final newline, and files with different record
counts.  Every chunk goes through one context as ONE batch and again sliced into ragged batches with carry-over
(per-file end-of-file flags); counters, per-bin streams, consumed bytes and skipped counts must equal the oracle's.
Headers without a name token are generated too: both sides must refuse them (IndexError in the reference).
"""
import os

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

from oracle import oracle as O

pytestmark = pytest.mark.gpu

SPANS = [(0, 4), (1, 5), (3, 6), (4, 6)]
KEYS = ["ACAGCAGT", "CTTGTTGA", "ANNTACGN", "GGGGCCCC"]
BASES = st.sampled_from(list("ACGTacgtNn") + ["R", "-"])
QUALS = st.sampled_from(list("!+5:;?@II#"))
EOL = st.sampled_from(["\n", "\n", "\n", "\r\n", " \n", "\t\n"])


@st.composite
def record(draw, index_read):
    name = draw(st.sampled_from(["r%d" % draw(st.integers(0, 99)), "a:b/1", "x" * 40, "q\x01z"] * 12 + ["", " ", "\t"]))
    lead = draw(st.sampled_from(["@", "@", "@", "X", "@ ", "@\t"]))
    descr = draw(st.sampled_from(["", " 1:N:0", "\tdesc", "  "]))
    n = draw(st.integers(0, 9) if index_read else st.integers(0, 30))
    seq = "".join(draw(st.lists(BASES, min_size=n, max_size=n)))
    qn = n if draw(st.integers(0, 19)) else max(0, n + draw(st.sampled_from([-1, 1])))      # 5 %: invalid record
    qual = "".join(draw(st.lists(QUALS, min_size=qn, max_size=qn)))
    plus = draw(st.sampled_from(["+", "+", "+" + name, "", "@plus"]))
    return (lead + name + descr + draw(EOL) + seq + draw(EOL) + plus + draw(EOL) + qual + draw(EOL)).encode("latin-1")


@st.composite
def chunk(draw):
    streams = []
    for s in range(4):
        recs = draw(st.lists(record(index_read=s >= 2), min_size=draw(st.sampled_from([0, 3, 5, 8])), max_size=24))
        text = b"".join(recs)
        if text and draw(st.booleans()):
            text = text.rstrip(b"\n")                       # final line without '\n' still is a line (FastqReader.py:37)
        if draw(st.integers(0, 9)) == 0:
            text += draw(st.sampled_from([b"\n", b"@partial\nACGT\n", b"\n\n\n"]))
        streams.append(text)
    return streams


def _ctx():
    from quade_b200 import DemuxContext
    return DemuxContext(True, SPANS, 25, KEYS, max_stream_bytes=1 << 16, max_pairs=256, n_slots=1)


_CTX = {}


def _shared_ctx():
    if "c" not in _CTX:
        _CTX["c"] = _ctx()
    return _CTX["c"]


def _same(r, ref, tag):
    assert r.n_pairs == ref.n_pairs, tag
    assert (r.counters == ref.counters).all(), tag
    assert r.consumed[:4] == ref.consumed and r.n_skipped[:4] == ref.n_skipped, tag
    assert r.chunk_end == ref.chunk_end, tag
    for k in range(2):
        assert (r.bin_off[k] == ref.bin_off[k]).all(), tag
        assert r.arena(k).tobytes() == ref.out[k], tag


@settings(max_examples=int(os.environ.get("QUADE_FUZZ_EXAMPLES", "300")), deadline=None, suppress_health_check=list(HealthCheck), derandomize=True)
@given(chunk(), st.lists(st.integers(1, 400), min_size=4, max_size=4))
def test_generated_fastq_matches_the_oracle(streams, steps):
    from quade_b200 import QuadeError
    ctx = _shared_ctx()
    p = O.OracleParams(True, SPANS, 25, KEYS)
    # (1) the whole chunk as one batch
    try:
        ref = O.demux_batch(p, streams, last=True)
    except O.OracleError as e:
        assert e.code == O.ERR_BLANK_HEADER
        with pytest.raises(QuadeError) as g:
            ctx.run(streams, last=True)
        assert g.value.code == -1
        return
    _same(ctx.run(streams, last=True), ref, "one batch")
    # (2) ragged batches with carry-over: file s is fed `steps[s]` new bytes at a time
    pos, buf = [0] * 4, [b""] * 4
    total = 0
    for _ in range(400):
        for s in range(4):
            buf[s] += streams[s][pos[s]:pos[s] + steps[s]]
            pos[s] += steps[s]
        eof = [pos[s] >= len(streams[s]) for s in range(4)]
        try:
            rb = O.demux_batch(p, buf, last=eof)
        except O.OracleError:
            with pytest.raises(QuadeError):
                ctx.run(buf, last=eof)
            return
        r = ctx.run(buf, last=eof)
        _same(r, rb, "batched")
        total += r.n_pairs
        buf = [buf[s][r.consumed[s]:] for s in range(4)]
        if r.chunk_end:
            break
    else:
        raise AssertionError("the batched run did not end")
    assert total == ref.n_pairs
