"""Pins the C oracle: (1) the reference's own golden vector, (2) reference-generated adversarial vectors,
(3) when /root/reference is present, a fresh differential run against the reference source itself."""
import numpy as np
import pytest

from oracle import oracle as O
from tests import util


def test_oracle_matches_reference_golden_dataset():
    case, chunks, exp, report = util.quade_dataset()
    p = util.oracle_params(case)
    acc, counters = util.run_oracle_chunks(p, chunks)
    names = O.bin_names([s[0] for s in case["samples"]])
    got = util.files_from_bins(names, acc)
    assert sorted(got) == sorted(exp)           # same file set: no *_fail_* files
    for f in exp:
        assert got[f] == exp[f], f
    assert int(counters.sum()) == 299           # C1_R1 record 0 is malformed -> desync, 99 pairs in chunk 1
    assert O.report_lines([s[0] for s in case["samples"]], counters) == report


@pytest.mark.parametrize("case", util.load_cases(), ids=lambda c: c["name"])
def test_oracle_matches_reference_adversarial(case):
    p = util.oracle_params(case)
    acc, counters = util.run_oracle_chunks(p, util.case_chunks(case))
    names = O.bin_names([s[0] for s in case["samples"]])
    got = util.files_from_bins(names, acc)
    exp = util.expected_files(case)
    assert sorted(got) == sorted(exp)
    for f in exp:
        assert got[f] == exp[f], f
    assert O.report_lines([s[0] for s in case["samples"]], counters) == case["expect"]["report"]


def test_oracle_batching_is_invariant():
    """Feeding a chunk in arbitrary byte batches with carry-over gives the same bins as one shot."""
    rng = np.random.default_rng(7)
    for case in util.load_cases():
        p = util.oracle_params(case)
        for ch in util.case_chunks(case):
            ref = O.demux_batch(p, ch, last=True)
            ns = len(ch)
            pos = [0] * ns
            acc = [[bytearray(), bytearray()] for _ in range(p.nbins)]
            counters = np.zeros(p.nbins, dtype=np.int64)
            avail = [0] * ns
            while True:
                last = True
                for s in range(ns):
                    avail[s] = min(len(ch[s]), max(avail[s], pos[s]) + int(rng.integers(1, 400)))
                    last = last and avail[s] == len(ch[s])
                r = O.demux_batch(p, [ch[s][pos[s]:avail[s]] for s in range(ns)], last=last)
                counters += r.counters
                for b in range(p.nbins):
                    acc[b][0] += r.bin_bytes(0, b)
                    acc[b][1] += r.bin_bytes(1, b)
                for s in range(ns):
                    pos[s] += r.consumed[s]
                if last:
                    break
            assert (counters == ref.counters).all(), case["name"]
            for b in range(p.nbins):
                assert bytes(acc[b][0]) == ref.bin_bytes(0, b) and bytes(acc[b][1]) == ref.bin_bytes(1, b), case["name"]


def test_blank_header_is_fatal():
    p = O.OracleParams(True, [(0, 4), (0, 4), (0, 0), (0, 0)], 0, ["ACAGACAG"])
    good = b"@a\nACGT\n+\nIIII\n"
    with pytest.raises(O.OracleError) as e:
        O.demux_batch(p, [b"@\nACGT\n+\nIIII\n", good, good, good])
    assert e.value.code == O.ERR_BLANK_HEADER
