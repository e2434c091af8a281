"""Shared helpers for the test-suite: golden-case loading and oracle drivers."""
import gzip
import json
import os

import numpy as np

from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")


def load_cases():
    with open(os.path.join(GOLDEN, "adversarial.json")) as fh:
        return json.load(fh)["cases"]


def spans_from_conf(c):
    """1-based inclusive config positions -> python slices, disabled sub-indexes -> 0:0
    (src/Quade.py:96-116)."""
    idx2 = bool(c["index2"])
    mol1 = bool(c["molecular1"])
    mol2 = idx2 and bool(c["molecular2"])
    def sp(on, key):
        return (c["pos"][key][0] - 1, c["pos"][key][1]) if on else (0, 0)
    return [sp(True, "index1"), sp(idx2, "index2"), sp(mol1, "molecular1"), sp(mol2, "molecular2")]


def sample_keys(case):
    dual = bool(case["conf"]["index2"])
    return [(s[1] + s[2]) if dual else s[1] for s in case["samples"]]


def oracle_params(case):
    c = case["conf"]
    return O.OracleParams(dual=c["index2"], spans=spans_from_conf(c), min_qual=c["minimal_qual"],
                          sample_keys=sample_keys(case), write_undetermined=c["write_undetermined"],
                          write_pass=c["write_pass"], write_fail=c["write_fail"])


def case_chunks(case):
    return [[s.encode("latin-1") for s in ch] for ch in case["chunks"]]


def expected_files(case):
    return {k: v.encode("latin-1") for k, v in case["expect"]["files"].items()}


def files_from_bins(names, per_bin):
    """per_bin: list over bins of (R1 bytes, R2 bytes) accumulated over chunks -> {filename: bytes} for the
    files the reference would have created (a bin file exists iff it received a record, FastqWriter.py:55-57)."""
    out = {}
    for n, (a, b) in zip(names, per_bin):
        if len(a) or len(b):
            out[n + "_R1.fastq.gz"] = bytes(a)
            out[n + "_R2.fastq.gz"] = bytes(b)
    return out


def run_oracle_chunks(params, chunks):
    nb = params.nbins
    acc = [[bytearray(), bytearray()] for _ in range(nb)]
    counters = np.zeros(nb, dtype=np.int64)
    for ch in chunks:
        r = O.demux_batch(params, ch, last=True)
        counters += r.counters
        for b in range(nb):
            acc[b][0] += r.bin_bytes(0, b)
            acc[b][1] += r.bin_bytes(1, b)
    return acc, counters


def quade_dataset():
    """The reference's golden vector: 3 chunks x (R1, R4, R2, R3) in the template's wiring
    (src/Conf_file.py:46-49), decompressed."""
    d = os.path.join(GOLDEN, "quade_dataset", "dataset")
    def rd(n):
        with gzip.open(os.path.join(d, n), "rb") as fh:
            return fh.read()
    chunks = [[rd("C%d_R1.fastq.gz" % i), rd("C%d_R4.fastq.gz" % i), rd("C%d_R2.fastq.gz" % i), rd("C%d_R3.fastq.gz" % i)]
              for i in (1, 2, 3)]
    res = os.path.join(GOLDEN, "quade_dataset", "result")
    exp = {}
    for f in os.listdir(res):
        if f.endswith(".fastq.gz"):
            with gzip.open(os.path.join(res, f), "rb") as fh:
                exp[f] = fh.read()
    with open(os.path.join(res, "Quade_report.csv"), newline="") as fh:
        report = fh.readlines()[1:]
    case = {"conf": {"minimal_qual": 25, "index2": True, "molecular1": True, "molecular2": True,
                     "pos": {"index1": [1, 4], "index2": [1, 4], "molecular1": [4, 6], "molecular2": [4, 6]},
                     "write_pass": True, "write_fail": True, "write_undetermined": True},
            "samples": [["S1", "ACAG", "ACAG"], ["S2", "CTTG", "CTTG"]]}
    return case, chunks, exp, report


class OracleEngine:
    """Stand-in for DemuxContext in CPU-only tests of the HOST logic (pipeline, carry-over, writers, sharding):
    the same submit / peek / collect surface, answers computed by the oracle.  Test infrastructure only: installed
    by monkeypatching Quade._open_engine / Quade._alloc_staging (see patch_host_engine)."""

    class _Res:
        def __init__(self, r, ticket):
            self.ticket = ticket
            self.n_pairs, self.consumed, self.n_skipped = r.n_pairs, list(r.consumed) + [0] * (4 - len(r.consumed)), list(r.n_skipped) + [0] * (4 - len(r.n_skipped))
            self.chunk_end = r.chunk_end
            self.counters, self.bin_off, self._r = r.counters, r.bin_off, r
            self._arena = [np.frombuffer(r.out[k], dtype=np.uint8) if len(r.out[k]) else np.zeros(1, dtype=np.uint8) for k in range(2)]
            self.out_ptr = [a.ctypes.data for a in self._arena]

        def bin_view(self, read, b):
            return memoryview(self._r.bin_bytes(read, b))

    def __init__(self, quade, device=0):
        from quade_b200.sample import Sample
        self.params = O.OracleParams(quade.idx2, quade.spans(), quade.minimal_qual, [s.index for s in Sample.SAMPLE_LIST],
                                     Sample.WRITE_UNDETERMINED, Sample.WRITE_PASS, Sample.WRITE_FAIL)
        self.batches = 0
        self._q = []

    @property
    def in_flight(self):
        return len(self._q)

    def submit(self, streams, last=True, ticket=0):
        self.batches += 1
        try:
            r = O.demux_batch(self.params, [bytes(s) for s in streams], last=list(last) if isinstance(last, (list, tuple)) else last)
            self._q.append(self._Res(r, ticket))
        except O.OracleError as e:
            err = RuntimeError("oracle error %d" % e.code)
            err.code = e.code
            self._q.append(err)

    def peek(self):
        if isinstance(self._q[0], Exception):
            raise self._q[0]
        return self._q[0]

    def collect(self):
        r = self._q.pop(0)
        if isinstance(r, Exception):
            raise r
        return r

    def collect_discard(self):
        if self._q:
            self._q.pop(0)

    def run(self, streams, last=True):
        self.submit(streams, last)
        return self.collect()

    def close(self):
        pass


class _PlainBuffer:
    def __init__(self, n):
        self.array = np.empty(n, dtype=np.uint8)

    def free(self):
        self.array = None


def patch_host_engine(Quade, monkeypatch=None):
    """Replace the device of Quade by the oracle (and the pinned staging memory by numpy arrays) for CPU-only tests of
    the host pipeline.  The product class has no switch for this: the methods are patched from here."""
    def _open_engine(self):
        if self._engine is None:
            self._engine = OracleEngine(self, self.device)
        return self._engine

    def _alloc_staging(self, nbytes):
        return _PlainBuffer(nbytes)
    if monkeypatch is not None:
        monkeypatch.setattr(Quade, "_open_engine", _open_engine)
        monkeypatch.setattr(Quade, "_alloc_staging", _alloc_staging)
    else:
        Quade._open_engine = _open_engine
        Quade._alloc_staging = _alloc_staging
