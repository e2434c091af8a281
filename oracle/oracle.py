"""ctypes front-end of the C oracle (oracle/quade_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs.
The product package never imports it.  Parity status: pinned (see quade_oracle.h).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

SPAN_IDX1, SPAN_IDX2, SPAN_MOL1, SPAN_MOL2 = range(4)
ERR_BLANK_HEADER = -1


class _Params(C.Structure):
    _fields_ = [("dual", C.c_int32), ("span", (C.c_int32 * 2) * 4), ("min_qual", C.c_int32),
                ("write_undetermined", C.c_int32), ("write_pass", C.c_int32), ("write_fail", C.c_int32),
                ("n_samples", C.c_int32), ("keys", C.c_void_p), ("key_off", C.c_void_p)]


class _Result(C.Structure):
    _fields_ = [("out", C.c_void_p * 2), ("bin_off", C.c_void_p * 2), ("counters", C.c_void_p),
                ("n_pairs", C.c_uint64), ("consumed", C.c_uint64 * 4), ("n_skipped", C.c_uint64 * 4),
                ("nbins", C.c_int32), ("err", C.c_int32), ("chunk_end", C.c_int32), ("reserved", C.c_int32)]


def build(force=False):
    so = os.path.join(_HERE, "libquade_oracle.so")
    src = os.path.join(_HERE, "quade_oracle.c")
    if force or not os.path.exists(so) or (os.path.exists(src) and os.path.getmtime(src) > os.path.getmtime(so)):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = C.CDLL(build())
        _LIB.qo_demux_batch.restype = C.c_int
        _LIB.qo_demux_batch.argtypes = [C.POINTER(_Params), C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.c_int,
                                        C.POINTER(_Result)]
        _LIB.qo_result_free.argtypes = [C.POINTER(_Result)]
        _LIB.qo_demux_chunks_mt.restype = C.c_int64
        _LIB.qo_demux_chunks_mt.argtypes = [C.POINTER(_Params), C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64),
                                            C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
    return _LIB


class OracleParams:
    """Plain description of one demultiplexing configuration (the values Quade.__init__ derives,
    src/Quade.py:96-139)."""

    def __init__(self, dual, spans, min_qual, sample_keys, write_undetermined=True, write_pass=True, write_fail=True):
        # spans: 4 python slices (start, end) for idx1, idx2, mol1, mol2; disabled = (0, 0)
        self.dual = bool(dual)
        self.spans = [tuple(int(x) for x in s) for s in spans]
        self.min_qual = int(min_qual)
        self.sample_keys = [k.upper() if isinstance(k, str) else k.decode("latin-1").upper() for k in sample_keys]
        self.write_undetermined, self.write_pass, self.write_fail = bool(write_undetermined), bool(write_pass), bool(write_fail)

    @property
    def nbins(self):
        return 1 + 2 * len(self.sample_keys)

    def _c(self):
        p = _Params()
        p.dual = int(self.dual)
        for i, (a, b) in enumerate(self.spans):
            p.span[i][0], p.span[i][1] = a, b
        p.min_qual = self.min_qual
        p.write_undetermined, p.write_pass, p.write_fail = int(self.write_undetermined), int(self.write_pass), int(self.write_fail)
        p.n_samples = len(self.sample_keys)
        blob = "".join(self.sample_keys).encode("latin-1")
        offs = np.zeros(len(self.sample_keys) + 1, dtype=np.uint32)
        np.cumsum([len(k) for k in self.sample_keys], out=offs[1:])
        keep = (C.create_string_buffer(blob, len(blob) + 1), offs)
        p.keys = C.cast(keep[0], C.c_void_p)
        p.key_off = offs.ctypes.data_as(C.c_void_p)
        return p, keep


class OracleResult:
    def __init__(self, out, bin_off, counters, n_pairs, consumed, n_skipped, chunk_end=False):
        self.chunk_end = chunk_end
        self.out = out              # [bytes R1 arena, bytes R2 arena]
        self.bin_off = bin_off      # [np.uint64[nbins+1]] * 2
        self.counters = counters    # np.int64[nbins]
        self.n_pairs = n_pairs
        self.consumed = consumed
        self.n_skipped = n_skipped

    def bin_bytes(self, read, b):
        o = self.bin_off[read]
        return self.out[read][int(o[b]):int(o[b + 1])]


class OracleError(RuntimeError):
    def __init__(self, code):
        super().__init__("oracle error %d" % code)
        self.code = code


def _as_buf(b):
    a = np.frombuffer(b, dtype=np.uint8) if not isinstance(b, np.ndarray) else b
    return a


def demux_batch(params, streams, last=True):
    """streams: list of 3 or 4 bytes-like (R1, R2, I1[, I2])."""
    L = lib()
    p, keep = params._c()
    arrs = [_as_buf(s) for s in streams]
    while len(arrs) < 4:
        arrs.append(np.zeros(0, dtype=np.uint8))
    ptrs = (C.c_void_p * 4)(*[a.ctypes.data if a.size else None for a in arrs])
    nb = (C.c_uint64 * 4)(*[a.size for a in arrs])
    r = _Result()
    if isinstance(last, (list, tuple)):   # per-stream end-of-file flags
        lastm = sum(1 << (8 + s) for s, e in enumerate(last) if e)
    else:
        lastm = int(bool(last))
    rc = L.qo_demux_batch(C.byref(p), ptrs, nb, lastm, C.byref(r))
    try:
        if rc != 0:
            raise OracleError(rc)
        nbins = r.nbins
        bin_off = [np.ctypeslib.as_array(C.cast(r.bin_off[k], C.POINTER(C.c_uint64)), (nbins + 1,)).copy() for k in range(2)]
        out = [C.string_at(r.out[k], int(bin_off[k][nbins])) for k in range(2)]
        counters = np.ctypeslib.as_array(C.cast(r.counters, C.POINTER(C.c_int64)), (nbins,)).copy()
        return OracleResult(out, bin_off, counters, int(r.n_pairs), [int(x) for x in r.consumed], [int(x) for x in r.n_skipped], bool(r.chunk_end))
    finally:
        L.qo_result_free(C.byref(r))
        del keep


def demux_chunks_mt(params, chunks, n_threads):
    """chunks: list of [R1, R2, I1(, I2)] numpy uint8 arrays. Returns (pairs, seconds, out_bytes)."""
    L = lib()
    p, keep = params._c()
    n = len(chunks)
    ptrs = (C.c_void_p * (4 * n))()
    nb = (C.c_uint64 * (4 * n))()
    hold = []
    for c, ch in enumerate(chunks):
        for s in range(4):
            if s < len(ch):
                a = _as_buf(ch[s])
                hold.append(a)
                ptrs[4 * c + s] = a.ctypes.data if a.size else None
                nb[4 * c + s] = a.size
    secs = C.c_double(0)
    ob = C.c_uint64(0)
    pairs = L.qo_demux_chunks_mt(C.byref(p), n, ptrs, nb, int(n_threads), C.byref(secs), C.byref(ob))
    del keep, hold
    if pairs < 0:
        raise OracleError(int(pairs))
    return int(pairs), float(secs.value), int(ob.value)


def bin_names(sample_names):
    """Bin order used everywhere: Undetermined, then <name>_pass, <name>_fail per sample
    (src/Sample.py:44,147-148)."""
    names = ["Undetermined"]
    for n in sample_names:
        names += [n + "_pass", n + "_fail"]
    return names


def report_lines(sample_names, counters):
    """Sample.REPORT (src/Sample.py:104-128) + the writer loop of src/Quade.py:189-190, from line 2 on
    (line 1 carries the date). counters: int64[nbins]."""
    c = [int(x) for x in counters]
    und = c[0]
    pas = sum(c[1::2])
    fai = sum(c[2::2])
    tot = und + pas + fai
    rep = [["Total pair", tot], ["Pair pass quality", pas], ["Pair fail quality", fai], ["Pair Undetermined", und]]
    if tot - und > 0:
        rep.append(["Percent Pair pass quality (wo Undetermined)", pas * 100 // (tot - und)])
        rep.append(["Percent Pair fail quality (wo Undetermined)", fai * 100 // (tot - und)])
        rep.append(["Percent Pair Undetermined", und * 100 // tot])
    for i, n in enumerate(sample_names):
        p_, f_ = c[1 + 2 * i], c[2 + 2 * i]
        rep += [[" ", " "], ["Sample Name", n], ["Total pair", p_ + f_], ["Pair pass quality", p_], ["Pair fail quality", f_]]
        if p_ + f_ > 0:
            rep.append(["Percent of total pair", (p_ + f_) * 100 // (tot - und)])
            rep.append(["Percent Pair pass quality", p_ * 100 // (p_ + f_)])
            rep.append(["Percent Pair fail quality", f_ * 100 // (p_ + f_)])
    return ["\n"] + ["{}\t{}\n".format(d, v) for d, v in rep]
