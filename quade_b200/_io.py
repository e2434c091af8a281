"""ctypes binding of libquade_io.so (include/quade_io.h): the parallel gzip-member writer behind FastqWriter."""
import ctypes as C
import os

import numpy as np

from . import build as _build

SYMBOLS = ["qio_sink_create", "qio_sink_destroy", "qio_sink_set_paths", "qio_sink_append", "qio_sink_wait",
           "qio_sink_stats", "qio_last_error", "qio_reader_open", "qio_reader_read", "qio_reader_close"]
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(_build.build_io())
        vp, u64p = C.c_void_p, C.POINTER(C.c_uint64)
        L.qio_last_error.restype = C.c_char_p
        L.qio_sink_create.argtypes = [C.c_int, C.c_int, C.POINTER(vp)]
        L.qio_sink_destroy.argtypes = [vp]
        L.qio_sink_set_paths.argtypes = [vp, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p)]
        L.qio_sink_append.argtypes = [vp, vp, vp, vp, vp, u64p]
        L.qio_sink_wait.argtypes = [vp, C.c_uint64]
        L.qio_sink_stats.argtypes = [vp, u64p, u64p]
        L.qio_reader_open.argtypes = [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(vp)]
        L.qio_reader_read.argtypes = [vp, vp, C.c_uint64, u64p]
        L.qio_reader_close.argtypes = [vp]
        for n in SYMBOLS:
            if n != "qio_last_error":
                getattr(L, n).restype = C.c_int
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise IOError("libquade_io error %d: %s" % (rc, lib().qio_last_error().decode("utf-8", "replace")))


class GzipSink:
    """One gzip member per non-empty bin and read of a batch, compressed on a pool of host threads and appended to
    the bin's files in submission order (src/FastqWriter.py:83-90)."""

    def __init__(self, n_threads=0, level=1):
        self._L = lib()
        self._s = C.c_void_p(None)
        _check(self._L.qio_sink_create(int(n_threads), int(level), C.byref(self._s)))
        self._keep = {}              # ticket -> objects that own the memory of the batch

    def set_paths(self, r1, r2):
        n = len(r1)
        a1 = (C.c_char_p * n)(*[p.encode() for p in r1])
        a2 = (C.c_char_p * n)(*[p.encode() for p in r2])
        _check(self._L.qio_sink_set_paths(self._s, n, a1, a2))
        self._keep.clear()

    def append(self, out1_ptr, off1, out2_ptr, off2, keep=None):
        """off1 / off2: numpy uint64 arrays of nbins + 1 offsets into the arenas at out1_ptr / out2_ptr (addresses).
        Returns a ticket; the arenas must stay untouched until wait(ticket)."""
        o1 = np.ascontiguousarray(off1, dtype=np.uint64)
        o2 = np.ascontiguousarray(off2, dtype=np.uint64)
        t = C.c_uint64(0)
        _check(self._L.qio_sink_append(self._s, C.c_void_p(out1_ptr), o1.ctypes.data_as(C.c_void_p), C.c_void_p(out2_ptr),
                                       o2.ctypes.data_as(C.c_void_p), C.byref(t)))
        self._keep[t.value] = keep
        return t.value

    def wait(self, ticket):
        _check(self._L.qio_sink_wait(self._s, int(ticket)))
        for k in [k for k in self._keep if k <= ticket]:
            del self._keep[k]

    def stats(self):
        a, b = C.c_uint64(0), C.c_uint64(0)
        _check(self._L.qio_sink_stats(self._s, C.byref(a), C.byref(b)))
        return a.value, b.value

    def close(self):
        if self._s:
            self._L.qio_sink_destroy(self._s)
            self._s = C.c_void_p(None)
            self._keep.clear()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class GzipReader:
    """A gzip file inflated ahead by a host thread of libquade_io.so (no interpreter in the loop); readinto(arr) drains
    the decompressed bytes into a numpy uint8 view.  src/pyFastq/FastqReader.py:30."""

    def __init__(self, path, block_bytes=8 << 20, depth=4):
        self._L = lib()
        self._r = C.c_void_p(None)
        _check(self._L.qio_reader_open(os.fsencode(path), int(block_bytes), int(depth), C.byref(self._r)))

    def readinto(self, arr):
        """Fills arr as far as the stream allows; fewer bytes than arr.size means the end of the stream."""
        want, done, got = int(arr.size), 0, C.c_uint64(0)
        base = arr.ctypes.data
        while done < want:
            _check(self._L.qio_reader_read(self._r, C.c_void_p(base + done), want - done, C.byref(got)))
            if got.value == 0:
                break
            done += got.value
        return done

    def close(self):
        if self._r:
            self._L.qio_reader_close(self._r)
            self._r = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
