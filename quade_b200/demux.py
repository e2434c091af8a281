"""DemuxContext: thin object over the C-ABI context (one per GPU).  It carries what Quade.__init__ derives from
the config (src/Quade.py:96-139) and runs batches of a chunk through the CUDA path."""
import ctypes as C

import numpy as np

from . import _lib


class DemuxResult:
    """One collected batch: per-bin byte ranges of the R1 / R2 arenas + counters (bin order: Undetermined,
    then <sample>_pass, <sample>_fail per sample, src/Sample.py:44,147-148)."""

    def __init__(self, raw, nbins, device_mode):
        self.ticket = int(raw.ticket)
        self.n_pairs = int(raw.n_pairs)
        self.chunk_end = bool(raw.chunk_end)
        self.consumed = [int(x) for x in raw.consumed]
        self.n_skipped = [int(x) for x in raw.n_skipped]
        self.nbins = nbins
        self.device_mode = device_mode
        self.bin_off = [np.ctypeslib.as_array(C.cast(raw.bin_off[k], C.POINTER(C.c_uint64)), (nbins + 1,)).copy()
                        for k in range(2)]
        self.counters = np.ctypeslib.as_array(C.cast(raw.counters, C.POINTER(C.c_int64)), (nbins,)).copy()
        self.out_ptr = [raw.out[0], raw.out[1]]
        self.ms_h2d, self.ms_kernels, self.ms_d2h = float(raw.ms_h2d), float(raw.ms_kernels), float(raw.ms_d2h)
        self.ms_kernel = {n: float(raw.ms_kernel[i]) for i, n in enumerate(_lib.K_NAMES)}
        self.in_bytes, self.out_bytes = int(raw.in_bytes), int(raw.out_bytes)
        self.gpu_launches = int(raw.gpu_launches)
        self.reframed_tiles = int(raw.reframed_tiles)

    def arena(self, read):
        """Host view (no copy) of one output arena; valid until the slot is reused."""
        if self.device_mode:
            raise RuntimeError("outputs of a device-mode batch stay on the device")
        n = int(self.bin_off[read][self.nbins])
        if n == 0:
            return np.zeros(0, dtype=np.uint8)
        return np.ctypeslib.as_array(C.cast(self.out_ptr[read], C.POINTER(C.c_uint8)), (n,))

    def bin_view(self, read, b):
        """memoryview (no copy) of bin b of the R1 (read=0) / R2 (read=1) arena; valid until the slot is reused."""
        o = self.bin_off[read]
        lo, hi = int(o[b]), int(o[b + 1])
        if hi == lo:
            return memoryview(b"")
        return memoryview(self.arena(read))[lo:hi]

    def bin_bytes(self, read, b):
        o = self.bin_off[read]
        return self.arena(read)[int(o[b]):int(o[b + 1])].tobytes()


class DemuxContext:
    def __init__(self, dual, spans, min_qual, sample_keys, write_undetermined=True, write_pass=True, write_fail=True,
                 device=0, max_stream_bytes=0, max_pairs=0, n_slots=0, max_records=0, device_only=False):
        """spans: python slices (start, end) of idx1, idx2, mol1, mol2 (disabled = (0, 0)); sample_keys: fused
        index strings exactly as in the config (index1_seq + index2_seq)."""
        self._L = _lib.lib()
        self._ctx = C.c_void_p(None)
        p = _lib.Params()
        p.dual = int(bool(dual))
        for i, (a, b) in enumerate(spans):
            p.span[i][0], p.span[i][1] = int(a), int(b)
        p.min_qual = int(min_qual)
        p.write_undetermined, p.write_pass, p.write_fail = int(bool(write_undetermined)), int(bool(write_pass)), int(bool(write_fail))
        keys = [k.encode("latin-1") if isinstance(k, str) else bytes(k) for k in sample_keys]
        p.n_samples = len(keys)
        p.max_stream_bytes, p.max_pairs, p.n_slots = int(max_stream_bytes), int(max_pairs), int(n_slots)
        p.max_records = int(max_records)
        p.flags = _lib.CTX_DEVICE_ONLY if device_only else 0
        self.n_streams = 4 if dual else 3
        self.nbins = 1 + 2 * len(keys)
        self.n_slots = int(n_slots) or 2
        _lib.check(self._L.quade_ctx_create(int(device), C.byref(p), C.byref(self._ctx)))
        try:
            blob = b"".join(keys)
            offs = np.zeros(len(keys) + 1, dtype=np.uint32)
            if keys:
                np.cumsum([len(k) for k in keys], out=offs[1:])
            buf = C.create_string_buffer(blob, len(blob) + 1)
            _lib.check(self._L.quade_ctx_set_samples(self._ctx, C.cast(buf, C.c_void_p), offs.ctypes.data_as(C.c_void_p)))
        except Exception:
            self.close()
            raise
        self._pending = []   # (device_mode, keepalive) per batch in flight

    def close(self):
        if self._ctx:
            self._L.quade_ctx_destroy(self._ctx)
            self._ctx = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    @property
    def in_flight(self):
        return len(self._pending)

    @staticmethod
    def _flags(last):
        """last: True/False for all streams, or one bool per stream (end-of-file reached for that file only)."""
        if isinstance(last, (list, tuple)):
            f = 0
            for s, e in enumerate(last):
                if e:
                    f |= 1 << (8 + s)
            return f
        return _lib.BATCH_LAST if last else 0

    def submit(self, streams, last=True, ticket=0):
        """streams: 3 or 4 host buffers (bytes / numpy uint8 / pinned views): R1, R2, I1[, I2]."""
        bt = _lib.Batch()
        keep = []
        for s in range(self.n_streams):
            a = streams[s]
            if not isinstance(a, np.ndarray):
                a = np.frombuffer(a, dtype=np.uint8)
            keep.append(a)
            bt.stream[s] = a.ctypes.data if a.size else None
            bt.nbytes[s] = a.size
        bt.flags = self._flags(last)
        bt.ticket = int(ticket)
        _lib.check(self._L.quade_submit(self._ctx, C.byref(bt)))
        self._pending.append((False, keep))

    def submit_device(self, ptrs, nbytes, last=True, ticket=0, keepalive=None):
        """ptrs: 3 or 4 device pointers (16-byte aligned, readable up to nbytes rounded up to 16)."""
        bt = _lib.Batch()
        for s in range(self.n_streams):
            bt.stream[s] = int(ptrs[s]) if nbytes[s] else None
            bt.nbytes[s] = int(nbytes[s])
        bt.flags = self._flags(last) | _lib.BATCH_DEVICE
        bt.ticket = int(ticket)
        _lib.check(self._L.quade_submit(self._ctx, C.byref(bt)))
        self._pending.append((True, keepalive))

    def peek(self):
        """The oldest batch in flight, described once its kernels are done (consumed, n_pairs, bin offsets, counters),
        without copying its bins back and without releasing it."""
        r = _lib.Result()
        _lib.check(self._L.quade_peek(self._ctx, C.byref(r)))
        return DemuxResult(r, self.nbins, True)

    def collect(self):
        r = _lib.Result()
        device_mode, _ = self._pending.pop(0) if self._pending else (False, None)
        _lib.check(self._L.quade_collect(self._ctx, C.byref(r)))
        return DemuxResult(r, self.nbins, device_mode)

    def collect_discard(self):
        """Release the oldest batch in flight whatever its outcome."""
        r = _lib.Result()
        if self._pending:
            self._pending.pop(0)
        self._L.quade_collect(self._ctx, C.byref(r))

    def run(self, streams, last=True):
        self.submit(streams, last=last)
        return self.collect()

    def counters(self):
        out = np.zeros(self.nbins, dtype=np.int64)
        _lib.check(self._L.quade_counters(self._ctx, out.ctypes.data_as(C.c_void_p)))
        return out

    def reset_counters(self):
        _lib.check(self._L.quade_counters_reset(self._ctx))

    def compute_stream(self):
        s = C.c_void_p(None)
        _lib.check(self._L.quade_compute_stream(self._ctx, C.byref(s)))
        return s.value or 0


class PinnedBuffer:
    """Page-locked host staging memory owned by the library (quade_host_alloc)."""

    def __init__(self, nbytes):
        self._L = _lib.lib()
        self._p = C.c_void_p(None)
        self.nbytes = int(nbytes)
        _lib.check(self._L.quade_host_alloc(C.byref(self._p), self.nbytes))
        self.array = np.ctypeslib.as_array(C.cast(self._p, C.POINTER(C.c_uint8)), (max(self.nbytes, 1),))[:self.nbytes]

    def free(self):
        if self._p:
            self.array = None
            self._L.quade_host_free(self._p)
            self._p = C.c_void_p(None)

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def allreduce_counters(contexts):
    """Sum the cumulative counters of several DemuxContext objects (one per GPU of this process) in place:
    quade_allreduce_counters, NCCL over NVLink."""
    arr = (C.c_void_p * len(contexts))(*[c._ctx for c in contexts])
    _lib.check(_lib.lib().quade_allreduce_counters(arr, len(contexts)))
