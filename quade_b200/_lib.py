"""ctypes binding of libquade_b200.so (include/quade_b200.h).  No CPU fallback: if the library is missing it is
built with nvcc; if that fails, or no CUDA device is present at run time, the calls raise."""
import ctypes as C
import os

from . import build as _build

QUADE_OK = 0
ERR_BLANK_HEADER, ERR_NOMEM, ERR_BAD_SAMPLE, ERR_CUDA, ERR_ARG, ERR_CAPACITY, ERR_UNSUPPORTED, ERR_STATE = range(-1, -9, -1)
BATCH_LAST, BATCH_DEVICE = 1, 2
CTX_DEVICE_ONLY = 1
# stage 1 (QUADE_K_VALIDATE in the header) = k_tilescan + redo pass + k_locate_rare
K_NAMES = ["frame", "tilescan", "pairing", "classify", "binscan", "offsets", "compact"]
K_COUNT = len(K_NAMES)
SYMBOLS = ["quade_version", "quade_last_error", "quade_device_count", "quade_host_alloc", "quade_host_free",
           "quade_ctx_create", "quade_ctx_destroy", "quade_ctx_set_samples", "quade_submit", "quade_peek", "quade_collect",
           "quade_counters", "quade_counters_reset", "quade_allreduce_counters", "quade_compute_stream",
           "quade_bin_offsets_static"]


class Params(C.Structure):
    _fields_ = [("dual", C.c_int32), ("span", (C.c_int32 * 2) * 4), ("min_qual", C.c_int32),
                ("write_undetermined", C.c_int32), ("write_pass", C.c_int32), ("write_fail", C.c_int32),
                ("n_samples", C.c_int32), ("max_stream_bytes", C.c_uint64), ("max_pairs", C.c_uint64),
                ("n_slots", C.c_int32), ("flags", C.c_uint32), ("max_records", C.c_uint64)]


class Batch(C.Structure):
    _fields_ = [("stream", C.c_void_p * 4), ("nbytes", C.c_uint64 * 4), ("flags", C.c_uint32),
                ("reserved", C.c_uint32), ("ticket", C.c_uint64)]


class Result(C.Structure):
    _fields_ = [("ticket", C.c_uint64), ("n_pairs", C.c_uint64), ("consumed", C.c_uint64 * 4),
                ("n_skipped", C.c_uint64 * 4), ("out", C.c_void_p * 2), ("bin_off", C.c_void_p * 2),
                ("counters", C.c_void_p), ("nbins", C.c_int32), ("err", C.c_int32), ("chunk_end", C.c_int32),
                ("reserved", C.c_int32),
                ("ms_h2d", C.c_float), ("ms_kernels", C.c_float), ("ms_d2h", C.c_float),
                ("ms_kernel", C.c_float * K_COUNT), ("in_bytes", C.c_uint64), ("out_bytes", C.c_uint64),
                ("gpu_launches", C.c_uint64), ("reframed_tiles", C.c_uint64)]


class QuadeError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libquade_b200 error %d: %s" % (code, msg))
        self.code = code


_LIB = None


def lib_path():
    return _build.LIB


def lib():
    """Load (building if needed) the shared library. Raises if it cannot be built or loaded."""
    global _LIB
    if _LIB is not None:
        return _LIB
    path = _build.LIB
    if _build.is_stale():
        _build.build()
    L = C.CDLL(path)
    vp = C.c_void_p
    L.quade_version.restype = C.c_int
    L.quade_last_error.restype = C.c_char_p
    L.quade_device_count.argtypes = [C.POINTER(C.c_int)]
    L.quade_host_alloc.argtypes = [C.POINTER(vp), C.c_uint64]
    L.quade_host_free.argtypes = [vp]
    L.quade_ctx_create.argtypes = [C.c_int, C.POINTER(Params), C.POINTER(vp)]
    L.quade_ctx_destroy.argtypes = [vp]
    L.quade_ctx_set_samples.argtypes = [vp, vp, vp]
    L.quade_submit.argtypes = [vp, C.POINTER(Batch)]
    L.quade_collect.argtypes = [vp, C.POINTER(Result)]
    L.quade_peek.argtypes = [vp, C.POINTER(Result)]
    L.quade_counters.argtypes = [vp, vp]
    L.quade_counters_reset.argtypes = [vp]
    L.quade_allreduce_counters.argtypes = [C.POINTER(vp), C.c_int]
    L.quade_compute_stream.argtypes = [vp, C.POINTER(vp)]
    L.quade_bin_offsets_static.argtypes = [vp, vp, vp, C.c_uint64, C.c_int32, vp, vp, vp, vp, vp]
    for name in SYMBOLS:
        if name not in ("quade_version", "quade_last_error"):
            getattr(L, name).restype = C.c_int
    _LIB = L
    return L


def check(rc):
    if rc != QUADE_OK:
        raise QuadeError(rc, lib().quade_last_error().decode("utf-8", "replace"))
    return rc


def device_count():
    n = C.c_int(0)
    rc = lib().quade_device_count(C.byref(n))
    return n.value if rc == QUADE_OK else 0
