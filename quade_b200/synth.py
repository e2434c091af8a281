"""Synthetic dual/single-indexed paired-end FASTQ: ctypes binding of libquade_synth.so
(include/quade_synth.h, csrc/quade_synth.cu).  Test and benchmark infrastructure: its own library, the product
library is neither needed nor loaded here.  BASELINE.json configs 2-5 are named here."""
import ctypes as C
import os

import numpy as np

from . import build as _build


class SynthParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("dual", C.c_int32), ("read_len", C.c_int32), ("index_len", C.c_int32),
                ("mol_len", C.c_int32), ("n_samples", C.c_int32), ("floor_mode", C.c_int32),
                ("undetermined", C.c_double)]


SYMBOLS = ["quade_synth_barcodes", "quade_synth_sizes", "quade_synth_fill", "quade_synth_fill_device"]
_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(_build.build_synth())
        vp, u64p = C.c_void_p, C.POINTER(C.c_uint64)
        L.quade_synth_barcodes.argtypes = [C.POINTER(SynthParams), vp]
        L.quade_synth_sizes.argtypes = [C.POINTER(SynthParams), vp, C.c_uint64, C.c_uint64, u64p]
        L.quade_synth_fill.argtypes = [C.POINTER(SynthParams), vp, C.c_uint64, C.c_uint64, C.POINTER(vp), C.c_int]
        L.quade_synth_fill_device.argtypes = [C.POINTER(SynthParams), vp, C.c_uint64, C.c_uint64, C.POINTER(vp), u64p, vp]
        for name in SYMBOLS:
            getattr(L, name).restype = C.c_int
        _LIB = L
    return _LIB


def _check(rc):
    if rc != 0:
        raise RuntimeError("libquade_synth error %d" % rc)

# name -> (dual, index_len, mol_len, n_samples, min_qual, undetermined, floor_mode)
CONFIGS = {
    "cfg2": dict(dual=False, index_len=8, mol_len=0, n_samples=96, min_qual=25, undetermined=0.05, floor_mode=0),
    "cfg3": dict(dual=True, index_len=8, mol_len=6, n_samples=384, min_qual=25, undetermined=0.05, floor_mode=0),
    "cfg4": dict(dual=True, index_len=10, mol_len=0, n_samples=384, min_qual=25, undetermined=0.05, floor_mode=0),
    "cfg5": dict(dual=True, index_len=8, mol_len=6, n_samples=384, min_qual=25, undetermined=0.30, floor_mode=1),
}
CONFIG_SEED = {"cfg2": 0x5EED0002, "cfg3": 0x5EED0003, "cfg4": 0x5EED0004, "cfg5": 0x5EED0005}


class Synth:
    def __init__(self, dual=True, read_len=150, index_len=8, mol_len=6, n_samples=384, undetermined=0.05,
                 floor_mode=0, seed=0x5EED0003, min_qual=25):
        self._L = lib()
        self.p = SynthParams(seed=int(seed), dual=int(bool(dual)), read_len=int(read_len), index_len=int(index_len),
                                  mol_len=int(mol_len), n_samples=int(n_samples), floor_mode=int(floor_mode),
                                  undetermined=float(undetermined))
        self.dual, self.index_len, self.mol_len, self.n_samples = bool(dual), int(index_len), int(mol_len), int(n_samples)
        self.min_qual = int(min_qual)
        ni = 2 if dual else 1
        self.barcodes = np.zeros(max(1, n_samples * ni * index_len), dtype=np.uint8)
        _check(self._L.quade_synth_barcodes(C.byref(self.p), self.barcodes.ctypes.data_as(C.c_void_p)))

    @classmethod
    def from_config(cls, name, read_len=150, **over):
        kw = dict(CONFIGS[name])
        kw["seed"] = CONFIG_SEED[name]
        kw.update(over)
        return cls(read_len=read_len, **kw)

    @property
    def n_streams(self):
        return 4 if self.dual else 3

    def sample_names(self):
        return ["S%03d" % (i + 1) for i in range(self.n_samples)]

    def sample_keys(self):
        """Fused index per sample (index1_seq + index2_seq), in config order."""
        ni = 2 if self.dual else 1
        b = self.barcodes[:self.n_samples * ni * self.index_len].reshape(self.n_samples, ni * self.index_len)
        return [bytes(r).decode("ascii") for r in b]

    def spans(self):
        il, ml = self.index_len, self.mol_len
        idx = (0, il)
        mol = (il, il + ml) if ml else (0, 0)
        if self.dual:
            return [idx, idx, mol, mol]
        return [idx, (0, 0), (0, 0), (0, 0)]   # single index + molecular is a reference defect

    def sizes(self, first, count):
        nb = (C.c_uint64 * 4)()
        _check(self._L.quade_synth_sizes(C.byref(self.p), None, int(first), int(count), nb))
        return [int(x) for x in nb][:self.n_streams]

    def fill(self, first, count, arrays, n_threads=None):
        """arrays: numpy uint8 buffers sized by sizes()."""
        ptrs = (C.c_void_p * 4)()
        for s in range(self.n_streams):
            ptrs[s] = arrays[s].ctypes.data
        nt = n_threads or min(32, os.cpu_count() or 1)
        _check(self._L.quade_synth_fill(C.byref(self.p), self.barcodes.ctypes.data_as(C.c_void_p), int(first),
                                            int(count), ptrs, int(nt)))

    def generate(self, first, count, n_threads=None):
        arrays = [np.empty(n, dtype=np.uint8) for n in self.sizes(first, count)]
        self.fill(first, count, arrays, n_threads)
        return arrays

    def fill_device(self, first, count, ptrs, stream=0):
        """ptrs: device pointers of buffers sized by sizes(); the bytes are those of fill().  Returns the sizes."""
        dp = (C.c_void_p * 4)()
        for s in range(self.n_streams):
            dp[s] = int(ptrs[s])
        nb = (C.c_uint64 * 4)()
        _check(self._L.quade_synth_fill_device(C.byref(self.p), self.barcodes.ctypes.data_as(C.c_void_p), int(first),
                                               int(count), dp, nb, C.c_void_p(int(stream) or None)))
        return [int(x) for x in nb][:self.n_streams]

    def conf_text(self, files, min_qual=None):
        """A Quade configuration file (grammar of src/Conf_file.py:18-104) for this data set.  files: one list of
        paths per stream (seq_R1, seq_R2, index_R1[, index_R2]), one entry per chunk."""
        il, ml = self.index_len, self.mol_len
        keys = ["seq_R1", "seq_R2", "index_R1", "index_R2"][:self.n_streams]
        L = ["[quality]", "minimal_qual : %d" % (self.min_qual if min_qual is None else min_qual), "", "[fastq]"]
        L += ["%s : %s" % (k, " ".join(f)) for k, f in zip(keys, files)]
        L += ["", "[index]", "index2 : %s" % self.dual, "molecular1 : %s" % bool(ml and self.dual),
              "molecular2 : %s" % bool(ml and self.dual), "index1_start : 1", "index1_end : %d" % il]
        if self.dual:
            L += ["index2_start : 1", "index2_end : %d" % il]
            if ml:
                L += ["molecular1_start : %d" % (il + 1), "molecular1_end : %d" % (il + ml),
                      "molecular2_start : %d" % (il + 1), "molecular2_end : %d" % (il + ml)]
        L += ["", "[output]", "write_pass : True", "write_fail : True", "write_undetermined : True", ""]
        for i, (n, k) in enumerate(zip(self.sample_names(), self.sample_keys())):
            L += ["[sample%d]" % (i + 1), "name : %s" % n, "index1_seq : %s" % k[:il]]
            if self.dual:
                L.append("index2_seq : %s" % k[il:])
            L.append("")
        return "\n".join(L)
