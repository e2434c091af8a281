"""FastqWriter: the reference's per-bin output object (src/FastqWriter.py:19-90) fed with device-built bytes.

The reference appends ":INDEX[:MOLECULAR]" to both read names, serialises the pair with FastqSeq.fastqstr and
buffers 20 records before appending a gzip member (FastqWriter.py:48-74).  Here the suffix and the serialisation
are done by k_compact3 on the GPU; a writer receives the already serialised R1 / R2 byte ranges of its bin for one
batch and keeps only the reference's host duties: lazy creation of `<name>_R1.fastq.gz` / `<name>_R2.fastq.gz` in
the current working directory on the first record (:55-57, :76-81) and appending gzip members in arrival order
(:83-90).  Parity is on the decompressed stream (gzip members carry mtime).
"""
import os
import threading
import zlib
from concurrent.futures import ThreadPoolExecutor
from gzip import open as gopen

import numpy as np

_POOL = None
_POOL_LOCK = threading.Lock()


def _pool():
    global _POOL
    with _POOL_LOCK:
        if _POOL is None:
            _POOL = ThreadPoolExecutor(max_workers=max(2, min(32, os.cpu_count() or 2)))
        return _POOL


def _gzip_member(chunks, level):
    """One gzip member (RFC 1952) holding the concatenation of `chunks`; zlib releases the GIL."""
    co = zlib.compressobj(level, zlib.DEFLATED, 31)
    out = [co.compress(c) for c in chunks]
    out.append(co.flush())
    return b"".join(out)


class FastqWriter(object):
    """Handle creation of fastq files and appending through byte buffers (same fields as the reference)."""

    #: zlib level of the appended members (the reference's gzip.open default is 9; decompressed bytes are equal)
    COMPRESSLEVEL = 1
    #: pending decompressed bytes per file that trigger an asynchronous flush
    FLUSH_BYTES = 4 << 20
    #: when set (multi-GPU runs), files are written as `<name>.part<chunk>` and merged in chunk order at the end
    PART = None

    def __init__(self, name="Unknown"):
        """Init the object with counters, empty buffers, and names of fastq files (FastqWriter.py:27-34)"""
        self.buffer_size = 20          # kept for interface parity; members are cut by FLUSH_BYTES here
        self.R1_fastq_name = name + "_R1.fastq.gz"
        self.R2_fastq_name = name + "_R2.fastq.gz"
        self.counter = -1              # -1 until the first record, then records waiting in the buffers
        self.R1_buffer = []
        self.R2_buffer = []
        self._pending_bytes = 0
        self._created = set()          # paths already created (one per part in multi-GPU runs)
        self._last = None              # future of the previous asynchronous append (keeps member order)

    def __str__(self):
        msg = "FASTQ_WRITER CLASS\n"
        for key, value in self.__dict__.items():
            msg += "\t\t{}\t{}\n".format(key, value)
        return msg

    def __repr__(self):
        return "<Instance of {} from {} >\n".format(self.__class__.__name__, self.__module__)

    # ~~~~~~~PUBLIC METHODS~~~~~~~#

    def _paths(self):
        sfx = "" if FastqWriter.PART is None else ".part{:06d}".format(FastqWriter.PART)
        return self.R1_fastq_name + sfx, self.R2_fastq_name + sfx

    def __call__(self, read1, read2, index=None, molecular="", n_records=1):
        """Append the serialised R1 / R2 records of this bin for one batch.

        read1 / read2 are bytes-like (the bin's range of the R1 / R2 output arena, suffix already appended on the
        device); `index` / `molecular` are accepted for signature parity with FastqWriter.py:48 and ignored.
        Per-record FastqSeq objects are not accepted: there is no CPU serialisation path.
        """
        if not isinstance(read1, (bytes, bytearray, memoryview)) or not isinstance(read2, (bytes, bytearray, memoryview)):
            raise TypeError("FastqWriter takes the device-built byte ranges of its bin (bytes-like); "
                            "per-record FastqSeq serialisation on the CPU is not part of quade_b200")
        if self.counter == -1:
            self.counter = 0
        p1, _ = self._paths()
        if p1 not in self._created:
            self.init_files()
        self.counter += int(n_records)
        self.R1_buffer.append(bytes(read1))   # the arena slot is reused by the next batch: take a copy
        self.R2_buffer.append(bytes(read2))
        self._pending_bytes += len(read1) + len(read2)
        if self._pending_bytes >= self.FLUSH_BYTES:
            self.flush_buffers(wait=False)

    def init_files(self):
        """Init empty files for R1 and R2.fastq.gz (FastqWriter.py:76-81)"""
        self.wait()
        p1, p2 = self._paths()
        with gopen(p1, "wb"):
            print("\tCreate {} file".format(self.R1_fastq_name))
        with gopen(p2, "wb"):
            print("\tCreate {} file".format(self.R2_fastq_name))
        self._created.add(p1)

    def flush_buffers(self, wait=True):
        """Append to R1 and R2 fastq.gz files (FastqWriter.py:83-90); compression runs on the writer pool."""
        if self.R1_buffer or self.R2_buffer:
            b1, b2 = self.R1_buffer, self.R2_buffer
            self.R1_buffer, self.R2_buffer = [], []
            self._pending_bytes = 0
            self.counter = 0
            p1, p2 = self._paths()
            prev, level = self._last, self.COMPRESSLEVEL

            def job():
                if prev is not None:
                    prev.result()
                for path, chunks in ((p1, b1), (p2, b2)):
                    member = _gzip_member(chunks, level)
                    with open(path, "ab") as fh:
                        fh.write(member)
            self._last = _pool().submit(job)
        if wait:
            self.wait()

    def wait(self):
        if self._last is not None:
            self._last.result()
            self._last = None


class BinSink(object):
    """The batch side of the writers: one call per demultiplexed batch instead of one per record.

    A batch's bins are byte ranges of the two pinned output arenas (DemuxResult.out_ptr / bin_off).  The sink creates
    the files of every bin that receives its first record (FastqWriter.init_files: lazy creation, the reference's
    "Create ... file" lines, src/FastqWriter.py:55-57,76-81) and hands all non-empty ranges to libquade_io.so in ONE
    native call: gzip members are cut, compressed on all host cores and appended in order while the GPU works on the
    next batch (src/FastqWriter.py:83-90).  No per-bin Python loop, no copy of the arena.
    """

    def __init__(self, writers, level=None, n_threads=0):
        from ._io import GzipSink
        self.writers = list(writers)
        self.sink = GzipSink(n_threads, FastqWriter.COMPRESSLEVEL if level is None else level)
        self.created = np.zeros(len(self.writers), dtype=bool)
        self._paths = None

    def begin_chunk(self):
        """(Re)bind the bins to their files: `<name>_R?.fastq.gz`, or the chunk's `.partNNNNNN` files in a multi-GPU
        run (FastqWriter.PART)."""
        paths = [w._paths() for w in self.writers]
        if paths != self._paths:
            self.sink.set_paths([p[0] for p in paths], [p[1] for p in paths])
            self._paths = paths
            self.created[:] = [p[0] in w._created for w, p in zip(self.writers, paths)]

    def append(self, res):
        """res: a collected DemuxResult (host arenas).  Returns the ticket to wait on before the arenas are reused."""
        off1, off2 = res.bin_off[0], res.bin_off[1]
        nz = np.flatnonzero(off1[1:] > off1[:-1])
        for b in nz[~self.created[nz]]:
            w = self.writers[int(b)]
            w.init_files()
            w.counter = 0
            self.created[b] = True
        return self.sink.append(res.out_ptr[0] or 0, off1, res.out_ptr[1] or 0, off2, keep=res)

    def wait(self, ticket):
        self.sink.wait(ticket)

    def close(self):
        self.sink.close()
