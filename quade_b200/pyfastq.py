"""pyFastq's role in the B200 build: staging, not parsing.

The reference's `FastqReader(path)` (src/pyFastq/FastqReader.py:25-58) is a generator that frames 4 lines per
record and builds one FastqSeq per record on the CPU.  Here the framing, validation and slicing run on the GPU
(k_frame6 / k_classify2), so `FastqReader(path)` only does what is left for the host: open the file with the
reference's rule (gzip iff the path ends in "gz", FastqReader.py:30) and move raw bytes into page-locked
buffers that `quade_submit` copies to the device.  There is no per-record Python code and no CPU parser.
"""
import gzip
import os
import stat
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_READ_POOL = None
_READ_POOL_LOCK = threading.Lock()


def _read_pool():
    """Threads that pread() slices of plain files into the staging buffers (the copy out of the page cache releases the
    GIL: one thread moves ~2.5 GB/s, a 64 MiB read split over the pool several times that)."""
    global _READ_POOL
    with _READ_POOL_LOCK:
        if _READ_POOL is None:
            _READ_POOL = ThreadPoolExecutor(max_workers=max(4, min(16, os.cpu_count() or 4)))
        return _READ_POOL


class FastqReader:
    """Byte stager over one FASTQ file.

    readinto(arr) fills a numpy uint8 view (normally a slice of a PinnedBuffer) and returns the number of bytes
    stored; `eof` turns True once the file is exhausted.  A gzip file is inflated ahead by a host thread of
    libquade_io.so (`depth` blocks of BLOCK bytes), so that inflating overlaps the GPU work on the current batch and
    the files of the chunks opened ahead inflate side by side; prefetch=False keeps Python's gzip module in the
    calling thread (the form the native reader is tested against).
    """

    BLOCK = 8 << 20
    SLICE = 4 << 20     # plain files: bytes per pread() task

    def __init__(self, fastq_file, prefetch=True, depth=4):
        self.path = fastq_file
        # same rule as the reference: FastqReader.py:30
        self._plain = fastq_file[-2:] != "gz"
        self._native = None
        if not self._plain and prefetch:
            from ._io import GzipReader
            self._native = GzipReader(fastq_file, self.BLOCK, depth)
            self._fh = None
        else:
            self._fh = open(fastq_file, "rb", buffering=0) if self._plain else gzip.open(fastq_file, "rb")
        self._size = None          # regular plain files: size known, slices are pread() in parallel
        self._pos = 0
        if self._plain:
            prefetch = False       # plain files are read straight into the caller's (pinned) buffer: no copy to hide
            st = os.fstat(self._fh.fileno())
            if stat.S_ISREG(st.st_mode):
                self._size = st.st_size
        self.eof = False
        self.bytes_read = 0
        self._pending = None       # (memoryview, offset) of a block partly handed out

    def _next_block(self):
        return self._fh.read(self.BLOCK)

    def readinto(self, arr):
        """Fill `arr` (numpy uint8, C-contiguous) as far as the file allows; returns bytes written."""
        want = int(arr.size)
        done = 0
        if self._native is not None:
            done = self._native.readinto(arr) if not self.eof else 0
            if done < want:
                self.eof = True
            self.bytes_read += done
            return done
        if self._plain and self._size is not None:
            want = min(want, self._size - self._pos)
            mv = memoryview(arr)
            fd, pos = self._fh.fileno(), self._pos

            def piece(a, b):
                got = a
                while got < b:
                    n = os.preadv(fd, [mv[got:b]], pos + got)
                    if n <= 0:
                        break
                    got += n
                return got - a
            step = self.SLICE
            if want <= step:
                done = piece(0, want)
            else:
                futs = [_read_pool().submit(piece, a, min(a + step, want)) for a in range(0, want, step)]
                parts = [f.result() for f in futs]
                done = 0
                for a, n in zip(range(0, want, step), parts):      # a short slice (file truncated meanwhile) ends the data
                    done = a + n
                    if n < min(step, want - a):
                        break
            self._pos += done
            if self._pos >= self._size or done < want:
                self.eof = True
            self.bytes_read += done
            return done
        if self._plain:
            mv = memoryview(arr)
            while done < want:
                n = self._fh.readinto(mv[done:])
                if not n:
                    self.eof = True
                    break
                done += n
            self.bytes_read += done
            return done
        while done < want and not self.eof:
            if self._pending is None:
                blk = self._next_block()
                if not blk:
                    self.eof = True
                    break
                self._pending = (memoryview(blk), 0)
            mv, off = self._pending
            n = min(want - done, len(mv) - off)
            arr[done:done + n] = np.frombuffer(mv[off:off + n], dtype=np.uint8)
            done += n
            off += n
            self._pending = (mv, off) if off < len(mv) else None
        self.bytes_read += done
        return done

    def close(self):
        try:
            if self._native is not None:
                self._native.close()
            else:
                self._fh.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()
