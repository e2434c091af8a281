"""Quade: the reference's driver (src/Quade.py:36-293) over the CUDA demultiplexing path.

Same command line (`-c CONF [-i] [--version]`), same configuration grammar and attribute names, same messages,
exit codes, output files and report.  What changes is the body of the hot loop: instead of pulling one record
per file per iteration (Quade.py:210-221) the host stages raw bytes of each file of a chunk into page-locked
buffers and calls quade_submit / quade_collect; framing, pairing, index slicing, sample lookup, quality filter,
name suffix and per-bin serialisation happen in the sm_100a kernels.  There is no CPU fallback.
"""
import configparser
import optparse
import os
import sys
from datetime import datetime
from time import time

import numpy as np

from . import parallel
from .conf_file import write_example_conf
from .fastq_writer import FastqWriter
from .pyfastq import FastqReader
from .sample import Sample


class Quade(object):
    """Fastq file demultiplexer, handling double indexing, molecular indexing and filtering based on index
    quality"""

    VERSION = "Quade 0.3.2"
    USAGE = "Usage: %prog -c Conf.txt [-i -h]"

    #: bytes of the largest file in a batch (the other files in proportion to their record size); env
    #: QUADE_BATCH_BYTES.  A device slot takes SLOT_BATCHES times that per file: windows grow when no complete pair fits.
    BATCH_BYTES = int(os.environ.get("QUADE_BATCH_BYTES", 64 << 20))
    SLOT_BATCHES = 1.25
    #: batches in flight on the device: H2D + kernels of batch N+1 beside the D2H of batch N and the gzip of N-1
    N_SLOTS = 3
    #: a file's page-locked buffer is two halves of HALF_BATCHES x BATCH_BYTES of data behind BATCH_BYTES / 2 of room for
    #: the rest moved over from the other half (quade_b200/feeder.py).  One batch per half: the window of the largest
    #: file is the whole half, so that what moves over is a partial record
    HALF_BATCHES = 1.0
    #: chunks whose files are opened ahead of the one being parsed: a gzip file is inflated by its own thread, so the
    #: files of the next chunks inflate on idle cores while this chunk is on the device (FastqReader.py:30)
    CHUNKS_AHEAD = 3
    #: decompressed bytes a file opened ahead may hold ready (8 MiB blocks): 3 chunks x 4 files x 384 MiB at most
    AHEAD_BLOCKS = 48

    @classmethod
    def class_init(cls):
        """init class method for instantiation from command line (Quade.py:50-66)"""
        optparser = optparse.OptionParser(usage=cls.USAGE, version=cls.VERSION)
        optparser.add_option('-c', dest="conf_file", help="Path to the configuration file [Mandatory]")
        optparser.add_option('-i', dest="init_conf", action='store_true',
                             help="Generate an example configuration file and exit [Facultative]")
        options, args = optparser.parse_args()
        return cls(options.conf_file, options.init_conf)

    def __init__(self, conf_file=None, init_conf=None, device=None):
        """Parse options from the configuration file and verify their values (Quade.py:70-153)."""
        if init_conf:
            print("Create an example configuration file in the current folder")
            write_example_conf()
            sys.exit(0)

        print("Initialize Quade")
        self.rank, self.world, local_rank = parallel.env_world()
        self.device = local_rank if device is None else int(device)
        try:
            assert conf_file, "A path to the configuration file is mandatory"
            self._is_readable_file(conf_file)
            self.conf = conf_file

            # Py2 RawConfigParser(allow_no_value=True): duplicate sections/options merge silently, no
            # interpolation, ' ;' starts an inline comment
            cp = configparser.RawConfigParser(allow_no_value=True, strict=False, inline_comment_prefixes=(";",))
            cp.read(self.conf)

            self.minimal_qual = cp.getint("quality", "minimal_qual")

            self.idx1 = True
            self.idx2 = cp.getboolean("index", "index2")
            self.mol1 = self.idx1 and cp.getboolean("index", "molecular1")
            self.mol2 = self.idx2 and cp.getboolean("index", "molecular2")

            # 1-based inclusive positions -> python slices; unused sub-indexes are 0:0 (Quade.py:105-116)
            def pos(on, key):
                return {"start": 0 if not on else cp.getint("index", key + "_start") - 1,
                        "end": 0 if not on else cp.getint("index", key + "_end")}
            self.idx1_pos = pos(True, "index1")
            self.idx2_pos = pos(self.idx2, "index2")
            self.mol1_pos = pos(self.mol1, "molecular1")
            self.mol2_pos = pos(self.mol2, "molecular2")

            self.seq_R1 = cp.get("fastq", "seq_R1").split()
            self.seq_R2 = cp.get("fastq", "seq_R2").split()
            self.index_R1 = cp.get("fastq", "index_R1").split()
            self.index_R2 = [] if not self.idx2 else cp.get("fastq", "index_R2").split()

            Sample.CLASS_INIT(
                write_undetermined=cp.getboolean("output", "write_undetermined"),
                write_pass=cp.getboolean("output", "write_pass"),
                write_fail=cp.getboolean("output", "write_fail"),
                min_qual=self.minimal_qual)

            for sample in [i for i in cp.sections() if i.startswith("sample")]:
                if self.idx2:
                    Sample(name=cp.get(sample, "name"), index=(cp.get(sample, "index1_seq") + cp.get(sample, "index2_seq")))
                else:
                    Sample(name=cp.get(sample, "name"), index=cp.get(sample, "index1_seq"))

            self._test_values()

        except (configparser.NoOptionError, configparser.NoSectionError) as E:
            print("Option or section missing. Report to the template configuration file\n" + str(E))
            sys.exit(1)
        except (ValueError, AssertionError) as E:
            print("One of the value in the configuration file is not correct\n" + str(E))
            sys.exit(1)
        except (IOError) as E:
            print("One of the file is incorrect or unreadable\n" + str(E))
            sys.exit(1)
        self._engine = None
        self._stage = []               # two sets of page-locked file buffers (this chunk's, the next chunk's)
        self._fed = {}                 # chunk -> (readers, feeders) started ahead
        self._bpr_hint = None          # bytes per record of every file, from the last batch that had pairs
        self._tickets = []             # per submitted batch of the run: ticket of its bins in the sink (or None)
        import collections
        self._prof = collections.defaultdict(float)

    def __str__(self):
        msg = "QUADE CLASS\n\tParameters list\n"
        for key in sorted(self.__dict__.keys()):
            msg += "\t{}\t{}\n".format(key, self.__dict__[key])
        return msg

    def __repr__(self):
        return "<Instance of {} from {} >\n".format(self.__class__.__name__, self.__module__)

    # ~~~~~~~PUBLIC METHODS~~~~~~~#

    def __call__(self):
        """Main function of the script (Quade.py:169-193)"""
        start_time = time()
        self.rank, self.world, local_rank = parallel.init_process_group()
        FastqWriter.PART = None
        if self.world > 1 and not os.environ.get("QUADE_NO_BIND"):
            from . import affinity
            affinity.bind_to_gpu(local_rank, self.world)      # threads and pinned pages of this rank next to its GPU

        print("Start parsing files: {} chunks to be parsed".format(len(self.seq_R1)))
        failure = None
        self._tickets = []
        # the main loop shares the interpreter with ~20 reader threads that take the GIL for a moment after every read:
        # with the default 5 ms switch interval a lock or a thread start in the loop can wait that long
        switch_interval = sys.getswitchinterval()
        sys.setswitchinterval(min(switch_interval, 2e-4))
        try:
            # the readers of the first chunks start now: gzip files inflate (libquade_io.so threads) while the CUDA
            # context and the pinned buffers are set up
            chunks, mine = self._chunk_plan()
            self._open_ahead(chunks, mine[:1 + self.CHUNKS_AHEAD], deep_first=True)
            t0 = time()
            self._open_engine()
            Sample.SINK = self._make_sink()
            self._prof["open engine + sink"] += time() - t0
            t0 = time()
            if self.idx2:
                self.double_index_parser()
            else:
                self.simple_index_parser()
            self._prof["all chunks"] += time() - t0
            t0 = time()
            for t in reversed(self._tickets):      # the sink is in order: the last ticket waits for all (and reports errors)
                if t is not None:
                    Sample.SINK.wait(t)
                    break
            self._prof["wait sink"] += time() - t0
            Sample.FLUSH_ALL()
        except BaseException as E:                  # noqa: B902 -- re-raised below, after the ranks have agreed
            failure = E
        finally:
            t0 = time()
            self._close_engine()
            self._prof["close"] += time() - t0
            sys.setswitchinterval(switch_interval)
        if self.world > 1:
            # a rank that failed (blank header, capacity, I/O) must not leave the others waiting in the all-reduce
            others = parallel.any_rank_failed(failure is not None, self.device)
            if failure is None and others:
                raise RuntimeError("another rank of this run failed: see its output")
        if failure is not None:
            raise failure

        if self.world > 1:
            # the path's only collective: the report counters; then restore bin order by chunk index
            Sample.SET_COUNTS(parallel.allreduce_counters(np.array(Sample.COUNTS(), dtype=np.int64), self.device))
            parallel.barrier()
            names = []
            for w in Sample.BIN_WRITERS():
                names += [w.R1_fastq_name, w.R2_fastq_name]
            parallel.merge_parts(names, self.rank, self.world)
            FastqWriter.PART = None
            parallel.barrier()

        if self.rank == 0:
            print("Generate_a csv report")
            with open("Quade_report.csv", "w") as report:
                report.write("Program {}\tDate {}\n\n".format(self.VERSION, str(datetime.today())))
                for descr, value in Sample.REPORT():
                    report.write("{}\t{}\n".format(descr, value))

        if os.environ.get("QUADE_PROFILE"):
            print("profile (s): " + ", ".join("{} {:.3f}".format(k, v) for k, v in sorted(self._prof.items())))
        print("Done in {}s".format(round(time() - start_time, 3)))
        return 0

    def double_index_parser(self):
        """Quade.py:195-224: chunk n = n-th file of seq_R1, seq_R2, index_R1, index_R2"""
        chunks, mine = self._chunk_plan()
        for i, n in enumerate(mine):
            print("Start parsing chunk {}".format(n + 1))
            self._open_ahead(chunks, mine[i:i + 1 + self.CHUNKS_AHEAD])
            t0 = time()
            self._feed(n, chunks[n], i)
            if i + 1 < len(mine):
                # the next chunk's first windows fill while this chunk is parsed (its buffers are page-locked once, on
                # this thread: the library allocates on the calling thread's current device)
                self._feed(mine[i + 1], chunks[mine[i + 1]], i + 1)
            self._prof["start feeders"] += time() - t0
            self._parse_chunk(n, chunks[n])
            print("\tEnd of chunk {}".format(n + 1))

    def simple_index_parser(self):
        """Quade.py:226-254"""
        chunks, mine = self._chunk_plan()
        for i, n in enumerate(mine):
            print("Start parsing chunk {}/{}".format(n + 1, len(self.seq_R1)))
            self._open_ahead(chunks, mine[i:i + 1 + self.CHUNKS_AHEAD])
            t0 = time()
            self._feed(n, chunks[n], i)
            if i + 1 < len(mine):
                # the next chunk's first windows fill while this chunk is parsed (its buffers are page-locked once, on
                # this thread: the library allocates on the calling thread's current device)
                self._feed(mine[i + 1], chunks[mine[i + 1]], i + 1)
            self._prof["start feeders"] += time() - t0
            self._parse_chunk(n, chunks[n])
            print("\t{} sequences parsed".format(self._chunk_pairs))     # the reader's line, FastqReader.py:57-58
            print("\tEnd of chunk {}".format(n + 1))

    # ~~~~~~~PRIVATE METHODS~~~~~~~#

    def spans(self):
        """python slices of idx1, idx2, mol1, mol2 in the order of the C-ABI (QUADE_SPAN_*)"""
        return [(p["start"], p["end"]) for p in (self.idx1_pos, self.idx2_pos, self.mol1_pos, self.mol2_pos)]

    def _slot_bytes(self):
        return (int(self.SLOT_BATCHES * self.BATCH_BYTES) + 15) & ~15

    def _open_engine(self):
        """The C-ABI context of this process's GPU (loads libquade_b200.so; raises without a CUDA device: there is
        no CPU path).  A slot takes SLOT_BATCHES x BATCH_BYTES of every file (the longest window _parse_chunk submits)."""
        if self._engine is None:
            from .demux import DemuxContext
            self._engine = DemuxContext(self.idx2, self.spans(), self.minimal_qual,
                                        [s.index for s in Sample.SAMPLE_LIST],
                                        write_undetermined=Sample.WRITE_UNDETERMINED, write_pass=Sample.WRITE_PASS,
                                        write_fail=Sample.WRITE_FAIL, device=self.device,
                                        max_stream_bytes=self._slot_bytes(), n_slots=self.N_SLOTS)
        return self._engine

    def _chunk_plan(self):
        """(file tuples of all chunks, indices of the chunks of this rank): chunk n = n-th file of every list
        (Quade.py:195-198, :226-229)"""
        if self.idx2:
            chunks = list(zip(self.seq_R1, self.seq_R2, self.index_R1, self.index_R2))
        else:
            chunks = list(zip(self.seq_R1, self.seq_R2, self.index_R1))
        return chunks, parallel.chunks_of_rank(len(chunks), self.rank, self.world)

    def _open_ahead(self, chunks, which, deep_first=False):
        """FastqReader objects of the chunks in `which` (the current one first), opened now.  The current chunk is
        drained as it inflates and needs little read-ahead, unless nobody drains it yet (deep_first: start-up)."""
        if not hasattr(self, "_opened"):
            self._opened = {}
        for n in which:
            if n not in self._opened:
                deep = deep_first or n != which[0]
                self._opened[n] = [FastqReader(f, depth=self.AHEAD_BLOCKS if deep else 4) for f in chunks[n]]

    def _make_sink(self):
        from .fastq_writer import BinSink
        return BinSink(Sample.BIN_WRITERS())

    def _alloc_staging(self, nbytes):
        """One page-locked staging buffer (quade_host_alloc): object with .array (numpy uint8) and .free()."""
        from .demux import PinnedBuffer
        return PinnedBuffer(nbytes)

    def _close_engine(self):
        sink, Sample.SINK = Sample.SINK, None
        try:
            if sink is not None:
                sink.close()
        finally:
            try:
                if self._engine is not None:
                    self._engine.close()
            finally:
                self._engine = None
                for readers, feeders in self._fed.values():
                    for f in feeders:
                        f.close()
                    for r in readers:
                        r.close()
                self._fed = {}
                for st in self._stage:
                    for b, _ in st:
                        b.free()
                self._stage = []
                for rs in getattr(self, "_opened", {}).values():
                    for r in rs:
                        r.close()
                self._opened = {}

    def _staging(self, files, which):
        """Set `which` (0 / 1: this chunk's, the next chunk's) of page-locked buffers, one per file of a chunk, allocated
        on first use (two halves each, quade_b200/feeder.py).  Returns [(buffer, head bytes)].  A file much smaller than
        the largest of its chunk (an index file beside the insert files) holds as many records in a quarter or half
        of the bytes, and page-locking memory costs ~0.4 s per GiB at start-up."""
        while len(self._stage) <= which:
            self._stage.append([])
        st = self._stage[which]
        sizes = []
        for f in files:
            try:
                sizes.append(os.path.getsize(f))
            except OSError:
                sizes.append(0)
        top = max(sizes + [1])
        while len(st) < len(files):
            share = sizes[len(st)] / top
            scale = 0.25 if share <= 0.2 else (0.5 if share <= 0.4 else 1.0)
            b = max(512, int(scale * self.BATCH_BYTES)) & ~15
            st.append((self._alloc_staging(2 * (b // 2 + int(self.HALF_BATCHES * b))), b // 2))
        return st

    def _feed(self, n, files, i):
        """Readers and feeder threads of chunk n (the i-th chunk of this rank), started now."""
        from .feeder import Feeder
        if n in self._fed:
            return
        readers = self._opened.pop(n, None) if hasattr(self, "_opened") else None
        readers = readers or [FastqReader(f) for f in files]
        B = self.BATCH_BYTES
        chunk = max(256, min(16 << 20, B // 2))
        stage = self._staging(files, i % 2)
        feeders = []
        self._fed[n] = (readers, feeders)
        for s in range(len(files)):
            buf, head = stage[s]
            feeders.append(Feeder(readers[s], buf.array, head, min(chunk, max(256, head))))

    def _parse_chunk(self, n, files):
        """The body of the reference's `while True` loop for one chunk (Quade.py:210-224 / :240-254), as a pipeline.

        Every file is streamed into its own page-locked buffer by a Feeder thread, ahead of the batches.  A batch is
        a window of every buffer: the bytes a previous batch left unconsumed and as many more as hold about the same
        number of records in every file.  While batch N is on the device the feeders keep reading; as soon as its
        kernels are done (quade_peek) the windows move on and batch N+1 is submitted; only then batch N is collected
        -- its D2H copy runs beside the H2D copy and the kernels of N+1 -- and its bins go to the native gzip sink,
        which compresses and appends them on all host cores while the loop goes on.  The chunk ends when a file has
        run dry (:223, :252).
        """
        eng = self._engine
        ns = len(files)
        B = self.BATCH_BYTES
        cap = self._slot_bytes()                   # bytes of one file a device slot takes (max_stream_bytes)
        if self.world > 1:
            Sample.FLUSH_ALL()                     # nothing of the previous chunk may land in this chunk's parts
            FastqWriter.PART = n
        if Sample.SINK is not None:
            Sample.SINK.begin_chunk()
        self._feed(n, files, 0)                    # (a caller that did not start them: set 0)
        readers, feeders = self._fed.pop(n)
        self._chunk_pairs = 0
        sink_tickets = self._tickets               # batches of the whole run: a slot's arenas outlive the chunk
        first_ticket = len(sink_tickets)
        ok = False
        prof = self._prof                          # seconds per pipeline stage (QUADE_PROFILE=1 prints them)
        clock = time

        def finish():
            """collect the oldest batch in flight and hand its bins to the writers"""
            try:
                res = eng.collect()
            except RuntimeError as E:
                if getattr(E, "code", None) == -1:              # QUADE_ERR_BLANK_HEADER
                    # name = line.rstrip()[1:].split()[0] on a header without a token (FastqReader.py:42)
                    raise IndexError("list index out of range") from E
                raise
            for s in range(ns):
                for _ in range(int(res.n_skipped[s])):         # FastqReader.py:48-50
                    print("Sequence length and quality string length are not equal.")
                    print("Skipping the sequence")
            sink_tickets.append(Sample.FINDER(res))

        try:
            # first batch of the first chunk: a probe that tells the bytes per record of every file; later chunks
            # start from the previous chunk's figures
            bpr = self._bpr_hint if self._bpr_hint is not None and len(self._bpr_hint) == ns else None
            limit = [min(B, 8 << 20)] * ns if bpr is None else [int(min(cap, B / max(bpr) * bpr[s] + 4096)) for s in range(ns)]
            stalled = 0
            k = 0
            while True:
                t0 = clock()
                for s in range(ns):
                    feeders[s].wait_for(limit[s])
                wins = [feeders[s].window(limit[s], cap) for s in range(ns)]
                prof["wait readers"] += clock() - t0
                views = [w[0] for w in wins]
                eof = [w[1] for w in wins]
                t0 = clock()
                K = first_ticket + k                # batches submitted since the start of the run
                if K >= self.N_SLOTS and sink_tickets[K - self.N_SLOTS] is not None:
                    Sample.SINK.wait(sink_tickets[K - self.N_SLOTS])   # the slot's output arenas are about to be reused
                prof["wait sink"] += clock() - t0
                t0 = clock()
                eng.submit(views, last=eof, ticket=k)
                prof["submit"] += clock() - t0
                t0 = clock()
                if k >= 1:
                    finish()                       # batch k-1: D2H beside the H2D and the kernels of batch k
                prof["collect + sink append"] += clock() - t0
                t0 = clock()
                try:
                    res = eng.peek()               # kernels of batch k done (so is its H2D copy): the windows may move
                except RuntimeError as E:
                    if getattr(E, "code", None) == -1:
                        eng.collect_discard()
                        raise IndexError("list index out of range") from E
                    raise
                prof["peek"] += clock() - t0
                prof["batches"] += 1
                k += 1
                self._chunk_pairs += res.n_pairs
                t0 = clock()
                for s in range(ns):
                    feeders[s].consume(int(res.consumed[s]))
                prof["move windows"] += clock() - t0
                if res.chunk_end:
                    finish()
                    ok = True
                    break
                # bytes per record of every file -> windows that hold about the same number of records; a file that is
                # ahead of the others keeps its surplus in its buffer
                if res.n_pairs:
                    bpr = [max(1.0, int(res.consumed[s]) / res.n_pairs) for s in range(ns)]
                    self._bpr_hint = bpr
                progressed = res.n_pairs > 0 or any(int(c) > 0 for c in res.consumed[:ns])
                stalled = 0 if progressed else stalled + 1
                if bpr is None or stalled:
                    # no complete pair in the windows: longer ones (the end of a feeder's half also lands here, once)
                    if stalled > 8:
                        raise IOError("a FASTQ record of chunk {} is larger than the staging buffer "
                                      "(QUADE_BATCH_BYTES={})".format(n + 1, B))
                    limit = [min(cap, max(w, 1 << 10) * 2) for w in limit]
                else:
                    pairs = B / max(bpr)
                    limit = [int(min(cap, pairs * bpr[s] + 4096)) for s in range(ns)]
        finally:
            t0 = clock()
            for f in feeders:
                f.close()
            prof["stop feeders"] += clock() - t0
            while eng.in_flight:                   # an exception above: leave no batch in flight
                try:
                    eng.collect()
                except Exception:
                    pass
            if not ok:                             # an exception above: the sink may not outlive the arenas either
                for t in sink_tickets[first_ticket:]:
                    if t is not None and Sample.SINK is not None:
                        try:
                            Sample.SINK.wait(t)
                        except Exception:
                            pass
            for r in readers:
                r.close()

    def _test_values(self):
        """Test the validity of options in the configuration file (Quade.py:258-279)"""
        assert 0 <= self.minimal_qual <= 40, "Authorized values for minimal_qual : 0 to 40"
        if self.idx2:
            assert len(self.seq_R1) == len(self.seq_R2) == len(self.index_R1) == len(self.index_R2) > 0, \
                "seq_R1, seq_R2, index_R1 and index_R2 are mandatory and have to contain the same number of files"
            for fp in (self.seq_R1 + self.seq_R2 + self.index_R1 + self.index_R2):
                self._is_readable_file(fp)
        else:
            assert len(self.seq_R1) == len(self.seq_R2) == len(self.index_R1) > 0, \
                "seq_R1, seq_R2 and index_R1 are mandatory and have to contain the same number of files"
            for fp in (self.seq_R1 + self.seq_R2 + self.index_R1):
                self._is_readable_file(fp)
        for pos in [self.idx1_pos, self.idx2_pos, self.mol1_pos, self.mol2_pos]:
            assert pos["start"] >= 0
            assert pos["end"] >= pos["start"]
        # The reference crashes on the first written pair when a molecular index is asked for without a second
        # sample index (Quade.py:247 passes a str where FastqWriter.py:62 wants a FastqSeq):
        # refuse the combination at configuration time instead.
        assert self.idx2 or not self.mol1, \
            "molecular1 without index2 is not supported (the reference fails on it, src/Quade.py:247)"
        assert Sample.SAMPLE_LIST, "At least one [sample] section is required"

    def _is_readable_file(self, fp):
        """Verify the readability of a file (Quade.py:281-284)"""
        if not os.access(fp, os.R_OK):
            raise IOError("{} is not a valid file".format(fp))


def main():
    quade = Quade.class_init()
    return quade()
