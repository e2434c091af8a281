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

    #: bytes of each file staged per batch (pinned host buffer and device slot size), env QUADE_BATCH_BYTES
    BATCH_BYTES = int(os.environ.get("QUADE_BATCH_BYTES", 64 << 20))
    #: callable(quade, device) -> engine with run(streams, last) / counters(); tests inject a stand-in here
    ENGINE_FACTORY = None

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
        self.rank, self.world, _ = parallel.init_process_group()
        FastqWriter.PART = None

        print("Start parsing files: {} chunks to be parsed".format(len(self.seq_R1)))
        try:
            if self.idx2:
                self.double_index_parser()
            else:
                self.simple_index_parser()
            Sample.FLUSH_ALL()
        finally:
            self._close_engine()

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

        print("Done in {}s".format(round(time() - start_time, 3)))
        return 0

    def double_index_parser(self):
        """Quade.py:195-224: chunk n = n-th file of seq_R1, seq_R2, index_R1, index_R2"""
        chunks = list(zip(self.seq_R1, self.seq_R2, self.index_R1, self.index_R2))
        for n in parallel.chunks_of_rank(len(chunks), self.rank, self.world):
            print("Start parsing chunk {}".format(n + 1))
            self._parse_chunk(n, chunks[n])
            print("\tEnd of chunk {}".format(n + 1))

    def simple_index_parser(self):
        """Quade.py:226-254"""
        chunks = list(zip(self.seq_R1, self.seq_R2, self.index_R1))
        for n in parallel.chunks_of_rank(len(chunks), self.rank, self.world):
            print("Start parsing chunk {}/{}".format(n + 1, len(self.seq_R1)))
            self._parse_chunk(n, chunks[n])
            print("\tEnd of chunk {}".format(n + 1))

    # ~~~~~~~PRIVATE METHODS~~~~~~~#

    def spans(self):
        """python slices of idx1, idx2, mol1, mol2 in the order of the C-ABI (QUADE_SPAN_*)"""
        return [(p["start"], p["end"]) for p in (self.idx1_pos, self.idx2_pos, self.mol1_pos, self.mol2_pos)]

    def _open_engine(self):
        if self._engine is None:
            if Quade.ENGINE_FACTORY is not None:
                self._engine = Quade.ENGINE_FACTORY(self, self.device)
            else:
                from .demux import DemuxContext   # loads libquade_b200.so; raises without a CUDA device
                self._engine = DemuxContext(self.idx2, self.spans(), self.minimal_qual,
                                            [s.index for s in Sample.SAMPLE_LIST],
                                            write_undetermined=Sample.WRITE_UNDETERMINED, write_pass=Sample.WRITE_PASS,
                                            write_fail=Sample.WRITE_FAIL, device=self.device,
                                            max_stream_bytes=self.BATCH_BYTES, n_slots=1)
        return self._engine

    def _close_engine(self):
        if self._engine is not None:
            try:
                self._engine.close()
            finally:
                self._engine = None
        for b in getattr(self, "_stage", []):
            b.free()
        self._stage = []

    def _staging(self, ns):
        """Page-locked staging buffers, one per file of a chunk (allocated once, reused by every chunk)."""
        if len(getattr(self, "_stage", [])) < ns:
            if Quade.ENGINE_FACTORY is not None:
                class _Plain:
                    def __init__(self, n):
                        self.array = np.empty(n, dtype=np.uint8)

                    def free(self):
                        self.array = None
                self._stage = [_Plain(self.BATCH_BYTES) for _ in range(ns)]
            else:
                from .demux import PinnedBuffer
                self._stage = [PinnedBuffer(self.BATCH_BYTES) for _ in range(ns)]
        return self._stage

    def _parse_chunk(self, n, files):
        """The body of the reference's `while True` loop for one chunk (Quade.py:210-224 / :240-254), batched.

        Every file is staged `BATCH_BYTES` at a time; a batch pairs as many records as all buffers hold
        completely, the unconsumed tail of each buffer (whole records of the longer files and the trailing
        partial line) is carried into the next batch.  The chunk ends when a file is exhausted (:223, :252).
        """
        eng = self._open_engine()
        ns = len(files)
        cap = self.BATCH_BYTES
        if self.world > 1:
            Sample.FLUSH_ALL()             # nothing of the previous chunk may land in this chunk's parts
            FastqWriter.PART = n
        readers = [FastqReader(f) for f in files]
        bufs = [b.array for b in self._staging(ns)]
        fill = [0] * ns
        try:
            while True:
                for s in range(ns):
                    if not readers[s].eof and fill[s] < cap:
                        fill[s] += readers[s].readinto(bufs[s][fill[s]:cap])
                eof = [r.eof for r in readers]
                try:
                    res = eng.run([bufs[s][:fill[s]] for s in range(ns)], last=eof)
                except RuntimeError as E:
                    if getattr(E, "code", None) == -1:          # QUADE_ERR_BLANK_HEADER
                        # name = line.rstrip()[1:].split()[0] on a header without a token (FastqReader.py:42)
                        raise IndexError("list index out of range") from E
                    raise
                for s in range(ns):
                    for _ in range(int(res.n_skipped[s])):     # FastqReader.py:48-50
                        print("Sequence length and quality string length are not equal.")
                        print("Skipping the sequence")
                Sample.FINDER(res)
                for s in range(ns):
                    c = int(res.consumed[s])
                    left = fill[s] - c
                    if c and left:
                        bufs[s][:left] = bufs[s][c:fill[s]]      # numpy handles the overlap
                    fill[s] = left
                if all(eof):
                    break
                if res.n_pairs == 0:
                    if any(eof):
                        break                                   # StopIteration of an exhausted file
                    raise IOError("a FASTQ record of chunk {} is larger than the staging buffer "
                                  "(QUADE_BATCH_BYTES={})".format(n + 1, cap))
        finally:
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
