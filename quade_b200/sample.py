"""Sample: the reference's class-level sample registry, counters and report (src/Sample.py:19-175).

Same class fields and methods.  The per-pair classification of `Sample.FINDER` (Sample.py:56-91: dict lookup of
the upper-cased fused index, min(phred) >= MIN_QUAL, counters, writer dispatch) runs on the GPU in k_classify2; here
FINDER takes one collected batch (a DemuxResult) and applies what the kernels decided: it adds the per-bin pair
counts to the class and sample counters and hands every non-empty bin range to that bin's FastqWriter.
"""
from .fastq_writer import FastqWriter


class Sample(object):
    # ~~~~~~~CLASS FIELDS~~~~~~~ (Sample.py:32-44)
    TOTAL = FAIL_QUAL = PASS_QUAL = UNDETERMINED = 0
    WRITE_UNDETERMINED = WRITE_PASS = WRITE_FAIL = False
    NAME_TO_SAMPLE = {}
    INDEX_TO_SAMPLE = {}
    SAMPLE_LIST = []
    DNA = ["A", "T", "C", "G", "N"]
    MIN_QUAL = 0
    UNDETERMINED_WRITER = FastqWriter(name="Undetermined")

    # ~~~~~~~CLASS METHODS~~~~~~~#

    @classmethod
    def CLASS_INIT(cls, write_undetermined=False, write_pass=False, write_fail=False, min_qual=0):
        """Generic class method to define all class fields at once (Sample.py:48-54)"""
        cls.WRITE_UNDETERMINED = write_undetermined
        cls.WRITE_PASS = write_pass
        cls.WRITE_FAIL = write_fail
        cls.MIN_QUAL = min_qual

    @classmethod
    def RESET(cls):
        """Forget samples, counters and writers: the reference keeps all of it in class fields, one run per
        process (Sample.py:32-44); tests and long-lived hosts need a way back to the import-time state."""
        cls.TOTAL = cls.FAIL_QUAL = cls.PASS_QUAL = cls.UNDETERMINED = 0
        cls.WRITE_UNDETERMINED = cls.WRITE_PASS = cls.WRITE_FAIL = False
        cls.NAME_TO_SAMPLE = {}
        cls.INDEX_TO_SAMPLE = {}
        cls.SAMPLE_LIST = []
        cls.MIN_QUAL = 0
        cls.UNDETERMINED_WRITER = FastqWriter(name="Undetermined")
        cls.SINK = None

    @classmethod
    def BIN_WRITERS(cls):
        """Writers in device bin order: Undetermined, then <sample>_pass, <sample>_fail in config order."""
        w = [cls.UNDETERMINED_WRITER]
        for s in cls.SAMPLE_LIST:
            w += [s.pass_writer, s.fail_writer]
        return w

    @classmethod
    def ADD_COUNTS(cls, counters):
        """counters: pair counts per bin (device order).  Counters move whatever the write flags say
        (Sample.py:62,71-72,79-80,88)."""
        c = [int(x) for x in counters]
        if len(c) != 1 + 2 * len(cls.SAMPLE_LIST):
            raise ValueError("expected {} bin counters, got {}".format(1 + 2 * len(cls.SAMPLE_LIST), len(c)))
        cls.UNDETERMINED += c[0]
        cls.TOTAL += c[0]
        for i, s in enumerate(cls.SAMPLE_LIST):
            p, f = c[1 + 2 * i], c[2 + 2 * i]
            s.pass_qual += p
            s.fail_qual += f
            cls.PASS_QUAL += p
            cls.FAIL_QUAL += f
            cls.TOTAL += p + f

    @classmethod
    def SET_COUNTS(cls, counters):
        """Replace all counters by `counters` (used after the all-reduce over GPUs)."""
        cls.TOTAL = cls.FAIL_QUAL = cls.PASS_QUAL = cls.UNDETERMINED = 0
        for s in cls.SAMPLE_LIST:
            s.pass_qual = s.fail_qual = 0
        cls.ADD_COUNTS(counters)

    @classmethod
    def COUNTS(cls):
        c = [cls.UNDETERMINED]
        for s in cls.SAMPLE_LIST:
            c += [s.pass_qual, s.fail_qual]
        return c

    #: BinSink of the running Quade (native gzip-member writer pool); None: FINDER writes through the FastqWriters
    SINK = None

    @classmethod
    def FINDER(cls, batch, *per_record_args):
        """Apply one demultiplexed batch (Sample.py:56-91, batched).

        `batch` is a DemuxResult: the GPU already looked every fused index up, checked min phred against MIN_QUAL
        and serialised the records per bin.  The counters move by the per-bin pair counts; every non-empty bin range
        goes to its files -- through the native sink (one call per batch; returns its ticket) when a Quade run has
        installed one, else through the bins' FastqWriter objects.  The reference's per-record form
        FINDER(read1, read2, index, molecular) has no CPU implementation here.
        """
        if per_record_args or not hasattr(batch, "bin_off"):
            raise TypeError("Sample.FINDER takes one DemuxResult per batch; the per-record CPU form of the "
                            "reference (Sample.py:57) is replaced by the CUDA path and is not provided")
        cls.ADD_COUNTS(batch.counters)
        if cls.SINK is not None:
            return cls.SINK.append(batch)
        writers = cls.BIN_WRITERS()
        for b, w in enumerate(writers):
            r1 = batch.bin_view(0, b)
            if len(r1) == 0:
                continue                      # write flag off (Sample.py:74,82,90) or no pair in this bin
            w(r1, batch.bin_view(1, b), n_records=int(batch.counters[b]))
        return None

    @classmethod
    def FLUSH_ALL(cls):
        """Flush all fastq buffers at once (Sample.py:93-102)"""
        ws = [w for w in cls.BIN_WRITERS() if w.counter > 0 or w._last is not None]
        for w in ws:
            w.flush_buffers(wait=False)
        for w in ws:
            w.wait()

    @classmethod
    def REPORT(cls):
        """Output a report under the form of a list (Sample.py:104-128; Python-2 '/' on ints is '//')"""
        report = []
        report.append(["Total pair", cls.TOTAL])
        report.append(["Pair pass quality", cls.PASS_QUAL])
        report.append(["Pair fail quality", cls.FAIL_QUAL])
        report.append(["Pair Undetermined", cls.UNDETERMINED])
        if cls.TOTAL - cls.UNDETERMINED > 0:
            report.append(["Percent Pair pass quality (wo Undetermined)", cls.PASS_QUAL * 100 // (cls.TOTAL - cls.UNDETERMINED)])
            report.append(["Percent Pair fail quality (wo Undetermined)", cls.FAIL_QUAL * 100 // (cls.TOTAL - cls.UNDETERMINED)])
            report.append(["Percent Pair Undetermined", cls.UNDETERMINED * 100 // cls.TOTAL])
        for sample in cls.SAMPLE_LIST:
            report.append([" ", " "])
            report.append(["Sample Name", sample.name])
            report.append(["Total pair", sample.total])
            report.append(["Pair pass quality", sample.pass_qual])
            report.append(["Pair fail quality", sample.fail_qual])
            if sample.total > 0:
                report.append(["Percent of total pair", sample.total * 100 // (cls.TOTAL - cls.UNDETERMINED)])
                report.append(["Percent Pair pass quality", sample.pass_qual * 100 // sample.total])
                report.append(["Percent Pair fail quality", sample.fail_qual * 100 // sample.total])
        return report

    # ~~~~~~~FUNDAMENTAL METHODS~~~~~~~#

    def __init__(self, name, index):
        """Sample.py:132-153: same assertions, same messages."""
        self.name = name
        self.index = index.upper()
        assert self.name not in self.NAME_TO_SAMPLE, "{} : Name is not unique".format(self.name)
        assert self.index not in self.INDEX_TO_SAMPLE, "{} : Index is not unique".format(self.name)
        assert self._is_dna(index), "{} : Non canonical DNA base in index".format(self.name)
        # the reference accepts an empty index and later dies in min() of an empty sequence
        assert len(index) > 0, "{} : Empty index".format(self.name)
        self.pass_qual = self.fail_qual = 0
        self.pass_writer = FastqWriter(name="{}_pass".format(self.name))
        self.fail_writer = FastqWriter(name="{}_fail".format(self.name))
        self.INDEX_TO_SAMPLE[self.index] = self
        self.NAME_TO_SAMPLE[self.name] = self
        self.SAMPLE_LIST.append(self)

    @property
    def total(self):
        return self.pass_qual + self.fail_qual

    def __str__(self):
        msg = "SAMPLE CLASS\n"
        for key, value in self.__dict__.items():
            msg += "\t{}\t{}\n".format(key, value)
        return msg

    def __repr__(self):
        return "<Instance of {} from {} >\n".format(self.__class__.__name__, self.__module__)

    def _is_dna(self, sequence):
        """Checked on the string as written in the config, so lower case is rejected (Sample.py:141,171-175)."""
        for base in sequence:
            if base not in self.DNA:
                return False
        return True
