"""quade_b200 -- B200-native demultiplexing hot path of a-slide/Quade behind Quade's own interfaces.

Python host (config file, Sample / FastqWriter object model, report) over hand-written sm_100a CUDA kernels
called through a ctypes C-ABI (include/quade_b200.h).  There is no CPU fallback."""
__version__ = "0.3.2+b200.1"

from ._lib import QuadeError, device_count, lib_path  # noqa: F401
from .demux import DemuxContext, DemuxResult, PinnedBuffer  # noqa: F401
