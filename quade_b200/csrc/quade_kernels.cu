// quade_kernels.cu -- sm_100a kernels of the demultiplexing hot path (see DESIGN.md for the data flow).
//
//   k_frame<fast>   newline scan, record framing, field extraction, index digest  (FastqReader.py:37-44, Quade.py:217-218)
//   k_tilescan      line index of every tile, check of the guessed record phase, record -> scratch position table
//   k_frame<redo>   the tiles the fast pass could not frame (malformed text, end of buffer, long records)
//   k_locate_rare   files with len(seq) != len(qual) records: pairing ordinal = rank among the valid ones
//                                                                       (FastqSeq.py:55, FastqReader.py:48-50)
//   k_pairing       n_pairs = min valid, consumed bytes, fatal checks    (Quade.py:210-224)
//   k_classify      fuse the index digests, perfect-hash lookup, quality filter, class, suffix, lengths
//                                                                       (Sample.py:65-91)
//   k_binscan_*     per-bin exclusive scan over pair blocks              (FastqWriter.py:68-69 append order)
//   k_offsets       stable in-block ranks -> output offset of every record
//   k_compact       '@' name suffix seq "\n+\n" qual '\n' into the bin-segmented arenas
//                                                                       (FastqWriter.py:61-69, FastqSeq.py:76-82)
// All of it is byte / integer work bound by HBM bandwidth: TMA bulk copies into shared memory, 128-bit
// conflict-free shared loads, SWAR byte tests, warp ballots and shuffles; no tensor cores.
#include "quade_dev.h"
#include "quade_ptx.cuh"

namespace quade {
namespace {

typedef unsigned long long u64;

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 0x20u || (c - 9u) <= 4u; } // Py2 str whitespace

} // namespace
} // namespace quade

#include "quade_frame.cuh"
#include "quade_classify.cuh"
#include "quade_bins.cuh"
#include "quade_compact.cuh"
#include "quade_launch.cuh"
