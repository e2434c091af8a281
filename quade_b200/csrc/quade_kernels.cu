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

// Self-made bounds checks (compute-sanitizer is closed on this pool): a build with -DQUADE_CHECKS=1 tests every
// shared / global store index and every window read of the hot kernels against its buffer and counts violations in
// g_qcheck {violations, code of the first one}; quade_ctx_destroy prints the tally (tools/gpu_checks.sh).
#if QUADE_CHECKS
__device__ unsigned int g_qcheck[2];
#define QCHK(cond, code)                                                                                           \
    do {                                                                                                           \
        if (!(cond)) {                                                                                             \
            if (atomicAdd(&g_qcheck[0], 1u) == 0u) g_qcheck[1] = (code);                                           \
        }                                                                                                          \
    } while (0)
#else
#define QCHK(cond, code) do { } while (0)
#endif

} // namespace
} // namespace quade

#include "quade_frame.cuh"
#include "quade_classify.cuh"
#include "quade_bins.cuh"
#include "quade_compact.cuh"
#include "quade_compact2.cuh"
#include "quade_launch.cuh"

namespace quade {
// violations counted by a QUADE_CHECKS build since the last call (0 0 in a normal build)
void debug_check_tally(unsigned int out[2])
{
    out[0] = out[1] = 0u;
#if QUADE_CHECKS
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_qcheck, sizeof(unsigned int) * 2);
    const unsigned int z[2] = {0u, 0u};
    cudaMemcpyToSymbol(g_qcheck, z, sizeof(z));
#endif
}
} // namespace quade
