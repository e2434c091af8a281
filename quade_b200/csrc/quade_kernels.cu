// quade_kernels.cu -- sm_100a kernels of the demultiplexing hot path (see DESIGN.md for the data flow).
//
//   K1 k_frame     newline scan + line framing + field extraction      (FastqReader.py:37-44)
//   K2 k_validate  len(seq)==len(qual) check, per-stream valid ordinals (FastqSeq.py:55, FastqReader.py:48-50)
//   K2b k_pairing  n_pairs = min valid, consumed bytes, fatal checks    (Quade.py:210-224)
//   K3 k_classify  slice + fuse index key, perfect-hash lookup, min phred, class, suffix, lengths
//                                                                       (Quade.py:217-218, Sample.py:65-91)
//   K4 k_binscan*  per-bin exclusive scan over pair blocks              (FastqWriter.py:68-69 append order)
//   K4b k_offsets  stable in-block ranks -> output offset of every record
//   K5 k_compact   '@' name suffix seq "\n+\n" qual '\n' into the bin-segmented arenas
//                                                                       (FastqWriter.py:61-69, FastqSeq.py:76-82)
// All of it is byte / integer work bound by HBM bandwidth: 128-bit coalesced loads, warp ballots and shuffles,
// shared-memory staging; no tensor cores.
#include "quade_dev.h"
#include "quade_ptx.cuh"

namespace quade {
namespace {

typedef unsigned long long u64;

constexpr u64 ST_AGG = 1ull << 62, ST_INC = 2ull << 62, ST_VAL = (1ull << 62) - 1;

__device__ __forceinline__ bool is_ws(uint32_t c) { return c == 0x20u || (c - 9u) <= 4u; } // Py2 str whitespace

__device__ __forceinline__ u64 ld_status(const u64 *p)
{
    u64 v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_status(u64 *p, u64 v)
{
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Decoupled look-back: exclusive prefix of `total` over the tiles of one sequence.  Called by warp 0 only.
__device__ __forceinline__ uint32_t lookback(u64 *status, int tile, uint32_t total, int lane)
{
    if (tile == 0) {
        if (lane == 0) st_status(status, ST_INC | total);
        return 0;
    }
    if (lane == 0) st_status(status + tile, ST_AGG | total);
    uint32_t prefix = 0;
    int look = tile - 1;
    for (;;) {
        int idx = look - lane;
        u64 sv = idx >= 0 ? ld_status(status + idx) : ST_INC;
        uint32_t flag = (uint32_t)(sv >> 62);
        unsigned inc = __ballot_sync(0xffffffffu, flag == 2u);
        unsigned empty = __ballot_sync(0xffffffffu, flag == 0u);
        unsigned upto = inc ? ((2u << (__ffs(inc) - 1)) - 1u) : 0xffffffffu; // lanes up to the first inclusive one
        if (empty & upto) continue;                                             // a predecessor is not there yet
        uint32_t v = ((1u << lane) & upto) ? (uint32_t)(sv & ST_VAL) : 0u;
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        prefix += v;
        if (inc) break;
        look -= 32;
    }
    if (lane == 0) st_status(status + tile, ST_INC | (u64)(prefix + total));
    return prefix;
}

// block-wide exclusive scan of one value per thread (256 threads); returns exclusive prefix, total in *total
__device__ __forceinline__ uint32_t block_scan256(uint32_t v, uint32_t *s_warp /*[9]*/, uint32_t *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < 8 ? s_warp[lane] : 0u, wi = w;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += t;
        }
        if (lane < 8) s_warp[lane] = wi - w;
        if (lane == 7) s_warp[8] = wi;
    }
    __syncthreads();
    uint32_t r = s_warp[warp] + inc - v;
    *total = s_warp[8];
    return r;
}

} // namespace
} // namespace quade

#include "quade_frame2.cuh"
#include "quade_frame6.cuh"
#include "quade_kernels3.cuh"
#include "quade_compact3.cuh"
#include "quade_launch.cuh"
