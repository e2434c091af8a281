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

__device__ __forceinline__ int4 ld_stream_v4(const void *p)
{
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
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

// SWAR byte classification, exact for all 256 byte values: 0x80 in every byte of w that is '\n' (nl) / Python-2
// whitespace " \t\n\v\f\r" (ws).  The 7-bit adds cannot carry across bytes.
__device__ __forceinline__ void word_flags(uint32_t w, uint32_t &nl, uint32_t &ws)
{
    const uint32_t a = w & 0x7f7f7f7fu, lo = ~w & 0x80808080u;
    const uint32_t ge9 = a + 0x77777777u, ge10 = a + 0x76767676u, ge11 = a + 0x75757575u, ge14 = a + 0x72727272u;
    const uint32_t ge32 = a + 0x60606060u, ge33 = a + 0x5f5f5f5fu;
    nl = ge10 & ~ge11 & lo;
    ws = ((ge9 & ~ge14) | (ge32 & ~ge33)) & lo;
}
// the eight 0x80 flags of two words -> 8 bits (byte i of f0 -> bit i, byte i of f1 -> bit 4+i); the partial
// products of the multiply land on distinct bit positions, so there are no carries
__device__ __forceinline__ uint32_t gather8(uint32_t f0, uint32_t f1)
{
    return ((((f0 >> 7) | (f1 >> 3)) * 0x00204081u) >> 21) & 0xffu;
}
__device__ __forceinline__ void vector_masks(const int4 q, uint32_t &nlm, uint32_t &wsm)
{
    uint32_t n0, n1, n2, n3, w0, w1, w2, w3;
    word_flags((uint32_t)q.x, n0, w0);
    word_flags((uint32_t)q.y, n1, w1);
    word_flags((uint32_t)q.z, n2, w2);
    word_flags((uint32_t)q.w, n3, w3);
    nlm = gather8(n0, n1) | gather8(n2, n3) << 8;
    wsm = gather8(w0, w1) | gather8(w2, w3) << 8;
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

// ------------------------------------------------------------------------------------------------------------
// K1: frame.  One 16 KiB tile per CTA iteration (dynamic ticket so that look-back predecessors always run).
// ------------------------------------------------------------------------------------------------------------
constexpr int FT = 256;
constexpr int FVEC = kFrameTile / 16;            // 1024 vectors of 16 bytes
constexpr int FHV = kFrameHalo / 16;             // halo vectors
constexpr int FW = (FVEC + FHV) / 4;             // 64-bit bitmap words
constexpr int FBITS = (FVEC + FHV) * 16;

struct FrameTile {
    const u64 *ws, *nl;   // bitmaps in shared memory, bit i = byte tbase+i
    const uint8_t *data;
    uint32_t tbase, n;
};

// first tile-local index >= from whose bit is set in (STOP ? ~ws|nl : ws); continues in global memory past the
// bitmap; returns a global position, n when nothing is found before the end of the buffer
template <bool STOP> __device__ __forceinline__ uint32_t find_fwd(const FrameTile &t, uint32_t from)
{
    uint32_t w = from >> 6;
    if (w < (uint32_t)FW) {
        u64 m = (STOP ? (~t.ws[w] | t.nl[w]) : t.ws[w]) & (~0ull << (from & 63));
        for (;;) {
            if (m) {
                uint32_t p = t.tbase + w * 64 + __ffsll((long long)m) - 1;
                return p < t.n ? p : t.n;
            }
            if (++w >= (uint32_t)FW) break;
            m = STOP ? (~t.ws[w] | t.nl[w]) : t.ws[w];
        }
        from = FBITS;
    }
    for (u64 p = (u64)t.tbase + from; p < t.n; ++p) {
        uint32_t c = t.data[p];
        if (STOP ? (c == '\n' || !is_ws(c)) : is_ws(c)) return (uint32_t)p;
    }
    return t.n;
}

// rstrip: end of the line content for the line terminated at tile-local index q (Py2 str.rstrip())
__device__ __forceinline__ uint32_t rstrip_end(const FrameTile &t, uint32_t q)
{
    int w = (int)(q >> 6);
    u64 m = (~t.ws[w] | t.nl[w]) & ((1ull << (q & 63)) - 1ull);
    for (;;) {
        if (m) return t.tbase + w * 64 + 64 - __clzll((long long)m);
        if (--w < 0) break;
        m = ~t.ws[w] | t.nl[w];
    }
    uint32_t p = t.tbase;
    while (p > 0) {
        uint32_t c = t.data[p - 1];
        if (c == '\n' || !is_ws(c)) break;
        --p;
    }
    return p;
}

// name = line.rstrip()[1:].split()[0] for the header line starting at global position hs (FastqReader.py:42)
__device__ __forceinline__ void parse_header(const FrameTile &t, uint32_t hs, uint32_t *noff, uint32_t *nlen)
{
    *noff = hs < t.n ? hs : t.n;
    *nlen = 0;
    if (hs >= t.n) return;
    uint32_t hq = hs - t.tbase;
    bool first_is_nl = hq < (uint32_t)FBITS ? ((t.nl[hq >> 6] >> (hq & 63)) & 1ull) : (t.data[hs] == '\n');
    if (first_is_nl) return;                       // empty line: no token
    uint32_t x = find_fwd<true>(t, hq + 1);        // first non-blank byte or newline after the dropped first byte
    if (x >= t.n) return;
    uint32_t xq = x - t.tbase;
    bool x_is_nl = xq < (uint32_t)FBITS ? ((t.nl[xq >> 6] >> (xq & 63)) & 1ull) : (t.data[x] == '\n');
    if (x_is_nl) return;
    uint32_t e = find_fwd<false>(t, xq + 1);       // the token ends at the next whitespace (or end of buffer)
    *noff = x;
    *nlen = e - x;
}

__global__ void __launch_bounds__(FT) k_frame(const __grid_constant__ BatchDev b, uint32_t total_tiles)
{
    __shared__ __align__(16) uint16_t s_nl[FVEC + FHV];
    __shared__ __align__(16) uint16_t s_ws[FVEC + FHV];
    __shared__ uint32_t s_warp[9];
    __shared__ uint32_t s_tile, s_base;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    for (;;) {
        if (tid == 0) s_tile = atomicAdd(&b.state->ticket_frame, 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= total_tiles) return;
        int s = 0;
        while (s + 1 < b.n_streams && tile >= b.frame_tile_start[s + 1]) ++s;
        const StreamDev &st = b.st[s];
        const int lt = (int)(tile - b.frame_tile_start[s]);
        const uint32_t n = st.n;
        const uint32_t tbase = (uint32_t)lt * (uint32_t)kFrameTile;
        const uint8_t *data = st.data;

        // 1. coalesced 128-bit loads -> per-vector newline / whitespace bit masks in shared memory
        if ((u64)tbase + FBITS + 16u <= (u64)n) {        // tile and halo entirely inside the buffer
#pragma unroll
            for (int i = 0; i < FVEC / FT; ++i) {
                const int v = i * FT + tid;
                uint32_t nlm, wsm;
                vector_masks(ld_stream_v4(data + (size_t)tbase + (size_t)v * 16u), nlm, wsm);
                s_nl[v] = (uint16_t)nlm;
                s_ws[v] = (uint16_t)wsm;
            }
            if (tid < FHV) {
                uint32_t nlm, wsm;
                vector_masks(ld_stream_v4(data + (size_t)tbase + (size_t)(FVEC + tid) * 16u), nlm, wsm);
                s_nl[FVEC + tid] = (uint16_t)nlm;
                s_ws[FVEC + tid] = (uint16_t)wsm;
            }
        } else {
            const bool vnl = ((b.last >> s) & 1) && n > 0 && data[n - 1] != '\n'; // final line without '\n' still is a line
            for (int v = tid; v < FVEC + FHV; v += FT) {
                const u64 pos = (u64)tbase + (u64)v * 16u;
                uint32_t nlm = 0, wsm = 0xffffu;
                if (pos < (((u64)n + 15u) & ~15ull)) {
                    vector_masks(ld_stream_v4(data + pos), nlm, wsm);
                    if (pos + 16u > n) { // bytes past the end count as blanks
                        uint32_t keep = (1u << (uint32_t)(n - pos)) - 1u;
                        nlm &= keep;
                        wsm |= ~keep & 0xffffu;
                    }
                }
                if (vnl && (u64)n >= pos && (u64)n < pos + 16u) nlm |= 1u << (uint32_t)(n - pos);
                s_nl[v] = (uint16_t)nlm;
                s_ws[v] = (uint16_t)wsm;
            }
        }
        __syncthreads();

        // 2. thread t owns bytes [64t, 64t+64): count, block scan, look-back for the global line index
        const u64 *NL = reinterpret_cast<const u64 *>(s_nl), *WS = reinterpret_cast<const u64 *>(s_ws);
        u64 mine = NL[tid];
        uint32_t total;
        uint32_t excl = block_scan256((uint32_t)__popcll(mine), s_warp, &total);
        if (warp == 0) {
            uint32_t base = lookback(st.status_frame, lt, total, lane);
            if (lane == 0) {
                s_base = base;
                if ((u64)tbase + kFrameTile > (u64)n) b.state->n_lines[s] = base + total; // the tile holding position n
            }
        }
        __syncthreads();
        uint32_t k = s_base + excl;

        // 3. per line end: newline position, rstripped end; per record start: name token
        FrameTile ft{WS, NL, data, tbase, n};
        if (lt == 0 && tid == 0 && st.cap_rec > 0) {
            uint32_t o, l;
            parse_header(ft, 0u, &o, &l);
            st.noff[0] = o;
            st.nlen[0] = l;
        }
        while (mine) {
            const uint32_t bit = (uint32_t)__ffsll((long long)mine) - 1u;
            mine &= mine - 1ull;
            const uint32_t q = (uint32_t)tid * 64u + bit, pos = tbase + q;
            const uint32_t j = k >> 2, r = k & 3u;
            if (j < st.cap_rec) {
                st.nl[k] = pos;
                if (r == 1u) st.e1[j] = rstrip_end(ft, q);
                if (r == 3u) {
                    st.e3[j] = rstrip_end(ft, q);
                    if (j + 1 < st.cap_rec) {
                        uint32_t o, l;
                        parse_header(ft, pos + 1u, &o, &l);
                        st.noff[j + 1] = o;
                        st.nlen[j + 1] = l;
                    }
                }
            } else {
                b.state->err_capacity = 1u;
            }
            ++k;
        }
        __syncthreads(); // bitmaps are rewritten by the next iteration
    }
}

} // namespace
} // namespace quade

#include "quade_kernels2.cuh"
#include "quade_frame2.cuh"
#include "quade_frame3.cuh"
#include "quade_frame4.cuh"
#include "quade_frame5.cuh"
#include "quade_frame6.cuh"
#include "quade_kernels3.cuh"
#include "quade_compact.cuh"
#include "quade_compact3.cuh"
#include "quade_launch.cuh"
