// quade_compact.cuh -- K5 compact, TMA-staged (included by quade_kernels.cu)
//
//   '@' name ':' index [':' molecular] '\n' seq '\n' '+' '\n' qual '\n'      (FastqWriter.py:61-69, FastqSeq.py:76-82)
//
// One warp (= one CTA) owns tiles of `tp` (<= 8) consecutive pairs = 2*tp records.  The raw bytes of the tile's
// R1 records and of its R2 records are each ONE contiguous span of the input (pairs are in input order), so
// they are staged with two cp.async.bulk (TMA) copies per tile into a two-stage shared-memory ring, signalled
// through an mbarrier; the copy of tile t+1 is in flight while tile t is assembled.
//
// Phase 1 (dense, lane = record x role): the variable part of the record -- '@' name suffix and the first
// bytes of the sequence up to the next 16-byte boundary of the DESTINATION -- is composed word-wise into a
// per-record HEAD slot at destination alignment.
// Phase 2 (one record per iteration, lane = 16-byte vector of the output): HEAD vectors are copied from the
// slot, the rest (seq "\n+\n" qual "\n", contiguous in a clean input record) is funnel-shifted from the staged
// span; aligned 128-bit stores, the first / last partial vector by bytes.
// Records that are not clean (CRLF, decorated '+' line, trailing blanks, no final newline), too long for the
// stage or with an oversized name take the generic byte path straight from global memory.
#pragma once

namespace quade {
namespace {

constexpr int C3_TP = kCompactTilePairs;                  // pairs per tile (upper bound; the batch's value is in BatchState)
static_assert(C3_TP >= 1 && C3_TP <= 8, "k_compact: one record lane per read of the tile, 16 record lanes");
constexpr int C3_CAP = kCompactStageBytes;              // staged bytes per stream per stage
#ifndef QUADE_C_PAIRED
#define QUADE_C_PAIRED 1        /* phase 2 copies a pair (R1 | R2 on the two half warps) per iteration, one packed parameter block per record */
#endif
#ifndef QUADE_C_UNROLL
#define QUADE_C_UNROLL 8         /* phase 2 unrolled over the pairs of a tile: 1 / 2 / 4 / 8 measured 0.609 / 0.607 / 0.612 / 0.599 ms */
#endif
#ifndef QUADE_C3_HS
#define QUADE_C3_HS 144         /* name + suffix up to 98 bytes always take the fast path (176: 130); longer ones the generic one */
#endif
#ifndef QUADE_C3_SUFV
#define QUADE_C3_SUFV 4
#endif
constexpr int C3_UNROLL = QUADE_C_UNROLL;
constexpr int C3_HS = QUADE_C3_HS;                // HEAD slot stride: 16-byte multiple (slots are read back as vectors)
constexpr int C3_SUFV = QUADE_C3_SUFV;                // staged suffix: up to 64 bytes per pair
#ifndef QUADE_C3_STAGES
#define QUADE_C3_STAGES 1
#endif
constexpr int C3_STAGES = QUADE_C3_STAGES;    // 1: more resident warps hide the copy latency; 2: per-warp prefetch
constexpr int C3_IN_BYTES = C3_STAGES * 2 * (C3_CAP + 32);
constexpr int C3_PAR = QUADE_C_PAIRED ? 16 : 32;  // phase-2 parameter bytes per record
constexpr int C3_SMEM = C3_IN_BYTES + 2 * C3_TP * C3_HS + C3_TP * C3_SUFV * 16 + 2 * C3_TP * C3_PAR + 32;

// shared -> shared copy of n bytes, any alignment on both sides, full destination words by funnel shift
__device__ __forceinline__ void seg_copy(uint32_t dst, uint32_t src, uint32_t n)
{
    uint32_t head = (4u - (dst & 3u)) & 3u;
    head = head < n ? head : n;
    for (uint32_t i = 0; i < head; ++i) sts8(dst + i, lds8(src + i));
    dst += head;
    src += head;
    n -= head;
    const uint32_t nw = n >> 2, sh = (src & 3u) * 8u;
    uint32_t sa = src & ~3u;
    uint32_t w0 = nw ? lds32(sa) : 0u;
    for (uint32_t i = 0; i < nw; ++i) {
        sa += 4u;
        const uint32_t w1 = lds32(sa);
        sts32(dst, __funnelshift_r(w0, w1, sh));
        w0 = w1;
        dst += 4u;
    }
    src += nw * 4u;
    for (uint32_t i = 0; i < (n & 3u); ++i) sts8(dst + i, lds8(src + i));
}

// byte path straight from global memory: any record shape (the whole warp works on one record)
__device__ __forceinline__ void emit_generic(const BatchDev &b, int k, uint32_t p, uint32_t cls, uint32_t ooff, int lane)
{
    const uint32_t v = b.st[k].loc[p];
    const uint4 rc = b.st[k].rec[v];
    const uint32_t nlen = b.st[k].nv[v].x, L = rc.w, suf_len = (cls >> 16) & 0xffu;
    const uint8_t *sufp = b.suffix + (size_t)p * b.suf_cap, *data = b.st[k].data;
    uint8_t *gout = b.out[k] + (size_t)ooff;
    const uint32_t total = nlen + suf_len + 2u * L + 5u, D1 = 1u + nlen + suf_len, D2 = D1 + L + 3u;
#pragma unroll 1
    for (uint32_t x = lane; x < total; x += 32) {
        uint32_t ch;
        if (x == 0u) ch = '@';
        else if (x < 1u + nlen) ch = data[rc.x + x - 1u];
        else if (x < D1) ch = sufp[x - 1u - nlen];
        else if (x < D1 + L) ch = data[rc.y + x - D1];
        else if (x < D2) ch = (x - D1 - L == 1u) ? '+' : '\n';
        else if (x < D2 + L) ch = data[rc.z + x - D2];
        else ch = '\n';
        QCHK((size_t)ooff + x < b.state->out_total[k], 206u);
        gout[x] = (uint8_t)ch;
    }
}

// 16 bytes starting at byte r (0..15) of the 32 bytes {A,B}; r is warp-uniform
__device__ __forceinline__ uint4 shift_bytes(const uint4 A, const uint4 B, uint32_t r)
{
    const uint32_t sh = (r & 3u) * 8u;
    uint32_t w0, w1, w2, w3, w4;
    switch (r >> 2) {
    case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
    case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
    case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
    default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
    }
    return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                      __funnelshift_r(w3, w4, sh));
}

// 16 bytes starting at byte r (0..15) of the 32 bytes {A,B}; r is warp-uniform.  Two levels of uniform selects
// (r & 4, r & 8) pick the five source words, four funnel shifts align them: 16 instructions, no branch
__device__ __forceinline__ uint4 shift_bytes_sel(const uint4 A, const uint4 B, uint32_t r)
{
    const bool b0 = (r & 4u) != 0u, b1 = (r & 8u) != 0u;
    const uint32_t u0 = b0 ? A.y : A.x, u1 = b0 ? A.z : A.y, u2 = b0 ? A.w : A.z, u3 = b0 ? B.x : A.w, u4 = b0 ? B.y : B.x,
                   u5 = b0 ? B.z : B.y, u6 = b0 ? B.w : B.z;
    const uint32_t w0 = b1 ? u2 : u0, w1 = b1 ? u3 : u1, w2 = b1 ? u4 : u2, w3 = b1 ? u5 : u3, w4 = b1 ? u6 : u4;
    const uint32_t sh = (r & 3u) * 8u;
    return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                      __funnelshift_r(w3, w4, sh));
}

struct TileMeta {
    uint4 rc;          // this lane's record {name_off, seq_off, qual_off, seq_len}
    uint32_t nlen, cls, ooff;
    uint4 suf0, suf1;             // first two 16-byte vectors of the pair's suffix (R1 lanes of role 0)
    uint32_t start0, start1, size0, size1;   // staged span per stream (uniform); size 0 = nothing staged
    bool staged;                  // uniform: the spans fit the stage
};

// C3_WPC independent warps per CTA, each with its own C3_SMEM bytes and tiles: the 1 KiB of shared memory the
// system reserves per CTA is paid once per C3_WPC warps (with the 144-byte HEAD slots: 24 warps of 9568 bytes in one CTA per SM)
#ifndef QUADE_C3_WPC
#define QUADE_C3_WPC 24        // 6 warps per scheduler; measured 20 (1 per CTA) / 22 / 23 / 24 warps per SM = 0.674 / 0.656 / 0.62 / 0.601 ms
#endif
constexpr int C3_WPC = QUADE_C3_WPC;
static_assert(C3_SMEM % 16 == 0, "per-warp shared memory regions stay 16-byte aligned (bulk copies, 128-bit accesses)");

__global__ void __launch_bounds__(32 * C3_WPC) k_compact(const __grid_constant__ BatchDev b)
{
    extern __shared__ __align__(128) uint8_t c3_smem_all[];
    const BatchState *St = b.state;
    if (St->err_capacity) return;
    const int lane = threadIdx.x & 31;
    const uint32_t warp_in_cta = threadIdx.x >> 5;
    uint8_t *c3_smem = c3_smem_all + warp_in_cta * (uint32_t)C3_SMEM;
    const uint32_t first_tile = blockIdx.x * (uint32_t)C3_WPC + warp_in_cta, tile_stride = gridDim.x * (uint32_t)C3_WPC;
    const uint32_t n_pairs = St->n_pairs;
    uint32_t tp = St->compact_tp;
    tp = tp < 1u ? 1u : (tp > (uint32_t)C3_TP ? (uint32_t)C3_TP : tp);
    const uint32_t n_tiles = (n_pairs + tp - 1u) / tp;

    const uint32_t sm0 = smem_u32(c3_smem);
    const uint32_t s_in = sm0;                                         // [stage][stream][C3_CAP + 32]
    const uint32_t s_head = s_in + C3_IN_BYTES;                        // [2*TP][C3_HS]
    const uint32_t s_suf = s_head + 2 * C3_TP * C3_HS;                 // [TP][64]
    const uint32_t s_par = s_suf + C3_TP * C3_SUFV * 16;               // [2*TP] phase-2 parameters
    const uint32_t s_bar = s_par + 2 * C3_TP * C3_PAR;                 // 2 mbarriers
    if (lane == 0) {
        mbar_init(s_bar, 1);
        mbar_init(s_bar + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const uint32_t r = lane & 15, role = lane >> 4;
    const uint32_t k = r >= tp ? 1u : 0u, pr = r - k * tp;             // stream and pair slot of this lane's record
    const bool rec_lane = r < 2u * tp;
    const uint32_t sufv = b.suf_cap >> 4;                              // suffix slot = sufv 16-byte vectors

    // per-pair metadata of a tile.  load_raw only ISSUES the global loads (nothing consumes them), so that the
    // loads of the next tile fly during this tile's phases; finish_meta turns them into the staged spans.
    // the scratch position of a record is read one tile earlier still: two dependent loads, each a tile ahead
    auto load_loc = [&](uint32_t tile) -> uint32_t {
        const uint32_t p = tile * tp + pr;
        return rec_lane && tile < n_tiles && p < n_pairs ? b.st[k].loc[p] : 0u;
    };
    auto load_raw = [&](uint32_t tile, TileMeta &m, uint32_t v) {
        const uint32_t p = tile * tp + pr;
        m.rc = make_uint4(0, 0, 0, 0);
        m.nlen = m.cls = m.ooff = 0u;
        m.suf0 = m.suf1 = make_uint4(0, 0, 0, 0);
        if (rec_lane && tile < n_tiles && p < n_pairs) {
            m.cls = b.cls[p];
            m.rc = b.st[k].rec[v];
            m.nlen = b.st[k].nv[v].x;
            // ONE load through a selected pointer: written as `k ? out2[p] : out1[p]` it compiled to two predicated loads
            // into the same register, and the second waited for the first (a quarter of the kernel's stall samples)
            // Written like this, ptxas emits two predicated loads into one register and the second waits for the first: the
            // warp sits out one DRAM latency here, at the top of every tile (22 % of the kernel's stall samples).  The
            // obvious form -- one load through a selected pointer -- removes that wait and was measured SLOWER
            // (0.665 vs 0.607 ms, A B A B on one box, with and without a sleep / time hint in the mbarrier poll): the wait
            // overlaps the TMA copy of this tile, which is polled for right after; without it the warps spin there.
#if QUADE_C_OOFF1
            const uint32_t *const op = k ? b.out2 : b.out1;
            m.ooff = op[p];
#else
            m.ooff = k ? b.out2[p] : b.out1[p];
#endif
            if (role == 0u && k == 0u) {                                // the pair's suffix, staged by this lane
                const uint4 *sp = reinterpret_cast<const uint4 *>(b.suffix + (size_t)p * b.suf_cap);
                m.suf0 = sp[0];
                if (sufv > 1u) m.suf1 = sp[1];
            }
        }
    };
    auto finish_meta = [&](TileMeta &m) {
        if (!(m.cls >> 31)) {                                           // the bin's writer is off: nothing to emit
            m.rc = make_uint4(0, 0, 0, 0);
            m.nlen = m.ooff = 0u;
        }
        // span of each stream: first .. last written record of the tile (lanes of role 0 vote)
        m.staged = true;
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
            const unsigned mine = __ballot_sync(0xffffffffu, role == 0u && rec_lane && k == (uint32_t)kk && (m.cls >> 31));
            uint32_t start = 0, size = 0;
            if (mine) {
                const int lo = __ffs(mine) - 1, hi = 31 - __clz(mine);
                const uint32_t first = __shfl_sync(0xffffffffu, m.rc.x, lo);
                const uint32_t lz = __shfl_sync(0xffffffffu, m.rc.z, hi), lw = __shfl_sync(0xffffffffu, m.rc.w, hi);
                start = first & ~15u;
                uint32_t end = lz + lw + 1u;                            // one byte past the quality string
                const uint32_t nmax = (b.st[kk].n + 15u) & ~15u;
                end = (end + 15u) & ~15u;
                end = end < nmax ? end : nmax;
                size = end - start;
                if (size > (uint32_t)C3_CAP) { m.staged = false; size = 0; }
            }
            if (kk == 0) { m.start0 = start; m.size0 = size; }
            else { m.start1 = start; m.size1 = size; }
        }
        if (!m.staged) m.size0 = m.size1 = 0;
    };
    auto issue = [&](const TileMeta &m, uint32_t stage) {
        if (lane == 0 && (m.size0 | m.size1)) {
            QCHK(m.size0 <= (uint32_t)C3_CAP && m.size1 <= (uint32_t)C3_CAP && (m.size0 & 15u) == 0u && (m.size1 & 15u) == 0u &&
                     (m.start0 & 15u) == 0u && (m.start1 & 15u) == 0u, 205u);
            const uint32_t bar = s_bar + 8u * stage;
            fence_proxy_async();
            mbar_expect_tx(bar, m.size0 + m.size1);
            if (m.size0) bulk_g2s(s_in + (stage * 2u + 0u) * (C3_CAP + 32) + 16u, b.st[0].data + m.start0, m.size0, bar);
            if (m.size1) bulk_g2s(s_in + (stage * 2u + 1u) * (C3_CAP + 32) + 16u, b.st[1].data + m.start1, m.size1, bar);
        }
    };

    uint32_t phase_bits = 0;
    uint32_t tile = first_tile, it = 0;
    TileMeta cur, nxt;
    uint32_t v_next = load_loc(tile + tile_stride);
    if (tile < n_tiles) {
        load_raw(tile, cur, load_loc(tile));
        finish_meta(cur);
        issue(cur, 0);
    }
#pragma unroll 1
    for (; tile < n_tiles; tile += tile_stride, ++it) {
        const uint32_t stage = C3_STAGES == 2 ? (it & 1u) : 0u;
        const uint32_t ntile = tile + tile_stride;
        load_raw(ntile, nxt, v_next);                                  // loads only; consumed at the end of the iteration
        v_next = load_loc(ntile + tile_stride);
        if (C3_STAGES == 2 && ntile < n_tiles) {
            finish_meta(nxt);
            issue(nxt, stage ^ 1u);
        }
        if (cur.size0 | cur.size1) {
            const uint32_t bar = s_bar + 8u * stage;
#ifdef QUADE_C_SLEEP
            while (!mbar_try_wait(bar, (phase_bits >> stage) & 1u)) __nanosleep(QUADE_C_SLEEP);
#else
            while (!mbar_try_wait(bar, (phase_bits >> stage) & 1u)) {}
#endif
            phase_bits ^= 1u << stage;
        }

        // ---------------- phase 1: HEAD slots ----------------
        const uint32_t p = tile * tp + pr;
        const bool written = rec_lane && (cur.cls >> 31);
        const uint32_t L = cur.rc.w, nlen = cur.nlen, suf_len = (cur.cls >> 16) & 0xffu;
        const uint32_t total = nlen + suf_len + 2u * L + 5u, D1 = 1u + nlen + suf_len;
        const uint32_t a = cur.ooff & 15u;
        uint32_t J = ((a + D1 + 15u) & ~15u) - a;
        J = J < total ? J : total;
        // smem address of global byte g of this lane's stream: sbase + g
        const uint32_t sbase = s_in + (stage * 2u + k) * (C3_CAP + 32) + 16u - (k ? cur.start1 : cur.start0);
        // srcb + y = smem address, in the staged span, of byte y of the record's 16-byte aligned output frame
        // (y = a + x for output byte x >= D1).  The HEAD slot is laid out with the SAME residue mod 16
        // (hb + y, hb = slot + (srcb & 15)), so that phase 2 shifts every vector of the record alike.
        const uint32_t srcb = sbase + cur.rc.y - D1 - a;
        const uint32_t hb = s_head + r * C3_HS + (srcb & 15u);
        bool fast = written && cur.staged && (cur.rc.z - cur.rc.y == L + 3u) && (15u + a + J <= (uint32_t)C3_HS) &&
                    total < (1u << 20) - 16u && suf_len <= (sufv < (uint32_t)C3_SUFV ? sufv : (uint32_t)C3_SUFV) * 16u && (cur.rc.z + L < b.st[k].n);
        if (fast) fast = lds8(sbase + cur.rc.y + L + 1u) == '+' && lds8(sbase + cur.rc.z + L) == '\n';
        // stage the pair's suffix (one copy per pair, by the R1 lanes of role 0)
        if (role == 0u && k == 0u && written) {
            sts128(s_suf + (pr * C3_SUFV + 0u) * 16u, cur.suf0);    // prefetched with the tile's metadata
            sts128(s_suf + (pr * C3_SUFV + 1u) * 16u, cur.suf1);
            const uint4 *sp = reinterpret_cast<const uint4 *>(b.suffix + (size_t)p * b.suf_cap);
            for (uint32_t v = 2; v < sufv && v < (uint32_t)C3_SUFV; ++v) sts128(s_suf + (pr * C3_SUFV + v) * 16u, sp[v]);
        }
        __syncwarp();
        if (fast) {
            uint32_t d1, s1, n1, d2 = 0, s2 = 0, n2 = 0;
            if (role == 0u) {                      // '@' name
                sts8(hb + a, '@');
                d1 = hb + a + 1u; s1 = sbase + cur.rc.x; n1 = nlen;
            } else {                               // suffix, then the sequence bytes up to the vector boundary
                d1 = hb + a + 1u + nlen; s1 = s_suf + pr * C3_SUFV * 16u; n1 = suf_len;
                d2 = hb + a + D1; s2 = sbase + cur.rc.y; n2 = J - D1;
            }
            seg_copy(d1, s1, n1);
            seg_copy(d2, s2, n2);
        }
        // everything phase 2 needs per record is derived HERE, once, by the record's lane (phase 2 runs these
        // values through all 32 lanes 2*tp times per tile): destination pointer of the record's first 16-byte
        // frame, number of full vectors, byte ranges of the partial first / last vector, source addresses
        if (role == 0u && rec_lane) {
#if QUADE_C_PAIRED
            // ONE 16-byte parameter block per record: a broadcast 128-bit shared load costs a wavefront per quarter warp
            // whatever the addresses, so two blocks per record were 8 of the ~22 load wavefronts a record took
            if (fast) {
                const uint32_t endb = a + total, F1 = endb >> 4;
                const uint32_t h0 = a, h1 = a ? (endb < 16u ? endb : 16u) : 0u;     // first partial vector [h0, h1)
                const uint32_t t1 = (F1 > 0u || a == 0u) ? endb : 0u;               // last partial vector [16 F1, t1)
                sts128(s_par + r * 16u, make_uint4(cur.ooff - a,                                             // frame start in the arena
                                                   F1 | (a + J) << 16 | (srcb & 15u) << 24 | (a ? 1u : 0u) << 28,
                                                   (hb - s_in) | (srcb - s_in + 1024u) << 16, h0 | h1 << 4 | t1 << 9));
            } else {
                sts128(s_par + r * 16u, make_uint4(cur.ooff, 0u, 1024u << 16, 0u));    // inactive in phase 2
            }
#else
            if (fast) {
                const uint32_t endb = a + total, F1 = endb >> 4;
                const u64 gp = (u64)(uintptr_t)(b.out[k] + (size_t)(cur.ooff - a));
                const uint32_t h0 = a, h1 = a ? (endb < 16u ? endb : 16u) : 0u;     // first partial vector [h0, h1)
                const uint32_t t1 = (F1 > 0u || a == 0u) ? endb : 0u;               // last partial vector [16 F1, t1)
                sts128(s_par + r * 32u, make_uint4((uint32_t)gp, (uint32_t)(gp >> 32), F1 | (a ? 1u : 0u) << 16, a + J));
                sts128(s_par + r * 32u + 16u, make_uint4(hb, srcb, (srcb & 15u) | h0 << 8 | h1 << 16, t1));
            } else {
                sts128(s_par + r * 32u, make_uint4(cur.ooff, 0u, 0u, 0u));          // inactive in phase 2
                sts128(s_par + r * 32u + 16u, make_uint4(s_in, s_in, 0u, 0u));
            }
#endif
        }
        const unsigned slow = __ballot_sync(0xffffffffu, role == 0u && written && !fast);
        __syncwarp();

#if QUADE_C_PAIRED
        // ---------------- phase 2: one PAIR per iteration, branch-free ----------------
        // lanes 0-15 copy the pair's R1 record, lanes 16-31 its R2 record, two 16-byte vectors per lane (records up to
        // 512 bytes in one step): half the iterations, half the parameter loads and bookkeeping per record
        {
            const uint32_t hw = (uint32_t)lane >> 4, l16 = (uint32_t)lane & 15u;
            uint8_t *const obase = b.out[hw];
#pragma unroll C3_UNROLL
            for (uint32_t it = 0; it < tp; ++it) {
                const uint4 P = lds128(s_par + (it + hw * tp) * 16u);
                uint8_t *const gdst = obase + P.x;
                const uint32_t F1 = P.y & 0xffffu, hend = (P.y >> 16) & 0xffu, sr = (P.y >> 24) & 15u, vlo = (P.y >> 28) & 1u;
                const uint32_t hbi = s_in + (P.z & 0xffffu), sb = s_in + (P.z >> 16) - 1024u;
                const uint32_t h0 = P.w & 15u, h1 = (P.w >> 4) & 31u, t1 = P.w >> 9;
#pragma unroll
                for (uint32_t j = 0; j < 2; ++j) {
                    const uint32_t v = l16 + 16u * j, y0 = 16u * v;
                    const uint32_t sa = ((y0 < hend ? hbi : sb) + y0) & ~15u;
                    const uint4 val = shift_bytes_sel(lds128(sa), lds128(sa + 16u), sr);
                    QCHK(!(v >= vlo && v < F1) || (sa >= s_in && sa + 32u <= s_bar), 201u);
                    QCHK(!(v >= vlo && v < F1) || (gdst + y0 >= obase && gdst + y0 + 16u <= obase + St->out_total[hw]), 202u);
                    if (v >= vlo && v < F1) *reinterpret_cast<uint4 *>(gdst + y0) = val;
                }
                if (F1 > 32u) {                                                    // records longer than 512 bytes
#pragma unroll 1
                    for (uint32_t v = l16 + 32u; v < F1; v += 16u) {
                        const uint32_t sa = (sb + 16u * v) & ~15u;
                        *reinterpret_cast<uint4 *>(gdst + 16u * v) = shift_bytes_sel(lds128(sa), lds128(sa + 16u), sr);
                    }
                }
                // first partial vector: bytes [h0, h1); last one: [16 F1, t1); one byte per lane of the record's half warp
#pragma unroll
                for (uint32_t j = 0; j < 2; ++j) {
                    const uint32_t y = j ? 16u * F1 + l16 : l16;
                    const bool okb = j ? y < t1 : (y >= h0 && y < h1);
                    const uint32_t bsrc = (y < hend ? hbi : sb) + y;
                    QCHK(!okb || (bsrc >= s_in && bsrc < s_bar), 203u);
                    QCHK(!okb || (gdst + y >= obase && gdst + y < obase + St->out_total[hw]), 204u);
                    if (okb) gdst[y] = (uint8_t)lds8(bsrc);
                }
            }
        }
#else
        // ---------------- phase 2: one record per iteration, branch-free ----------------
        const uint32_t nrec = 2u * tp;
#pragma unroll 2
        for (uint32_t i = 0; i < nrec; ++i) {
            const uint4 A = lds128(s_par + i * 32u), B = lds128(s_par + i * 32u + 16u);
            uint8_t *gdst = reinterpret_cast<uint8_t *>((uintptr_t)((u64)A.y << 32 | A.x));
            const uint32_t F1 = A.z & 0xffffu, vlo = A.z >> 16, hend = A.w;        // hend: end of the HEAD part
            const uint32_t hbi = B.x, sb = B.y, sr = B.z & 0xffu, h0 = (B.z >> 8) & 0xffu, h1 = B.z >> 16, t1 = B.w;
            {
                const uint32_t v = lane, y0 = 16u * v;
                const uint32_t sa = ((y0 < hend ? hbi : sb) + y0) & ~15u;
                const uint4 val = shift_bytes_sel(lds128(sa), lds128(sa + 16u), sr);
                QCHK(!(v >= vlo && v < F1) || (sa >= s_in && sa + 32u <= s_bar), 201u);
                QCHK(!(v >= vlo && v < F1) || (gdst + y0 >= b.out[i >= tp ? 1 : 0] && gdst + y0 + 16u <= b.out[i >= tp ? 1 : 0] + St->out_total[i >= tp ? 1 : 0]), 202u);
                if (v >= vlo && v < F1) *reinterpret_cast<uint4 *>(gdst + y0) = val;
            }
            if (F1 > 32u) {                                                        // records longer than 512 bytes
#pragma unroll 1
                for (uint32_t v = lane + 32u; v < F1; v += 32) {
                    const uint32_t sa = (sb + 16u * v) & ~15u;
                    *reinterpret_cast<uint4 *>(gdst + 16u * v) = shift_bytes(lds128(sa), lds128(sa + 16u), sr);
                }
            }
            // first partial vector: bytes [h0, h1) by lanes 0-15; last one: [16*F1, t1) by lanes 16-31
            const uint32_t y = lane < 16 ? (uint32_t)lane : 16u * F1 + (uint32_t)(lane - 16);
            const bool okb = lane < 16 ? (y >= h0 && y < h1) : y < t1;
            const uint32_t bsrc = (y < hend ? hbi : sb) + y;
            QCHK(!okb || (bsrc >= s_in && bsrc < s_bar), 203u);
            QCHK(!okb || (gdst + y >= b.out[i >= tp ? 1 : 0] && gdst + y < b.out[i >= tp ? 1 : 0] + St->out_total[i >= tp ? 1 : 0]), 204u);
            if (okb) gdst[y] = (uint8_t)lds8(bsrc);
        }
#endif
        if (slow) {                                                                // records off the fast path
#pragma unroll 1
            for (unsigned rem = slow; rem; rem &= rem - 1u) {
                const uint32_t i = (uint32_t)__ffs((int)rem) - 1u;                 // record slot = its role-0 lane
                const uint32_t kk = i >= tp ? 1u : 0u, pp = tile * tp + (i - kk * tp);
                emit_generic(b, (int)kk, pp, b.cls[pp], lds32(s_par + i * (QUADE_C_PAIRED ? 16u : 32u)), lane);
            }
        }
        __syncwarp();
        if (C3_STAGES == 1) {
            finish_meta(nxt);
            if (ntile < n_tiles) issue(nxt, 0u);
        }
        cur = nxt;
    }
}

} // namespace
} // namespace quade
