// quade_compact2.cuh -- K5 compact, stream ping-pong form (included by quade_kernels.cu after quade_compact.cuh)
//
// Same output and the same two phases per record as k_compact (quade_compact.cuh), another schedule.  k_compact stages
// the R1 span and the R2 span of a tile, waits for both, then works on both: the copy's round trip (a DRAM latency) is
// exposed once per tile and only other warps hide it -- and a second stage per warp would halve the warps.  Here a
// warp alternates between the two read streams of its tiles:
//
//      wait R1(t) | HEAD + copy of the R1 records of tile t | issue R1(t+1) | wait R2(t) | R2 records of t | issue R2(t+1)
//
// so each stream's copy flies while the warp is busy with the other stream: the two spans of k_compact's one stage ARE
// the two stages, no more shared memory.  A half step has at most 8 records (tp <= 8 pairs per tile), so phase 1 uses
// four lanes per record ('@' + half the name | suffix | first sequence bytes | rest of the name) and keeps 8 HEAD
// slots instead of 16.
#pragma once

namespace quade {
namespace {

#ifndef QUADE_C4_WPC
#define QUADE_C4_WPC 28
#endif
constexpr int C4_WPC = QUADE_C4_WPC;
constexpr int C4_SMEM = 2 * (C3_CAP + 32) + C3_TP * C3_HS + C3_TP * C3_SUFV * 16 + C3_TP * 16 + 32;
static_assert(C4_SMEM % 16 == 0, "per-warp shared memory regions stay 16-byte aligned");
static_assert(C3_TP <= 8, "k_compact2: 8 record slots x 4 roles");

struct HalfMeta {
    uint4 rc;                 // this lane's record {name_off, seq_off, qual_off, seq_len}
    uint32_t nlen, cls, ooff;
    uint32_t start, size;     // staged span of the stream (uniform); size 0 = nothing staged
    bool staged;              // uniform: the span fits the stage
};

__global__ void __launch_bounds__(32 * C4_WPC) k_compact2(const __grid_constant__ BatchDev b)
{
    extern __shared__ __align__(128) uint8_t c4_smem_all[];
    const BatchState *St = b.state;
    if (St->err_capacity) return;
    const int lane = threadIdx.x & 31;
    const uint32_t warp_in_cta = threadIdx.x >> 5;
    const uint32_t first_tile = blockIdx.x * (uint32_t)C4_WPC + warp_in_cta, tile_stride = gridDim.x * (uint32_t)C4_WPC;
    const uint32_t n_pairs = St->n_pairs;
    uint32_t tp = St->compact_tp;
    tp = tp < 1u ? 1u : (tp > (uint32_t)C3_TP ? (uint32_t)C3_TP : tp);
    const uint32_t n_tiles = (n_pairs + tp - 1u) / tp;

    const uint32_t s_in = smem_u32(c4_smem_all + warp_in_cta * (uint32_t)C4_SMEM);     // [stream][C3_CAP + 32]
    const uint32_t s_head = s_in + 2 * (C3_CAP + 32);                   // [TP][C3_HS]
    const uint32_t s_suf = s_head + C3_TP * C3_HS;                      // [TP][64]
    const uint32_t s_par = s_suf + C3_TP * C3_SUFV * 16;                // [TP] uint4: phase-2 parameters
    const uint32_t s_bar = s_par + C3_TP * 16;                          // 2 mbarriers, one per stream
    if (lane == 0) {
        mbar_init(s_bar, 1);
        mbar_init(s_bar + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();

    const uint32_t pr = lane & 7, role = lane >> 3;                     // pair slot of the tile, role of this lane
    const bool rec_lane = pr < tp;
    const uint32_t sufv = b.suf_cap >> 4;                               // suffix slot = sufv 16-byte vectors

    auto load_loc = [&](int k, uint32_t tile) -> uint32_t {
        const uint32_t p = tile * tp + pr;
        return rec_lane && tile < n_tiles && p < n_pairs ? b.st[k].loc[p] : 0u;
    };
    // only ISSUES the loads; finish_meta consumes them half a step later
    auto load_raw = [&](int k, uint32_t tile, HalfMeta &m, uint32_t v) {
        const uint32_t p = tile * tp + pr;
        m.rc = make_uint4(0, 0, 0, 0);
        m.nlen = m.cls = m.ooff = 0u;
        if (rec_lane && tile < n_tiles && p < n_pairs) {
            m.cls = b.cls[p];
            m.rc = b.st[k].rec[v];
            m.nlen = b.st[k].nv[v].x;
            m.ooff = (k ? b.out2 : b.out1)[p];
        }
    };
    auto finish_meta = [&](int k, HalfMeta &m) {
        if (!(m.cls >> 31)) {                                           // the bin's writer is off: nothing to emit
            m.rc = make_uint4(0, 0, 0, 0);
            m.nlen = m.ooff = 0u;
        }
        // span of the stream: first .. last written record of the tile (lanes of role 0 vote)
        const unsigned mine = __ballot_sync(0xffffffffu, role == 0u && rec_lane && (m.cls >> 31));
        m.staged = true;
        m.start = m.size = 0u;
        if (mine) {
            const int lo = __ffs(mine) - 1, hi = 31 - __clz(mine);
            const uint32_t first = __shfl_sync(0xffffffffu, m.rc.x, lo);
            const uint32_t lz = __shfl_sync(0xffffffffu, m.rc.z, hi), lw = __shfl_sync(0xffffffffu, m.rc.w, hi);
            m.start = first & ~15u;
            uint32_t end = (lz + lw + 1u + 15u) & ~15u;                 // one byte past the quality string, rounded up
            const uint32_t nmax = (b.st[k].n + 15u) & ~15u;
            end = end < nmax ? end : nmax;
            m.size = end - m.start;
            if (m.size > (uint32_t)C3_CAP) { m.staged = false; m.size = 0u; }
        }
    };
    auto issue = [&](int k, const HalfMeta &m) {
        if (lane == 0 && m.size) {
            QCHK(m.size <= (uint32_t)C3_CAP && (m.size & 15u) == 0u && (m.start & 15u) == 0u, 205u);
            const uint32_t bar = s_bar + 8u * (uint32_t)k;
            fence_proxy_async();
            mbar_expect_tx(bar, m.size);
            bulk_g2s(s_in + (uint32_t)k * (C3_CAP + 32) + 16u, b.st[k].data + m.start, m.size, bar);
        }
    };

    // one half step: the records of stream k of one tile, staged by an earlier issue(k, ...)
    auto half_step = [&](int k, const HalfMeta &cur, const uint4 suf0, const uint4 suf1, uint32_t tile, uint32_t &phase_bits) {
        if (cur.size) {
            const uint32_t bar = s_bar + 8u * (uint32_t)k;
            while (!mbar_try_wait(bar, (phase_bits >> k) & 1u)) {}
            phase_bits ^= 1u << k;
        }
        // ---------------- phase 1: HEAD slots, four lanes per record ----------------
        const uint32_t p = tile * tp + pr;
        const bool written = rec_lane && (cur.cls >> 31);
        const uint32_t L = cur.rc.w, nlen = cur.nlen, suf_len = (cur.cls >> 16) & 0xffu;
        const uint32_t total = nlen + suf_len + 2u * L + 5u, D1 = 1u + nlen + suf_len;
        const uint32_t a = cur.ooff & 15u;
        uint32_t J = ((a + D1 + 15u) & ~15u) - a;
        J = J < total ? J : total;
        const uint32_t sbase = s_in + (uint32_t)k * (C3_CAP + 32) + 16u - cur.start;   // + g = address of global byte g of the stream
        // srcb + y = address, in the staged span, of byte y of the record's 16-byte aligned output frame (y = a + x for
        // output byte x >= D1); the HEAD slot has the SAME residue mod 16, so phase 2 shifts every vector alike
        const uint32_t srcb = sbase + cur.rc.y - D1 - a;
        const uint32_t hb = s_head + pr * C3_HS + (srcb & 15u);
        bool fast = written && cur.staged && (cur.rc.z - cur.rc.y == L + 3u) && (15u + a + J <= (uint32_t)C3_HS) &&
                    total < (1u << 20) - 16u && suf_len <= (sufv < (uint32_t)C3_SUFV ? sufv : (uint32_t)C3_SUFV) * 16u &&
                    (cur.rc.z + L < b.st[k].n);
        if (k == 0 && role == 1u && written) {                          // the pair's suffix: staged once, used by both reads
            sts128(s_suf + (pr * C3_SUFV + 0u) * 16u, suf0);
            sts128(s_suf + (pr * C3_SUFV + 1u) * 16u, suf1);
            const uint4 *sp = reinterpret_cast<const uint4 *>(b.suffix + (size_t)p * b.suf_cap);
            for (uint32_t v = 2; v < sufv && v < (uint32_t)C3_SUFV; ++v) sts128(s_suf + (pr * C3_SUFV + v) * 16u, sp[v]);
        }
        // the record's own lane (role 0) also looks at the text; the other roles copy on the register checks alone
        // (in bounds; a HEAD slot of a record that turns out not to be clean is never read)
        if (role == 0u && fast) fast = lds8(sbase + cur.rc.y + L + 1u) == '+' && lds8(sbase + cur.rc.z + L) == '\n';
        __syncwarp();
        if (fast) {
            // ONE copy per lane, the same call for all four roles (three calls in three branches ran one after the
            // other): '@' + first half of the name | suffix | first sequence bytes | second half of the name
            const uint32_t h = (nlen >> 1) & ~3u;
            uint32_t d = hb + a + 1u, s = sbase + cur.rc.x, n = h;
            if (role == 1u) { d += nlen; s = s_suf + pr * C3_SUFV * 16u; n = suf_len; }
            if (role == 2u) { d = hb + a + D1; s = sbase + cur.rc.y; n = J - D1; }
            if (role == 3u) { d += h; s += h; n = nlen - h; }
            if (role == 0u) sts8(hb + a, '@');
            seg_copy(d, s, n);
        }
        if (role == 0u) {                                               // all 8 slots: phase 2 reads slot tp when tp is odd
            if (fast) {
                const uint32_t endb = a + total, F1 = endb >> 4;
                const uint32_t h0 = a, h1 = a ? (endb < 16u ? endb : 16u) : 0u;     // first partial vector [h0, h1)
                const uint32_t t1 = (F1 > 0u || a == 0u) ? endb : 0u;               // last partial vector [16 F1, t1)
                sts128(s_par + pr * 16u, make_uint4(cur.ooff - a, F1 | (a + J) << 16 | (srcb & 15u) << 24 | (a ? 1u : 0u) << 28,
                                                    (hb - s_in) | (srcb - s_in + 1024u) << 16, h0 | h1 << 4 | t1 << 9));
            } else {
                sts128(s_par + pr * 16u, make_uint4(cur.ooff, 0u, 1024u << 16, 0u));   // inactive in phase 2
            }
        }
        const unsigned slow = __ballot_sync(0xffffffffu, role == 0u && written && !fast);
        __syncwarp();

        // ---------------- phase 2: two records per iteration (one per half warp), branch-free ----------------
        {
            const uint32_t hw = (uint32_t)lane >> 4, l16 = (uint32_t)lane & 15u;
            uint8_t *const obase = b.out[k];
            const uint32_t n_it = (tp + 1u) >> 1;
#pragma unroll 2
            for (uint32_t it = 0; it < n_it; ++it) {
                const uint4 P = lds128(s_par + (2u * it + hw) * 16u);
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
                    QCHK(!(v >= vlo && v < F1) || (gdst + y0 >= obase && gdst + y0 + 16u <= obase + St->out_total[k]), 202u);
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
                    QCHK(!okb || (gdst + y >= obase && gdst + y < obase + St->out_total[k]), 204u);
                    if (okb) gdst[y] = (uint8_t)lds8(bsrc);
                }
            }
        }
        if (slow) {                                                                // records off the fast path
#pragma unroll 1
            for (unsigned rem = slow; rem; rem &= rem - 1u) {
                const uint32_t i = (uint32_t)__ffs((int)rem) - 1u;                 // pair slot = its role-0 lane
                const uint32_t pp = tile * tp + i;
                emit_generic(b, k, pp, b.cls[pp], lds32(s_par + i * 16u), lane);
            }
        }
        __syncwarp();
    };

    uint32_t phase_bits = 0;
    uint32_t tile = first_tile;
    HalfMeta m[2];
    uint32_t vn[2];
    uint4 suf0 = make_uint4(0, 0, 0, 0), suf1 = suf0;                   // suffix vectors of the CURRENT tile (role-1 lanes)
    auto load_suffix = [&](uint32_t t, uint4 &s0, uint4 &s1) {          // issued with the R1 metadata, a tile ahead
        const uint32_t p = t * tp + pr;
        s0 = s1 = make_uint4(0, 0, 0, 0);
        if (role == 1u && rec_lane && t < n_tiles && p < n_pairs) {
            const uint4 *sp = reinterpret_cast<const uint4 *>(b.suffix + (size_t)p * b.suf_cap);
            s0 = sp[0];
            if (sufv > 1u) s1 = sp[1];
        }
    };
    if (tile < n_tiles) {
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            load_raw(k, tile, m[k], load_loc(k, tile));
            finish_meta(k, m[k]);
            issue(k, m[k]);
        }
        load_suffix(tile, suf0, suf1);
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) vn[k] = load_loc(k, tile + tile_stride);
#pragma unroll 1
    for (; tile < n_tiles; tile += tile_stride) {
        const uint32_t ntile = tile + tile_stride;
        uint4 nsuf0, nsuf1;
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            HalfMeta n;
            load_raw(k, ntile, n, vn[k]);                              // loads only; consumed after this half step
            vn[k] = load_loc(k, ntile + tile_stride);
            if (k == 0) load_suffix(ntile, nsuf0, nsuf1);
            half_step(k, m[k], suf0, suf1, tile, phase_bits);
            finish_meta(k, n);                                         // the stage of stream k is free: next tile's span
            if (ntile < n_tiles) issue(k, n);
            m[k] = n;
        }
        suf0 = nsuf0;
        suf1 = nsuf1;
    }
}

} // namespace
} // namespace quade
