// quade_frame6.cuh -- K1 k_frame6: dependency-free framing (included by quade_kernels.cu after quade_frame2/3.cuh)
//
// What every single-pass variant before it (k_frame2 .. k_frame5) was bound by, measured with clock64 probes
// (tools/probe_f3prof.py): the decoupled look-back for a tile's global line index k0 takes ~5 us per tile -- not
// the L2 round trips but the wait for the slowest of the several hundred predecessor tiles in flight to publish
// its line count (a running maximum over a jittery ticket -> copy -> scan latency).  Scan warps sat idle 79 % of
// the time.  k0 is needed for two things only:
//   (a) k0 mod 4: which line ends close a record.  GUESSED from the text -- the first line end i in 0..3 whose
//       next line starts with '@' and whose third next line starts with '+' (exact for well-formed FASTQ: a
//       quality line may start with '@', but the line two after it is then a sequence line);
//   (b) the record index the outputs are stored at.  DEFERRED: a tile reserves room in per-stream scratch arrays
//       with one atomic and stores its records there in tile order of arrival.
// So tiles are independent: no look-back, no ordering between CTAs, nothing to wait for.  Afterwards
//   k_tilescan  one CTA per stream: exclusive scan of the tiles' line counts -> true k0, record base and record
//               count per tile; a tile whose guess was wrong (or that found no scratch room) goes to a redo list;
//   k_densify   one warp per tile: scratch -> the dense per-record arrays at the true record index; counts records
//               with len(seq) != len(qual) and finds the first blank header (FastqReader.py:42,48-50);
//   k_frame6<true> parses the redo tiles again with the true k0, straight into the dense arrays.
// The result never depends on the guess; on well-formed input the redo list is empty.
#pragma once

namespace quade {
namespace {

constexpr int F6T = 256;
#ifndef F6_MINB
#define F6_MINB 8        // 32 registers (a few spills in the parse path) and 64 resident warps: 0.653 -> 0.631 ms vs 6 CTAs at 40
#endif
// shared memory of a CTA: window | newline bitmap | line-end list | misc (no candidate bitmap: 24.4 KiB, 8 CTAs / SM)
constexpr int F6_PCAP = 2048;                  // listed line ends per window (else: redo pass)
constexpr int F6_OFF_BM = F2_WIN + 16;
constexpr int F6_OFF_POS = F6_OFF_BM + F2_WORDS * 8;            // u16[F6_PCAP]
constexpr int F6_OFF_MISC = F6_OFF_POS + F6_PCAP * 2;
constexpr int F6_SMEM = F6_OFF_MISC + 128;
constexpr uint32_t F6_BLANK = 1u << 25;        // scratch digest: the header had no name token

// ---- record parser without the candidate bitmap ------------------------------------------------------------
// k_frame2..5 keep a second bitmap (bytes <= 0x20) for every input byte so that the record threads can hop over
// name tokens and test line ends for trailing blanks; building it costs ~19 instructions per 16 input bytes for
// ALL bytes, while only the ~40 header bytes of a record are ever searched.  Here the record thread scans its
// own header with the same SWAR test, 4 bytes per shared-memory load.

// first position in [p, e) holding whitespace (Python-2 str set), e if none
__device__ __forceinline__ uint32_t f6_find_ws(const F2Tile &t, uint32_t p, uint32_t e)
{
    if (p >= e) return e;
    if (p >= t.tbase && e - t.tbase <= t.win) {                 // the whole range is staged
        const uint32_t a = t.sdata + (p - t.tbase), end = t.sdata + (e - t.tbase);
        uint32_t a4 = a & ~3u;
        uint32_t w = lds32(a4);
        uint32_t f = f2_c_flags(w) & (0xffffffffu << ((a & 3u) * 8u));   // bytes <= 0x20 at or after a
        for (;;) {
            while (f) {
                const uint32_t i = ((uint32_t)__ffs((int)f) - 1u) >> 3;
                const uint32_t at = a4 + i;
                if (at >= end) return e;
                if (is_ws((w >> (8u * i)) & 0xffu)) return p + (at - a);
                f &= f - 1u;                                    // a control character inside the token
            }
            a4 += 4u;
            if (a4 >= end) return e;
            w = lds32(a4);
            f = f2_c_flags(w);
        }
    }
    for (; p < e; ++p)
        if (is_ws(f2_byte(t, p))) return p;
    return e;
}

__device__ __forceinline__ uint32_t f6_rstrip(const F2Tile &t, uint32_t s, uint32_t e)
{
    while (e > s && is_ws(f2_byte(t, e - 1u))) --e;
    return e;
}

// f2_parse without t.bc (FastqReader.py:37-44, FastqSeq.py:55)
__device__ __forceinline__ F2Rec f6_parse(const F2Tile &t, uint32_t hs, const uint16_t *pos, uint32_t li, uint32_t npos)
{
    F2Rec r;
    r.complete = r.valid = false;
    r.name_off = r.nlen = r.seq_off = r.qual_off = r.seq_len = r.end = 0u;
    if (hs > t.n) return r;
    const uint32_t nl0 = li < npos ? t.tbase + pos[li] : f2_next_nl(t, hs);
    if (nl0 == F2_NONE || nl0 >= t.n) return r;           // a line ending at n is the file's last line
    const uint32_t nl1 = li + 1u < npos ? t.tbase + pos[li + 1u] : f2_next_nl(t, nl0 + 1u);
    if (nl1 == F2_NONE || nl1 >= t.n) return r;
    const uint32_t nl2 = li + 2u < npos ? t.tbase + pos[li + 2u] : f2_next_nl(t, nl1 + 1u);
    if (nl2 == F2_NONE || nl2 >= t.n) return r;
    const uint32_t nl3 = li + 3u < npos ? t.tbase + pos[li + 3u] : f2_next_nl(t, nl2 + 1u);
    if (nl3 == F2_NONE) return r;
    r.complete = true;
    r.end = nl3 < t.n ? nl3 + 1u : t.n;
    // name = line.rstrip()[1:].split()[0]: drop the first byte whatever it is, skip blanks, take the token
    uint32_t p = hs + 1u;
    while (p < nl0 && is_ws(f2_byte(t, p))) ++p;
    r.name_off = p < nl0 ? p : nl0;
    r.nlen = p < nl0 ? f6_find_ws(t, p, nl0) - p : 0u;
    r.seq_off = nl0 + 1u;
    r.qual_off = nl2 + 1u;
    r.seq_len = f6_rstrip(t, r.seq_off, nl1) - r.seq_off;
    r.valid = r.seq_len == f6_rstrip(t, r.qual_off, nl3) - r.qual_off;
    return r;
}
// f2_digest with the bytes taken straight from the staged window when the whole record lies inside it (always,
// except for records longer than the 2 KiB halo): no per-byte window check
template <bool INWIN>
__device__ __forceinline__ void f6_digest_t(const BatchDev &b, const StreamDev &st, int s, const F2Tile &t, const F2Rec &r, uint32_t v)
{
    auto byte_at = [&](uint32_t pos) -> uint32_t { return INWIN ? lds8(t.sdata + (pos - t.tbase)) : f2_byte(t, pos); };
    uint32_t ia, in, ma, mn;
    f2_clamp(b.span[s - 2], r.seq_len, &ia, &in);          // QUADE_SPAN_IDX1 / IDX2
    f2_clamp(b.span[s], r.seq_len, &ma, &mn);              // QUADE_SPAN_MOL1 / MOL2
    u64 key = 0;
    uint32_t nm = 0, bad = in > 32u ? 1u : 0u, minq = 255u;
    uint8_t *slot = st.sfx + (size_t)v * st.sfx_slot;
    uint32_t word = 0;
    for (uint32_t i = 0; i < in; ++i) {
        uint32_t c = byte_at(r.seq_off + ia + i);
        const uint32_t q = byte_at(r.qual_off + ia + i);
        minq = q < minq ? q : minq;                         // phred = q - 33 (FastqSeq.py:44), compared later
        word |= c << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == in) {
            *reinterpret_cast<uint32_t *>(slot + (i & ~3u)) = word;
            word = 0;
        }
        // index.seq.upper() (Sample.py:65) and the base class in 4 instructions: (c >> 1) & 3 maps A,C,T,G (either
        // case) to 0,1,2,3; c is one of them iff c with bit 5 cleared equals the letter of its code (clearing bit 5
        // turns only 'a'..'z' into 'A'..'Z' among the bytes that can end up equal to a capital letter)
        const uint32_t code = (c >> 1) & 3u, up = c & 0xdfu;
        const bool acgt = up == ((0x47544341u >> (8u * code)) & 0xffu);
        const bool isn = up == 'N';
        if (!acgt && !isn) bad = 1u;
        if (i < 32u) {
            if (acgt) key |= (u64)code << (2u * i);
            else if (isn) nm |= 1u << i;
        }
    }
    uint8_t *mslot = slot + ((st.wi + 3u) & ~3u);
    for (uint32_t i = 0; i < mn; ++i) {
        word |= byte_at(r.seq_off + ma + i) << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == mn) {
            *reinterpret_cast<uint32_t *>(mslot + (i & ~3u)) = word;
            word = 0;
        }
    }
    st.dig[v] = make_uint4((uint32_t)key, (uint32_t)(key >> 32), nm, in | mn << 8 | minq << 16 | bad << 24);
}
__device__ __forceinline__ void f6_digest(const BatchDev &b, const StreamDev &st, int s, const F2Tile &t, const F2Rec &r, uint32_t v)
{
    if (r.seq_off >= t.tbase && r.qual_off + r.seq_len - t.tbase <= t.win) f6_digest_t<true>(b, st, s, t, r, v);
    else f6_digest_t<false>(b, st, s, t, r, v);
}

template <bool REDO>
__global__ void __launch_bounds__(F6T, F6_MINB) k_frame6(const __grid_constant__ BatchDev b, uint32_t total_tiles)
{
    extern __shared__ __align__(128) uint8_t f6_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f6_smem);
    const u64 *s_bm = reinterpret_cast<const u64 *>(f6_smem + F6_OFF_BM);
    uint16_t *s_pos = reinterpret_cast<uint16_t *>(f6_smem + F6_OFF_POS);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f6_smem + F6_OFF_MISC);
    const uint32_t s_bar = sm0 + F6_OFF_MISC + 64, a_pos = sm0 + F6_OFF_POS;
    BatchState *S = b.state;
    const uint32_t n_todo = REDO ? min(S->n_redo, b.redo_cap) : total_tiles;

    // thread 0, early in a tile: the NEXT ticket (tiles are independent: nothing orders the tickets), its geometry,
    // and an L2 prefetch of its window, so that the TMA copy issued once this tile is done hits L2
    auto ticket_next = [&]() {
        const uint32_t t = atomicAdd(REDO ? &S->ticket_redo : &S->ticket_frame, 1u);
        s_misc[13] = t;
        if (t >= n_todo) return;
        int s = 0;
        uint32_t lt;
        if (REDO) {
            const uint32_t packed = b.redo_list[t];
            s = (int)(packed >> 28);
            lt = packed & 0x0fffffffu;
        } else {
#pragma unroll
            for (int i = 1; i < kMaxStreams; ++i) s += (i < b.n_streams && t >= b.frame_tile_start[i]) ? 1 : 0;
            lt = t - b.frame_tile_start[s];
        }
        s_misc[14] = (uint32_t)s;
        s_misc[15] = lt;
        const uint32_t tbase = lt * (uint32_t)F2_TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)F2_WIN, n16 - tbase) : 0u;
        if (bytes)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b.st[s].data + tbase), "r"(bytes) : "memory");
    };
    // thread 0, when the window is free: make the prefetched ticket current and copy its window
    auto issue_next = [&]() {
        const uint32_t t = s_misc[13];
        s_misc[10] = t;
        if (t >= n_todo) return;
        const int s = (int)s_misc[14];
        const uint32_t lt = s_misc[15];
        s_misc[11] = (uint32_t)s;
        s_misc[12] = lt;
        const uint32_t tbase = lt * (uint32_t)F2_TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)F2_WIN, n16 - tbase) : 0u;
        if (bytes) {
            fence_proxy_async();
            mbar_expect_tx(s_bar, bytes);
            bulk_g2s(sm0, b.st[s].data + tbase, bytes, s_bar);
        }
    };
    if (tid == 0) {
        mbar_init(s_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        ticket_next();
        issue_next();
    }
    __syncthreads();
    uint32_t phase = 0;

    for (;;) {
        const uint32_t tile = s_misc[10];
        if (tile >= n_todo) break;
        const int s = (int)s_misc[11];
        const uint32_t lt = s_misc[12];
        const StreamDev &st = b.st[s];
        F2Tile T;
        T.data = st.data;
        T.sdata = sm0;
        T.bm = s_bm;
        T.bc = nullptr;                                         // no candidate bitmap here: f6_parse
        T.tbase = lt * (uint32_t)F2_TILE;
        T.cap = (uint32_t)F2_WIN;
        T.n = st.n;
        T.win = T.n > T.tbase ? min((uint32_t)F2_WIN, T.n - T.tbase) : 0u;
        T.last = ((b.last >> s) & 1) != 0;
        if (((T.n + 15u) & ~15u) > T.tbase) {                   // a copy was issued for this tile
            while (!mbar_try_wait(s_bar, phase)) {}
            phase ^= 1u;
        }

        // 1. newline / candidate masks: thread t takes the 16-byte chunks t, t+256, ... (conflict-free 128-bit
        //    loads); the 16-bit masks stored per chunk ARE the 64-bit bitmap words
        {
            const bool edge = T.win < (uint32_t)F2_WIN;         // the window reaches the end of the buffer
            uint32_t vbit = F2_NONE;                            // window offset of a virtual last line end
            if (edge && T.last && T.n > 0u && T.n - T.tbase < (uint32_t)F2_WIN) vbit = T.n - T.tbase;
            if (!edge) {                                        // the usual tile: nothing to clip
#pragma unroll
                for (uint32_t i = 0; i < 5; ++i) {
                    const uint32_t c = (uint32_t)tid + (uint32_t)F6T * i;
                    if (i == 4 && tid >= F2_HALO / 16) break;   // halo: 128 chunks, warps 0-3
                    const uint4 q = lds128(sm0 + 16u * c);
                    sts16(sm0 + F6_OFF_BM + 2u * c,
                          f2_gather16(f2_nl_flags(q.x), f2_nl_flags(q.y), f2_nl_flags(q.z), f2_nl_flags(q.w)));
                }
            } else {
#pragma unroll 1
                for (uint32_t i = 0; i < 5; ++i) {
                    const uint32_t c = (uint32_t)tid + (uint32_t)F6T * i, off = 16u * c;
                    if (i == 4 && tid >= F2_HALO / 16) break;
                    const uint4 q = lds128(sm0 + off);
                    uint32_t m = f2_gather16(f2_nl_flags(q.x), f2_nl_flags(q.y), f2_nl_flags(q.z), f2_nl_flags(q.w));
                    const uint32_t keep = off >= T.win ? 0u : (T.win - off >= 16u ? 0xffffu : (1u << (T.win - off)) - 1u);
                    m &= keep;
                    if ((vbit >> 4) == c && f2_byte(T, T.n - 1u) != '\n') m |= 1u << (vbit & 15u);   // no final '\n'
                    sts16(sm0 + F6_OFF_BM + 2u * c, m);
                }
            }
        }
        __syncthreads();                                        // B1: bitmaps complete

        // 2. line ends per thread (64 bytes = bitmap word tid), warp scans; the halo's 32 words by warp 1
        const u64 mine = s_bm[tid];
        const uint32_t cnt = (uint32_t)__popcll(mine);
        uint32_t inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) s_misc[warp] = inc;
        u64 hm = 0;
        uint32_t hinc = 0, hcnt = 0;
        if (warp == 1) {
            hm = s_bm[F6T + lane];
            hcnt = (uint32_t)__popcll(hm);
            hinc = hcnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, hinc, o);
                if (lane >= o) hinc += x;
            }
            if (lane == 31) s_misc[8] = hinc;
        }
        __syncthreads();                                        // B2: warp counts
        const uint4 wa = *reinterpret_cast<const uint4 *>(s_misc), wb = *reinterpret_cast<const uint4 *>(s_misc + 4);
        const uint32_t c1 = wa.x, c2 = c1 + wa.y, c3 = c2 + wa.z, c4 = c3 + wa.w, c5 = c4 + wb.x, c6 = c5 + wb.y,
                       c7 = c6 + wb.z, total = c7 + wb.w;
        uint32_t before = 0u;
        before = warp == 1 ? c1 : before;
        before = warp == 2 ? c2 : before;
        before = warp == 3 ? c3 : before;
        before = warp == 4 ? c4 : before;
        before = warp == 5 ? c5 : before;
        before = warp == 6 ? c6 : before;
        before = warp == 7 ? c7 : before;
        const uint32_t npos = min(total + s_misc[8], (uint32_t)F6_PCAP);
        const uint32_t first = lt == 0u ? 1u : 0u;              // slot 0 of a file's first tile starts at byte 0
        if (tid == 0) {
            // scratch room for the tile's records, in tile order of arrival: an upper bound over the four possible
            // record phases (at most one slot too many), so that the atomic flies during the list build and parse
            if (!REDO) {
                const uint32_t nres = first + ((total + 3u) >> 2);
                uint32_t sb = atomicAdd(&S->scratch_cursor[s], nres);
                if ((u64)sb + nres > (u64)st.cap_rec || total + s_misc[8] > (uint32_t)F6_PCAP) sb = F2_NONE;   // -> redo
                s_misc[9] = sb;
            }
            if (!REDO) atomicAdd(&S->chunk_lines[s][lt / (uint32_t)kTileChunk], total);   // for k_tilescan
            ticket_next();
        }

        // 3. dense list of the window's line ends (tile, then halo: list order = byte order)
        {
            uint32_t i = before + inc - cnt;
            uint32_t lo = (uint32_t)mine, hi = (uint32_t)(mine >> 32);
            const uint32_t o0 = (uint32_t)tid * 64u;
            // both halves of the word advance in the same iteration: the warp loops max(popc(lo), popc(hi)) times
            // over its lanes instead of max popc(lo) + max popc(hi)
            uint32_t ih = i + (uint32_t)__popc(lo);
            while (lo | hi) {
                if (lo) {
                    const uint32_t bit = (uint32_t)__ffs((int)lo) - 1u;
                    lo &= lo - 1u;
                    if (i < (uint32_t)F6_PCAP) s_pos[i] = (uint16_t)(o0 + bit);
                    ++i;
                }
                if (hi) {
                    const uint32_t bit = (uint32_t)__ffs((int)hi) - 1u;
                    hi &= hi - 1u;
                    if (ih < (uint32_t)F6_PCAP) s_pos[ih] = (uint16_t)(o0 + 32u + bit);
                    ++ih;
                }
            }
            if (warp == 1) {
                i = total + hinc - hcnt;
                const uint32_t h0 = (uint32_t)(F6T + lane) * 64u;
                while (hm) {
                    const uint32_t bit = (uint32_t)__ffsll((long long)hm) - 1u;
                    hm &= hm - 1ull;
                    if (i < (uint32_t)F6_PCAP) s_pos[i] = (uint16_t)(h0 + bit);
                    ++i;
                }
            }
        }
        __syncthreads();                                        // B3: list complete

        // 4. which list entries close a record: the true k0 (redo) or the guess from the text.  Only the warps that
        //    can own a record of this tile take part (insert reads: 47 records per tile = 2 of 8 warps).
        const bool warp_has = (uint32_t)warp * 32u < first + ((total + 3u) >> 2);
        uint32_t t0 = 3u, k0 = 0u, nslot = 0u, rbase = 0u;      // t0: list index of the first record-closing line
        if (warp_has || warp == 0) {
            if (REDO) {
                k0 = st.tile_tab[lt].x;
                t0 = (3u - k0) & 3u;
            } else if (lt != 0u) {
                bool ok = false;
                if (lane < 4 && (uint32_t)lane + 2u < npos)
                    ok = lds8(sm0 + lds16(a_pos + 2u * lane) + 1u) == '@' && lds8(sm0 + lds16(a_pos + 2u * (lane + 2)) + 1u) == '+';
                const unsigned g = __ballot_sync(0xffffffffu, ok);
                t0 = g ? (uint32_t)__ffs((int)g) - 1u : 0u;
            }
            nslot = first + (total > t0 ? (total - t0 + 3u) >> 2 : 0u);
            rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;             // REDO: record index of slot 0
            if (!REDO && tid == 0) st.tile_tab[lt] = make_uint4(total, t0, nslot, s_misc[9]);
        }

        // record of slot i (FastqReader.py:37-44); false: only the redo pass can find it (line-end list overflow)
        auto parse_slot = [&](uint32_t slot, F2Rec &R) -> bool {
            if (first && slot == 0u) { R = f6_parse(T, 0u, s_pos, 0u, npos); return true; }
            const uint32_t ti = t0 + 4u * (slot - first);
            if (ti < npos) { R = f6_parse(T, T.tbase + s_pos[ti] + 1u, s_pos, ti + 1u, npos); return true; }
            R.complete = R.valid = false;
            R.name_off = R.nlen = R.seq_off = R.qual_off = R.seq_len = R.end = 0u;
            if (!REDO) return false;
            const uint32_t want = (rbase + slot) * 4u - 1u;     // more than F6_PCAP lines in 18 KiB: bitmap search
            uint32_t k = k0, ppos = T.tbase, e = F2_NONE;
            while (k <= want) {
                e = f2_next_nl(T, ppos);
                if (e == F2_NONE) break;
                ppos = e + 1u;
                ++k;
            }
            if (e != F2_NONE) R = f6_parse(T, e + 1u, s_pos, 0u, 0u);
            return true;
        };
        // redo pass: outputs at the raw record index, as k_frame2.  First pass: the same fields at scratch position
        // v; an index read with a blank header carries F6_BLANK in its digest (insert reads: name length 0).
        auto store = [&](uint32_t v, const F2Rec &R) {
            if (!R.complete) return;
            const uint32_t inv = R.valid ? 0u : 0x80000000u;
            if (REDO) {
                if (v >= st.cap_rec) { S->err_capacity = 1u; return; }
                if (R.nlen == 0u) atomicMax(&S->first_blank[s], ~v);           // keeps the LOWEST blank record index
                if (inv) atomicAdd(&S->n_invalid[s], 1u);
                st.vend[v] = R.end;
                if (s < 2) {
                    st.rec[v] = make_uint4(R.name_off, R.seq_off, R.qual_off, R.seq_len | inv);
                    st.vnlen[v] = R.nlen;
                } else if (inv) {
                    st.dig[v] = make_uint4(0u, 0u, 0u, inv);
                } else {
                    f6_digest(b, st, s, T, R, v);
                }
            } else {
                st.vend_t[v] = R.end;
                if (s < 2) {
                    st.rec_t[v] = make_uint4(R.name_off, R.seq_off, R.qual_off, R.seq_len | inv);
                    st.vnlen_t[v] = R.nlen;
                } else if (inv) {
                    st.dig_t[v] = make_uint4(0u, 0u, 0u, inv | (R.nlen == 0u ? F6_BLANK : 0u));
                } else {
                    StreamDev stt = st;
                    stt.dig = st.dig_t;
                    stt.sfx = st.sfx_t;
                    f6_digest(b, stt, s, T, R, v);
                    if (R.nlen == 0u) st.dig_t[v].w |= F6_BLANK;
                }
            }
        };

        // 5. one thread per record; the first one is parsed while the scratch reservation is in flight
        F2Rec R;
        bool have = false;
        if ((uint32_t)tid < nslot) have = parse_slot((uint32_t)tid, R);
        __syncthreads();                                        // B4: scratch base known
        const uint32_t base = REDO ? rbase : s_misc[9];
        if (base != F2_NONE) {
            if (have) store(base + (uint32_t)tid, R);
            for (uint32_t slot = (uint32_t)tid + F6T; slot < nslot; slot += F6T)
                if (parse_slot(slot, R)) store(base + slot, R);
        }
        __syncthreads();                                        // B5: window, bitmaps and list are free
        if (tid == 0) issue_next();
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// k_tilescan: one CTA per chunk of kTileChunk tiles of one stream, one tile per thread.  The line ends of the
// chunks before it were summed by k_frame6 (chunk_lines), so every CTA is independent.
// tile_tab[t] = {line ends, guessed t0, records, scratch base} becomes {k0, ok, records, scratch base}; n_lines;
// tiles to parse again are appended to the redo list.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTileChunk) k_tilescan(const __grid_constant__ BatchDev b)
{
    __shared__ uint32_t s_warp[33];
    __shared__ uint32_t s_base;
    const int s = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const StreamDev &st = b.st[s];
    BatchState *S = b.state;
    const uint32_t n_t = st.n ? st.n / (uint32_t)F2_TILE + 1u : 0u, chunk = blockIdx.x;
    if (chunk * (uint32_t)kTileChunk >= n_t) return;
    if (tid == 0) s_base = 0u;
    __syncthreads();
    {                                                           // line ends in the chunks before this one
        uint32_t v = (uint32_t)tid < chunk && tid < kMaxChunks ? S->chunk_lines[s][tid] : 0u;
#pragma unroll
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0 && v) atomicAdd(&s_base, v);
    }
    const uint32_t t = chunk * (uint32_t)kTileChunk + (uint32_t)tid;
    const uint4 e = t < n_t ? st.tile_tab[t] : make_uint4(0u, 0u, 0u, 0u);
    const uint32_t total = e.x;
    uint32_t inc = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += x;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        const uint32_t w = s_warp[lane];
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += x;
        }
        s_warp[lane] = wi - w;
    }
    __syncthreads();
    if (t >= n_t) return;
    const uint32_t k0 = s_base + s_warp[warp] + inc - total, first = t == 0u ? 1u : 0u;
    const uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + first;
    const bool ok = nslot == 0u || (e.y == ((3u - k0) & 3u) && e.w != F2_NONE);
    st.tile_tab[t] = make_uint4(k0, ok ? 1u : 0u, nslot, e.w);
    if (!ok) {
        const uint32_t idx = atomicAdd(&S->n_redo, 1u);
        if (idx < b.redo_cap) b.redo_list[idx] = (uint32_t)s << 28 | t;
    }
    if (t == n_t - 1u) S->n_lines[s] = k0 + total;
}

// ------------------------------------------------------------------------------------------------------------
// k_densify: one warp per tile, scratch -> dense per-record arrays at the true record index (the pairing ordinal
// of a file without invalid records).  Counts the records with len(seq) != len(qual), finds the first blank header.
// ------------------------------------------------------------------------------------------------------------
#ifndef F6_DENSIFY_MINB
#define F6_DENSIFY_MINB 3   // 80 registers, no spills: 2 / 3 / 4 / 6 / 8 CTAs per SM measured 0.179 / 0.142 / 0.148 / 0.155 / 0.188 ms for the stage
#endif
__global__ void __launch_bounds__(256, F6_DENSIFY_MINB) k_densify(const __grid_constant__ BatchDev b, uint32_t total_tiles)
{
    const int lane = threadIdx.x & 31;
    BatchState *S = b.state;
    const uint32_t nwarps = gridDim.x * (blockDim.x >> 5);
    auto locate = [&](uint32_t tile, int &s, uint32_t &lt) -> uint4 {
        s = 0;
#pragma unroll
        for (int i = 1; i < kMaxStreams; ++i) s += (i < b.n_streams && tile >= b.frame_tile_start[i]) ? 1 : 0;
        lt = tile - b.frame_tile_start[s];
        return b.st[s].tile_tab[lt];
    };
    uint32_t tile = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (tile >= total_tiles) return;
    int s, s_next = 0;
    uint32_t lt, lt_next = 0u;
    uint4 e = locate(tile, s, lt), e_next = make_uint4(0u, 0u, 0u, 0u);
    for (;; tile += nwarps, e = e_next, s = s_next, lt = lt_next) {
        if (tile >= total_tiles) break;
        if (tile + nwarps < total_tiles) e_next = locate(tile + nwarps, s_next, lt_next);   // a tile ahead
        const StreamDev &st = b.st[s];
        if (!(e.y & 1u) || e.z == 0u) continue;
        const uint32_t rbase = lt == 0u ? 0u : (e.x >> 2) + 1u, n_rec = S->n_lines[s] >> 2, sv = st.sfx_slot >> 4;
        // Two records per lane and pass; every load of a pass is issued before its first store (the arrays may
        // alias as far as the compiler knows, and a load -> store -> load chain costs one DRAM round trip per link).
        for (uint32_t i0 = lane; i0 < e.z; i0 += 64) {
            uint32_t j[2], v[2];
            bool act[2];
            uint32_t ve[2];
            uint4 r[2], x0[2], x1[2];
            uint32_t nl[2];
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                const uint32_t i = i0 + 32u * u;
                j[u] = rbase + i;
                v[u] = e.w + i;
                act[u] = i < e.z && j[u] < n_rec;               // trailing partial record (FastqReader.py:57-58)
                if (act[u] && j[u] >= st.cap_rec) { S->err_capacity = 1u; act[u] = false; }
                nl[u] = 0u;
                x0[u] = x1[u] = make_uint4(0u, 0u, 0u, 0u);
                if (act[u]) {
                    ve[u] = st.vend_t[v[u]];
                    if (s < 2) {
                        r[u] = st.rec_t[v[u]];
                        nl[u] = st.vnlen_t[v[u]];
                    } else {
                        r[u] = st.dig_t[v[u]];
                        const uint4 *src = reinterpret_cast<const uint4 *>(st.sfx_t + (size_t)v[u] * st.sfx_slot);
                        if (sv > 0u) x0[u] = src[0];
                        if (sv > 1u) x1[u] = src[1];
                    }
                }
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (!act[u]) continue;
                st.vend[j[u]] = ve[u];
                bool blank;
                const bool inv = (r[u].w >> 31) != 0u;
                if (s < 2) {
                    st.rec[j[u]] = r[u];
                    st.vnlen[j[u]] = nl[u];
                    blank = nl[u] == 0u;
                } else {
                    blank = (r[u].w & F6_BLANK) != 0u;
                    r[u].w &= ~F6_BLANK;
                    st.dig[j[u]] = r[u];
                    if (!inv) {
                        const uint4 *src = reinterpret_cast<const uint4 *>(st.sfx_t + (size_t)v[u] * st.sfx_slot);
                        uint4 *dst = reinterpret_cast<uint4 *>(st.sfx + (size_t)j[u] * st.sfx_slot);
                        if (sv > 0u) dst[0] = x0[u];
                        if (sv > 1u) dst[1] = x1[u];
                        for (uint32_t q = 2; q < sv; ++q) dst[q] = src[q];
                    }
                }
                if (blank) atomicMax(&S->first_blank[s], ~j[u]);    // keeps the LOWEST blank record index
                if (inv) atomicAdd(&S->n_invalid[s], 1u);
            }
        }
    }
}

} // namespace
} // namespace quade
