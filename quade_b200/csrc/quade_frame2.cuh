// quade_frame2.cuh -- K1' k_frame2: newline scan + record framing + validation + index digest in ONE pass
// (included by quade_kernels.cu)
//
// Replaces k_frame + k_validate and the raw-byte half of k_classify.  Per 16 KiB tile of one input file:
//   1. the tile and a 2 KiB halo are staged with one cp.async.bulk (TMA) copy into a two-stage shared-memory ring
//      (the copy of the next tile is in flight while this one is parsed);
//   2. every thread scans 64 bytes from shared memory: SWAR newline flags, gathered to a 64-bit mask with IDP.4A;
//      popcount, block scan and a decoupled look-back over the tiles give the global line index;
//   3. lines k = 3 (mod 4) end a record, so the byte after them starts the next one (FastqReader.py:37): the
//      record starts of the tile are written to a dense list;
//   4. one thread per record start: the 4 line ends (bitmap search), name token, rstripped sequence / quality
//      (FastqReader.py:41-44), valid <=> len(seq) == len(qual) (FastqSeq.py:55); for index reads the sample-index
//      and molecular spans are sliced with python clamping (Quade.py:217-218, FastqSeq.py:91-93) and digested:
//      2-bit packed upper-cased key, N mask, min quality byte, raw bytes for the name suffix;
//   5. second block scan + look-back: rank among the VALID records of the file = pairing ordinal
//      (FastqReader.py:48-50 skips a bad record in its own file only), outputs written at that ordinal.
#pragma once

namespace quade {
namespace {

constexpr int F2T = 256;                       // threads per CTA
constexpr int F2_TILE = kFrameTile;            // 16384 bytes owned by a tile
constexpr int F2_HALO = 2048;                  // bytes after the tile that are staged as well
constexpr int F2_WIN = F2_TILE + F2_HALO;
constexpr int F2_WORDS = F2_WIN / 64;          // 64-bit bitmap words (288)
constexpr int F2_PCAP = 4096;                  // line ends of a window kept as a dense list (else: bitmap search)
constexpr uint32_t F2_NONE = 0xffffffffu;

constexpr int F2_OFF_DATA = 0;                                   // [F2_STAGES][F2_WIN + 16]
constexpr int F2_STAGES = 1;                   // 1: ticket taken just in time, more CTAs per SM hide the copy latency
constexpr int F2_OFF_BM = F2_STAGES * (F2_WIN + 16);             // u64[F2_WORDS] newlines
constexpr int F2_OFF_BC = F2_OFF_BM + F2_WORDS * 8;              // u64[F2_WORDS] bytes <= 0x20 (blank candidates)
constexpr int F2_OFF_POS = F2_OFF_BC + F2_WORDS * 8;             // u16[F2_PCAP] window offsets of the line ends
constexpr int F2_OFF_MISC = F2_OFF_POS + F2_PCAP * 2;
constexpr int F2_SMEM = F2_OFF_MISC + 128;

struct F2Tile {
    const uint8_t *data;    // global bytes of the file
    uint32_t sdata;         // shared address of byte `tbase`
    const u64 *bm;          // newline bitmap, bit i = byte tbase + i (i < win)
    const u64 *bc;          // bitmap of the bytes <= 0x20: every blank is in there
    uint32_t tbase, n, win; // win: staged bytes that exist (<= cap)
    uint32_t cap;           // window capacity in bytes (multiple of 64)
    bool last;              // last batch of the file: if it ends without '\n', position n counts as a line end
};

__device__ __forceinline__ uint32_t f2_byte(const F2Tile &t, uint32_t pos)
{
    const uint32_t q = pos - t.tbase;
    return q < t.win ? lds8(t.sdata + q) : (uint32_t)t.data[pos];
}

// position of the first line end at or after global position `from` (the virtual one at n included), F2_NONE if
// the file has no further line
__device__ __forceinline__ uint32_t f2_next_nl(const F2Tile &t, uint32_t from)
{
    uint32_t q = from - t.tbase;
    if (q < t.cap) {
        uint32_t w = q >> 6;
        u64 m = t.bm[w] & (~0ull << (q & 63u));
        for (;;) {
            if (m) return t.tbase + w * 64u + (uint32_t)__ffsll((long long)m) - 1u;
            if (++w >= (t.cap >> 6)) break;
            m = t.bm[w];
        }
        from = t.tbase + t.cap;
    }
    for (u64 p = from; p < t.n; ++p)
        if (t.data[p] == '\n') return (uint32_t)p;
    return (t.last && t.n > 0u && from <= t.n && t.data[t.n - 1u] != '\n') ? t.n : F2_NONE;
}

// is the byte at global position pos <= 0x20 ?
__device__ __forceinline__ bool f2_cand(const F2Tile &t, uint32_t pos)
{
    const uint32_t q = pos - t.tbase;
    return q < t.win ? ((t.bc[q >> 6] >> (q & 63u)) & 1ull) != 0ull : t.data[pos] <= 0x20u;
}
// first position in [from, e) whose byte is <= 0x20, e if none
__device__ __forceinline__ uint32_t f2_next_cand(const F2Tile &t, uint32_t from, uint32_t e)
{
    uint32_t q = from - t.tbase;
    if (q < t.cap) {
        uint32_t w = q >> 6;
        u64 m = t.bc[w] & (~0ull << (q & 63u));
        for (;;) {
            if (m) {
                const uint32_t p = t.tbase + w * 64u + (uint32_t)__ffsll((long long)m) - 1u;
                return p < e ? p : e;
            }
            if (++w >= (t.cap >> 6) || t.tbase + w * 64u >= e) break;
            m = t.bc[w];
        }
        from = t.tbase + t.cap;
    }
    for (u64 p = from; p < e; ++p)
        if (t.data[p] <= 0x20u) return (uint32_t)p;
    return e;
}

// end of line.rstrip() for the line [s, e) whose terminator sits at e (Python-2 str whitespace)
__device__ __forceinline__ uint32_t f2_rstrip(const F2Tile &t, uint32_t s, uint32_t e)
{
    if (e > s && f2_cand(t, e - 1u))
        while (e > s && is_ws(f2_byte(t, e - 1u))) --e;
    return e;
}

// first position in [p, e) holding whitespace, e if none: hop over the candidate bytes
__device__ __forceinline__ uint32_t f2_find_ws(const F2Tile &t, uint32_t p, uint32_t e)
{
    for (;;) {
        const uint32_t c = f2_next_cand(t, p, e);
        if (c >= e) return e;
        if (is_ws(f2_byte(t, c))) return c;
        p = c + 1u;                                           // a control character inside the token
    }
}

struct F2Rec {
    uint32_t name_off, nlen, seq_off, qual_off, seq_len, end;   // end: position after the record's last line
    bool complete, valid;
};

// one record starting at global position hs (FastqReader.py:37-44, FastqSeq.py:55).  pos[li .. li+3] are the
// window offsets of its four line ends as far as the list (npos entries) reaches; the rest is searched.
__device__ __forceinline__ F2Rec f2_parse(const F2Tile &t, uint32_t hs, const uint16_t *pos, uint32_t li, uint32_t npos)
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
    if (p < nl0 && f2_cand(t, p))
        while (p < nl0 && is_ws(f2_byte(t, p))) ++p;
    r.name_off = p < nl0 ? p : nl0;
    r.nlen = p < nl0 ? f2_find_ws(t, p, nl0) - p : 0u;
    r.seq_off = nl0 + 1u;
    r.qual_off = nl2 + 1u;
    r.seq_len = f2_rstrip(t, r.seq_off, nl1) - r.seq_off;
    r.valid = r.seq_len == f2_rstrip(t, r.qual_off, nl3) - r.qual_off;
    return r;
}

// newline flags of one word: 0x80 in every byte equal to '\n' (exact for all byte values)
__device__ __forceinline__ uint32_t f2_nl_flags(uint32_t w)
{
    // (w ^ 0x0a..) & 0x7f.. as ONE lop3 (ptxas keeps 0x7f7f7f7f in a register; written as C it is an AND and a XOR
    // with an immediate each); bit 7 of every byte is the same in w and w ^ 0x0a0a0a0a.  0.638 -> 0.624 ms for k_frame6
    uint32_t t;
    asm("lop3.b32 %0, %1, 0x0a0a0a0a, 0x7f7f7f7f, 0x28;" : "=r"(t) : "r"(w));
    return ~((t + 0x7f7f7f7fu) | w) & 0x80808080u;
}
// 0x80 in every byte <= 0x20
__device__ __forceinline__ uint32_t f2_c_flags(uint32_t w) { return ~(((w & 0x7f7f7f7fu) + 0x5f5f5f5fu) | w) & 0x80808080u; }
// 16 bytes -> 16 mask bits: IDP.4A weighs the 0x80 flags of two words into 128 x (8-bit mask)
__device__ __forceinline__ uint32_t f2_gather16(uint32_t f0, uint32_t f1, uint32_t f2, uint32_t f3)
{
    const uint32_t lo = __dp4a(f1, 0x80402010u, __dp4a(f0, 0x08040201u, 0u));
    const uint32_t hi = __dp4a(f3, 0x80402010u, __dp4a(f2, 0x08040201u, 0u));
    return (lo >> 7) | (hi << 1);
}
// newline / candidate masks of the 64 bytes at shared address a (16-byte aligned); the four vectors are read in
// a rotated order so that the 8 lanes of a quarter warp hit 8 different bank groups
__device__ __forceinline__ void f2_mask64(uint32_t a, int lane, u64 &nl, u64 &cd)
{
    const uint32_t rot = ((uint32_t)lane >> 1) & 3u;
    nl = cd = 0;
#pragma unroll
    for (uint32_t i = 0; i < 4; ++i) {
        const uint32_t v = (i + rot) & 3u;
        const uint4 q = lds128(a + 16u * v);
        nl |= (u64)f2_gather16(f2_nl_flags(q.x), f2_nl_flags(q.y), f2_nl_flags(q.z), f2_nl_flags(q.w)) << (16u * v);
        cd |= (u64)f2_gather16(f2_c_flags(q.x), f2_c_flags(q.y), f2_c_flags(q.z), f2_c_flags(q.w)) << (16u * v);
    }
}

// python slice [sp0:sp1] of a string of length L -> offset and length (FastqSeq.py:91-93)
__device__ __forceinline__ void f2_clamp(const int sp[2], uint32_t L, uint32_t *a, uint32_t *n)
{
    uint32_t s = (uint32_t)sp[0], e = (uint32_t)sp[1];
    s = s < L ? s : L;
    e = e < L ? e : L;
    *a = s;
    *n = e > s ? e - s : 0u;
}

// index read: digest of the sample-index span + raw bytes of both spans for the name suffix
__device__ __forceinline__ void f2_digest(const BatchDev &b, const StreamDev &st, int s, const F2Tile &t, const F2Rec &r,
                                          uint32_t v)
{
    uint32_t ia, in, ma, mn;
    f2_clamp(b.span[s - 2], r.seq_len, &ia, &in);          // QUADE_SPAN_IDX1 / IDX2
    f2_clamp(b.span[s], r.seq_len, &ma, &mn);              // QUADE_SPAN_MOL1 / MOL2
    u64 key = 0;
    uint32_t nm = 0, bad = in > 32u ? 1u : 0u, minq = 255u;
    uint8_t *slot = st.sfx + (size_t)v * st.sfx_slot;
    uint32_t word = 0;
    for (uint32_t i = 0; i < in; ++i) {
        uint32_t c = f2_byte(t, r.seq_off + ia + i);
        const uint32_t q = f2_byte(t, r.qual_off + ia + i);
        minq = q < minq ? q : minq;                         // phred = q - 33 (FastqSeq.py:44), compared later
        word |= c << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == in) {
            *reinterpret_cast<uint32_t *>(slot + (i & ~3u)) = word;
            word = 0;
        }
        if (c - 'a' < 26u) c -= 32u;                        // index.seq.upper() (Sample.py:65)
        const bool acgt = c == 'A' || c == 'C' || c == 'G' || c == 'T';
        if (!acgt && c != 'N') bad = 1u;
        if (i < 32u) {
            if (acgt) key |= (u64)((c >> 1) & 3u) << (2u * i);
            else if (c == 'N') nm |= 1u << i;
        }
    }
    uint8_t *mslot = slot + ((st.wi + 3u) & ~3u);
    for (uint32_t i = 0; i < mn; ++i) {
        word |= f2_byte(t, r.seq_off + ma + i) << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == mn) {
            *reinterpret_cast<uint32_t *>(mslot + (i & ~3u)) = word;
            word = 0;
        }
    }
    st.dig[v] = make_uint4((uint32_t)key, (uint32_t)(key >> 32), nm, in | mn << 8 | minq << 16 | bad << 24);
}

__global__ void __launch_bounds__(F2T, 6) k_frame2(const __grid_constant__ BatchDev b, uint32_t total_tiles)
{
    extern __shared__ __align__(128) uint8_t f2_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f2_smem);
    u64 *s_bm = reinterpret_cast<u64 *>(f2_smem + F2_OFF_BM);
    u64 *s_bc = reinterpret_cast<u64 *>(f2_smem + F2_OFF_BC);
    uint16_t *s_pos = reinterpret_cast<uint16_t *>(f2_smem + F2_OFF_POS);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f2_smem + F2_OFF_MISC);   // [0..8] scan, [9] base, [10..11] tickets
    const uint32_t s_bar = sm0 + F2_OFF_MISC + 64;                            // 2 mbarriers
    BatchState *S = b.state;

    auto tile_geom = [&](uint32_t tile, int &s, uint32_t &lt) {
        s = 0;
        while (s + 1 < b.n_streams && tile >= b.frame_tile_start[s + 1]) ++s;
        lt = tile - b.frame_tile_start[s];
    };
    auto issue = [&](uint32_t tile, uint32_t stage) {       // thread 0: TMA copy of the tile's window
        int s;
        uint32_t lt;
        tile_geom(tile, s, lt);
        const uint32_t tbase = lt * (uint32_t)F2_TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)F2_WIN, n16 - tbase) : 0u;
        if (bytes) {
            fence_proxy_async();
            mbar_expect_tx(s_bar + 8u * stage, bytes);
            bulk_g2s(sm0 + F2_OFF_DATA + stage * (F2_WIN + 16), b.st[s].data + tbase, bytes, s_bar + 8u * stage);
        }
    };

    if (tid == 0) {
        mbar_init(s_bar, 1);
        mbar_init(s_bar + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        const uint32_t t0 = atomicAdd(&S->ticket_frame, 1u);
        s_misc[10] = t0;
        if (t0 < total_tiles) issue(t0, 0);
    }
    __syncthreads();
    uint32_t tile = s_misc[10], phase_bits = 0;

    for (uint32_t it = 0; tile < total_tiles; ++it) {
        const uint32_t stage = F2_STAGES == 2 ? (it & 1u) : 0u;
        if (F2_STAGES == 2 && tid == 0) {                    // next ticket, its copy flies while this tile is parsed
            const uint32_t tn = atomicAdd(&S->ticket_frame, 1u);
            s_misc[10 + ((it + 1u) & 1u)] = tn;
            if (tn < total_tiles) issue(tn, stage ^ 1u);
        }
        int s;
        uint32_t lt;
        tile_geom(tile, s, lt);
        const StreamDev &st = b.st[s];
        const uint32_t n = st.n, tbase = lt * (uint32_t)F2_TILE, n16 = (n + 15u) & ~15u;
        const uint32_t staged = n16 > tbase ? min((uint32_t)F2_WIN, n16 - tbase) : 0u;
        const uint32_t sdata = sm0 + F2_OFF_DATA + stage * (F2_WIN + 16);
        if (staged) {
            while (!mbar_try_wait(s_bar + 8u * stage, (phase_bits >> stage) & 1u)) {}
            phase_bits ^= 1u << stage;
        }
        F2Tile T;
        T.data = st.data;
        T.sdata = sdata;
        T.bm = s_bm;
        T.bc = s_bc;
        T.tbase = tbase;
        T.cap = (uint32_t)F2_WIN;
        T.n = n;
        T.win = n > tbase ? min((uint32_t)F2_WIN, n - tbase) : 0u;
        T.last = ((b.last >> s) & 1) != 0;

        // 1. newline / candidate bitmaps of the window: 64 bytes per thread (+ the halo by warp 0)
        u64 mine = 0, halo = 0;
        for (int c = tid; c < F2_WORDS; c += F2T) {
            const uint32_t c0 = (uint32_t)c * 64u;            // window offset of this chunk
            u64 m = 0, cd = 0;
            if (c0 < T.win) {
                f2_mask64(sdata + c0, lane, m, cd);
                if (c0 + 64u > T.win) {                       // bytes past the end of the file
                    m &= (1ull << (T.win - c0)) - 1ull;
                    cd &= (1ull << (T.win - c0)) - 1ull;
                }
            }
            if (T.last && n > 0u && n - tbase >= c0 && n - tbase < c0 + 64u && f2_byte(T, n - 1u) != '\n')
                m |= 1ull << (n - tbase - c0);                // the file's last line has no '\n'
            s_bm[c] = m;
            s_bc[c] = cd;
            if (c < F2T) mine = m; else halo = m;
        }
        // 2. line index: block scan + look-back over the tiles of this file
        uint32_t total;
        const uint32_t excl = block_scan256((uint32_t)__popcll(mine), s_misc, &total);
        uint32_t hexcl = 0, htotal = 0;                       // halo line ends (warp 0), listed after the tile's
        if (warp == 0) {
            const uint32_t hc = (uint32_t)__popcll(halo);
            uint32_t inc = hc;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += x;
            }
            hexcl = inc - hc;
            htotal = __shfl_sync(0xffffffffu, inc, 31);
            const uint32_t base = lookback(st.status_frame, (int)lt, total, lane);
            if (lane == 0) {
                s_misc[9] = base;
                s_misc[12] = htotal;
                if ((u64)tbase + F2_TILE > (u64)n) S->n_lines[s] = base + total;
            }
        }
        // 3. dense list of the window's line ends; the byte after every line k = 3 (mod 4) starts a record
        {
            uint32_t i = excl;
            u64 m = mine;
            while (m) {
                const uint32_t bit = (uint32_t)__ffsll((long long)m) - 1u;
                m &= m - 1ull;
                if (i < (uint32_t)F2_PCAP) s_pos[i] = (uint16_t)((uint32_t)tid * 64u + bit);
                ++i;
            }
            if (warp == 0) {
                i = total + hexcl;
                m = halo;
                while (m) {
                    const uint32_t bit = (uint32_t)__ffsll((long long)m) - 1u;
                    m &= m - 1ull;
                    if (i < (uint32_t)F2_PCAP) s_pos[i] = (uint16_t)((uint32_t)(F2T + lane) * 64u + bit);
                    ++i;
                }
            }
        }
        __syncthreads();
        const uint32_t k0 = s_misc[9];
        const uint32_t npos = min(total + s_misc[12], (uint32_t)F2_PCAP);
        const uint32_t rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;                // record index of slot 0
        const uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + (lt == 0u ? 1u : 0u);
        const uint32_t t0i = (3u - k0) & 3u;                  // list index of the first line k = 3 (mod 4)
        // record of slot i: starts after list entry ti = t0i + 4 (i - first), its line ends are entries ti+1 .. ti+4
        auto parse_slot = [&](uint32_t slot) -> F2Rec {
            uint32_t hs, li;
            if (lt == 0u && slot == 0u) { hs = 0u; li = 0u; }
            else {
                const uint32_t ti = t0i + 4u * (slot - (lt == 0u ? 1u : 0u));
                li = ti + 1u;
                hs = ti < npos ? tbase + s_pos[ti] + 1u : F2_NONE;
            }
            if (hs == F2_NONE) {                              // list overflow: find the terminator by bitmap search
                uint32_t k = k0, ppos = tbase;                // (pathological tiles only: > F2_PCAP lines in 18 KiB)
                const uint32_t want = (rbase + slot) * 4u - 1u;
                uint32_t e = F2_NONE;
                while (k <= want) {
                    e = f2_next_nl(T, ppos);
                    if (e == F2_NONE) break;
                    ppos = e + 1u;
                    ++k;
                }
                F2Rec z;
                z.complete = z.valid = false;
                z.name_off = z.nlen = z.seq_off = z.qual_off = z.seq_len = z.end = 0u;
                if (e == F2_NONE) return z;
                return f2_parse(T, e + 1u, s_pos, 0u, 0u);
            }
            return f2_parse(T, hs, s_pos, li, npos);
        };
        // 4. one thread per record start, outputs at the RAW record index j.  Records with len(seq) != len(qual)
        //    are flagged (bit 31) and counted; they are rare, and only when a file has some does k_fix_compact
        //    squeeze them out afterwards (pairing ordinal = rank among the valid records, FastqReader.py:48-50).
        for (uint32_t slot = tid; slot < nslot; slot += F2T) {
            const F2Rec R = parse_slot(slot);
            if (!R.complete) continue;
            const uint32_t j = rbase + slot;
            if (j >= st.cap_rec) { S->err_capacity = 1u; continue; }
            if (R.nlen == 0u) atomicMax(&S->first_blank[s], ~j);               // keeps the LOWEST blank record index
            const uint32_t inv = R.valid ? 0u : 0x80000000u;
            if (inv) atomicAdd(&S->n_invalid[s], 1u);
            st.vend[j] = R.end;
            if (s < 2) {
                st.rec[j] = make_uint4(R.name_off, R.seq_off, R.qual_off, R.seq_len | inv);
                st.vnlen[j] = R.nlen;
            } else if (inv) {
                st.dig[j] = make_uint4(0u, 0u, 0u, inv);
            } else {
                f2_digest(b, st, s, T, R, j);
            }
        }
        if (F2_STAGES == 2) {
            tile = s_misc[10 + ((it + 1u) & 1u)];
            __syncthreads();                                  // bitmap, list and the stage are reused
        } else {
            // A ticket held while another tile is parsed would make every successor spin in its look-back for a
            // whole tile time: take it just in time; the other CTAs of the SM cover the copy latency.
            __syncthreads();
            if (tid == 0) {
                const uint32_t tn = atomicAdd(&S->ticket_frame, 1u);
                s_misc[10] = tn;
                if (tn < total_tiles) issue(tn, 0u);
            }
            __syncthreads();
            tile = s_misc[10];
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// K2': rare path.  A file that holds records with len(seq) != len(qual) gets them squeezed out: stable
// compaction of its per-record arrays (tiles of 1024 records, decoupled look-back) into scratch arrays, then a
// copy back.  Files without such records (the normal case) return at once: raw record index = pairing ordinal.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool f2_is_invalid(const StreamDev &st, int s, uint32_t j)
{
    return ((s < 2 ? st.rec[j].w : st.dig[j].w) >> 31) != 0u;
}

__global__ void __launch_bounds__(256) k_fix_compact(const __grid_constant__ BatchDev b)
{
    __shared__ uint32_t s_warp[9];
    __shared__ uint32_t s_tile, s_base;
    const int s = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    BatchState *S = b.state;
    if (S->n_invalid[s] == 0u) return;
    const StreamDev &st = b.st[s];
    const uint32_t n_rec = min(S->n_lines[s] >> 2, st.cap_rec);
    const uint32_t n_tiles = (n_rec + kValidateTile - 1) / kValidateTile, sv = st.sfx_slot >> 4;
    for (;;) {
        if (tid == 0) s_tile = atomicAdd(&S->ticket_validate[s], 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) return;
        const uint32_t j0 = tile * kValidateTile + (uint32_t)tid * 4u;
        uint32_t vmask = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (j0 + i < n_rec && !f2_is_invalid(st, s, j0 + i)) vmask |= 1u << i;
        uint32_t total;
        const uint32_t excl = block_scan256((uint32_t)__popc(vmask), s_warp, &total);
        if (warp == 0) {
            const uint32_t base = lookback(st.status_validate, (int)tile, total, lane);
            if (lane == 0) s_base = base;
        }
        __syncthreads();
        uint32_t v = s_base + excl;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (!(vmask & (1u << i))) continue;
            const uint32_t j = j0 + i;
            st.vrec[v] = j;
            st.vend_t[v] = st.vend[j];
            if (s < 2) {
                st.rec_t[v] = st.rec[j];
                st.vnlen_t[v] = st.vnlen[j];
            } else {
                st.dig_t[v] = st.dig[j];
                const uint4 *src = reinterpret_cast<const uint4 *>(st.sfx + (size_t)j * st.sfx_slot);
                uint4 *dst = reinterpret_cast<uint4 *>(st.sfx_t + (size_t)v * st.sfx_slot);
                for (uint32_t q = 0; q < sv; ++q) dst[q] = src[q];
            }
            ++v;
        }
        __syncthreads();
    }
}

__global__ void __launch_bounds__(256) k_fix_copyback(const __grid_constant__ BatchDev b)
{
    const int s = blockIdx.y;
    const BatchState *S = b.state;
    if (S->n_invalid[s] == 0u) return;
    const StreamDev &st = b.st[s];
    const uint32_t n_rec = min(S->n_lines[s] >> 2, st.cap_rec), nv = n_rec - min(n_rec, S->n_invalid[s]), sv = st.sfx_slot >> 4;
    for (uint32_t v = blockIdx.x * 256u + threadIdx.x; v < nv; v += gridDim.x * 256u) {
        st.vend[v] = st.vend_t[v];
        if (s < 2) {
            st.rec[v] = st.rec_t[v];
            st.vnlen[v] = st.vnlen_t[v];
        } else {
            st.dig[v] = st.dig_t[v];
            const uint4 *src = reinterpret_cast<const uint4 *>(st.sfx_t + (size_t)v * st.sfx_slot);
            uint4 *dst = reinterpret_cast<uint4 *>(st.sfx + (size_t)v * st.sfx_slot);
            for (uint32_t q = 0; q < sv; ++q) dst[q] = src[q];
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// K2b': pairing on the outputs of k_frame2 (Quade.py:210-224): n_pairs = min over files of the valid count.
// ------------------------------------------------------------------------------------------------------------
__global__ void k_pairing2(const __grid_constant__ BatchDev b)
{
    if (threadIdx.x != 0) return;
    BatchState *S = b.state;
    uint32_t np = 0xffffffffu;
    for (int s = 0; s < b.n_streams; ++s) {
        const uint32_t n_rec = min(S->n_lines[s] >> 2, b.st[s].cap_rec);
        S->n_valid[s] = n_rec - min(n_rec, S->n_invalid[s]);
        np = min(np, S->n_valid[s]);
    }
    if (np > b.cap_pairs) {
        S->err_capacity = 1u;
        np = 0;
    }
    if (S->err_capacity) np = 0;
    int sx = 0;
    while (S->n_valid[sx] != np && sx + 1 < b.n_streams) ++sx;   // first file (in pull order) that runs dry
    for (int s = 0; s < b.n_streams && !S->err_capacity; ++s) {
        const StreamDev &st = b.st[s];
        const bool sq = S->n_invalid[s] != 0u;                  // squeezed: vrec[v] = raw index, else v itself
        auto vrec = [&](uint32_t v) { return sq ? st.vrec[v] : v; };
        if (np > 0) {
            S->consumed[s] = st.vend[np - 1];
            S->n_skipped[s] = vrec(np - 1) + 1u - np;
        }
        if (S->first_blank[s]) {
            const uint32_t jb = ~S->first_blank[s];
            bool parsed;
            if (s < sx) parsed = jb <= vrec(np);             // pulled once more in the failing iteration
            else if (s == sx) parsed = true;                     // parsed to the end looking for a valid record
            else parsed = np > 0 && jb <= vrec(np - 1);
            if (parsed) S->err_blank = 1u;
        }
    }
    uint32_t tp = kCompactTilePairs;                                            // k_compact3 tile size from the mean record size
    if (np > 0) {
        uint32_t avg = max(S->consumed[0], S->consumed[1]) / np + 1u;
        avg += avg / (uint32_t)QUADE_C3_MARGIN;
        tp = ((uint32_t)kCompactStageBytes - 48u) / avg;
        tp = tp > (uint32_t)kCompactTilePairs ? (uint32_t)kCompactTilePairs : (tp < 1u ? 1u : tp);
    }
    S->compact_tp = tp;
    S->n_pairs = np;
    S->n_blocks = (np + b.pairs_per_block - 1) / b.pairs_per_block;
}

// ------------------------------------------------------------------------------------------------------------
// K3': classify from the index digests (Sample.py:56-91).  One thread per pair: fuse the two digests into the
// key of index1 + index2, probe the perfect-hash table (full-key verify), min quality against the threshold,
// bin, the name suffix ':' index [':' molecular] '\n' (FastqWriter.py:61-66), output record lengths, per-block
// per-bin byte histogram, counters.
// ------------------------------------------------------------------------------------------------------------
constexpr int C2T = 256;

__global__ void __launch_bounds__(C2T) k_classify2(const __grid_constant__ BatchDev b, int table_in_smem)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t n_pairs = b.state->n_pairs, n_blocks = b.state->n_blocks;
    if (blockIdx.x >= n_blocks) return;
    const int tid = threadIdx.x, nb = b.n_bins;
    uint32_t *h_b1 = reinterpret_cast<uint32_t *>(smem), *h_b2 = h_b1 + nb, *h_cnt = h_b2 + nb;
    const uint32_t n_slots = b.table.n_slots_mask + 1u, n_buckets = b.table.n_buckets_mask + 1u;
    const uint4 *entries = b.table.entries;
    const uint16_t *disp = b.table.disp;
    for (int i = tid; i < nb; i += C2T) h_cnt[i] = 0u;
    if (table_in_smem) {
        uint4 *se = reinterpret_cast<uint4 *>(smem + (((size_t)3 * nb * 4 + 15) & ~(size_t)15));
        uint16_t *sd = reinterpret_cast<uint16_t *>(se + n_slots);
        for (uint32_t i = tid; i < n_slots; i += C2T) se[i] = entries[i];
        for (uint32_t i = tid; i < n_buckets; i += C2T) sd[i] = disp[i];
        entries = se;
        disp = sd;
    }
    const StreamDev &i1 = b.st[2], &i2 = b.st[3];
    const uint32_t mo1 = (i1.wi + 3u) & ~3u, mo2 = b.dual ? ((i2.wi + 3u) & ~3u) : 0u;
    const uint32_t minbyte = (uint32_t)b.min_qual + 33u;

    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        for (int i = tid; i < 2 * nb; i += C2T) h_b1[i] = 0u;
        __syncthreads();
        const uint32_t p0 = blk * b.pairs_per_block;
        uint32_t end = p0 + b.pairs_per_block;
        if (end > n_pairs) end = n_pairs;
        for (uint32_t p = p0 + tid; p < end; p += C2T) {
            const uint4 d1 = i1.dig[p];
            const uint4 d2 = b.dual ? i2.dig[p] : make_uint4(0, 0, 0, 255u << 16);
            const uint32_t in1 = d1.w & 0xffu, mn1 = (d1.w >> 8) & 0xffu, in2 = d2.w & 0xffu, mn2 = (d2.w >> 8) & 0xffu;
            const uint32_t klen = in1 + in2, mlen = mn1 + mn2;
            const uint32_t bad = (d1.w >> 24) | (d2.w >> 24) | (klen > 32u ? 1u : 0u);
            const uint32_t minq = min((d1.w >> 16) & 0xffu, (d2.w >> 16) & 0xffu);
            const uint64_t k1 = (uint64_t)d1.x | (uint64_t)d1.y << 32, k2 = (uint64_t)d2.x | (uint64_t)d2.y << 32;
            const uint64_t key = k1 | (in1 < 32u ? k2 << (2u * in1) : 0ull);
            const uint32_t nm = d1.z | (in1 < 32u ? d2.z << in1 : 0u);
            int sid = -1;
            if (!bad && klen > 0u) {
                const uint64_t h = key_hash(key, nm, klen, b.table.seed);
                const uint32_t dsp = disp[(uint32_t)h & b.table.n_buckets_mask];
                const uint4 e = entries[key_slot(h, dsp, b.table.n_slots_mask)];
                if (e.x == (uint32_t)key && e.y == (uint32_t)(key >> 32) && e.z == nm && (e.w & 0xffu) == klen && (e.w >> 8))
                    sid = (int)(e.w >> 8) - 1;
            }
            uint32_t bin, wr;
            if (sid >= 0) {
                if (minq >= minbyte) { bin = 1u + 2u * sid; wr = b.wr_pass; }   // min(index.qual) >= MIN_QUAL, Sample.py:70
                else                 { bin = 2u + 2u * sid; wr = b.wr_fail; }
            } else { bin = 0u; wr = b.wr_und; }

            // ':' index [':' molecular] '\n', raw case, assembled 4 bytes at a time
            const uint32_t suf_len = 1u + klen + (mlen ? 1u + mlen : 0u) + 1u;
            const uint8_t *s1 = i1.sfx + (size_t)p * i1.sfx_slot, *s2 = b.dual ? i2.sfx + (size_t)p * i2.sfx_slot : s1;
            uint32_t *suf = reinterpret_cast<uint32_t *>(b.suffix + (size_t)p * b.suf_cap);
            if (in1 <= 8u && in2 <= 8u && mn1 <= 8u && mn2 <= 8u) {
                // every piece fits a 64-bit word (the usual 6..10-base indexes): pieces are appended to a 64-bit
                // accumulator by shifts and flushed with 8-byte stores -- ~80 instructions instead of ~670 for the
                // byte-wise assembly below, and 4 independent 8-byte loads instead of ~30 dependent byte loads
                auto piece = [](const uint8_t *q, uint32_t n) -> u64 {          // n <= 8 bytes at a 4-byte aligned address
                    const uint32_t *w = reinterpret_cast<const uint32_t *>(q);
                    const u64 v = n ? ((u64)w[0] | (n > 4u ? (u64)w[1] << 32 : 0ull)) : 0ull;
                    return n >= 8u ? v : v & ((1ull << (8u * n)) - 1ull);
                };
                const u64 pa = piece(s1, in1), pb = b.dual ? piece(s2, in2) : 0ull;
                const u64 pc = piece(s1 + mo1, mn1), pd = b.dual ? piece(s2 + mo2, mn2) : 0ull;
                u64 *suf64 = reinterpret_cast<u64 *>(suf);
                u64 acc = ':';
                uint32_t fill = 1, wq = 0;                                       // bytes in acc, 8-byte words flushed
                auto append = [&](u64 v, uint32_t n) {
                    if (!n) return;
                    acc |= v << (8u * fill);                                     // fill < 8 here
                    if (fill + n >= 8u) {
                        suf64[wq++] = acc;
                        acc = fill ? v >> (8u * (8u - fill)) : 0ull;
                        fill = fill + n - 8u;
                    } else {
                        fill += n;
                    }
                };
                append(pa, in1);
                append(pb, in2);
                if (mlen) {
                    append((u64)':', 1u);
                    append(pc, mn1);
                    append(pd, mn2);
                }
                append((u64)'\n', 1u);
                if (fill) suf64[wq] = acc;
            } else
            for (uint32_t w0 = 0; w0 < suf_len; w0 += 4) {
                uint32_t word = 0;
#pragma unroll
                for (uint32_t j = 0; j < 4; ++j) {
                    uint32_t x = w0 + j, c = 0;
                    if (x < suf_len) {
                        if (x == 0u) c = ':';
                        else if (x == suf_len - 1u) c = '\n';
                        else {
                            x -= 1u;
                            if (x < in1) c = s1[x];
                            else if (x < klen) c = s2[x - in1];
                            else if (x == klen) c = ':';
                            else {
                                x -= klen + 1u;
                                c = x < mn1 ? s1[mo1 + x] : s2[mo2 + (x - mn1)];
                            }
                        }
                    }
                    word |= c << (8u * j);
                }
                suf[w0 >> 2] = word;
            }
            uint32_t l1 = 0, l2 = 0;
            if (wr) {
                l1 = b.st[0].vnlen[p] + suf_len + 2u * b.st[0].rec[p].w + 5u;
                l2 = b.st[1].vnlen[p] + suf_len + 2u * b.st[1].rec[p].w + 5u;
                atomicAdd(&h_b1[bin], l1);
                atomicAdd(&h_b2[bin], l2);
            }
            atomicAdd(&h_cnt[bin], 1u);                      // counters ignore the write flags (Sample.py:62-88)
            b.cls[p] = bin | suf_len << 16 | (wr ? 0x80000000u : 0u);
            b.len1[p] = l1;
            b.len2[p] = l2;
        }
        __syncthreads();
        uint32_t *hg = b.hist + (size_t)blk * 2u * nb;
        for (int i = tid; i < 2 * nb; i += C2T) hg[i] = h_b1[i];
        __syncthreads();
    }
    for (int i = tid; i < nb; i += C2T) {
        const uint32_t c = h_cnt[i];
        if (c) atomicAdd(reinterpret_cast<u64 *>(b.counters) + i, (u64)c);
    }
}

} // namespace
} // namespace quade
