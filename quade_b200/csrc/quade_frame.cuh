// quade_frame.cuh -- k_frame: newline scan + record framing + field extraction + index digest in ONE pass over the
// input, dependency-free across tiles (included by quade_kernels.cu).
//
// A tile is FT_TILE bytes of one input file (+ a 2 KiB halo), staged with one cp.async.bulk (TMA) copy; the next
// tile's ticket is taken early and its window prefetched into L2.  Per tile:
//   1. SWAR newline flags over conflict-free 128-bit shared loads, gathered with IDP.4A into 16-bit masks that ARE
//      the 64-bit bitmap words (one per thread);
//   2. popcount, warp scans, cross-warp prefix: number of line ends before every word;
//   3. dense list of the window's line ends (u16 window offsets);
//   4. the global line index k0 of a tile is needed for two things only.  (a) k0 mod 4, i.e. which line ends close
//      a record (FastqReader.py:37 takes 4 lines blindly): GUESSED from the text -- the first line end i in 0..3
//      whose next line starts with '@' and whose third next line starts with '+' (exact for well-formed FASTQ: a
//      quality line may start with '@', but the line two after it is then a sequence line) -- and verified by
//      k_tilescan.  (b) the record index the outputs are stored at: DEFERRED -- a tile reserves room in per-stream
//      scratch arrays with one atomic (tile order of arrival); k_tilescan writes the record -> position table.
//   5. one thread per record: name token, rstripped sequence / quality, len(seq) == len(qual)
//      (FastqReader.py:41-44, FastqSeq.py:55); index reads: python-clamped spans (Quade.py:217-218,
//      FastqSeq.py:91-93), 2-bit packed upper-cased key, N mask, quality filter bit, raw bytes for the name suffix.
// The FAST instantiation handles what real FASTQ looks like with straight-line SWAR code on the staged window and
// flags every tile it cannot be sure about (window clipped by the end of the buffer, a record that does not end
// inside the window, more line ends than the list holds); k_tilescan adds the tiles whose guess was wrong, and the
// REDO instantiation frames all of those again with the true k0 and generic byte-wise code (bitmap search, global
// memory beyond the window).  The result never depends on the guess.
#pragma once

namespace quade {
namespace {

constexpr int FT_TILE = kFrameTile;            // bytes owned by a tile
#ifndef QUADE_F_WPT
#define QUADE_F_WPT 2
#endif
constexpr int FT_WPT = QUADE_F_WPT;            // 64-byte bitmap words per thread
constexpr int FT_NT = FT_TILE / 64 / FT_WPT;   // threads per CTA
constexpr int FT_NW = FT_NT / 32;
constexpr int FT_HALO = 2048;                  // bytes after the tile that are staged as well
constexpr int FT_WIN = FT_TILE + FT_HALO;
#ifndef QUADE_F_SPLIT
#define QUADE_F_SPLIT FT_WIN
#endif
constexpr int FT_SPLIT = QUADE_F_SPLIT;        // bytes of the first of two window copies (FT_WIN: one copy).  FT_TILE / 2 -- the threads
                                               // start on the first half while the second is in flight -- measured no gain (0.3494 vs 0.3476 ms)
constexpr int FT_WORDS = FT_WIN / 64;          // 64-bit bitmap words
#ifndef QUADE_F_PCAP
#define QUADE_F_PCAP (FT_TILE * 3 / 32)
#endif
constexpr int FT_PCAP = QUADE_F_PCAP;          // listed line ends per window: files whose lines average >= 10.7 bytes (records >= 43) stay
                                               // on the fast path, and 5 CTAs fit an SM: 0.377 -> 0.349 ms against twice the list at 4 CTAs
constexpr uint32_t FT_NONE = 0xffffffffu;
static_assert(FT_HALO / 64 == 32, "the halo's bitmap words are scanned by one warp");
static_assert(FT_NW >= 2 && FT_NW <= 32 && FT_WIN < 65536, "line ends are listed as u16 window offsets");
constexpr int FT_NPK = (FT_WPT + 1) / 2;       // registers that carry the line-end counts of a thread's words through the scans
static_assert(FT_WPT >= 1 && FT_WPT <= 4 && FT_NPK * FT_NW <= 32, "warp counts live in the first misc words");
// measured and rejected: one lane per warp spinning on the copy's mbarrier (0.504 vs 0.477 ms)
#ifndef QUADE_F_GUESSW
#define QUADE_F_GUESSW 0                       // only the warps that can own a record slot guess the record phase: measured slower (0.476 vs 0.456 ms)
#endif
#ifndef QUADE_F_IDXTOK
#define QUADE_F_IDXTOK 0                       // index reads: full name token scan (only "is there a token" is needed)
#endif
#ifndef QUADE_F_MINB
#define QUADE_F_MINB 5                         // as many CTAs as the shared memory of an SM holds (32 KiB tiles): 40 warps, 48 registers
#endif

// shared memory of a CTA: window (+ slack for the 64-byte header scan) | newline bitmap | line-end list | misc
constexpr int FT_OFF_POS = FT_WIN + 128;                        // u16[FT_PCAP] + a dump entry
constexpr int FT_OFF_MISC = FT_OFF_POS + FT_PCAP * 2 + 16;
constexpr int FT_OFF_BM = FT_OFF_MISC + 256;                    // redo pass only
constexpr int FT_SMEM_FAST = FT_OFF_BM, FT_SMEM_REDO = FT_OFF_BM + FT_WORDS * 8;
// misc (u32 index): [0..FT_NW) warp line-end counts, then
enum { FM_HALO = 32, FM_BASE, FM_NINV, FM_BLANK, FM_BAD, FM_TK = 40 /* [2][4]: ticket, stream, tile in stream */, FM_BAR = 56 /* byte 224 */ };

struct FWin {
    const uint8_t *data;    // global bytes of the file
    uint32_t sdata;         // shared address of byte `tbase`
    const u64 *bm;          // newline bitmap, bit i = byte tbase + i (i < win)
    uint32_t tbase, n, win; // win: staged bytes that exist (<= FT_WIN)
    bool last;              // last batch of the file: if it ends without '\n', position n counts as a line end
};

struct FRec {
    uint32_t name_off, nlen, seq_off, qual_off, seq_len, end;   // end: position after the record's last line
    bool complete, valid;
};

// newline flags of one word: 0x80 in every byte equal to '\n' (exact for all byte values).  (w ^ 0x0a..) & 0x7f..
// is ONE lop3 (written as C it is an AND and a XOR with an immediate each); bit 7 of a byte is the same in w and
// w ^ 0x0a0a0a0a
__device__ __forceinline__ uint32_t f_nl_flags(uint32_t w)
{
    uint32_t t;
    asm("lop3.b32 %0, %1, 0x0a0a0a0a, 0x7f7f7f7f, 0x28;" : "=r"(t) : "r"(w));
    return ~((t + 0x7f7f7f7fu) | w) & 0x80808080u;
}
// 0x80 in every byte <= 0x20 (every Python-2 whitespace byte is one; so are the other control characters)
__device__ __forceinline__ uint32_t f_c_flags(uint32_t w) { return ~(((w & 0x7f7f7f7fu) + 0x5f5f5f5fu) | w) & 0x80808080u; }
// 0x80 in every zero byte of x (exact)
__device__ __forceinline__ uint32_t f_z_flags(uint32_t x) { return ~(((x & 0x7f7f7f7fu) + 0x7f7f7f7fu) | x) & 0x80808080u; }
// 16 bytes -> 16 mask bits: IDP.4A weighs the 0x80 flags of two words into 128 x (8-bit mask)
__device__ __forceinline__ uint32_t f_gather16(uint32_t f0, uint32_t f1, uint32_t f2, uint32_t f3)
{
    const uint32_t lo = __dp4a(f1, 0x80402010u, __dp4a(f0, 0x08040201u, 0u));
    const uint32_t hi = __dp4a(f3, 0x80402010u, __dp4a(f2, 0x08040201u, 0u));
    return (lo >> 7) | (hi << 1);
}
// the four 0x80 flags of a word -> 4 bits
__device__ __forceinline__ uint32_t f_gather4(uint32_t f) { return __dp4a(f, 0x08040201u, 0u) >> 7; }

// python slice [sp0:sp1] of a string of length L -> offset and length (FastqSeq.py:91-93)
__device__ __forceinline__ void f_clamp(const int sp[2], uint32_t L, uint32_t *a, uint32_t *n)
{
    uint32_t s = (uint32_t)sp[0], e = (uint32_t)sp[1];
    s = s < L ? s : L;
    e = e < L ? e : L;
    *a = s;
    *n = e > s ? e - s : 0u;
}

// ---- generic (REDO) path: any text, any record length ------------------------------------------------------

__device__ __forceinline__ uint32_t g_byte(const FWin &t, uint32_t pos)
{
    const uint32_t q = pos - t.tbase;
    return q < t.win ? lds8(t.sdata + q) : (uint32_t)t.data[pos];
}

// position of the first line end at or after global position `from` (the virtual one at n included), FT_NONE if
// the file has no further line
__device__ __noinline__ uint32_t g_next_nl(const FWin &t, uint32_t from)
{
    uint32_t q = from - t.tbase;
    if (q < (uint32_t)FT_WIN) {
        uint32_t w = q >> 6;
        u64 m = t.bm[w] & (~0ull << (q & 63u));
        for (;;) {
            if (m) return t.tbase + w * 64u + (uint32_t)__ffsll((long long)m) - 1u;
            if (++w >= (uint32_t)FT_WORDS) break;
            m = t.bm[w];
        }
        from = t.tbase + (uint32_t)FT_WIN;
    }
    for (u64 p = from; p < t.n; ++p)
        if (t.data[p] == '\n') return (uint32_t)p;
    return (t.last && t.n > 0u && from <= t.n && t.data[t.n - 1u] != '\n') ? t.n : FT_NONE;
}

__device__ __forceinline__ uint32_t g_rstrip(const FWin &t, uint32_t s, uint32_t e)
{
    while (e > s && is_ws(g_byte(t, e - 1u))) --e;
    return e;
}

// one record starting at global position hs (FastqReader.py:37-44, FastqSeq.py:55).  pos[li .. li+3] are the
// window offsets of its four line ends as far as the list (npos entries) reaches; the rest is searched.
__device__ __noinline__ FRec g_parse(const FWin &t, uint32_t hs, const uint16_t *pos, uint32_t li, uint32_t npos)
{
    FRec r;
    r.complete = r.valid = false;
    r.name_off = r.nlen = r.seq_off = r.qual_off = r.seq_len = r.end = 0u;
    if (hs > t.n) return r;
    const uint32_t nl0 = li < npos ? t.tbase + pos[li] : g_next_nl(t, hs);
    if (nl0 == FT_NONE || nl0 >= t.n) return r;           // a line ending at n is the file's last line
    const uint32_t nl1 = li + 1u < npos ? t.tbase + pos[li + 1u] : g_next_nl(t, nl0 + 1u);
    if (nl1 == FT_NONE || nl1 >= t.n) return r;
    const uint32_t nl2 = li + 2u < npos ? t.tbase + pos[li + 2u] : g_next_nl(t, nl1 + 1u);
    if (nl2 == FT_NONE || nl2 >= t.n) return r;
    const uint32_t nl3 = li + 3u < npos ? t.tbase + pos[li + 3u] : g_next_nl(t, nl2 + 1u);
    if (nl3 == FT_NONE) return r;
    r.complete = true;
    r.end = nl3 < t.n ? nl3 + 1u : t.n;
    // name = line.rstrip()[1:].split()[0]: drop the first byte whatever it is, skip blanks, take the token
    uint32_t p = hs + 1u;
    while (p < nl0 && is_ws(g_byte(t, p))) ++p;
    r.name_off = p < nl0 ? p : nl0;
    uint32_t e = p;
    while (e < nl0 && !is_ws(g_byte(t, e))) ++e;
    r.nlen = e - p;
    r.seq_off = nl0 + 1u;
    r.qual_off = nl2 + 1u;
    r.seq_len = g_rstrip(t, r.seq_off, nl1) - r.seq_off;
    r.valid = r.seq_len == g_rstrip(t, r.qual_off, nl3) - r.qual_off;
    return r;
}

// index read: digest of the sample-index span + raw bytes of both spans for the name suffix, byte by byte.
//   dig = {key lo, key hi, N mask, in | mn << 8 | lowq << 16 | bad << 24}: 2 bits per upper-cased base (A,C,T,G = 0..3,
//   index.seq.upper(), Sample.py:65), bit i of the N mask for an 'N' at base i, bad = some base is none of ACGTN (the
//   read can never match, Sample.py:65-67) or the span is longer than 32 bases; lowq = some quality byte of the span is
//   below min_qual + 33, i.e. min(index.qual) < MIN_QUAL (FastqSeq.py:44, Sample.py:70).
__device__ __noinline__ uint4 g_digest(const BatchDev &b, const StreamDev &st, int s, const FWin &t, const FRec &r, uint32_t v)
{
    uint32_t ia, in, ma, mn;
    f_clamp(b.span[s - 2], r.seq_len, &ia, &in);           // QUADE_SPAN_IDX1 / IDX2
    f_clamp(b.span[s], r.seq_len, &ma, &mn);               // QUADE_SPAN_MOL1 / MOL2
    const uint32_t thr = (uint32_t)b.min_qual + 33u;
    u64 key = 0;
    uint32_t nm = 0, bad = in > 32u ? 1u : 0u, lowq = 0u;
    uint8_t *slot = st.sfx + (size_t)v * st.sfx_slot;
    uint32_t word = 0;
    for (uint32_t i = 0; i < in; ++i) {
        const uint32_t c = g_byte(t, r.seq_off + ia + i);
        if (g_byte(t, r.qual_off + ia + i) < thr) lowq = 1u;
        word |= c << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == in) {
            *reinterpret_cast<uint32_t *>(slot + (i & ~3u)) = word;
            word = 0;
        }
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
        word |= g_byte(t, r.seq_off + ma + i) << ((i & 3u) * 8u);
        if ((i & 3u) == 3u || i + 1u == mn) {
            *reinterpret_cast<uint32_t *>(mslot + (i & ~3u)) = word;
            word = 0;
        }
    }
    return make_uint4((uint32_t)key, (uint32_t)(key >> 32), nm, in | mn << 8 | lowq << 16 | bad << 24);
}

// ---- fast path: the whole record lies inside the staged window ---------------------------------------------
// Positions here are WINDOW offsets x: byte tbase + x of the file sits at shared address sm0 + x.

// 64 mask bits of the bytes <= 0x20 among the 64 bytes at the 16-byte aligned shared address a
__device__ __forceinline__ u64 f_cand64(uint32_t a)
{
#if QUADE_CHECKS
    {
        extern __shared__ __align__(128) uint8_t f_smem_chk[];
        const uint32_t lo = smem_u32(f_smem_chk);
        QCHK(a >= lo && a + 64u <= lo + (uint32_t)FT_WIN + 128u && (a & 15u) == 0u, 106u);
    }
#endif
    const uint4 q0 = lds128(a), q1 = lds128(a + 16u), q2 = lds128(a + 32u), q3 = lds128(a + 48u);
    const uint32_t lo = f_gather16(f_c_flags(q0.x), f_c_flags(q0.y), f_c_flags(q0.z), f_c_flags(q0.w)) |
                        f_gather16(f_c_flags(q1.x), f_c_flags(q1.y), f_c_flags(q1.z), f_c_flags(q1.w)) << 16;
    const uint32_t hi = f_gather16(f_c_flags(q2.x), f_c_flags(q2.y), f_c_flags(q2.z), f_c_flags(q2.w)) |
                        f_gather16(f_c_flags(q3.x), f_c_flags(q3.y), f_c_flags(q3.z), f_c_flags(q3.w)) << 16;
    return (u64)hi << 32 | lo;
}

// name = line.rstrip()[1:].split()[0] (FastqReader.py:42) of the header line [hs, nl0], nl0 = its '\n': drop the
// first byte whatever it is, skip blanks, take the bytes up to the next whitespace.  The '\n' at nl0 is whitespace,
// so the search always ends there at the latest; 64 header bytes are tested per step.
__device__ __forceinline__ void f_token(uint32_t sm0, uint32_t hs, uint32_t nl0, uint32_t &noff, uint32_t &nlen)
{
    uint32_t p = hs + 1u;
    while (p < nl0 && is_ws(lds8(sm0 + p))) ++p;
    if (p >= nl0) {
        noff = nl0;
        nlen = 0u;
        return;
    }
    uint32_t off = p & ~15u;                                   // sm0 is 16-byte aligned
    u64 m = f_cand64(sm0 + off) & (~0ull << (p & 15u));
    uint32_t e;
    for (;;) {
        if (m) {
            e = off + (uint32_t)__ffsll((long long)m) - 1u;
            if (e >= nl0) { e = nl0; break; }                  // nl0 may be the virtual end of a buffer without a final '\n'
            if (is_ws(lds8(sm0 + e))) break;
            m &= m - 1ull;                                     // a control character inside the token
        } else if (off + 64u > nl0) {
            e = nl0;
            break;
        } else {
            off += 64u;
            m = f_cand64(sm0 + off);
        }
    }
    noff = p;
    nlen = e - p;
}

__device__ __forceinline__ uint32_t f_rstrip(uint32_t sm0, uint32_t s, uint32_t e)
{
    while (e > s && is_ws(lds8(sm0 + e - 1u))) --e;
    return e;
}

// index read whose spans are at most 16 bases wide: the digest of g_digest, four bases per step.  sq / qq: window
// offsets of the sequence / quality line
__device__ __forceinline__ uint4 f_digest(const BatchDev &b, const StreamDev &st, int s, uint32_t sm0, uint32_t sq, uint32_t qq,
                                          uint32_t L, uint32_t v)
{
    uint32_t ia, in, ma, mn;
    f_clamp(b.span[s - 2], L, &ia, &in);                   // QUADE_SPAN_IDX1 / IDX2
    f_clamp(b.span[s], L, &ma, &mn);                       // QUADE_SPAN_MOL1 / MOL2
    const uint32_t W = (st.wi + 3u) >> 2, WM = (st.wm + 3u) >> 2;   // uniform: words per span (<= 4)
    const uint32_t thr4 = ((uint32_t)b.min_qual + 33u) * 0x01010101u;
    uint32_t *slot = reinterpret_cast<uint32_t *>(st.sfx + (size_t)v * st.sfx_slot);
    QCHK(v < st.cap_rec && 4u * (W + WM) <= st.sfx_slot && sq + ia + 4u * W + 4u <= (uint32_t)FT_WIN + 128u &&
             qq + ia + 4u * W + 4u <= (uint32_t)FT_WIN + 128u && sq + ma + 4u * WM + 4u <= (uint32_t)FT_WIN + 128u, 107u);
    u64 key = 0;
    uint32_t nm = 0, badf = 0, lowf = 0;
    {
        const uint32_t a = sm0 + sq + ia, a4 = a & ~3u, sh = (a & 3u) * 8u;
        const uint32_t c = sm0 + qq + ia, c4 = c & ~3u, shq = (c & 3u) * 8u;
        uint32_t xs = lds32(a4), xq = lds32(c4);
#pragma unroll
        for (uint32_t k = 0; k < 4; ++k) {
            if (k < W) {
                const uint32_t xs1 = lds32(a4 + 4u * k + 4u), xq1 = lds32(c4 + 4u * k + 4u);
                const uint32_t w = __funnelshift_r(xs, xs1, sh), qw = __funnelshift_r(xq, xq1, shq);
                xs = xs1;
                xq = xq1;
                const uint32_t rem = in > 4u * k ? in - 4u * k : 0u;
                const uint32_t msk = rem >= 4u ? 0xffffffffu : (1u << (8u * rem)) - 1u;     // bases of this word inside the span
                slot[k] = w & msk;
                // index.seq.upper() (Sample.py:65): clearing bit 5 turns only 'a'..'z' into 'A'..'Z' among the bytes that
                // can end up equal to a capital letter; (c >> 1) & 3 maps A,C,T,G (either case) to 0,1,2,3 and a byte is
                // one of them iff its upper-cased value equals the letter of its code (PRMT as a 4-entry table)
                const uint32_t up = w & 0xdfdfdfdfu, codes = (w >> 1) & 0x03030303u;
                uint32_t sel = (codes | codes >> 4) & 0x00ff00ffu;
                sel = (sel | sel >> 8) & 0xffffu;
                const uint32_t want = __byte_perm(0x47544341u, 0u, sel);
                const uint32_t m80 = msk & 0x80808080u;
                uint32_t t1, t2;
                asm("lop3.b32 %0, %1, %2, 0x7f7f7f7f, 0x28;" : "=r"(t1) : "r"(up), "r"(want));          // (up ^ want) & 0x7f..
                asm("lop3.b32 %0, %1, 0x4e4e4e4e, 0x7f7f7f7f, 0x28;" : "=r"(t2) : "r"(up));             // (up ^ 'N'..) & 0x7f..
                const uint32_t acgt = ~((t1 + 0x7f7f7f7fu) | up) & m80, isn = ~((t2 + 0x7f7f7f7fu) | up) & m80;
                badf |= ~(acgt | isn) & m80;
                const uint32_t kc = codes & ((acgt >> 7) * 3u);
                key |= (u64)((kc * 0x01041040u) >> 24) << (8u * k);
                nm |= f_gather4(isn) << (4u * k);
                // phred = q - 33 (FastqSeq.py:44): q < thr for some byte of the span <=> min(index.qual) < MIN_QUAL
                lowf |= ~((((qw | 0x80808080u) - thr4) | qw)) & m80;
            }
        }
    }
    {
        const uint32_t a = sm0 + sq + ma, a4 = a & ~3u, sh = (a & 3u) * 8u;
        uint32_t *mslot = slot + W;
        uint32_t xs = lds32(a4);
#pragma unroll
        for (uint32_t k = 0; k < 4; ++k) {
            if (k < WM) {
                const uint32_t xs1 = lds32(a4 + 4u * k + 4u);
                const uint32_t w = __funnelshift_r(xs, xs1, sh);
                xs = xs1;
                const uint32_t rem = mn > 4u * k ? mn - 4u * k : 0u;
                mslot[k] = w & (rem >= 4u ? 0xffffffffu : (1u << (8u * rem)) - 1u);
            }
        }
    }
    return make_uint4((uint32_t)key, (uint32_t)(key >> 32), nm, in | mn << 8 | (lowf ? 1u << 16 : 0u) | (badf ? 1u << 24 : 0u));
}

// ------------------------------------------------------------------------------------------------------------
// k_frame.  Persistent CTAs, dynamic tile tickets; 4 block barriers per tile.
// tile_tab[2 * lt]     = {line ends of the tile, guessed t0 (| 0x100: could not frame), record slots, scratch base}
// tile_tab[2 * lt + 1] = {records with len(seq) != len(qual), first slot with a blank header or FT_NONE, 0, 0}
// ------------------------------------------------------------------------------------------------------------
template <bool REDO>
__global__ void __launch_bounds__(FT_NT, REDO ? 1 : QUADE_F_MINB) k_frame(const __grid_constant__ BatchDev b, uint32_t total_tiles)
{
    extern __shared__ __align__(128) uint8_t f_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f_smem);
    const u64 *s_bm = reinterpret_cast<const u64 *>(f_smem + FT_OFF_BM);
    uint16_t *s_pos = reinterpret_cast<uint16_t *>(f_smem + FT_OFF_POS);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f_smem + FT_OFF_MISC);
    const uint32_t s_bar = sm0 + FT_OFF_MISC + 4u * FM_BAR, a_pos = sm0 + FT_OFF_POS;
    BatchState *S = b.state;
    const uint32_t n_todo = REDO ? min(S->n_redo, b.redo_cap) : total_tiles;

    // thread 0, early in a tile: the NEXT ticket (tiles are independent: nothing orders the tickets); later in the
    // tile, once the atomic has returned, its geometry goes into the other half of the ticket slots together with an
    // L2 prefetch of its window, so that the TMA copy issued once this tile is done hits L2
    auto ticket_take = [&]() -> uint32_t { return atomicAdd(REDO ? &S->ticket_redo : &S->ticket_frame, 1u); };
    auto ticket_put = [&](uint32_t half, uint32_t t) {
        uint32_t *tk = s_misc + FM_TK + 4u * half;
        tk[0] = t;
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
        tk[1] = (uint32_t)s;
        tk[2] = lt;
        const uint32_t tbase = lt * (uint32_t)FT_TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)FT_WIN, n16 - tbase) : 0u;
        if (bytes)
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(b.st[s].data + tbase), "r"(bytes) : "memory");
    };
    // thread 0, when the window is free: copy the window of the ticket in `half`
    auto issue = [&](uint32_t half) {
        const uint32_t *tk = s_misc + FM_TK + 4u * half;
        if (tk[0] >= n_todo) return;
        const int s = (int)tk[1];
        const uint32_t tbase = tk[2] * (uint32_t)FT_TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)FT_WIN, n16 - tbase) : 0u;
        if (bytes) {
            // two copies, two mbarriers: the threads start on the first half of the tile while the second is in flight
            const uint32_t first = min(bytes, (uint32_t)FT_SPLIT);
            fence_proxy_async();
            mbar_expect_tx(s_bar, first);
            bulk_g2s(sm0, b.st[s].data + tbase, first, s_bar);
            if (bytes > first) {
                mbar_expect_tx(s_bar + 8u, bytes - first);
                bulk_g2s(sm0 + first, b.st[s].data + tbase + first, bytes - first, s_bar + 8u);
            }
        }
    };
    if (tid == 0) {
        mbar_init(s_bar, 1);
        mbar_init(s_bar + 8u, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        s_misc[FM_NINV] = 0u;
        s_misc[FM_BLANK] = FT_NONE;
        s_misc[FM_BAD] = 0u;
        ticket_put(0u, ticket_take());
        issue(0u);
    }
    __syncthreads();
    uint32_t phase = 0;

    for (uint32_t it = 0;; ++it) {
        const uint32_t half = it & 1u;
        const uint32_t tile = s_misc[FM_TK + 4u * half];
        if (tile >= n_todo) break;
        const int s = (int)s_misc[FM_TK + 4u * half + 1u];
        const uint32_t lt = s_misc[FM_TK + 4u * half + 2u];
        const StreamDev &st = b.st[s];
        const uint32_t tbase = lt * (uint32_t)FT_TILE, n = st.n;
        const uint32_t win = n > tbase ? min((uint32_t)FT_WIN, n - tbase) : 0u;
        const bool lastbuf = ((b.last >> s) & 1) != 0;
        const bool edge = win < (uint32_t)FT_WIN;               // the window reaches the end of the buffer
        const uint32_t copied = ((n + 15u) & ~15u) > tbase ? min((uint32_t)FT_WIN, ((n + 15u) & ~15u) - tbase) : 0u;
        if (copied) {                                           // first copy of this tile
            while (!mbar_try_wait(s_bar, phase & 1u)) {}
            phase ^= 1u;
        }
        if (copied > (uint32_t)FT_SPLIT && (REDO || FT_WPT == 1)) {   // the redo pass (and one-word threads) wait for all of it here
            while (!mbar_try_wait(s_bar + 8u, (phase >> 1) & 1u)) {}
            phase ^= 2u;
        }

        // 1. newline masks.  Fast pass: thread t owns the 64 bytes of the bitmap words t, t + FT_NT, ...; the four
        //    16-byte vectors of a word are read in an order rotated by the lane so that the 8 lanes of a quarter warp
        //    hit 8 different bank groups, and the four 16-bit masks are put in place by ONE 64-bit rotation: the words
        //    never leave the registers.  Redo pass: the masks go to the shared bitmap that the generic parser searches.
        u64 mine[FT_WPT];
        if (!REDO) {
            const uint32_t rot = ((uint32_t)lane >> 1) & 3u, sh = (rot & 1u) * 16u;
#pragma unroll
            for (int w = 0; w < FT_WPT; ++w) {
                const uint32_t o0 = 64u * ((uint32_t)tid + (uint32_t)(w * FT_NT)), a = sm0 + o0;
                if (FT_WPT > 1 && w == FT_WPT / 2 && copied > (uint32_t)FT_SPLIT) {   // second copy: the upper word rows and the halo
                    while (!mbar_try_wait(s_bar + 8u, (phase >> 1) & 1u)) {}
                    phase ^= 2u;
                }
                uint32_t r[4];
#pragma unroll
                for (uint32_t i = 0; i < 4; ++i) {
                    const uint4 q = lds128(a + 16u * ((i + rot) & 3u));
                    r[i] = f_gather16(f_nl_flags(q.x), f_nl_flags(q.y), f_nl_flags(q.z), f_nl_flags(q.w));
                }
                uint32_t xl = r[0] | r[1] << 16, xh = r[2] | r[3] << 16;   // mask of vector (i + rot) & 3 at bits 16 i
                if (rot & 2u) { const uint32_t t = xl; xl = xh; xh = t; }  // ... rotated left by 16 rot: vector v at bits 16 v
                u64 m = (u64)__funnelshift_l(xl, xh, sh) << 32 | __funnelshift_l(xh, xl, sh);
                if (edge) {                                     // the window reaches the end of the buffer: clip
                    m = o0 >= win ? 0ull : (win - o0 >= 64u ? m : m & ((1ull << (win - o0)) - 1ull));
                    if (lastbuf && n > 0u && ((n - tbase) >> 6) == (o0 >> 6) && st.data[n - 1u] != '\n')
                        m |= 1ull << ((n - tbase) & 63u);        // no final '\n': a virtual line end AT n
                }
                mine[w] = m;
            }
        } else {
            uint32_t vbit = FT_NONE;                            // window offset of a virtual last line end
            if (lastbuf && n > 0u && n - tbase < (uint32_t)FT_WIN) vbit = n - tbase;
#pragma unroll 1
            for (uint32_t c = (uint32_t)tid; c < (uint32_t)FT_WIN / 16u; c += (uint32_t)FT_NT) {
                const uint32_t off = 16u * c;
                uint32_t m = 0u;
                if (off < win) {
                    const uint4 q = lds128(sm0 + off);
                    m = f_gather16(f_nl_flags(q.x), f_nl_flags(q.y), f_nl_flags(q.z), f_nl_flags(q.w));
                    if (win - off < 16u) m &= (1u << (win - off)) - 1u;
                }
                if ((vbit >> 4) == c && st.data[n - 1u] != '\n') m |= 1u << (vbit & 15u);   // no final '\n'
                sts16(sm0 + FT_OFF_BM + 2u * c, m);
            }
            __syncthreads();                                    // bitmap complete
#pragma unroll
            for (int w = 0; w < FT_WPT; ++w) mine[w] = s_bm[tid + w * FT_NT];
        }

        // 2. line ends per thread; the counts of a thread's words (<= 64 each, <= 2^14 per tile and word row) travel
        //    through the scans as 16-bit fields, two per register.  Redo pass: the halo's 32 words by warp 1
        uint32_t cnt[FT_NPK], inc[FT_NPK];
#pragma unroll
        for (int k = 0; k < FT_NPK; ++k) {
            cnt[k] = (uint32_t)__popcll(mine[2 * k]);
            if (2 * k + 1 < FT_WPT) cnt[k] |= (uint32_t)__popcll(mine[2 * k + 1 < FT_WPT ? 2 * k + 1 : 0]) << 16;
            inc[k] = cnt[k];
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
            for (int k = 0; k < FT_NPK; ++k) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, inc[k], o);
                if (lane >= o) inc[k] += x;
            }
        }
        if (lane == 31) {
#pragma unroll
            for (int k = 0; k < FT_NPK; ++k) s_misc[FT_NPK * warp + k] = inc[k];
        }
        u64 hm = 0;
        uint32_t hinc = 0, hcnt = 0;
        if (REDO && warp == 1) {
            hm = s_bm[FT_TILE / 64 + lane];
            hcnt = (uint32_t)__popcll(hm);
            hinc = hcnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, hinc, o);
                if (lane >= o) hinc += x;
            }
            if (lane == 31) s_misc[FM_HALO] = hinc;
        }
        __syncthreads();                                        // B2: warp counts
        uint32_t before[FT_WPT], total = 0u;                    // line ends of the tile before each of the thread's words
#pragma unroll
        for (int k = 0; k < FT_NPK; ++k) {                      // exclusive prefix over the warps, by every warp
            const uint32_t w = lane < FT_NW ? s_misc[FT_NPK * lane + k] : 0u;
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < FT_NW; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += x;
            }
            const uint32_t bw = __shfl_sync(0xffffffffu, wi - w, warp) + inc[k] - cnt[k], tw = __shfl_sync(0xffffffffu, wi, FT_NW - 1);
#pragma unroll
            for (int j = 0; j < 2; ++j) {                       // word row w2 lies after all of the rows before it
                const int w2 = 2 * k + j;
                if (w2 < FT_WPT) {
                    before[w2] = total + ((bw >> (16 * j)) & 0xffffu);
                    total += (tw >> (16 * j)) & 0xffffu;
                }
            }
        }
        // the fast pass needs the halo only for the record that starts last in the tile: 4 line ends at most
        uint32_t npos = REDO ? min(total + s_misc[FM_HALO], (uint32_t)FT_PCAP) : 0u;
        const uint32_t first = lt == 0u ? 1u : 0u;              // slot 0 of a file's first tile starts at byte 0
        const uint32_t nres = first + ((total + 3u) >> 2);      // upper bound of the tile's record slots over the 4 phases
        const bool listed = REDO || total + 4u <= (uint32_t)FT_PCAP;
        uint32_t sb_reg = 0u, tk_reg = 0u;
        if (tid == 0) {
            // scratch room for the tile's records, in tile order of arrival, and the next ticket: both atomics fly
            // during the list build (their results are first touched after it)
            if (!REDO) sb_reg = atomicAdd(&S->scratch_cursor[s], nres);
            tk_reg = ticket_take();
            if (!REDO) atomicAdd(&S->chunk_lines[s][lt / (uint32_t)kTileChunk], total);   // for k_tilescan
        }

        // 3. dense list of the window's line ends (tile, then halo: list order = byte order)
        if (listed) {
            // all 32-bit halves of the thread's words advance in the same iteration: the warp loops as often as its
            // fullest half has line ends
            uint32_t hb[2 * FT_WPT], hi_[2 * FT_WPT];          // bits left / list index, per half
            uint32_t any = 0u;
#pragma unroll
            for (int w = 0; w < FT_WPT; ++w) {
                hb[2 * w] = (uint32_t)mine[w];
                hb[2 * w + 1] = (uint32_t)(mine[w] >> 32);
                hi_[2 * w] = before[w];
                hi_[2 * w + 1] = before[w] + (uint32_t)__popc(hb[2 * w]);
                any |= hb[2 * w] | hb[2 * w + 1];
            }
            if (REDO) {
                while (any) {
                    any = 0u;
#pragma unroll
                    for (int h = 0; h < 2 * FT_WPT; ++h) {
                        if (hb[h]) {
                            const uint32_t bit = (uint32_t)__ffs((int)hb[h]) - 1u;
                            hb[h] &= hb[h] - 1u;
                            const uint32_t o0 = 64u * ((uint32_t)tid + (uint32_t)((h >> 1) * FT_NT)) + 32u * (uint32_t)(h & 1);
                            if (hi_[h] < (uint32_t)FT_PCAP) s_pos[hi_[h]] = (uint16_t)(o0 + bit);
                            ++hi_[h];
                            any |= hb[h];
                        }
                    }
                }
            } else {
                // branch-free body: a half without line ends stores to a dump slot behind the list (68 -> ~40
                // instructions per iteration against the branchy form)
                while (any) {
                    any = 0u;
#pragma unroll
                    for (int h = 0; h < 2 * FT_WPT; ++h) {
                        const uint32_t x = hb[h];
                        const uint32_t o0 = 64u * ((uint32_t)tid + (uint32_t)((h >> 1) * FT_NT)) + 32u * (uint32_t)(h & 1);
                        const uint32_t pos = o0 + (uint32_t)__ffs((int)x) - 1u;
                        QCHK(!x || hi_[h] < (uint32_t)FT_PCAP, 101u);
                        sts16(a_pos + 2u * (x ? hi_[h] : (uint32_t)FT_PCAP), pos);
                        hi_[h] += x ? 1u : 0u;
                        hb[h] = x & (x - 1u);
                        any |= hb[h];
                    }
                }
            }
            if (REDO && warp == 1) {
                uint32_t i = total + hinc - hcnt;
                const uint32_t h0 = (uint32_t)(FT_TILE / 64 + lane) * 64u;
                while (hm && i < npos) {
                    const uint32_t bit = (uint32_t)__ffsll((long long)hm) - 1u;
                    hm &= hm - 1ull;
                    s_pos[i] = (uint16_t)(h0 + bit);
                    ++i;
                }
            }
            if (!REDO && warp == 1) {
                // the first 4 line ends of the halo, 512 bytes (one vector per lane) at a time: 2 x 150 reads need one step
                uint32_t found = 0u;
                for (uint32_t h = 0; h < (uint32_t)FT_HALO / 512u && found < 4u; ++h) {
                    const uint32_t off = (uint32_t)FT_TILE + 512u * h + 16u * (uint32_t)lane;
                    uint32_t m = 0u;
                    if (off < win) {
                        const uint4 q = lds128(sm0 + off);
                        m = f_gather16(f_nl_flags(q.x), f_nl_flags(q.y), f_nl_flags(q.z), f_nl_flags(q.w));
                        if (win - off < 16u) m &= (1u << (win - off)) - 1u;
                    }
                    if (edge && lastbuf && n > 0u && ((n - tbase) >> 4) == (off >> 4) && st.data[n - 1u] != '\n')
                        m |= 1u << ((n - tbase) & 15u);
                    const uint32_t c = (uint32_t)__popc(m);
                    uint32_t ci = c;
#pragma unroll
                    for (int o = 1; o < 32; o <<= 1) {
                        const uint32_t x = __shfl_up_sync(0xffffffffu, ci, o);
                        if (lane >= o) ci += x;
                    }
                    uint32_t r = found + ci - c;
                    while (m && r < 4u) {
                        QCHK(total + r < (uint32_t)FT_PCAP, 102u);
                        s_pos[total + r] = (uint16_t)(off + (uint32_t)__ffs((int)m) - 1u);
                        m &= m - 1u;
                        ++r;
                    }
                    found += __shfl_sync(0xffffffffu, ci, 31);
                }
                if (lane == 0) s_misc[FM_HALO] = min(found, 4u);
            }
        }
        if (tid == 0) {
            ticket_put(half ^ 1u, tk_reg);
            if (!REDO) s_misc[FM_BASE] = (u64)sb_reg + nres > (u64)st.cap_rec ? FT_NONE : sb_reg;
        }
        __syncthreads();                                        // B3: list complete, next ticket and scratch base known
        if (!REDO) npos = total + s_misc[FM_HALO];

        // 4. which list entries close a record: the true k0 (redo) or the guess from the text
        uint32_t t0 = 3u, k0 = 0u;                              // t0: list index of the first record-closing line end
        if (REDO) {
            k0 = st.tile_tab[2u * lt].x;
            t0 = (3u - k0) & 3u;
        } else if (lt != 0u && listed && (!QUADE_F_GUESSW || (uint32_t)warp * 32u < nres)) {   // warps that can own a record slot
            bool ok = false;
            if (lane < 4 && (uint32_t)lane + 2u < npos)
                ok = lds8(sm0 + lds16(a_pos + 2u * lane) + 1u) == '@' && lds8(sm0 + lds16(a_pos + 2u * (lane + 2)) + 1u) == '+';
            const unsigned g = __ballot_sync(0xffffffffu, ok);
            t0 = g ? (uint32_t)__ffs((int)g) - 1u : 0u;
        }
        const uint32_t nslot = listed ? first + (total > t0 ? (total - t0 + 3u) >> 2 : 0u) : nres;
        const uint32_t rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;  // REDO: record index of slot 0

        if (REDO) {
            // ---- generic pass: outputs at the scratch position reserved by the fast pass -----------------------
            FWin T;
            T.data = st.data;
            T.sdata = sm0;
            T.bm = s_bm;
            T.tbase = tbase;
            T.n = n;
            T.win = win;
            T.last = lastbuf;
            const uint32_t sb = st.tile_tab[2u * lt].w;
            if (sb == FT_NONE) {
                if (tid == 0 && nslot) S->err_capacity = 1u;
            } else {
                for (uint32_t slot = (uint32_t)tid; slot < nslot; slot += FT_NT) {
                    FRec R;
                    if (first && slot == 0u) {
                        R = g_parse(T, 0u, s_pos, 0u, npos);
                    } else {
                        const uint32_t ti = t0 + 4u * (slot - first);
                        if (ti < npos) {
                            R = g_parse(T, tbase + s_pos[ti] + 1u, s_pos, ti + 1u, npos);
                        } else {                                // more than FT_PCAP lines in the window: bitmap search
                            const uint32_t want = (rbase + slot) * 4u - 1u;
                            uint32_t k = k0, ppos = tbase, e = FT_NONE;
                            while (k <= want) {
                                e = g_next_nl(T, ppos);
                                if (e == FT_NONE) break;
                                ppos = e + 1u;
                                ++k;
                            }
                            R.complete = false;
                            if (e != FT_NONE) R = g_parse(T, e + 1u, s_pos, 0u, 0u);
                        }
                    }
                    const uint32_t v = sb + slot;
                    if (!R.complete) {                          // trailing partial record (FastqReader.py:57-58): never located
                        st.nv[v] = make_uint2(1u, 0u);
                        st.rec[v] = make_uint4(0u, 0u, 0u, 0x80000000u);
                        continue;
                    }
                    const uint32_t inv = R.valid ? 0u : 0x80000000u;
                    if (R.nlen == 0u) atomicMax(&S->first_blank[s], ~(rbase + slot));   // keeps the LOWEST blank record index
                    if (inv) atomicAdd(&S->n_invalid[s], 1u);
                    st.nv[v] = make_uint2(R.nlen, R.end);
                    if (s < 2) st.rec[v] = make_uint4(R.name_off, R.seq_off, R.qual_off, R.seq_len | inv);
                    else if (inv) st.rec[v] = make_uint4(0u, 0u, 0u, inv);
                    else st.rec[v] = g_digest(b, st, s, T, R, v);
                }
            }
            __syncthreads();                                    // B4: window, bitmap and list are free
            if (tid == 0) issue(half ^ 1u);
            continue;
        }

        // ---- fast pass: one thread per record, everything inside the window ------------------------------------
        const bool fastdig = st.wi <= 16u && st.wm <= 16u;
        uint4 o_rec = make_uint4(0u, 0u, 0u, 0u);
        uint2 o_nv = make_uint2(0u, 0u);
        auto parse_slot = [&](uint32_t slot, uint32_t v) -> bool {
            // window offsets of the record's start and of its four line ends
            uint32_t hs, li;
            if (first && slot == 0u) { hs = 0u; li = 0u; }
            else {
                const uint32_t ti = t0 + 4u * (slot - first);
                hs = lds16(a_pos + 2u * ti) + 1u;
                li = ti + 1u;
            }
            if (li + 4u > npos) {
                // the record does not end inside the window: the trailing partial record of the buffer
                // (FastqReader.py:57-58; the caller carries it over), or a record for the redo pass
                if (!edge) s_misc[FM_BAD] = 1u;
                return false;
            }
            const uint32_t nl0 = lds16(a_pos + 2u * li), nl1 = lds16(a_pos + 2u * li + 2u), nl2 = lds16(a_pos + 2u * li + 4u),
                           nl3 = lds16(a_pos + 2u * li + 6u);
            QCHK(li + 3u < (uint32_t)FT_PCAP && hs <= nl0 && nl0 < nl1 && nl1 < nl2 && nl2 < nl3 && nl3 <= win, 103u);
            uint32_t noff, nlen;
            if (s < 2 || QUADE_F_IDXTOK) f_token(sm0, hs, nl0, noff, nlen);
            else {                                              // index reads: only whether there is a name token matters
                noff = hs + 1u;
                while (noff < nl0 && is_ws(lds8(sm0 + noff))) ++noff;
                nlen = noff < nl0 ? 1u : 0u;
            }
            const uint32_t sq = nl0 + 1u, qq = nl2 + 1u;
            const uint32_t L = f_rstrip(sm0, sq, nl1) - sq;
            const bool valid = L == f_rstrip(sm0, qq, nl3) - qq;
            o_nv = make_uint2(nlen, min(tbase + nl3 + 1u, n));   // a virtual last line end sits AT n
            if (nlen == 0u) atomicMin(&s_misc[FM_BLANK], slot);
            if (!valid) atomicAdd(&s_misc[FM_NINV], 1u);
            if (s < 2) {
                o_rec = make_uint4(tbase + noff, tbase + sq, tbase + qq, L | (valid ? 0u : 0x80000000u));
            } else if (!valid) {
                o_rec = make_uint4(0u, 0u, 0u, 0x80000000u);
            } else {                                            // the digest also stores the raw span bytes at v
                if (fastdig) o_rec = f_digest(b, st, s, sm0, sq, qq, L, v);
                else {
                    FWin T;
                    T.data = st.data; T.sdata = sm0; T.bm = s_bm; T.tbase = tbase; T.n = n; T.win = win; T.last = lastbuf;
                    FRec R;
                    R.seq_off = tbase + sq; R.qual_off = tbase + qq; R.seq_len = L;
                    o_rec = g_digest(b, st, s, T, R, v);
                }
            }
            return true;
        };
        const uint32_t base = s_misc[FM_BASE];
        const bool can = listed && base != FT_NONE;
        if (can) {
            for (uint32_t slot = (uint32_t)tid; slot < nslot; slot += FT_NT)     // more than one round: records < 64 bytes
                if (parse_slot(slot, base + slot)) {
                    QCHK(base + slot < st.cap_rec, 104u);
                    st.nv[base + slot] = o_nv;
                    st.rec[base + slot] = o_rec;
                }
        }
        __syncthreads();                                        // B4: window, bitmap and list are free; tile counters complete
        if (tid == 0) {
            const bool bad = !can || s_misc[FM_BAD] != 0u;
            st.tile_tab[2u * lt] = make_uint4(total, t0 | (bad ? 0x100u : 0u), nslot, base);
            st.tile_tab[2u * lt + 1u] = make_uint4(s_misc[FM_NINV], s_misc[FM_BLANK], 0u, 0u);
            s_misc[FM_NINV] = 0u;
            s_misc[FM_BLANK] = FT_NONE;
            s_misc[FM_BAD] = 0u;
            issue(half ^ 1u);
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// k_tilescan: one CTA per chunk of kTileChunk tiles of one stream.  The line ends of the chunks before it were
// summed by k_frame (chunk_lines), so every CTA is independent.
//   phase 1, one tile per thread: exclusive scan of the tiles' line counts -> true k0, record base and record count
//   per tile; tile_tab[2t] becomes {k0, ok, records, scratch base}; a tile the fast pass could not frame, or whose
//   guessed phase was wrong, goes to the redo list; the others add their invalid-record count and first blank header;
//   phase 2, one tile per warp: loc[record] = scratch position (the pairing ordinal of a file without invalid
//   records is the record index; k_locate_rare rewrites loc for the files that hold some).
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTileChunk) k_tilescan(const __grid_constant__ BatchDev b)
{
    __shared__ uint32_t s_warp[33];
    __shared__ uint32_t s_base, s_all;
    __shared__ uint32_t s_rb[kTileChunk], s_ns[kTileChunk], s_sb[kTileChunk];
    const int s = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const StreamDev &st = b.st[s];
    BatchState *S = b.state;
    const uint32_t n_t = st.n ? st.n / (uint32_t)FT_TILE + 1u : 0u, chunk = blockIdx.x;
    if (chunk * (uint32_t)kTileChunk >= n_t) return;
    if (tid == 0) s_base = s_all = 0u;
    __syncthreads();
    {                                                           // line ends in the chunks before this one, and in all
        uint32_t v = 0u, a = 0u;
        for (uint32_t c = (uint32_t)tid; c < (uint32_t)kMaxChunks; c += (uint32_t)kTileChunk) {
            const uint32_t x = S->chunk_lines[s][c];
            a += x;
            v += c < chunk ? x : 0u;
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            v += __shfl_xor_sync(0xffffffffu, v, o);
            a += __shfl_xor_sync(0xffffffffu, a, o);
        }
        if (lane == 0 && v) atomicAdd(&s_base, v);
        if (lane == 0 && a) atomicAdd(&s_all, a);
    }
    const uint32_t t = chunk * (uint32_t)kTileChunk + (uint32_t)tid;
    const uint4 e = t < n_t ? st.tile_tab[2u * t] : make_uint4(0u, 0u, 0u, 0u);
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
        const uint32_t w = lane < kTileChunk / 32 ? s_warp[lane] : 0u;
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= o) wi += x;
        }
        s_warp[lane] = wi - w;
    }
    __syncthreads();
    const uint32_t n_rec = min(s_all >> 2, st.cap_rec);         // complete records of the buffer (FastqReader.py:57-58)
    s_ns[tid] = 0u;
    if (t < n_t) {
        const uint32_t k0 = s_base + s_warp[warp] + inc - total, first = t == 0u ? 1u : 0u;
        const uint32_t rbase = t == 0u ? 0u : (k0 >> 2) + 1u;
        uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + first;
        const bool ok = nslot == 0u || (e.y == ((3u - k0) & 3u) && e.w != FT_NONE);
        st.tile_tab[2u * t] = make_uint4(k0, ok ? 1u : 0u, nslot, e.w);
        if (!ok) {
            const uint32_t idx = atomicAdd(&S->n_redo, 1u);
            if (idx < b.redo_cap) b.redo_list[idx] = (uint32_t)s << 28 | t;
            else S->err_capacity = 1u;
        } else if (nslot) {
            const uint4 x = st.tile_tab[2u * t + 1u];
            if (x.x) atomicAdd(&S->n_invalid[s], x.x);
            if (x.y != FT_NONE) atomicMax(&S->first_blank[s], ~(rbase + x.y));   // keeps the LOWEST blank record index
        }
        if (e.w == FT_NONE && nslot) S->err_capacity = 1u;       // no scratch room: the batch holds more records than cap_rec
        nslot = rbase < n_rec ? min(nslot, n_rec - rbase) : 0u;  // the slot of a trailing partial record is not a record
        s_rb[tid] = rbase;
        s_ns[tid] = e.w == FT_NONE ? 0u : nslot;
        s_sb[tid] = e.w;
        if (t == n_t - 1u) S->n_lines[s] = k0 + total;
    }
    __syncthreads();
    for (uint32_t i = (uint32_t)warp; i < (uint32_t)kTileChunk; i += kTileChunk / 32) {
        const uint32_t ns = s_ns[i], rb = s_rb[i], sb = s_sb[i];
        for (uint32_t j = (uint32_t)lane; j < ns; j += 32u) {
            QCHK(rb + j < st.cap_rec && sb + j < st.cap_rec, 105u);
            st.loc[rb + j] = sb + j;
        }
    }
}

// ------------------------------------------------------------------------------------------------------------
// k_locate_rare: only for a file that holds records with len(seq) != len(qual) (FastqReader.py:48-50 skips them in
// their own file only, so the pairing ordinal is the rank among the VALID records).  One CTA per stream walks the
// tiles in order, kTileChunk at a time: valid records per tile, block scan, loc / vraw rewritten.  Files without
// such records (the normal case) return at once.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kTileChunk) k_locate_rare(const __grid_constant__ BatchDev b)
{
    __shared__ uint32_t s_warp[33];
    __shared__ uint32_t s_run;
    const int s = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    BatchState *S = b.state;
    if (S->n_invalid[s] == 0u || S->err_capacity) return;
    const StreamDev &st = b.st[s];
    const uint32_t n_t = st.n ? st.n / (uint32_t)FT_TILE + 1u : 0u;
    const uint32_t n_rec = min(S->n_lines[s] >> 2, st.cap_rec);
    if (tid == 0) s_run = 0u;
    __syncthreads();
    for (uint32_t t0 = 0; t0 < n_t; t0 += (uint32_t)kTileChunk) {
        const uint32_t t = t0 + (uint32_t)tid;
        uint32_t rbase = 0u, ns = 0u, sb = 0u, nv = 0u;
        if (t < n_t) {
            const uint4 e = st.tile_tab[2u * t];
            rbase = t == 0u ? 0u : (e.x >> 2) + 1u;
            ns = rbase < n_rec ? min(e.z, n_rec - rbase) : 0u;
            sb = e.w;
            for (uint32_t j = 0; j < ns; ++j) nv += (st.rec[sb + j].w >> 31) ? 0u : 1u;
        }
        uint32_t inc = nv;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const uint32_t w = lane < kTileChunk / 32 ? s_warp[lane] : 0u;
            uint32_t wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t x = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += x;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_warp[32] = wi;
        }
        __syncthreads();
        uint32_t v = s_run + s_warp[warp] + inc - nv;
        for (uint32_t j = 0; j < ns; ++j) {
            if (st.rec[sb + j].w >> 31) continue;
            st.loc[v] = sb + j;
            st.vraw[v] = rbase + j;
            ++v;
        }
        __syncthreads();
        if (tid == 0) s_run += s_warp[32];
        __syncthreads();
    }
}

} // namespace
} // namespace quade
