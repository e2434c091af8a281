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

struct F2Rec {
    uint32_t name_off, nlen, seq_off, qual_off, seq_len, end;   // end: position after the record's last line
    bool complete, valid;
};

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
// python slice [sp0:sp1] of a string of length L -> offset and length (FastqSeq.py:91-93)
__device__ __forceinline__ void f2_clamp(const int sp[2], uint32_t L, uint32_t *a, uint32_t *n)
{
    uint32_t s = (uint32_t)sp[0], e = (uint32_t)sp[1];
    s = s < L ? s : L;
    e = e < L ? e : L;
    *a = s;
    *n = e > s ? e - s : 0u;
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
