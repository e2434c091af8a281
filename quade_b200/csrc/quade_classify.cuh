// quade_classify.cuh -- k_pairing and k_classify (included by quade_kernels.cu)
#pragma once

namespace quade {
namespace {

// ------------------------------------------------------------------------------------------------------------
// k_pairing (Quade.py:210-224): the loop pulls the next VALID record of R1, R2, I1[, I2] in that order and ends at
// the first file that is exhausted, so n_pairs = min over files of the valid count; one thread.
// A batch is a slice of the chunk: a file "runs dry" only if its buffer ends at end-of-file (b.last).  Without a dry
// file the caller carries the unconsumed tail into the next batch, so consumed[] may and does run over the invalid
// records that follow the last paired one (a buffer holding nothing but such records must make progress), but never
// over a blank header that the reference may not reach (IndexError of FastqReader.py:42 only when it is parsed).
// ------------------------------------------------------------------------------------------------------------
__device__ uint32_t raw_end(const BatchDev &b, const BatchState *S, int s, uint32_t j)
{
    const StreamDev &st = b.st[s];
    if (S->n_invalid[s] == 0u) return st.nv[st.loc[j]].y;   // record index = pairing ordinal
    // rare path: the tile holding raw record j (largest tile whose first record index is <= j)
    const uint32_t n_t = st.n ? st.n / (uint32_t)FT_TILE + 1u : 0u;
    uint32_t lo = 0u, hi = n_t;
    while (hi - lo > 1u) {
        const uint32_t mid = (lo + hi) >> 1;
        const uint32_t rb = (st.tile_tab[2u * mid].x >> 2) + 1u;
        if (rb <= j) lo = mid; else hi = mid;
    }
    const uint4 e = st.tile_tab[2u * lo];
    const uint32_t rbase = lo == 0u ? 0u : (e.x >> 2) + 1u;
    return st.nv[e.w + (j - rbase)].y;
}

__global__ void k_pairing(const __grid_constant__ BatchDev b)
{
    if (threadIdx.x != 0) return;
    BatchState *S = b.state;
    uint32_t n_rec[kMaxStreams], np_raw = 0xffffffffu;
    for (int s = 0; s < b.n_streams; ++s) {
        n_rec[s] = min(S->n_lines[s] >> 2, b.st[s].cap_rec);
        S->n_valid[s] = n_rec[s] - min(n_rec[s], S->n_invalid[s]);
        np_raw = min(np_raw, S->n_valid[s]);
    }
    uint32_t np = min(np_raw, b.cap_pairs);                     // the rest is carried into the next batch
    if (S->err_capacity) np = 0u;
    // iteration np of the loop fails at the first file (in pull order) that has no further valid record; the files
    // before it were pulled once more, the failing one was read to the end of its complete records.  It has run DRY
    // (end of the chunk, Quade.py:223) only if its buffer ends at end-of-file.
    int first = -1;
    if (np == np_raw && !S->err_capacity)
        for (int s = 0; s < b.n_streams && first < 0; ++s)
            if (S->n_valid[s] == np) first = s;
    const bool dry = first >= 0 && ((b.last >> first) & 1);
    S->chunk_end = dry ? 1u : 0u;
    for (int s = 0; s < b.n_streams && !S->err_capacity; ++s) {
        const StreamDev &st = b.st[s];
        const bool sq = S->n_invalid[s] != 0u;                  // vraw[v] = raw index, else v itself
        auto vraw = [&](uint32_t v) { return sq ? st.vraw[v] : v; };
        const uint32_t upto = np > 0u ? vraw(np - 1u) + 1u : 0u;                 // raw records up to the last paired one
        const uint32_t next = S->n_valid[s] > np ? vraw(np) : n_rec[s];          // raw index of the next valid record
        const uint32_t fb = S->first_blank[s], jb = ~fb;
        // raw records the reference has parsed when iteration np fails
        const uint32_t parsed = first < 0 ? upto : (s < first ? min(next + 1u, n_rec[s]) : (s == first ? n_rec[s] : upto));
        if (fb && jb < parsed) S->err_blank = 1u;
        uint32_t cons = upto;
        if (!dry) {
            // the caller carries the tails over: the invalid records in front of the next valid one stay consumed
            // (FastqReader.py:48-50 skips them in their own file whatever the other files hold), a blank header the
            // reference may never reach is not passed
            cons = next;
            if (fb && jb >= upto && jb < next) cons = jb;
        }
        S->n_skipped[s] = cons - np;
        S->consumed[s] = cons > 0u ? raw_end(b, S, s, cons - 1u) : 0u;
    }
    uint32_t tp = kCompactTilePairs;                                            // k_compact tile size from the mean record size
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
// k_classify (Sample.py:56-91).  One thread per pair: fuse the two digests into the key of index1 + index2, probe
// the perfect-hash table (full-key verify), quality filter, bin, the name suffix ':' index [':' molecular] '\n'
// (FastqWriter.py:61-66), output record lengths, per-block per-bin byte histogram, counters.
// ------------------------------------------------------------------------------------------------------------
constexpr int C2T = 256;

__global__ void __launch_bounds__(C2T) k_classify(const __grid_constant__ BatchDev b, int table_in_smem)
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

    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        for (int i = tid; i < 2 * nb; i += C2T) h_b1[i] = 0u;
        __syncthreads();
        const uint32_t p0 = blk * b.pairs_per_block;
        uint32_t end = p0 + b.pairs_per_block;
        if (end > n_pairs) end = n_pairs;
        // the scratch positions of the pair after this one are loaded while this one is classified: the chain
        // loc -> record is two dependent global loads
        uint32_t p = p0 + tid;
        uint32_t nv1 = 0u, nv2 = 0u, nr1 = 0u, nr2 = 0u;
        if (p < end) {
            nv1 = i1.loc[p];
            nv2 = b.dual ? i2.loc[p] : 0u;
            nr1 = b.st[0].loc[p];
            nr2 = b.st[1].loc[p];
        }
        for (; p < end; p += C2T) {
            const uint32_t v1 = nv1, v2 = nv2, r1 = nr1, r2 = nr2;
            QCHK(v1 < i1.cap_rec && (!b.dual || v2 < i2.cap_rec) && r1 < b.st[0].cap_rec && r2 < b.st[1].cap_rec && p < b.cap_pairs, 301u);
            if (p + C2T < end) {
                nv1 = i1.loc[p + C2T];
                nv2 = b.dual ? i2.loc[p + C2T] : 0u;
                nr1 = b.st[0].loc[p + C2T];
                nr2 = b.st[1].loc[p + C2T];
            }
            // insert-read fields for the output lengths (read even when the bin's writer is off: no dependent wait)
            const uint32_t a1 = b.st[0].nv[r1].x + 2u * b.st[0].rec[r1].w, a2 = b.st[1].nv[r2].x + 2u * b.st[1].rec[r2].w;
            const uint4 d1 = i1.rec[v1];
            const uint4 d2 = b.dual ? i2.rec[v2] : make_uint4(0, 0, 0, 0);
            const uint32_t in1 = d1.w & 0xffu, mn1 = (d1.w >> 8) & 0xffu, in2 = d2.w & 0xffu, mn2 = (d2.w >> 8) & 0xffu;
            const uint32_t klen = in1 + in2, mlen = mn1 + mn2;
            const uint32_t bad = ((d1.w | d2.w) >> 24) | (klen > 32u ? 1u : 0u);
            const uint32_t lowq = ((d1.w | d2.w) >> 16) & 1u;  // some quality byte of the index is below min_qual + 33
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
                if (!lowq) { bin = 1u + 2u * sid; wr = b.wr_pass; }                // min(index.qual) >= MIN_QUAL, Sample.py:70
                else                 { bin = 2u + 2u * sid; wr = b.wr_fail; }
            } else { bin = 0u; wr = b.wr_und; }

            // ':' index [':' molecular] '\n', raw case, assembled 4 bytes at a time
            const uint32_t suf_len = 1u + klen + (mlen ? 1u + mlen : 0u) + 1u;
            const uint8_t *s1 = i1.sfx + (size_t)v1 * i1.sfx_slot, *s2 = b.dual ? i2.sfx + (size_t)v2 * i2.sfx_slot : s1;
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
                l1 = a1 + suf_len + 5u;
                l2 = a2 + suf_len + 5u;
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
