// quade_kernels2.cuh -- K2 validate, K2b pairing, K3 classify (included by quade_kernels.cu)
#pragma once

namespace quade {
namespace {

// ------------------------------------------------------------------------------------------------------------
// K2: validate.  Thread t of a tile owns records 4t..4t+3 (stable order).  valid <=> len(seq)==len(qual)
// (FastqSeq.py:55); invalid records are skipped in THIS stream only (FastqReader.py:48-50), so the pairing
// ordinal of a record is its rank among the valid records of its stream.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_validate(const __grid_constant__ BatchDev b)
{
    __shared__ uint32_t s_warp[9];
    __shared__ uint32_t s_tile, s_base;
    const int s = blockIdx.y, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const StreamDev &st = b.st[s];
    uint32_t n_rec = b.state->n_lines[s] >> 2;
    if (n_rec > st.cap_rec) n_rec = st.cap_rec; // k_frame raised err_capacity already
    const uint32_t n_tiles = (n_rec + kValidateTile - 1) / kValidateTile;
    const uint4 *nl4 = reinterpret_cast<const uint4 *>(st.nl);

    for (;;) {
        if (tid == 0) s_tile = atomicAdd(&b.state->ticket_validate[s], 1u);
        __syncthreads();
        const uint32_t tile = s_tile;
        if (tile >= n_tiles) return;
        const uint32_t j0 = tile * kValidateTile + (uint32_t)tid * 4u;
        uint4 rec[4];
        uint32_t nlen[4];
        uint32_t vmask = 0, blank = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const uint32_t j = j0 + i;
            if (j < n_rec) {
                const uint4 l = nl4[j];
                const uint32_t seq_off = l.x + 1u, qual_off = l.z + 1u;
                const uint32_t sl = st.e1[j] - seq_off, ql = st.e3[j] - qual_off;
                nlen[i] = st.nlen[j];
                rec[i] = make_uint4(st.noff[j], seq_off, qual_off, sl);
                if (sl == ql) vmask |= 1u << i;
                if (nlen[i] == 0u) blank = max(blank, ~j); // keeps the LOWEST blank record index
            }
        }
        uint32_t total;
        const uint32_t excl = block_scan256((uint32_t)__popc(vmask), s_warp, &total);
        if (warp == 0) {
            const uint32_t base = lookback(st.status_validate, (int)tile, total, lane);
            if (lane == 0) {
                s_base = base;
                if (tile == n_tiles - 1) b.state->n_valid[s] = base + total;
            }
        }
        __syncthreads();
        uint32_t v = s_base + excl;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (vmask & (1u << i)) {
                st.rec[v] = rec[i];
                st.vnlen[v] = nlen[i];
                st.vrec[v] = j0 + i;
                ++v;
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) blank = max(blank, __shfl_xor_sync(0xffffffffu, blank, o));
        if (lane == 0 && blank) atomicMax(&b.state->first_blank[s], blank);
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// K2b: pairing.  The reference pulls one valid record from R1, R2, I1[, I2] per iteration and stops at the
// first exhausted stream (Quade.py:210-224): n_pairs = min over streams of the valid count.  Also decides
// whether a blank header was parsed (IndexError in the reference, FastqReader.py:42).
// ------------------------------------------------------------------------------------------------------------
__global__ void k_pairing(const __grid_constant__ BatchDev b)
{
    if (threadIdx.x != 0) return;
    BatchState *S = b.state;
    uint32_t np = 0xffffffffu;
    for (int s = 0; s < b.n_streams; ++s) np = min(np, S->n_valid[s]);
    int sx = 0;
    while (S->n_valid[sx] != np) ++sx;  // first stream (in pull order) that runs dry in iteration np
    for (int s = 0; s < b.n_streams; ++s) {
        const StreamDev &st = b.st[s];
        uint32_t n_rec = S->n_lines[s] >> 2;
        if (n_rec > st.cap_rec) n_rec = st.cap_rec;
        if (np > 0) {
            const uint32_t j = st.vrec[np - 1];
            const uint32_t end = st.nl[4u * j + 3u];
            S->consumed[s] = end < st.n ? end + 1u : st.n;
            S->n_skipped[s] = j + 1u - np;
        }
        if (S->first_blank[s]) {
            const uint32_t jb = ~S->first_blank[s];
            bool parsed;
            if (s < sx) parsed = jb <= st.vrec[np];             // pulled once more in the failing iteration
            else if (s == sx) parsed = true;                     // parsed to the end looking for a valid record
            else parsed = np > 0 && jb <= st.vrec[np - 1];
            if (parsed) S->err_blank = 1u;
        }
    }
    if (np > b.cap_pairs) {
        S->err_capacity = 1u;
        np = 0;
    }
    if (S->err_capacity) np = 0;
    // k_compact3 stages the R1 / R2 spans of a tile of tp pairs: size tp from the mean record size
    uint32_t tp = kCompactTilePairs;
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
// K3: classify.  One CTA per block of pairs_per_block pairs, 8 lanes per pair.
//   index = I1[i1s:i1e] + I2[i2s:i2e], molecular = I1[m1s:m1e] + I2[m2s:m2e]   (Quade.py:217-218, python
//   slice clamping FastqSeq.py:91-93); key = upper(index) packed 2 bits/base + N mask + length; probe of the
//   perfect-hash table with full-key verify (Sample.py:65-67); sub-warp min of phred over the index bytes
//   only (Sample.py:70); class -> bin; suffix bytes; output record lengths; per-block per-bin byte histogram.
// ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void clamp_span(const int sp[2], uint32_t L, uint32_t *a, uint32_t *n)
{
    uint32_t s = (uint32_t)sp[0], e = (uint32_t)sp[1];
    s = s < L ? s : L;
    e = e < L ? e : L;
    *a = s;
    *n = e > s ? e - s : 0u;
}

constexpr int CT = 1024;   // threads per classify CTA: 128 pairs in flight, full occupancy with 2 CTAs per SM

__global__ void __launch_bounds__(CT) k_classify(const __grid_constant__ BatchDev b, int table_in_smem)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const uint32_t n_pairs = b.state->n_pairs, n_blocks = b.state->n_blocks;
    if (blockIdx.x >= n_blocks) return;
    const int tid = threadIdx.x, nb = b.n_bins;
    uint32_t *h_b1 = reinterpret_cast<uint32_t *>(smem), *h_b2 = h_b1 + nb, *h_cnt = h_b2 + nb;
    const uint32_t n_slots = b.table.n_slots_mask + 1u, n_buckets = b.table.n_buckets_mask + 1u;
    const uint4 *entries = b.table.entries;
    const uint16_t *disp = b.table.disp;
    for (int i = tid; i < nb; i += CT) h_cnt[i] = 0u;
    if (table_in_smem) {
        uint4 *se = reinterpret_cast<uint4 *>(smem + (((size_t)3 * nb * 4 + 15) & ~(size_t)15));
        uint16_t *sd = reinterpret_cast<uint16_t *>(se + n_slots);
        for (uint32_t i = tid; i < n_slots; i += CT) se[i] = entries[i];
        for (uint32_t i = tid; i < n_buckets; i += CT) sd[i] = disp[i];
        entries = se;
        disp = sd;
    }

    const int gl = tid & 7, lane = tid & 31;
    const unsigned gmask = 0xffu << (lane & 24);
    const uint8_t *d1 = b.st[2].data, *d2 = b.dual ? b.st[3].data : nullptr;

    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
    for (int i = tid; i < 2 * nb; i += CT) h_b1[i] = 0u;
    __syncthreads();
    const uint32_t p0 = blk * b.pairs_per_block;
    uint32_t end = p0 + b.pairs_per_block;
    if (end > n_pairs) end = n_pairs;

    for (uint32_t p = p0 + (tid >> 3); p < end; p += CT / 8) {
        const uint4 r1 = b.st[2].rec[p];
        uint4 r2 = make_uint4(0, 0, 0, 0);
        if (b.dual) r2 = b.st[3].rec[p];
        uint32_t ia1, in1, ia2, in2, ma1, mn1, ma2, mn2;
        clamp_span(b.span[0], r1.w, &ia1, &in1);
        clamp_span(b.span[1], r2.w, &ia2, &in2);
        clamp_span(b.span[2], r1.w, &ma1, &mn1);
        clamp_span(b.span[3], r2.w, &ma2, &mn2);
        const uint32_t klen = in1 + in2, mlen = mn1 + mn2;

        // fused key + min phred: lane gl takes bases gl, gl+8, ...
        uint64_t key = 0;
        uint32_t nm = 0, bad = klen > 32u ? 1u : 0u;
        int minq = 0x7fffffff;
        for (uint32_t t = gl; t < klen; t += 8) {
            uint32_t c, q;
            if (t < in1) {
                c = d1[r1.y + ia1 + t];
                q = d1[r1.z + ia1 + t];
            } else {
                c = d2[r2.y + ia2 + (t - in1)];
                q = d2[r2.z + ia2 + (t - in1)];
            }
            minq = min(minq, (int)q - 33);                  // FastqSeq.py:44
            if (c - 'a' < 26u) c -= 32u;                     // index.seq.upper(), Sample.py:65
            const bool acgt = c == 'A' || c == 'C' || c == 'G' || c == 'T';
            const bool isn = c == 'N';
            if (!acgt && !isn) bad = 1u;
            if (t < 32u) {
                if (acgt) key |= (uint64_t)((c >> 1) & 3u) << (2u * t);
                if (isn) nm |= 1u << t;
            }
        }
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            key |= __shfl_xor_sync(gmask, key, o);
            nm |= __shfl_xor_sync(gmask, nm, o);
            bad |= __shfl_xor_sync(gmask, bad, o);
            minq = min(minq, __shfl_xor_sync(gmask, minq, o));
        }
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
            if (minq >= b.min_qual) { bin = 1u + 2u * sid; wr = b.wr_pass; }   // Sample.py:70-75
            else                    { bin = 2u + 2u * sid; wr = b.wr_fail; }   // Sample.py:78-83
        } else { bin = 0u; wr = b.wr_und; }                                       // Sample.py:86-91

        // ':' index [':' molecular] '\n' with the bytes as they appear in the file (FastqWriter.py:61-66)
        const uint32_t suf_len = 1u + klen + (mlen ? 1u + mlen : 0u) + 1u;
        uint8_t *suf = b.suffix + (size_t)p * b.suf_cap;
        for (uint32_t t = gl; t < suf_len; t += 8) {
            uint32_t c, x = t;
            if (x == 0u) c = ':';
            else if (x == suf_len - 1u) c = '\n';
            else {
                x -= 1u;
                if (x < in1) c = d1[r1.y + ia1 + x];
                else if (x < klen) c = d2[r2.y + ia2 + (x - in1)];
                else if (x == klen) c = ':';
                else {
                    x -= klen + 1u;
                    c = x < mn1 ? d1[r1.y + ma1 + x] : d2[r2.y + ma2 + (x - mn1)];
                }
            }
            suf[t] = (uint8_t)c;
        }
        if (gl == 0) {
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
    }
    __syncthreads();
    uint32_t *hg = b.hist + (size_t)blk * 2u * nb;
    for (int i = tid; i < 2 * nb; i += CT) hg[i] = h_b1[i];
    __syncthreads();
    } // blocks
    for (int i = tid; i < nb; i += CT) {
        const uint32_t c = h_cnt[i];
        if (c) atomicAdd(reinterpret_cast<u64 *>(b.counters) + i, (u64)c);
    }
}

} // namespace
} // namespace quade
