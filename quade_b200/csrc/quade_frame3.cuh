// quade_frame3.cuh -- K1'' k_frame3: k_frame2's single pass with the look-back taken off the critical path
// (included by quade_kernels.cu after quade_frame2.cuh, whose tile helpers, record parser and digest it reuses)
//
// ncu on k_frame2 (profiles/r01b_*): 42 % issue-slot use, 60 % of the stall samples at block barriers: seven warps
// wait while warp 0 looks back for the tile's global line index k0, and no record can be parsed before it is known.
// k0 is needed for two things only: its residue mod 4 (which line ends close a record) and the record index the
// outputs are stored at.  So here
//   * the residue is GUESSED from the text -- the first line end i in 0..3 whose next line starts with '@' and
//     whose third next line starts with '+' (exact for well-formed FASTQ: a quality line may start with '@', but
//     the line two after it is a sequence) -- and warps 1-7 parse their records into registers with that guess
//     WHILE warp 0 looks back;
//   * after the barrier the guess is checked against the true k0: equal (always, on well-formed input) -> the
//     registers are stored at the true record index; different -> the records are parsed again with the true
//     residue.  The result never depends on the guess.
// (A three-stage variant that prefetched two tiles ahead was measured first and rejected: holding tickets for
// two iterations leaves hundreds of tiles between "count published" and "prefix known", and every look-back then
// waits for the slowest CTA: 2.7 ms against 1.38 ms.)
// Instruction diet against k_frame2: conflict-free chunk-interleaved mask pass that stores 16-bit masks (no
// 64-bit variable shifts), stream / tile geometry resolved once by the ticket thread, one barrier less in the
// block scan, 32-bit list extraction.
#pragma once

namespace quade {
namespace {

constexpr int F3T = 256;
constexpr int F3_PCAP = 2048;                  // listed line ends per window (else: bitmap search)
constexpr int F3_OFF_BM = F2_WIN + 16;                          // u64[F2_WORDS] newlines
constexpr int F3_OFF_BC = F3_OFF_BM + F2_WORDS * 8;             // u64[F2_WORDS] bytes <= 0x20
constexpr int F3_OFF_POS = F3_OFF_BC + F2_WORDS * 8;            // u16[F3_PCAP]
constexpr int F3_OFF_MISC = F3_OFF_POS + F3_PCAP * 2;
constexpr int F3_SMEM = F3_OFF_MISC + 128;
// misc (u32 index): [0..7] warp line-end counts, [8] halo count, [9] k0, [10] tile, [11] stream, [12] tile in
// stream, byte 64: mbarrier

__device__ __forceinline__ void sts16(uint32_t a, uint32_t v) { asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory"); }
__device__ __forceinline__ uint32_t lds16(uint32_t a)
{
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}

// Decoupled look-back with a 256-tile fan-in per round trip.  A new 16 KiB tile starts every ~12 ns on the whole
// GPU while one L2 round trip takes ~0.5 us, so dozens of tiles always sit between "count published" and
// "prefix known"; walking them 32 per round trip (lookback()) caps the kernel at ~32 tiles per round trip --
// which is where k_frame2 sat (1.3 ms) whatever else was optimised.  Here each lane loads 8 status words at
// once, so one round covers 256 predecessors.  Called by one full warp.
__device__ __forceinline__ uint32_t lookback_wide(u64 *status, int tile, uint32_t total, int lane, uint32_t backoff_ns)
{
    if (tile == 0) {
        if (lane == 0) st_status(status, ST_INC | total);
        return 0;
    }
#ifdef QUADE_F3_PROF
    if (backoff_ns == 0xffffffffu) return 0;                    // timing experiment only: no dependency (wrong results)
#endif
    uint32_t mine = 0;                                          // per-lane partial sum, reduced once at the end
    int look = tile - 1;
    for (;;) {
        u64 sv[8];
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            const int idx = look - 32 * g - lane;
            sv[g] = idx >= 0 ? ld_status(status + idx) : ST_INC;
        }
        bool done = false, stop = false;
        int used = 0;                                           // groups consumed without meeting an inclusive word
#pragma unroll
        for (int g = 0; g < 8; ++g) {
            if (done || stop) continue;
            const uint32_t flag = (uint32_t)(sv[g] >> 62);
            const unsigned inc = __ballot_sync(0xffffffffu, flag == 2u);
            const unsigned empty = __ballot_sync(0xffffffffu, flag == 0u);
            const unsigned upto = inc ? ((2u << (__ffs(inc) - 1)) - 1u) : 0xffffffffu;   // lanes up to the first inclusive
            if (empty & upto) { stop = true; continue; }       // a predecessor is not there yet: reload from here
            if ((1u << lane) & upto) mine += (uint32_t)(sv[g] & ST_VAL);
            if (inc) done = true; else ++used;
        }
        if (done) break;
        look -= 32 * used;
        if (stop && backoff_ns) __nanosleep(backoff_ns);        // polling the same L2 lines from ~900 CTAs slows the writers
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, o);
    if (lane == 0) st_status(status + tile, ST_INC | (u64)(mine + total));
    return mine;
}

#ifdef QUADE_F3_PROF
__device__ unsigned long long g_f3_prof[16];
#define F3P(i) do { if ((tid & 31) == 0 && (tid >> 5) < 2) { const long long t_ = clock64(); prof[i] += t_ - tprev; tprev = t_; } } while (0)
#else
#define F3P(i) do { } while (0)
#endif

__global__ void __launch_bounds__(F3T, 6) k_frame3(const __grid_constant__ BatchDev b, uint32_t total_tiles, uint32_t backoff_ns)
{
    extern __shared__ __align__(128) uint8_t f3_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f3_smem);
    const u64 *s_bm = reinterpret_cast<const u64 *>(f3_smem + F3_OFF_BM);
    const u64 *s_bc = reinterpret_cast<const u64 *>(f3_smem + F3_OFF_BC);
    uint16_t *s_pos = reinterpret_cast<uint16_t *>(f3_smem + F3_OFF_POS);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f3_smem + F3_OFF_MISC);
    const uint32_t s_bar = sm0 + F3_OFF_MISC + 64, a_pos = sm0 + F3_OFF_POS;
    BatchState *S = b.state;

    auto take = [&]() {                                  // thread 0: next ticket, geometry, TMA copy of the window
        const uint32_t t = atomicAdd(&S->ticket_frame, 1u);
        s_misc[10] = t;
        if (t >= total_tiles) return;
        int s = 0;
#pragma unroll
        for (int i = 1; i < kMaxStreams; ++i) s += (i < b.n_streams && t >= b.frame_tile_start[i]) ? 1 : 0;
        const uint32_t lt = t - b.frame_tile_start[s];
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
        take();
    }
    __syncthreads();
    uint32_t phase = 0;
#ifdef QUADE_F3_PROF
    long long prof[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#endif

    for (;;) {
        const uint32_t tile = s_misc[10];
        if (tile >= total_tiles) break;
        const int s = (int)s_misc[11];
        const uint32_t lt = s_misc[12];
        const StreamDev &st = b.st[s];
        F2Tile T;
        T.data = st.data;
        T.sdata = sm0;
        T.bm = s_bm;
        T.bc = s_bc;
        T.tbase = lt * (uint32_t)F2_TILE;
        T.cap = (uint32_t)F2_WIN;
        T.n = st.n;
        T.win = T.n > T.tbase ? min((uint32_t)F2_WIN, T.n - T.tbase) : 0u;
        T.last = ((b.last >> s) & 1) != 0;
        if (((T.n + 15u) & ~15u) > T.tbase) {                   // a copy was issued for this tile
            while (!mbar_try_wait(s_bar, phase)) {}
            phase ^= 1u;
        }
        F3P(0);

        // 1. newline / candidate masks: thread t takes the 16-byte chunks t, t+256, ... (conflict-free 128-bit
        //    loads); the 16-bit masks stored per chunk ARE the 64-bit bitmap words
        {
            const bool edge = T.win < (uint32_t)F2_WIN;         // the window reaches the end of the buffer
            uint32_t vbit = F2_NONE;                            // window offset of a virtual last line end
            if (edge && T.last && T.n > 0u && T.n - T.tbase < (uint32_t)F2_WIN) vbit = T.n - T.tbase;
#pragma unroll
            for (uint32_t i = 0; i < 5; ++i) {
                const uint32_t c = (uint32_t)tid + (uint32_t)F3T * i, off = 16u * c;
                if (i == 4 && tid >= F2_HALO / 16) break;       // halo: 128 chunks, warps 0-3
                const uint4 q = lds128(sm0 + off);
                uint32_t m = f2_gather16(f2_nl_flags(q.x), f2_nl_flags(q.y), f2_nl_flags(q.z), f2_nl_flags(q.w));
                uint32_t cd = f2_gather16(f2_c_flags(q.x), f2_c_flags(q.y), f2_c_flags(q.z), f2_c_flags(q.w));
                if (edge) {
                    const uint32_t keep = off >= T.win ? 0u : (T.win - off >= 16u ? 0xffffu : (1u << (T.win - off)) - 1u);
                    m &= keep;
                    cd &= keep;
                    if ((vbit >> 4) == c && f2_byte(T, T.n - 1u) != '\n') m |= 1u << (vbit & 15u);   // no final '\n'
                }
                sts16(sm0 + F3_OFF_BM + 2u * c, m);
                sts16(sm0 + F3_OFF_BC + 2u * c, cd);
            }
        }
        __syncthreads();                                        // B1: bitmaps complete
        F3P(1);

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
            hm = s_bm[F3T + lane];
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
        F3P(2);
        const uint4 wa = *reinterpret_cast<const uint4 *>(s_misc), wb = *reinterpret_cast<const uint4 *>(s_misc + 4);
        const uint32_t c1 = wa.x, c2 = c1 + wa.y, c3 = c2 + wa.z, c4 = c3 + wa.w, c5 = c4 + wb.x, c6 = c5 + wb.y,
                       c7 = c6 + wb.z, total = c7 + wb.w;
        if (tid == 0) st_status(st.status_frame + lt, (lt ? ST_AGG : ST_INC) | total);   // publish before anything else
        uint32_t before = 0u;
        before = warp == 1 ? c1 : before;
        before = warp == 2 ? c2 : before;
        before = warp == 3 ? c3 : before;
        before = warp == 4 ? c4 : before;
        before = warp == 5 ? c5 : before;
        before = warp == 6 ? c6 : before;
        before = warp == 7 ? c7 : before;
        const uint32_t npos = min(total + s_misc[8], (uint32_t)F3_PCAP);

        // 3. dense list of the window's line ends (tile, then halo: list order = byte order)
        {
            uint32_t i = before + inc - cnt;
            uint32_t lo = (uint32_t)mine, hi = (uint32_t)(mine >> 32);
            const uint32_t o0 = (uint32_t)tid * 64u;
            while (lo) {
                const uint32_t bit = (uint32_t)__ffs((int)lo) - 1u;
                lo &= lo - 1u;
                if (i < (uint32_t)F3_PCAP) s_pos[i] = (uint16_t)(o0 + bit);
                ++i;
            }
            while (hi) {
                const uint32_t bit = (uint32_t)__ffs((int)hi) - 1u;
                hi &= hi - 1u;
                if (i < (uint32_t)F3_PCAP) s_pos[i] = (uint16_t)(o0 + 32u + bit);
                ++i;
            }
            if (warp == 1) {
                i = total + hinc - hcnt;
                const uint32_t h0 = (uint32_t)(F3T + lane) * 64u;
                while (hm) {
                    const uint32_t bit = (uint32_t)__ffsll((long long)hm) - 1u;
                    hm &= hm - 1ull;
                    if (i < (uint32_t)F3_PCAP) s_pos[i] = (uint16_t)(h0 + bit);
                    ++i;
                }
            }
        }
        __syncthreads();                                        // B3: list complete
        F3P(3);

        // 4. guess k0 mod 4 from the text (header '@', two lines on '+'); the file's first tile knows k0 = 0
        const uint32_t first = lt == 0u ? 1u : 0u;              // slot 0 of a file's first tile starts at byte 0
        uint32_t t0g = 3u;                                      // guessed list index of the first record-closing line
        if (lt != 0u) {
            bool ok = false;
            if (lane < 4 && (uint32_t)lane + 2u < npos)
                ok = lds8(sm0 + lds16(a_pos + 2u * lane) + 1u) == '@' && lds8(sm0 + lds16(a_pos + 2u * (lane + 2)) + 1u) == '+';
            const unsigned g = __ballot_sync(0xffffffffu, ok);
            t0g = g ? (uint32_t)__ffs((int)g) - 1u : 0u;
        }
        const uint32_t nslot_g = first + (total > t0g ? (total - t0g + 3u) >> 2 : 0u);

        // record of slot i under residue t0 (true = t0 is exact and k0 known: the list-overflow search may run)
        auto parse_slot = [&](uint32_t slot, uint32_t t0, bool exact, uint32_t k0, F2Rec &R) -> bool {
            if (first && slot == 0u) { R = f2_parse(T, 0u, s_pos, 0u, npos); return true; }
            const uint32_t ti = t0 + 4u * (slot - first);
            if (ti < npos) { R = f2_parse(T, T.tbase + s_pos[ti] + 1u, s_pos, ti + 1u, npos); return true; }
            R.complete = R.valid = false;
            R.name_off = R.nlen = R.seq_off = R.qual_off = R.seq_len = R.end = 0u;
            if (!exact) return false;
            // list overflow (> F3_PCAP lines in 18 KiB): find the terminator by bitmap search
            const uint32_t rb = lt == 0u ? 0u : (k0 >> 2) + 1u, want = (rb + slot) * 4u - 1u;
            uint32_t k = k0, ppos = T.tbase, e = F2_NONE;
            while (k <= want) {
                e = f2_next_nl(T, ppos);
                if (e == F2_NONE) break;
                ppos = e + 1u;
                ++k;
            }
            if (e != F2_NONE) R = f2_parse(T, e + 1u, s_pos, 0u, 0u);
            return true;
        };
        // outputs at the RAW record index j.  Records with len(seq) != len(qual) are flagged (bit 31) and counted;
        // only when a file has some does k_fix_compact squeeze them out afterwards (FastqReader.py:48-50).
        auto store = [&](uint32_t j, const F2Rec &R) {
            if (!R.complete) return;
            if (j >= st.cap_rec) { S->err_capacity = 1u; return; }
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
        };

        // 5. warp 0 looks back for k0 while warps 1-7 parse one record each under the guess
        F2Rec R;
        bool have = false;
        const uint32_t myslot = (uint32_t)tid - 32u;            // warps 1-7: slots 0..223
        if (warp == 0) {
            const uint32_t base = lookback_wide(st.status_frame, (int)lt, total, lane, backoff_ns);   // the count was published above
            if (lane == 0) {
                s_misc[9] = base;
                if ((u64)T.tbase + F2_TILE > (u64)T.n) S->n_lines[s] = base + total;
            }
        } else if (myslot < nslot_g) {
            have = parse_slot(myslot, t0g, false, 0u, R);
        }
        F3P(4);                                                 // warp 0: look-back, warp 1: parse
        __syncthreads();                                        // B4: k0 known
        F3P(5);
        const uint32_t k0 = s_misc[9];
        const uint32_t t0i = (3u - k0) & 3u;                    // true list index of the first line k = 3 (mod 4)
        const uint32_t rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;  // record index of slot 0
        const uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + first;
        if (t0i == t0g) {
            if (warp != 0 && myslot < nslot) {
                if (!have) parse_slot(myslot, t0i, true, k0, R);
                store(rbase + myslot, R);
            }
            for (uint32_t slot = 224u + (uint32_t)tid; slot < nslot; slot += F3T) {   // more than 224 records in 16 KiB
                parse_slot(slot, t0i, true, k0, R);
                store(rbase + slot, R);
            }
        } else {                                                // wrong guess (malformed text): parse again
            for (uint32_t slot = tid; slot < nslot; slot += F3T) {
                parse_slot(slot, t0i, true, k0, R);
                store(rbase + slot, R);
            }
        }
        __syncthreads();                                        // B5: window, bitmaps and list are free
        F3P(6);
        if (tid == 0) take();
        __syncthreads();
        F3P(7);
    }
#ifdef QUADE_F3_PROF
    if ((tid & 31) == 0 && (tid >> 5) < 2)
        for (int i = 0; i < 8; ++i) atomicAdd(&g_f3_prof[(tid >> 5) * 8 + i], (unsigned long long)prof[i]);
#endif
}

} // namespace
} // namespace quade
