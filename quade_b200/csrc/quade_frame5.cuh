// quade_frame5.cuh -- K1 k_frame5: single-pass frame + validate + index digest, three warp roles per CTA
// (included by quade_kernels.cu after quade_frame2/3/4.cuh: tile helpers, record parser, digest, wide look-back)
//
// clock64 probes on k_frame4 (tools/probe_f3prof.py): per tile a CTA spends 30 % of its time parsing records --
// one thread per record, so for insert reads (47 records per 16 KiB) two warps walk a dependent chain of
// shared-memory loads while seven warps idle at the barrier -- and 17-20 % waiting for the next TMA copy.
// Here the roles run concurrently on different tiles of a two-stage ring (window, bitmaps, line-end list):
//
//   scan warps (one 64-byte bitmap word per thread)      wait full[st] -> masks -> counts -> publish the line count
//                                                         -> line-end list -> listed[st]; go on with the other stage
//   look-back warp                                        go[st] -> 256-wide look-back -> k0 -> done[st]
//   parse warps (one record per thread and round)         listed[st] + done[st] -> records -> outputs at the raw
//                                                         record index -> take the NEXT ticket, TMA it into st
//
// The throughput-bound mask pass of tile j+1 overlaps the latency-bound parse of tile j; the copy of tile j+2
// is issued the moment the parse of tile j releases its stage.  A ticket is only ever taken when a stage is
// free, never across a wait for k0 (a held ticket delays a published count that every successor waits for).
#pragma once

namespace quade {
namespace {

template <int TILE, int PW> struct F5 {
    static constexpr int HALO = 2048;
    static constexpr int WIN = TILE + HALO;
    static constexpr int NW = WIN / 64;            // scan threads: one 64-byte bitmap word each
    static constexpr int NP = PW * 32;             // parse threads
    static constexpr int NT = NW + 32 + NP;        // scan warps, look-back warp, parse warps
    static constexpr int MAINW = TILE / 2048;      // scan warps whose words lie in the tile proper (the next: halo)
    static constexpr int PCAP = TILE / 8;          // listed line ends per window (else: bitmap search)
    static constexpr int DSTRIDE = WIN + 16;
    static constexpr int OFF_BM = 2 * DSTRIDE;                  // u64[2][NW]
    static constexpr int OFF_BC = OFF_BM + 2 * NW * 8;          // u64[2][NW]
    static constexpr int OFF_POS = OFF_BC + 2 * NW * 8;         // u16[2][PCAP]
    static constexpr int OFF_MISC = OFF_POS + 2 * PCAP * 2;
    static constexpr int SMEM = OFF_MISC + 256;
    // misc (u32 index): [0..23] warp counts; per stage st at 24 + 8 st: {tile, stream, tile in stream, total,
    // halo total, k0}; bytes 192..: mbarriers full[2], go[2], done[2], listed[2]
};

__device__ __forceinline__ void bar_named(int id, int n) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}

#ifdef QUADE_F3_PROF
#define F5P(cond, i) do { if (cond) { const long long t_ = clock64(); atomicAdd(&g_f3_prof[i], (unsigned long long)(t_ - tprev)); tprev = t_; } } while (0)
#else
#define F5P(cond, i) do { } while (0)
#endif

template <int TILE, int PW, int MINB>
__global__ void __launch_bounds__(F5<TILE, PW>::NT, MINB) k_frame5(const __grid_constant__ BatchDev b, uint32_t total_tiles, uint32_t backoff_ns)
{
    using G = F5<TILE, PW>;
    extern __shared__ __align__(128) uint8_t f5_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f5_smem);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f5_smem + G::OFF_MISC);
    const uint32_t bar_full = sm0 + G::OFF_MISC + 192, bar_go = bar_full + 16, bar_done = bar_full + 32, bar_listed = bar_full + 48;
    BatchState *S = b.state;

    // parse leader: next ticket into stage stg.  full[stg] always gets ONE arrival per use (plus the copy's bytes),
    // so its parity is the use count whatever the tile holds
    auto take = [&](uint32_t stg) {
        const uint32_t t = atomicAdd(&S->ticket_frame, 1u);
        uint32_t *m = s_misc + 24 + 8 * stg;
        m[0] = t;
        uint32_t bytes = 0;
        int s = 0;
        uint32_t lt = 0;
        if (t < total_tiles) {
            f4_geom<TILE>(b, t, s, lt);
            const uint32_t tbase = lt * (uint32_t)TILE, n16 = (b.st[s].n + 15u) & ~15u;
            bytes = n16 > tbase ? min((uint32_t)G::WIN, n16 - tbase) : 0u;
        }
        m[1] = (uint32_t)s;
        m[2] = lt;
        fence_proxy_async();
        mbar_arrive_tx(bar_full + 8u * stg, bytes);
        if (bytes) bulk_g2s(sm0 + stg * G::DSTRIDE, b.st[s].data + lt * (uint32_t)TILE, bytes, bar_full + 8u * stg);
    };
    if (tid == G::NW + 32) {                                    // parse leader
        for (uint32_t i = 0; i < 8; ++i) mbar_init(bar_full + 8u * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        take(0u);
        take(1u);
    }
    __syncthreads();
#ifdef QUADE_F3_PROF
    long long tprev = clock64();
#endif

    auto make_tile = [&](uint32_t stg) {
        const uint32_t *m = s_misc + 24 + 8 * stg;
        const int s = (int)m[1];
        F2Tile T;
        T.data = b.st[s].data;
        T.sdata = sm0 + stg * G::DSTRIDE;
        T.bm = reinterpret_cast<const u64 *>(f5_smem + G::OFF_BM) + stg * G::NW;
        T.bc = reinterpret_cast<const u64 *>(f5_smem + G::OFF_BC) + stg * G::NW;
        T.tbase = m[2] * (uint32_t)TILE;
        T.cap = (uint32_t)G::WIN;
        T.n = b.st[s].n;
        T.win = T.n > T.tbase ? min((uint32_t)G::WIN, T.n - T.tbase) : 0u;
        T.last = ((b.last >> s) & 1) != 0;
        return T;
    };

    // ---------------- look-back warp ----------------
    if (warp == G::NW / 32) {
        for (uint32_t j = 0;; ++j) {
            const uint32_t stg = j & 1u, par = (j >> 1) & 1u;
            mbar_wait(bar_go + 8u * stg, par);
            F5P(lane == 0, 2);
            const uint32_t *m = s_misc + 24 + 8 * stg;
            const uint32_t lt = m[2], total = m[3];
            if (m[0] >= total_tiles) return;
            const int s = (int)m[1];
            const uint32_t base = lookback_wide(b.st[s].status_frame, (int)lt, total, lane, backoff_ns);
            if (lane == 0) {
                s_misc[24 + 8 * stg + 5] = base;
                if ((u64)lt * TILE + TILE > (u64)b.st[s].n) S->n_lines[s] = base + total;   // the tile holding position n
                mbar_arrive(bar_done + 8u * stg);
            }
            F5P(lane == 0, 3);
            __syncwarp();
        }
    }

    // ---------------- parse warps ----------------
    if (warp > G::NW / 32) {
        const uint32_t ptid = (uint32_t)tid - (uint32_t)(G::NW + 32);
        for (uint32_t j = 0;; ++j) {
            const uint32_t stg = j & 1u, par = (j >> 1) & 1u;
            mbar_wait(bar_listed + 8u * stg, par);              // scanned: bitmaps, list, counts
            F5P(ptid == 0, 4);
            const uint32_t *m = s_misc + 24 + 8 * stg;
            if (m[0] >= total_tiles) return;
            mbar_wait(bar_done + 8u * stg, par);                // k0 from the look-back warp
            F5P(ptid == 0, 5);
            const int s = (int)m[1];
            const uint32_t lt = m[2], total = m[3], k0 = m[5];
            const StreamDev &st = b.st[s];
            const F2Tile T = make_tile(stg);
            const uint16_t *pos = reinterpret_cast<const uint16_t *>(f5_smem + G::OFF_POS) + stg * G::PCAP;
            const uint32_t npos = min(total + m[4], (uint32_t)G::PCAP);
            const uint32_t first = lt == 0u ? 1u : 0u;          // slot 0 of a file's first tile starts at byte 0
            const uint32_t rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;   // record index of slot 0
            const uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + first;
            const uint32_t t0i = (3u - k0) & 3u;                // list index of the first line k = 3 (mod 4)
            for (uint32_t slot = ptid; slot < nslot; slot += G::NP) {
                F2Rec R;
                if (first && slot == 0u) R = f2_parse(T, 0u, pos, 0u, npos);
                else {
                    const uint32_t ti = t0i + 4u * (slot - first);
                    if (ti < npos) R = f2_parse(T, T.tbase + pos[ti] + 1u, pos, ti + 1u, npos);
                    else {                                      // list overflow (> PCAP lines in a window): search
                        uint32_t k = k0, ppos = T.tbase, e = F2_NONE;
                        const uint32_t want = (rbase + slot) * 4u - 1u;
                        while (k <= want) {
                            e = f2_next_nl(T, ppos);
                            if (e == F2_NONE) break;
                            ppos = e + 1u;
                            ++k;
                        }
                        if (e == F2_NONE) continue;
                        R = f2_parse(T, e + 1u, pos, 0u, 0u);
                    }
                }
                if (!R.complete) continue;
                const uint32_t jr = rbase + slot;
                if (jr >= st.cap_rec) { S->err_capacity = 1u; continue; }
                if (R.nlen == 0u) atomicMax(&S->first_blank[s], ~jr);          // keeps the LOWEST blank record index
                const uint32_t inv = R.valid ? 0u : 0x80000000u;
                if (inv) atomicAdd(&S->n_invalid[s], 1u);
                st.vend[jr] = R.end;
                if (s < 2) {
                    st.rec[jr] = make_uint4(R.name_off, R.seq_off, R.qual_off, R.seq_len | inv);
                    st.vnlen[jr] = R.nlen;
                } else if (inv) {
                    st.dig[jr] = make_uint4(0u, 0u, 0u, inv);
                } else {
                    f2_digest(b, st, s, T, R, jr);
                }
            }
            bar_named(2, G::NP);                                // every parse thread is done with the stage
            F5P(ptid == 0, 6);
            if (ptid == 0) take(stg);
            F5P(ptid == 0, 7);
        }
    }

    // ---------------- scan warps ----------------
#pragma unroll 1
    for (uint32_t j = 0;; ++j) {
        const uint32_t stg = j & 1u, par = (j >> 1) & 1u;
        uint32_t *m = s_misc + 24 + 8 * stg;
        F5P(tid == 0, 1);
        mbar_wait(bar_full + 8u * stg, par);                    // ticket taken and its window copied
        F5P(tid == 0, 0);
        if (m[0] >= total_tiles) {                              // no tile left: release the other roles
            if (tid == 0) {
                mbar_arrive(bar_go + 8u * stg);
                mbar_arrive(bar_listed + 8u * stg);
            }
            return;
        }
        const int s = (int)m[1];
        const uint32_t lt = m[2];
        const F2Tile T = make_tile(stg);
        // 1. newline / candidate masks: thread t takes the 16-byte chunks t, t+NW, t+2NW, t+3NW (conflict-free
        //    128-bit loads); the 16-bit masks stored per chunk ARE the 64-bit bitmap words
        {
            const uint32_t a_bm = sm0 + G::OFF_BM + stg * G::NW * 8, a_bc = sm0 + G::OFF_BC + stg * G::NW * 8;
            const bool edge = T.win < (uint32_t)G::WIN;         // the window reaches the end of the buffer
            uint32_t vbit = F2_NONE;                            // window offset of a virtual last line end
            if (edge && T.last && T.n > 0u && T.n - T.tbase < (uint32_t)G::WIN) vbit = T.n - T.tbase;
#pragma unroll
            for (uint32_t i = 0; i < 4; ++i) {
                const uint32_t c = (uint32_t)tid + (uint32_t)G::NW * i, off = 16u * c;
                const uint4 q = lds128(T.sdata + off);
                uint32_t mk = f2_gather16(f2_nl_flags(q.x), f2_nl_flags(q.y), f2_nl_flags(q.z), f2_nl_flags(q.w));
                uint32_t cd = f2_gather16(f2_c_flags(q.x), f2_c_flags(q.y), f2_c_flags(q.z), f2_c_flags(q.w));
                if (edge) {
                    const uint32_t keep = off >= T.win ? 0u : (T.win - off >= 16u ? 0xffffu : (1u << (T.win - off)) - 1u);
                    mk &= keep;
                    cd &= keep;
                    if ((vbit >> 4) == c && f2_byte(T, T.n - 1u) != '\n') mk |= 1u << (vbit & 15u);   // no final '\n'
                }
                sts16(a_bm + 2u * c, mk);
                sts16(a_bc + 2u * c, cd);
            }
        }
        bar_named(1, G::NW);                                    // bitmaps complete
        // 2. line ends per thread (its 64 bytes = bitmap word tid), warp scans
        const u64 mine = T.bm[tid];
        const uint32_t cnt = (uint32_t)__popcll(mine);
        uint32_t inc = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += x;
        }
        if (lane == 31) s_misc[warp] = inc;
        bar_named(1, G::NW);                                    // warp counts
        uint32_t total = 0, before = 0;
#pragma unroll
        for (int w = 0; w < G::MAINW; ++w) {
            const uint32_t c = s_misc[w];
            before += w < warp ? c : 0u;
            total += c;
        }
        if (tid == 0) {                                         // publish the count, hand the tile to the look-back warp
            st_status(b.st[s].status_frame + lt, (lt ? ST_AGG : ST_INC) | total);
            m[3] = total;
            m[4] = s_misc[G::MAINW];
            mbar_arrive(bar_go + 8u * stg);
        }
        // 3. dense list of the window's line ends (tile, then halo: list order = byte order)
        {
            uint16_t *pos = reinterpret_cast<uint16_t *>(f5_smem + G::OFF_POS) + stg * G::PCAP;
            uint32_t i = before + inc - cnt;
            uint32_t lo = (uint32_t)mine, hi = (uint32_t)(mine >> 32);
            const uint32_t o0 = (uint32_t)tid * 64u;
            while (lo) {
                const uint32_t bit = (uint32_t)__ffs((int)lo) - 1u;
                lo &= lo - 1u;
                if (i < (uint32_t)G::PCAP) pos[i] = (uint16_t)(o0 + bit);
                ++i;
            }
            while (hi) {
                const uint32_t bit = (uint32_t)__ffs((int)hi) - 1u;
                hi &= hi - 1u;
                if (i < (uint32_t)G::PCAP) pos[i] = (uint16_t)(o0 + 32u + bit);
                ++i;
            }
        }
        bar_named(1, G::NW);                                    // list complete (also guards the warp counts)
        if (tid == 0) mbar_arrive(bar_listed + 8u * stg);       // hand the tile to the parse warps
    }
}

} // namespace
} // namespace quade
