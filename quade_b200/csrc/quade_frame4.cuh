// quade_frame4.cuh -- K1 k_frame4: single-pass frame + validate + index digest with the look-back DECOUPLED
// (included by quade_kernels.cu after quade_frame2/3.cuh, whose tile helpers, record parser and digest it reuses)
//
// Measured on k_frame2 / k_frame3 with clock64 probes (tools/probe_f3prof.py): the look-back for a tile's global
// line index k0 takes ~50 % of a CTA's time per tile -- not the L2 round trips, but waiting for the slowest of
// the hundreds of predecessor tiles in flight to publish its line count (running maximum of a jittery latency).
// A CTA cannot avoid that wait, but it can do something else meanwhile.  Here
//   * a dedicated look-back warp (the last warp of the CTA) is handed (tile, count) through an mbarrier, resolves
//     k0 with a 256-wide look-back and hands it back through a second mbarrier;
//   * the worker warps never wait for it in the same iteration: tile j is scanned (masks, counts, published
//     count, line-end list) in iteration j and its records are parsed in iteration j+1, after tile j+1 has been
//     scanned -- window, bitmaps and list are double-buffered -- so the look-back has a whole iteration of slack;
//   * the ticket -> publish latency stays short (ticket, TMA copy, masks, counts), which keeps the jitter that
//     every successor has to wait for small: the three-stage prefetching variant measured earlier held tickets
//     for two iterations and ran at 2.7 ms against 1.38 ms.
// Worker barriers are named barriers over the worker warps only (bar.sync 1, NW).
#pragma once

namespace quade {
namespace {

constexpr uint32_t F4_EXIT = 0xffffffffu;

template <int TILE> struct F4 {
    static constexpr int HALO = 2048;
    static constexpr int WIN = TILE + HALO;
    static constexpr int NW = WIN / 64;            // worker threads: one 64-byte bitmap word each
    static constexpr int NT = NW + 32;             // + the look-back warp
    static constexpr int MAINW = TILE / 2048;      // warps whose words lie in the tile proper (the next one: halo)
    static constexpr int PCAP = TILE / 8;          // listed line ends per window (else: bitmap search)
    static constexpr int DSTRIDE = WIN + 16;
    static constexpr int OFF_BM = 2 * DSTRIDE;                  // u64[2][NW]
    static constexpr int OFF_BC = OFF_BM + 2 * NW * 8;          // u64[2][NW]
    static constexpr int OFF_POS = OFF_BC + 2 * NW * 8;         // u16[2][PCAP]
    static constexpr int OFF_MISC = OFF_POS + 2 * PCAP * 2;
    static constexpr int SMEM = OFF_MISC + 256;
    static constexpr int MINB = TILE >= 32768 ? 2 : (TILE >= 16384 ? 4 : 7);   // CTAs per SM the register budget is cut for
    // misc (u32 index): [0..23] warp counts; per stage st = 0/1 at 24 + 8 st: {tile, stream, tile in stream, total,
    // halo total, k0}; bytes 192..: mbarriers full[2], go[2], done[2]
};

__device__ __forceinline__ void bar_workers(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    while (!mbar_try_wait(bar, parity)) {}
}

// tile -> stream and tile inside the stream, for tiles of TILE bytes (a stream of n bytes has n / TILE + 1 tiles)
template <int TILE> __device__ __forceinline__ void f4_geom(const BatchDev &b, uint32_t tile, int &s, uint32_t &lt)
{
    s = 0;
    lt = tile;
#pragma unroll
    for (int i = 0; i < kMaxStreams - 1; ++i) {
        const uint32_t cnt = (i < b.n_streams && b.st[i].n) ? b.st[i].n / (uint32_t)TILE + 1u : 0u;
        if (s == i && i + 1 < b.n_streams && lt >= cnt) {
            lt -= cnt;
            s = i + 1;
        }
    }
}

template <int TILE>
__global__ void __launch_bounds__(F4<TILE>::NT, F4<TILE>::MINB) k_frame4(const __grid_constant__ BatchDev b, uint32_t total_tiles, uint32_t backoff_ns)
{
    using G = F4<TILE>;
    extern __shared__ __align__(128) uint8_t f4_smem[];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t sm0 = smem_u32(f4_smem);
    uint32_t *s_misc = reinterpret_cast<uint32_t *>(f4_smem + G::OFF_MISC);
    const uint32_t bar_full = sm0 + G::OFF_MISC + 192, bar_go = bar_full + 16, bar_done = bar_full + 32;
    BatchState *S = b.state;

    auto issue = [&](uint32_t t, uint32_t stg) {         // thread 0: geometry of ticket t, TMA copy of its window
        uint32_t *m = s_misc + 24 + 8 * stg;
        m[0] = t;
        if (t >= total_tiles) return;
        int s;
        uint32_t lt;
        f4_geom<TILE>(b, t, s, lt);
        m[1] = (uint32_t)s;
        m[2] = lt;
        const uint32_t tbase = lt * (uint32_t)TILE, n16 = (b.st[s].n + 15u) & ~15u;
        const uint32_t bytes = n16 > tbase ? min((uint32_t)G::WIN, n16 - tbase) : 0u;
        if (bytes) {
            fence_proxy_async();
            mbar_expect_tx(bar_full + 8u * stg, bytes);
            bulk_g2s(sm0 + stg * G::DSTRIDE, b.st[s].data + tbase, bytes, bar_full + 8u * stg);
        }
    };
    if (tid == 0) {
        for (uint32_t i = 0; i < 6; ++i) mbar_init(bar_full + 8u * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        issue(atomicAdd(&S->ticket_frame, 1u), 0u);
    }
    __syncthreads();

    // ---------------- look-back warp ----------------
    if (warp == G::NW / 32) {
        for (uint32_t j = 0;; ++j) {
            const uint32_t stg = j & 1u, par = (j >> 1) & 1u;
            mbar_wait(bar_go + 8u * stg, par);
            const uint32_t *m = s_misc + 24 + 8 * stg;
            const uint32_t lt = m[2], total = m[3];
            if (lt == F4_EXIT) return;
            const int s = (int)m[1];
            const uint32_t base = lookback_wide(b.st[s].status_frame, (int)lt, total, lane, backoff_ns);
            if (lane == 0) {
                s_misc[24 + 8 * stg + 5] = base;
                if ((u64)lt * TILE + TILE > (u64)b.st[s].n) S->n_lines[s] = base + total;   // the tile holding position n
                mbar_arrive(bar_done + 8u * stg);
            }
            __syncwarp();
        }
    }

    // ---------------- worker warps ----------------
    auto make_tile = [&](uint32_t stg) {
        const uint32_t *m = s_misc + 24 + 8 * stg;
        const int s = (int)m[1];
        F2Tile T;
        T.data = b.st[s].data;
        T.sdata = sm0 + stg * G::DSTRIDE;
        T.bm = reinterpret_cast<const u64 *>(f4_smem + G::OFF_BM) + stg * G::NW;
        T.bc = reinterpret_cast<const u64 *>(f4_smem + G::OFF_BC) + stg * G::NW;
        T.tbase = m[2] * (uint32_t)TILE;
        T.cap = (uint32_t)G::WIN;
        T.n = b.st[s].n;
        T.win = T.n > T.tbase ? min((uint32_t)G::WIN, T.n - T.tbase) : 0u;
        T.last = ((b.last >> s) & 1) != 0;
        return T;
    };
    // records of the tile in stage stg (scanned in the previous iteration), outputs at the RAW record index
    auto parse = [&](uint32_t stg, uint32_t par) {
        mbar_wait(bar_done + 8u * stg, par);                    // k0 from the look-back warp
        const uint32_t *m = s_misc + 24 + 8 * stg;
        const int s = (int)m[1];
        const uint32_t lt = m[2], total = m[3], k0 = m[5];
        const StreamDev &st = b.st[s];
        const F2Tile T = make_tile(stg);
        const uint16_t *pos = reinterpret_cast<const uint16_t *>(f4_smem + G::OFF_POS) + stg * G::PCAP;
        const uint32_t npos = min(total + m[4], (uint32_t)G::PCAP);
        const uint32_t first = lt == 0u ? 1u : 0u;              // slot 0 of a file's first tile starts at byte 0
        const uint32_t rbase = lt == 0u ? 0u : (k0 >> 2) + 1u;  // record index of slot 0
        const uint32_t nslot = ((k0 + total) >> 2) - (k0 >> 2) + first;
        const uint32_t t0i = (3u - k0) & 3u;                    // list index of the first line k = 3 (mod 4)
        for (uint32_t slot = tid; slot < nslot; slot += G::NW) {
            F2Rec R;
            if (first && slot == 0u) R = f2_parse(T, 0u, pos, 0u, npos);
            else {
                const uint32_t ti = t0i + 4u * (slot - first);
                if (ti < npos) R = f2_parse(T, T.tbase + pos[ti] + 1u, pos, ti + 1u, npos);
                else {                                          // list overflow (> PCAP lines in a window): search
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
            if (R.nlen == 0u) atomicMax(&S->first_blank[s], ~jr);              // keeps the LOWEST blank record index
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
    };

    uint32_t full_phase = 0;
    uint32_t j = 0;
#ifdef QUADE_F3_PROF
    long long prof[8] = {0, 0, 0, 0, 0, 0, 0, 0}, tprev = clock64();
#endif
#pragma unroll 1
    for (;; ++j) {
        const uint32_t stg = j & 1u;
        uint32_t *m = s_misc + 24 + 8 * stg;
        if (m[0] >= total_tiles) break;
        const int s = (int)m[1];
        const uint32_t lt = m[2];
        const F2Tile T = make_tile(stg);
        if (((T.n + 15u) & ~15u) > T.tbase) {                   // a copy was issued for this tile
            mbar_wait(bar_full + 8u * stg, (full_phase >> stg) & 1u);
            full_phase ^= 1u << stg;
        }
        F3P(0);
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
        bar_workers(G::NW);                                     // bitmaps complete
        F3P(1);
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
        bar_workers(G::NW);                                     // warp counts
        F3P(2);
        uint32_t total = 0, before = 0;
#pragma unroll
        for (int w = 0; w < G::MAINW; ++w) {
            const uint32_t c = s_misc[w];
            before += w < warp ? c : 0u;
            total += c;
        }
        uint32_t tk = 0;
        if (tid == 0) {                                         // publish the count, hand the tile to the look-back warp
            st_status(b.st[s].status_frame + lt, (lt ? ST_AGG : ST_INC) | total);
            m[3] = total;
            m[4] = s_misc[G::MAINW];
            mbar_arrive(bar_go + 8u * stg);
        }
        // 3. dense list of the window's line ends (tile, then halo: list order = byte order)
        {
            uint16_t *pos = reinterpret_cast<uint16_t *>(f4_smem + G::OFF_POS) + stg * G::PCAP;
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
        F3P(3);
        // 4. the PREVIOUS tile's records: its look-back has had this whole iteration
        if (j > 0) {
            mbar_wait(bar_done + 8u * (stg ^ 1u), ((j - 1u) >> 1) & 1u);
            F3P(4);
        }
        // next ticket: taken only now, AFTER the wait for k0 -- a ticket held across a dependency wait delays this
        // CTA's next published count, which every successor's look-back then has to wait for (positive feedback)
        if (tid == 0) tk = atomicAdd(&S->ticket_frame, 1u);     // consumed after the parse: latency hidden
        if (j > 0) {
            parse(stg ^ 1u, ((j - 1u) >> 1) & 1u);
        }
        F3P(5);
        bar_workers(G::NW);                                     // list complete; the other stage is free
        F3P(6);
        if (tid == 0) issue(tk, stg ^ 1u);
        bar_workers(G::NW);                                     // next tile's geometry visible
        F3P(7);
    }
    if (j > 0) parse((j & 1u) ^ 1u, ((j - 1u) >> 1) & 1u);
#ifdef QUADE_F3_PROF
    if ((tid & 31) == 0 && (tid >> 5) < 2)
        for (int i = 0; i < 8; ++i) atomicAdd(&g_f3_prof[(tid >> 5) * 8 + i], (unsigned long long)prof[i]);
#endif
    if (tid == 0) {                                             // release the look-back warp
        s_misc[24 + 8 * (j & 1u) + 2] = F4_EXIT;
        mbar_arrive(bar_go + 8u * (j & 1u));
    }
}

} // namespace
} // namespace quade
