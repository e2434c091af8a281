// quade_bins.cuh -- k_binscan_* and k_offsets (included by quade_kernels.cu)
#pragma once

#ifndef QUADE_OFFSETS_BALLOT
#define QUADE_OFFSETS_BALLOT 1      /* peer groups from ballots over the bin's bits instead of MATCH.ANY (0.039 -> 0.033 ms) */
#endif

namespace quade {
namespace {

// ------------------------------------------------------------------------------------------------------------
// K4: per bin, exclusive scan of the block histograms over blocks (in place), then over bins.
// Output order inside a bin = pair order (FastqWriter.py:68-69), bins are laid out one after the other.
// ------------------------------------------------------------------------------------------------------------
// 32 columns (bin x read) per CTA, threadIdx.y = one of 32 contiguous slices of the block range: slice sums,
// a shared-memory scan over the slices, then each slice rewrites its rows as exclusive prefixes.
__global__ void __launch_bounds__(1024) k_binscan_blocks(const __grid_constant__ BatchDev b)
{
    __shared__ u64 s_part[32][33];
    const uint32_t width = 2u * (uint32_t)b.n_bins, col = blockIdx.x * 32u + threadIdx.x, sl = threadIdx.y;
    const uint32_t nblk = b.state->n_blocks, per = (nblk + 31u) / 32u;
    const uint32_t k0 = min(sl * per, nblk), k1 = min(k0 + per, nblk);
    const bool on = col < width;
    uint32_t *h = b.hist + col;
    u64 sum = 0;
    if (on) {
        for (uint32_t kb = k0; kb < k1; kb += 8u) {
            uint32_t v[8];
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) v[q] = kb + q < k1 ? h[(size_t)(kb + q) * width] : 0u;
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) sum += v[q];
        }
    }
    s_part[sl][threadIdx.x] = sum;
    __syncthreads();
    u64 run = 0;
    for (uint32_t i = 0; i < sl; ++i) run += s_part[i][threadIdx.x];
    if (on) {
        for (uint32_t kb = k0; kb < k1; kb += 8u) {           // 8 rows loaded at once (independent), then rewritten
            uint32_t v[8];
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) v[q] = kb + q < k1 ? h[(size_t)(kb + q) * width] : 0u;
#pragma unroll
            for (uint32_t q = 0; q < 8; ++q) {
                if (kb + q < k1) h[(size_t)(kb + q) * width] = (uint32_t)run;
                run += v[q];
            }
        }
        if (sl == 31) b.bin_total[col] = run;   // the last slice ends on the column total
    }
}

__global__ void __launch_bounds__(1024) k_binscan_bins(const __grid_constant__ BatchDev b)
{
    __shared__ u64 s_warp[33];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = b.n_bins;
    const int per = (nb + 1023) / 1024;
    for (int c = 0; c < 2; ++c) {
        const u64 *tot = b.bin_total + (size_t)c * nb;
        u64 *off = b.bin_off + (size_t)c * (nb + 1);
        const int lo = min(tid * per, nb), hi = min(lo + per, nb);
        u64 sum = 0;
        for (int i = lo; i < hi; ++i) sum += tot[i];
        u64 inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            u64 t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            u64 w = s_warp[lane], wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                u64 t = __shfl_up_sync(0xffffffffu, wi, o);
                if (lane >= o) wi += t;
            }
            s_warp[lane] = wi - w;
            if (lane == 31) s_warp[32] = wi;
        }
        __syncthreads();
        u64 run = s_warp[warp] + inc - sum;
        for (int i = lo; i < hi; ++i) {
            off[i] = run;
            run += tot[i];
        }
        if (tid == 0) {
            const u64 total = s_warp[32];
            off[nb] = total;
            b.state->out_total[c] = total;
            if (total > b.out_cap[c] || total >= (1ull << 32)) b.state->err_capacity = 1u;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------------------------
// K4b: offsets.  One CTA per block of pairs; the per-bin running offsets live in shared memory.  Inside a
// 256-pair chunk every warp ranks its 32 pairs with shuffles (stable: lower lane first), then the 8 warps
// apply their ranks to the running offsets one after the other.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_offsets(const __grid_constant__ BatchDev b)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const BatchState *S = b.state;
    if (blockIdx.x >= S->n_blocks || S->err_capacity) return;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nb = b.n_bins;
    uint32_t *run1 = reinterpret_cast<uint32_t *>(smem), *run2 = run1 + nb;
    const uint32_t *hg = b.hist + (size_t)blockIdx.x * 2u * nb;
    for (int i = tid; i < nb; i += 256) {
        run1[i] = (uint32_t)b.bin_off[i] + hg[i];
        run2[i] = (uint32_t)b.bin_off[nb + 1 + i] + hg[nb + i];
    }
    __syncthreads();
    const uint32_t bin_bits = 32u - (uint32_t)__clz(max(nb, 2) - 1);
    const uint32_t n_pairs = S->n_pairs, p0 = blockIdx.x * b.pairs_per_block;
    uint32_t end = p0 + b.pairs_per_block;
    if (end > n_pairs) end = n_pairs;
    // the inputs of the next 256 pairs are loaded while these are ranked: a round trip to L2 per chunk was a third of
    // the kernel's time
    uint32_t n_bin = 0xffffffffu, n_l1 = 0u, n_l2 = 0u;
    if (p0 + tid < end) {
        n_bin = b.cls[p0 + tid] & 0xffffu;
        n_l1 = b.len1[p0 + tid];
        n_l2 = b.len2[p0 + tid];
    }
    for (uint32_t c0 = p0; c0 < end; c0 += 256) {
        const uint32_t p = c0 + tid;
        const bool active = p < end;
        const uint32_t bin = n_bin, l1 = n_l1, l2 = n_l2;
        n_bin = 0xffffffffu;
        n_l1 = n_l2 = 0u;
        if (p + 256u < end) {
            n_bin = b.cls[p + 256u] & 0xffffu;
            n_l1 = b.len1[p + 256u];
            n_l2 = b.len2[p + 256u];
        }
        // stable rank inside the warp: output bytes of the lower lanes that go to the same bin.  match.any gives
        // the lanes sharing this lane's bin; groups are small (hundreds of bins), so the loop runs 1-3 times
        // instead of the 32 x 3 shuffles of an all-pairs comparison
#if QUADE_OFFSETS_BALLOT
        // lanes with this lane's bin, from one ballot per bit of the bin index (MATCH.ANY costs ~110 cycles of the ADU pipe)
        unsigned peers = __ballot_sync(0xffffffffu, active);
        if (!active) peers = ~peers;
        for (uint32_t bit = 0; bit < bin_bits; ++bit) {
            const unsigned m = __ballot_sync(0xffffffffu, (bin >> bit) & 1u);
            peers &= (bin >> bit) & 1u ? m : ~m;
        }
#else
        const unsigned peers = __match_any_sync(0xffffffffu, bin);
#endif
        const bool later = (peers >> lane) > 1u;                // a higher lane writes to the same bin
        unsigned rem = peers & ((1u << lane) - 1u);
        uint32_t pre1 = 0, pre2 = 0;
        while (__any_sync(0xffffffffu, rem != 0u)) {
            const int j = rem ? __ffs((int)rem) - 1 : 0;
            const uint32_t a1 = __shfl_sync(0xffffffffu, l1, j), a2 = __shfl_sync(0xffffffffu, l2, j);
            if (rem) {
                pre1 += a1;
                pre2 += a2;
                rem &= rem - 1u;
            }
        }
        for (int w = 0; w < 8; ++w) {
            if (warp == w) {
                uint32_t o1 = 0, o2 = 0;
                if (active) {
                    o1 = run1[bin] + pre1;
                    o2 = run2[bin] + pre2;
                    b.out1[p] = o1;
                    b.out2[p] = o2;
                }
                __syncwarp();
                if (active && !later) {
                    run1[bin] = o1 + l1;
                    run2[bin] = o2 + l2;
                }
            }
            __syncthreads();
        }
    }
}

} // namespace
} // namespace quade
