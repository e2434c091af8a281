// quade_static.cu -- quade_bin_offsets_static (include/quade_b200.h): the bin layout stage (K4 k_binscan_* + K4b
// k_offsets of quade_bins.cuh) as a stand-alone entry point that ALLOCATES NOTHING.
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The contexts of quade_ctx_create
// keep the stage's scratch -- the per-block per-bin histogram -- in their cudaMalloc'ed arena; here it lives in a
// file-scope `static __device__` array of a fixed, modest size, so the call needs no allocation at all.
//
//   * Every GPU that loads the library holds the array (kStaticScratchBytes = 6 MiB) until the process exits.
//   * Every call on a device shares it, and it keeps the last call's contents.  Nothing in it has to start at zero:
//     k_static_hist rewrites every column of every row the later kernels read (zeros included), so a call never sees
//     a predecessor's values.  Two calls that overlap on one GPU (different streams, no dependency) RACE on it.
//   * In host code the array's name is a host-side stand-in, not a device address: the address kernels get comes
//     from cudaGetSymbolAddress, asked once per device with that device current (scratch_of_current_device below).
#include "quade_dev.h"
#include "../../include/quade_b200.h"

#include <cstdio>
#include <mutex>

namespace quade {
namespace {

typedef unsigned long long u64;
static_assert(sizeof(u64) == sizeof(uint64_t), "u64 is the ABI's uint64_t");

constexpr uint32_t kStaticPPB = 4096;                                          // pairs per histogram block
constexpr uint32_t kStaticMaxBlocks = QUADE_STATIC_MAX_PAIRS / kStaticPPB;     // 256
constexpr uint32_t kStaticMaxCols = 3u * QUADE_STATIC_MAX_BINS;                // pairs, R1 bytes, R2 bytes per bin
constexpr size_t kStaticScratchWords = (size_t)kStaticMaxBlocks * kStaticMaxCols;
constexpr size_t kStaticScratchBytes = kStaticScratchWords * sizeof(u64);      // 6 MiB
static_assert(QUADE_STATIC_MAX_PAIRS % kStaticPPB == 0, "whole histogram blocks");

static __device__ __align__(256) u64 g_static_scratch[kStaticScratchWords];    // [block][3 * nbins], row stride 3 * nbins

// one CTA per block of kStaticPPB pairs: histogram of {pairs, R1 bytes, R2 bytes} per bin in shared memory, then the
// whole row (every column, zeros too) to the scratch.  Pairs whose bin is >= nbins are left out everywhere.
__global__ void __launch_bounds__(256) k_static_hist(const uint32_t *__restrict__ bin, const uint32_t *__restrict__ len1,
                                                     const uint32_t *__restrict__ len2, uint32_t n_pairs, uint32_t nbins,
                                                     u64 *__restrict__ hist)
{
    __shared__ u64 s_h[kStaticMaxCols];
    const uint32_t cols = 3u * nbins;
    for (uint32_t i = threadIdx.x; i < cols; i += 256) s_h[i] = 0;
    __syncthreads();
    const uint32_t p0 = blockIdx.x * kStaticPPB;
    const uint32_t end = min(p0 + kStaticPPB, n_pairs);
    for (uint32_t p = p0 + threadIdx.x; p < end; p += 256) {
        const uint32_t bn = bin[p];
        if (bn < nbins) {
            atomicAdd(reinterpret_cast<unsigned long long *>(&s_h[bn]), 1ull);
            atomicAdd(reinterpret_cast<unsigned long long *>(&s_h[nbins + bn]), (unsigned long long)len1[p]);
            atomicAdd(reinterpret_cast<unsigned long long *>(&s_h[2u * nbins + bn]), (unsigned long long)len2[p]);
        }
    }
    __syncthreads();
    u64 *row = hist + (size_t)blockIdx.x * cols;
    for (uint32_t i = threadIdx.x; i < cols; i += 256) row[i] = s_h[i];
}

// one CTA: per column an exclusive scan over the blocks (in place), then per read an exclusive scan of the column
// totals over the bins -> bin_off[2][nbins + 1], counters[nbins]
__global__ void __launch_bounds__(1024) k_static_scan(u64 *__restrict__ hist, uint32_t n_blocks, uint32_t nbins,
                                                      u64 *__restrict__ bin_off, int64_t *__restrict__ counters)
{
    __shared__ u64 s_tot[kStaticMaxCols];
    const uint32_t cols = 3u * nbins, tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    for (uint32_t col = tid; col < cols; col += 1024) {
        u64 run = 0;
        for (uint32_t kb = 0; kb < n_blocks; ++kb) {
            u64 *h = hist + (size_t)kb * cols + col;
            const u64 v = *h;
            *h = run;
            run += v;
        }
        s_tot[col] = run;
        if (col < nbins) counters[col] = (int64_t)run;
    }
    __syncthreads();
    if (warp < 2u) {                                            // warp 0: R1 bytes, warp 1: R2 bytes
        const u64 *tot = s_tot + (1u + warp) * nbins;
        u64 *off = bin_off + (size_t)warp * (nbins + 1u);
        u64 carry = 0;
        for (uint32_t base = 0; base < nbins; base += 32u) {
            const uint32_t i = base + lane;
            const u64 v = i < nbins ? tot[i] : 0;
            u64 inc = v;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const u64 t = __shfl_up_sync(0xffffffffu, inc, o);
                if ((int)lane >= o) inc += t;
            }
            if (i < nbins) off[i] = carry + inc - v;
            carry += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) off[nbins] = carry;
    }
}

// one CTA per block of pairs: the byte offset of every pair's R1 / R2 record in the arena = start of its bin + bytes
// of the bin's earlier blocks (scratch row, scanned) + bytes of the bin's earlier pairs in this block (input order
// inside a bin, FastqWriter.py:68-69).  A chunk of 256 pairs is ranked warp by warp: lanes of one bin are found with
// match.any, each adds the lengths of its lower peers, then the 8 warps apply their ranks one after the other.
__global__ void __launch_bounds__(256) k_static_offsets(const uint32_t *__restrict__ bin, const uint32_t *__restrict__ len1,
                                                        const uint32_t *__restrict__ len2, uint32_t n_pairs, uint32_t nbins,
                                                        const u64 *__restrict__ hist, const u64 *__restrict__ bin_off,
                                                        u64 *__restrict__ off1, u64 *__restrict__ off2)
{
    __shared__ u64 s_run[2u * QUADE_STATIC_MAX_BINS];
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5, cols = 3u * nbins;
    const u64 *row = hist + (size_t)blockIdx.x * cols;
    for (uint32_t i = tid; i < nbins; i += 256) {
        s_run[i] = bin_off[i] + row[nbins + i];
        s_run[nbins + i] = bin_off[nbins + 1u + i] + row[2u * nbins + i];
    }
    __syncthreads();
    const uint32_t p0 = blockIdx.x * kStaticPPB;
    const uint32_t end = min(p0 + kStaticPPB, n_pairs);
    for (uint32_t c0 = p0; c0 < end; c0 += 256) {
        const uint32_t p = c0 + tid;
        uint32_t bn = p < end ? bin[p] : 0xffffffffu;
        const bool active = bn < nbins;
        if (!active) bn = 0xffffffffu;
        const uint32_t l1 = active ? len1[p] : 0u, l2 = active ? len2[p] : 0u;
        const unsigned peers = __match_any_sync(0xffffffffu, bn);
        const bool last_of_bin = (peers >> lane) == 1u;          // no higher lane of this warp writes to the bin
        unsigned lower = peers & ((1u << lane) - 1u);
        u64 pre1 = 0, pre2 = 0;
        while (__any_sync(0xffffffffu, lower != 0u)) {
            const int j = lower ? __ffs((int)lower) - 1 : 0;
            const uint32_t a1 = __shfl_sync(0xffffffffu, l1, j), a2 = __shfl_sync(0xffffffffu, l2, j);
            if (lower) {
                pre1 += a1;
                pre2 += a2;
                lower &= lower - 1u;
            }
        }
        for (uint32_t w = 0; w < 8u; ++w) {
            if (warp == w) {
                u64 o1 = 0, o2 = 0;
                if (active) {
                    o1 = s_run[bn] + pre1;
                    o2 = s_run[nbins + bn] + pre2;
                    off1[p] = o1;
                    off2[p] = o2;
                }
                __syncwarp();
                if (active && last_of_bin) {
                    s_run[bn] = o1 + l1;
                    s_run[nbins + bn] = o2 + l2;
                }
            }
            __syncthreads();
        }
    }
}

// device address of the scratch on the CURRENT device.  The array's name in host code is a host-side stand-in, not a
// device address: cudaGetSymbolAddress gives the address, per device (each device that loads the module has its own
// copy), asked once with that device current and kept.
constexpr int kMaxDevices = 64;
std::mutex g_scratch_mu;
u64 *g_scratch_ptr[kMaxDevices];

cudaError_t scratch_of_current_device(u64 **out, int *dev_out)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[dev]) {
        u64 *ptr = nullptr;
        e = cudaGetSymbolAddress((void **)&ptr, g_static_scratch);
        if (e != cudaSuccess) return e;
        if (!ptr) return cudaErrorInvalidSymbol;
        g_scratch_ptr[dev] = ptr;
    }
    *out = g_scratch_ptr[dev];
    *dev_out = dev;
    return cudaSuccess;
}

int static_fail(int code, const char *fmt, unsigned long long a = 0, unsigned long long b = 0)
{
    char text[256];
    snprintf(text, sizeof(text), fmt, a, b);
    set_last_error(text);
    return code;
}

} // namespace
} // namespace quade

using namespace quade;

extern "C" int quade_bin_offsets_static(const uint32_t *d_bin, const uint32_t *d_len1, const uint32_t *d_len2, uint64_t n_pairs,
                                        int32_t nbins, uint64_t *d_bin_off, int64_t *d_counters, uint64_t *d_off1,
                                        uint64_t *d_off2, void *stream)
{
    if (nbins < 1 || nbins > QUADE_STATIC_MAX_BINS)
        return static_fail(QUADE_ERR_CAPACITY, "quade_bin_offsets_static: nbins = %llu, the static scratch holds 1..%llu bins",
                           (unsigned long long)(nbins < 0 ? 0 : nbins), (unsigned long long)QUADE_STATIC_MAX_BINS);
    if (n_pairs > QUADE_STATIC_MAX_PAIRS)
        return static_fail(QUADE_ERR_CAPACITY, "quade_bin_offsets_static: n_pairs = %llu, the static scratch holds at most %llu pairs",
                           (unsigned long long)n_pairs, (unsigned long long)QUADE_STATIC_MAX_PAIRS);
    if (!d_bin_off || !d_counters || (n_pairs && (!d_bin || !d_len1 || !d_len2)) || (!d_off1) != (!d_off2))
        return static_fail(QUADE_ERR_ARG, "quade_bin_offsets_static: NULL input / output (d_off1 and d_off2: both or neither)");
    u64 *scratch = nullptr;
    int dev = -1;
    cudaError_t e = scratch_of_current_device(&scratch, &dev);
    if (e != cudaSuccess) {
        set_last_error(cudaGetErrorString(e));
        return static_fail(QUADE_ERR_CUDA, "quade_bin_offsets_static: cudaGetSymbolAddress failed (see the CUDA error: code %llu)",
                           (unsigned long long)e);
    }
    cudaStream_t s = (cudaStream_t)stream;
    const uint32_t np = (uint32_t)n_pairs, nb = (uint32_t)nbins, n_blocks = (np + kStaticPPB - 1u) / kStaticPPB;
    if (n_blocks) k_static_hist<<<n_blocks, 256, 0, s>>>(d_bin, d_len1, d_len2, np, nb, scratch);
    k_static_scan<<<1, 1024, 0, s>>>(scratch, n_blocks, nb, reinterpret_cast<u64 *>(d_bin_off), d_counters);
    if (n_blocks && d_off1)
        k_static_offsets<<<n_blocks, 256, 0, s>>>(d_bin, d_len1, d_len2, np, nb, scratch, reinterpret_cast<const u64 *>(d_bin_off),
                                                  reinterpret_cast<u64 *>(d_off1), reinterpret_cast<u64 *>(d_off2));
    e = cudaGetLastError();
    if (e != cudaSuccess)
        return static_fail(QUADE_ERR_CUDA, "quade_bin_offsets_static: kernel launch failed on device %llu (CUDA error %llu)",
                           (unsigned long long)dev, (unsigned long long)e);
    return QUADE_OK;
}
