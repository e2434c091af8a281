// quade_kernels3.cuh -- K4 bin scans, K4b offsets, K5 compact and the launchers (included by quade_kernels.cu)
#pragma once

namespace quade {
namespace {

// ------------------------------------------------------------------------------------------------------------
// K4: per bin, exclusive scan of the block histograms over blocks (in place), then over bins.
// Output order inside a bin = pair order (FastqWriter.py:68-69), bins are laid out one after the other.
// ------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_binscan_blocks(const __grid_constant__ BatchDev b)
{
    const uint32_t idx = blockIdx.x * 256u + threadIdx.x, width = 2u * (uint32_t)b.n_bins;
    if (idx >= width) return;
    const uint32_t nblk = b.state->n_blocks;
    uint32_t *h = b.hist + idx;
    u64 run = 0;
    uint32_t k = 0;
    for (; k + 4 <= nblk; k += 4) {
        const uint32_t v0 = h[(size_t)(k + 0) * width], v1 = h[(size_t)(k + 1) * width];
        const uint32_t v2 = h[(size_t)(k + 2) * width], v3 = h[(size_t)(k + 3) * width];
        h[(size_t)(k + 0) * width] = (uint32_t)run; run += v0;
        h[(size_t)(k + 1) * width] = (uint32_t)run; run += v1;
        h[(size_t)(k + 2) * width] = (uint32_t)run; run += v2;
        h[(size_t)(k + 3) * width] = (uint32_t)run; run += v3;
    }
    for (; k < nblk; ++k) {
        const uint32_t v = h[(size_t)k * width];
        h[(size_t)k * width] = (uint32_t)run;
        run += v;
    }
    b.bin_total[idx] = run;
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
    const uint32_t n_pairs = S->n_pairs, p0 = blockIdx.x * b.pairs_per_block;
    uint32_t end = p0 + b.pairs_per_block;
    if (end > n_pairs) end = n_pairs;
    for (uint32_t c0 = p0; c0 < end; c0 += 256) {
        const uint32_t p = c0 + tid;
        const bool active = p < end;
        const uint32_t bin = active ? (b.cls[p] & 0xffffu) : 0xffffffffu;
        const uint32_t l1 = active ? b.len1[p] : 0u, l2 = active ? b.len2[p] : 0u;
        uint32_t pre1 = 0, pre2 = 0;
        bool later = false;
#pragma unroll 8
        for (int i = 0; i < 32; ++i) {
            const uint32_t bi = __shfl_sync(0xffffffffu, bin, i);
            const uint32_t a1 = __shfl_sync(0xffffffffu, l1, i), a2 = __shfl_sync(0xffffffffu, l2, i);
            if (bi == bin) {
                if (i < lane) { pre1 += a1; pre2 += a2; }
                else if (i > lane) later = true;
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

// ------------------------------------------------------------------------------------------------------------
// K5: compact.  One warp per 32 consecutive pairs, one record (R1 or R2 of a pair) at a time:
//   cp.async the raw record span [name_off, qual_off+len) into a 4-stage shared-memory ring (3 records ahead),
//   assemble '@' name suffix seq "\n+\n" qual '\n' in a staging buffer that has the alignment of the
//   destination (interior 16-byte vectors by funnel shift, segment edges / suffix / constants by bytes),
//   then store aligned 128-bit vectors (+ byte stores for the first/last partial vector).
//   FastqWriter.py:61-69, FastqSeq.py:76-82.  Qualities round-trip unchanged (chr(ord(x)-33+33)).
// ------------------------------------------------------------------------------------------------------------
constexpr int KT = 256;
constexpr int KSTAGES = 4;
constexpr int KCAPV = 48;                       // 16-byte vectors per staged span (768 B)
constexpr int KS_BYTES = (KCAPV + 1) * 16;      // one extra vector: the funnel reads sx/16 + 1
constexpr int KO_VEC = KCAPV + 2;
constexpr int KO_BYTES = KO_VEC * 16;

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc)
{
    unsigned sa = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// 16 bytes starting at byte r (0..15) of the 32 bytes {A,B}
__device__ __forceinline__ uint4 shift_bytes(const uint4 A, const uint4 B, uint32_t r)
{
    const uint32_t sh = (r & 3u) * 8u;
    uint32_t w0, w1, w2, w3, w4;
    switch (r >> 2) {
    case 0: w0 = A.x; w1 = A.y; w2 = A.z; w3 = A.w; w4 = B.x; break;
    case 1: w0 = A.y; w1 = A.z; w2 = A.w; w3 = B.x; w4 = B.y; break;
    case 2: w0 = A.z; w1 = A.w; w2 = B.x; w3 = B.y; w4 = B.z; break;
    default: w0 = A.w; w1 = B.x; w2 = B.y; w3 = B.z; w4 = B.w; break;
    }
    return make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                      __funnelshift_r(w3, w4, sh));
}

// copy `len` bytes S[so ..) -> O[d0 ..): full 16-byte vectors of O by funnel shift, the edges by bytes
__device__ __forceinline__ void copy_segment(const uint8_t *S, uint8_t *O, uint32_t d0, uint32_t so, uint32_t len, int lane)
{
    const uint32_t d1 = d0 + len, fv = (d0 + 15u) >> 4, lv = d1 >> 4;
    if (lv > fv) {
        const int delta = (int)so - (int)d0;
        const uint32_t r = (uint32_t)delta & 15u;
        const uint4 *S4 = reinterpret_cast<const uint4 *>(S);
        uint4 *O4 = reinterpret_cast<uint4 *>(O);
        for (uint32_t v = fv + lane; v < lv; v += 32) {
            const uint32_t sv = (uint32_t)(delta + (int)(16u * v)) >> 4;
            O4[v] = shift_bytes(S4[sv], S4[sv + 1], r);
        }
    }
    const uint32_t head_end = min(16u * fv, d1);
    const uint32_t t = lane & 15;
    if (lane < 16) {
        if (d0 + t < head_end) O[d0 + t] = S[so + t];
    } else if (lv >= fv) {
        const uint32_t x = 16u * lv + t;
        if (x < d1) O[x] = S[so + (x - d0)];
    }
}

__global__ void __launch_bounds__(KT) k_compact(const __grid_constant__ BatchDev b)
{
    __shared__ __align__(16) uint8_t s_S[KT / 32][KSTAGES][KS_BYTES];
    __shared__ __align__(16) uint8_t s_O[KT / 32][KO_BYTES];
    const BatchState *St = b.state;
    if (St->err_capacity) return;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t n_pairs = St->n_pairs, n_tasks = (n_pairs + 31u) >> 5;
    uint8_t *O = s_O[warp];

    for (uint32_t task = blockIdx.x * (KT / 32) + warp; task < n_tasks; task += gridDim.x * (KT / 32)) {
        const uint32_t p = task * 32u + lane;
        const bool active = p < n_pairs;
        uint32_t cls = 0, oo[2] = {0, 0}, nl[2] = {0, 0};
        uint4 rc[2] = {make_uint4(0, 0, 0, 0), make_uint4(0, 0, 0, 0)};
        if (active) {
            cls = b.cls[p];
            oo[0] = b.out1[p]; oo[1] = b.out2[p];
            rc[0] = b.st[0].rec[p]; rc[1] = b.st[1].rec[p];
            nl[0] = b.st[0].vnlen[p]; nl[1] = b.st[1].vnlen[p];
        }
        const int nrec = 2 * (int)min(32u, n_pairs - task * 32u);

        auto issue = [&](int r) {
            if (r < nrec) {
                const int i = r >> 1, k = r & 1;
                const uint32_t c = __shfl_sync(0xffffffffu, cls, i);
                const uint32_t s0 = __shfl_sync(0xffffffffu, k ? rc[1].x : rc[0].x, i);
                const uint32_t s1 = __shfl_sync(0xffffffffu, k ? rc[1].z + rc[1].w : rc[0].z + rc[0].w, i);
                const uint32_t al = s0 & ~15u, nv = (s1 - al + 15u) >> 4;
                if ((c >> 31) && nv <= (uint32_t)KCAPV) {
                    const uint8_t *src = b.st[k].data + al;
                    uint8_t *dst = s_S[warp][r & (KSTAGES - 1)];
                    for (uint32_t v = lane; v < nv; v += 32) cp_async16(dst + 16u * v, src + 16u * v);
                }
            }
            cp_async_commit();
        };
        issue(0); issue(1); issue(2);

        for (int r = 0; r < nrec; ++r) {
            issue(r + KSTAGES - 1);
            cp_async_wait<KSTAGES - 1>();
            __syncwarp();
            const int i = r >> 1, k = r & 1;
            const uint32_t c = __shfl_sync(0xffffffffu, cls, i);
            const uint32_t name_off = __shfl_sync(0xffffffffu, k ? rc[1].x : rc[0].x, i);
            const uint32_t seq_off = __shfl_sync(0xffffffffu, k ? rc[1].y : rc[0].y, i);
            const uint32_t qual_off = __shfl_sync(0xffffffffu, k ? rc[1].z : rc[0].z, i);
            const uint32_t L = __shfl_sync(0xffffffffu, k ? rc[1].w : rc[0].w, i);
            const uint32_t nlen = __shfl_sync(0xffffffffu, k ? nl[1] : nl[0], i);
            const uint32_t ooff = __shfl_sync(0xffffffffu, k ? oo[1] : oo[0], i);
            if (!(c >> 31)) continue;                        // the bin's writer is off (Sample.py:74,82,90)
            const uint32_t suf_len = (c >> 16) & 0xffu;
            const uint32_t total = nlen + suf_len + 2u * L + 5u;
            const uint32_t D1 = 1u + nlen + suf_len, D2 = D1 + L + 3u;
            const uint32_t al = name_off & ~15u, a = ooff & 15u;
            const uint32_t nv_src = (qual_off + L - al + 15u) >> 4, nv_out = (a + total + 15u) >> 4;
            const uint8_t *sufp = b.suffix + (size_t)(task * 32u + i) * b.suf_cap;
            const uint8_t *data = b.st[k].data;
            uint8_t *gout = b.out[k];

            if (nv_src > (uint32_t)KCAPV || nv_out > (uint32_t)KO_VEC) {
                // oversized record: byte-wise straight from global memory
                for (uint32_t x = lane; x < total; x += 32) {
                    uint32_t ch;
                    if (x == 0u) ch = '@';
                    else if (x < 1u + nlen) ch = data[name_off + x - 1u];
                    else if (x < D1) ch = sufp[x - 1u - nlen];
                    else if (x < D1 + L) ch = data[seq_off + x - D1];
                    else if (x < D2) ch = (x - D1 - L == 1u) ? '+' : '\n';
                    else if (x < D2 + L) ch = data[qual_off + x - D2];
                    else ch = '\n';
                    gout[(size_t)ooff + x] = (uint8_t)ch;
                }
                continue;
            }
            const uint8_t *S = s_S[warp][r & (KSTAGES - 1)];
            copy_segment(S, O, a + 1u, name_off - al, nlen, lane);
            if (qual_off - seq_off == L + 3u) {
                copy_segment(S, O, a + D1, seq_off - al, 2u * L + 3u, lane);   // seq "\n?\n" qual in one go
            } else {
                copy_segment(S, O, a + D1, seq_off - al, L, lane);
                copy_segment(S, O, a + D2, qual_off - al, L, lane);
            }
            for (uint32_t t = lane; t < suf_len; t += 32) O[a + 1u + nlen + t] = sufp[t];
            __syncwarp();
            if (lane == 0) O[a] = '@';
            else if (lane <= 3) O[a + D1 + L + (lane - 1)] = lane == 2 ? '+' : '\n';
            else if (lane == 4) O[a + total - 1u] = '\n';
            __syncwarp();

            // aligned 128-bit stores; first / last partial vector by bytes
            uint8_t *gdst = gout + (size_t)(ooff - a);
            const uint32_t endb = a + total, F0 = a ? 1u : 0u, F1 = endb >> 4;
            uint32_t head_end, tail_beg;
            if (F1 >= F0) {
                const uint4 *O4 = reinterpret_cast<const uint4 *>(O);
                uint4 *G4 = reinterpret_cast<uint4 *>(gdst);
                for (uint32_t v = F0 + lane; v < F1; v += 32) G4[v] = O4[v];
                head_end = 16u * F0;
                tail_beg = 16u * F1;
            } else {
                head_end = endb;
                tail_beg = endb;
            }
            const uint32_t t = lane & 15;
            if (lane < 16) {
                if (a + t < head_end) gdst[a + t] = O[a + t];
            } else {
                if (tail_beg + t < endb) gdst[tail_beg + t] = O[tail_beg + t];
            }
            __syncwarp();
        }
        cp_async_wait<0>();
        __syncwarp();
    }
}

} // namespace

// ------------------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------------------
static int g_sms = 0, g_occ_frame = 0, g_occ_validate = 0, g_occ_compact = 0;

static void init_launch_config()
{
    if (g_sms) return;
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame, k_frame, FT, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_validate, k_validate, 256, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact, k_compact, KT, 0);
    if (g_occ_frame < 1) g_occ_frame = 1;
    if (g_occ_validate < 1) g_occ_validate = 1;
    if (g_occ_compact < 1) g_occ_compact = 1;
}

void launch_frame(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (!total) return;
    uint32_t grid = (uint32_t)(g_sms * g_occ_frame);
    if (grid > total) grid = total;
    k_frame<<<grid, FT, 0, s>>>(b, total);
}

void launch_validate(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    uint32_t max_tiles = 1;
    for (int i = 0; i < b.n_streams; ++i) {
        uint32_t t = (b.st[i].cap_rec + kValidateTile - 1) / kValidateTile;
        if (t > max_tiles) max_tiles = t;
    }
    uint32_t gx = (uint32_t)(g_sms * g_occ_validate) / (uint32_t)b.n_streams;
    if (gx < 1) gx = 1;
    if (gx > max_tiles) gx = max_tiles;
    k_validate<<<dim3(gx, b.n_streams), 256, 0, s>>>(b);
}

void launch_pairing(const BatchDev &b, cudaStream_t s) { k_pairing<<<1, 32, 0, s>>>(b); }

size_t classify_smem_bytes(const BatchDev &b, bool *table_in_smem)
{
    size_t hist = (((size_t)3 * b.n_bins * 4) + 15) & ~(size_t)15;
    size_t table = (size_t)(b.table.n_slots_mask + 1) * 16 + (size_t)(b.table.n_buckets_mask + 1) * 2;
    bool in = hist + table <= 96 * 1024;
    if (table_in_smem) *table_in_smem = in;
    return in ? hist + table : hist;
}

void launch_classify(const BatchDev &b, cudaStream_t s)
{
    bool in_smem;
    size_t sm = classify_smem_bytes(b, &in_smem);
    cudaFuncSetAttribute(k_classify, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    k_classify<<<b.cap_blocks, 256, sm, s>>>(b, in_smem ? 1 : 0);
}

void launch_binscan(const BatchDev &b, cudaStream_t s)
{
    const uint32_t width = 2u * (uint32_t)b.n_bins;
    k_binscan_blocks<<<(width + 255) / 256, 256, 0, s>>>(b);
    k_binscan_bins<<<1, 1024, 0, s>>>(b);
}

void launch_offsets(const BatchDev &b, cudaStream_t s)
{
    size_t sm = (size_t)2 * b.n_bins * 4;
    cudaFuncSetAttribute(k_offsets, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    k_offsets<<<b.cap_blocks, 256, sm, s>>>(b);
}

void launch_compact(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    uint32_t grid = (uint32_t)(g_sms * g_occ_compact);
    uint32_t need = (b.cap_pairs + 255) / 256;
    if (need < 1) need = 1;
    if (grid > need) grid = need;
    k_compact<<<grid, KT, 0, s>>>(b);
}

int kernel_launch_count(int stage)
{
    switch (stage) {
    case 0: case 1: case 2: case 3: case 5: case 6: return 1;
    case 4: return 2;
    default: return 0;
    }
}

} // namespace quade
