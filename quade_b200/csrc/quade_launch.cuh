// quade_launch.cuh -- kernel launchers (included by quade_kernels.cu)
#pragma once

#ifndef QUADE_COMPACT2
#define QUADE_COMPACT2 0          /* 1: k_compact2 (stream ping-pong schedule, measured 0.643 ms) instead of k_compact (0.608 ms) */
#endif

namespace quade {
// ------------------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------------------
static int g_sms = 0, g_occ_frame = 0, g_occ_compact = 0, g_occ_compact2 = 0;
static int g_f_ctas = 0, g_c_ctas = 0;      // CTAs per SM when k_frame and k_compact share the SMs (0: all that fit)

static void init_launch_config()
{
    // function attributes (dynamic shared memory above 48 KiB) are per DEVICE: a process that drives several GPUs
    // (quade_allreduce_counters: one context per GPU) sets them on each device it launches on
    static unsigned long long done_mask = 0ull;
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 64 && (done_mask >> dev) & 1ull) return;
    if (dev < 64) done_mask |= 1ull << dev;
    if (const char *v = getenv("QUADE_F_CTAS")) g_f_ctas = atoi(v);
    if (const char *v = getenv("QUADE_C_CTAS")) g_c_ctas = atoi(v);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_frame<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FT_SMEM_FAST);
    cudaFuncSetAttribute(k_frame<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FT_SMEM_REDO);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame, k_frame<false>, FT_NT, FT_SMEM_FAST);
    if (g_occ_frame < 1) g_occ_frame = 1;
    cudaFuncSetAttribute(k_compact, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(C3_SMEM * C3_WPC));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact, k_compact, 32 * C3_WPC, C3_SMEM * C3_WPC);
    if (g_occ_compact < 1) g_occ_compact = 1;
    cudaFuncSetAttribute(k_compact2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(C4_SMEM * C4_WPC));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact2, k_compact2, 32 * C4_WPC, C4_SMEM * C4_WPC);
    if (g_occ_compact2 < 1) g_occ_compact2 = 1;
}

// fast pass: independent tiles into scratch
void launch_frame(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (!total) return;
    uint32_t g = (uint32_t)(g_sms * (g_f_ctas > 0 && g_f_ctas < g_occ_frame ? g_f_ctas : g_occ_frame));
    if (g > total) g = total;
    k_frame<false><<<g, FT_NT, FT_SMEM_FAST, s>>>(b, total);
}

// tile scan + record -> scratch table, redo pass, files with invalid records
void launch_fix(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (!total) return;
    uint32_t max_tiles = 0;
    for (int i = 0; i < b.n_streams; ++i) max_tiles = max(max_tiles, b.st[i].n ? b.st[i].n / (uint32_t)FT_TILE + 1u : 0u);
    k_tilescan<<<dim3((max_tiles + kTileChunk - 1) / kTileChunk, b.n_streams), kTileChunk, 0, s>>>(b);
    k_frame<true><<<(uint32_t)g_sms, FT_NT, FT_SMEM_REDO, s>>>(b, total);
    k_locate_rare<<<b.n_streams, kTileChunk, 0, s>>>(b);
}

void launch_pairing(const BatchDev &b, cudaStream_t s) { k_pairing<<<1, 32, 0, s>>>(b); }

void launch_classify(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    bool in_smem;
    size_t sm = classify_smem_bytes(b, &in_smem);
    cudaFuncSetAttribute(k_classify, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_classify, C2T, sm);
    if (occ < 1) occ = 1;
    uint32_t grid = (uint32_t)(g_sms * occ);
    if (grid > b.cap_blocks) grid = b.cap_blocks;
    k_classify<<<grid, C2T, sm, s>>>(b, in_smem ? 1 : 0);
}

size_t classify_smem_bytes(const BatchDev &b, bool *table_in_smem)
{
    size_t hist = (((size_t)3 * b.n_bins * 4) + 15) & ~(size_t)15;
    size_t table = (size_t)(b.table.n_slots_mask + 1) * 16 + (size_t)(b.table.n_buckets_mask + 1) * 2;
    bool in = hist + table <= 96 * 1024;
    if (table_in_smem) *table_in_smem = in;
    return in ? hist + table : hist;
}

void launch_binscan(const BatchDev &b, cudaStream_t s)
{
    const uint32_t width = 2u * (uint32_t)b.n_bins;
    k_binscan_blocks<<<(width + 31) / 32, dim3(32, 32), 0, s>>>(b);
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
#if QUADE_COMPACT2
    uint32_t grid = (uint32_t)(g_sms * g_occ_compact2);
    if (grid > b.cap_pairs) grid = b.cap_pairs;
    if (grid < 1) grid = 1;
    k_compact2<<<grid, 32 * C4_WPC, C4_SMEM * C4_WPC, s>>>(b);
#else
    uint32_t grid = (uint32_t)(g_sms * (g_c_ctas > 0 && g_c_ctas < g_occ_compact ? g_c_ctas : g_occ_compact));
    if (grid > b.cap_pairs) grid = b.cap_pairs;
    if (grid < 1) grid = 1;
    k_compact<<<grid, 32 * C3_WPC, C3_SMEM * C3_WPC, s>>>(b);
#endif
}

} // namespace quade
