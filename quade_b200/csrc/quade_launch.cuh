// quade_launch.cuh -- kernel launchers (included by quade_kernels.cu)
#pragma once

namespace quade {
// ------------------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------------------
static int g_sms = 0, g_occ_frame6 = 0, g_occ_compact3 = 0;
static int g_f6_ctas = 0, g_c3_ctas = 0;    // CTAs per SM when k_frame6 and k_compact3 share the SMs (0: all that fit)

static void init_launch_config()
{
    if (g_sms) return;
    if (const char *v = getenv("QUADE_F6_CTAS")) g_f6_ctas = atoi(v);
    if (const char *v = getenv("QUADE_C3_CTAS")) g_c3_ctas = atoi(v);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaFuncSetAttribute(k_frame6<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F6_SMEM);
    cudaFuncSetAttribute(k_frame6<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F6_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame6, k_frame6<false>, F6T, F6_SMEM);
    if (g_occ_frame6 < 1) g_occ_frame6 = 1;
    cudaFuncSetAttribute(k_compact3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(C3_SMEM * C3_WPC));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact3, k_compact3, 32 * C3_WPC, C3_SMEM * C3_WPC);
    if (g_occ_compact3 < 1) g_occ_compact3 = 1;
}

// k_frame6: independent tiles into scratch
void launch_frame(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (!total) return;
    uint32_t g6 = (uint32_t)(g_sms * (g_f6_ctas > 0 && g_f6_ctas < g_occ_frame6 ? g_f6_ctas : g_occ_frame6));
    if (g6 > total) g6 = total;
    k_frame6<false><<<g6, F6T, F6_SMEM, s>>>(b, total);
}

// tile scan, scratch -> dense arrays, redo pass, squeeze of files with invalid records
void launch_fix(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (total) {
        uint32_t max_tiles = 0;
        for (int i = 0; i < b.n_streams; ++i) max_tiles = max(max_tiles, b.st[i].n ? b.st[i].n / (uint32_t)F2_TILE + 1u : 0u);
        k_tilescan<<<dim3((max_tiles + kTileChunk - 1) / kTileChunk, b.n_streams), kTileChunk, 0, s>>>(b);
        k_densify<<<(uint32_t)g_sms * (uint32_t)F6_DENSIFY_MINB, 256, 0, s>>>(b, total);
        k_frame6<true><<<(uint32_t)g_sms, F6T, F6_SMEM, s>>>(b, total);
    }
    const uint32_t gx = (uint32_t)max(1, g_sms * 4 / b.n_streams);
    k_fix_compact<<<dim3(gx, b.n_streams), 256, 0, s>>>(b);
    k_fix_copyback<<<dim3(gx, b.n_streams), 256, 0, s>>>(b);
}

void launch_pairing2(const BatchDev &b, cudaStream_t s) { k_pairing2<<<1, 32, 0, s>>>(b); }

void launch_classify2(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    bool in_smem;
    size_t sm = classify_smem_bytes(b, &in_smem);
    cudaFuncSetAttribute(k_classify2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_classify2, C2T, sm);
    if (occ < 1) occ = 1;
    uint32_t grid = (uint32_t)(g_sms * occ);
    if (grid > b.cap_blocks) grid = b.cap_blocks;
    k_classify2<<<grid, C2T, sm, s>>>(b, in_smem ? 1 : 0);
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
    uint32_t grid3 = (uint32_t)(g_sms * (g_c3_ctas > 0 && g_c3_ctas < g_occ_compact3 ? g_c3_ctas : g_occ_compact3));
    if (grid3 > b.cap_pairs) grid3 = b.cap_pairs;
    if (grid3 < 1) grid3 = 1;
    k_compact3<<<grid3, 32 * C3_WPC, C3_SMEM * C3_WPC, s>>>(b);
}

} // namespace quade

