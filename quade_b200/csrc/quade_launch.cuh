// quade_launch.cuh -- kernel launchers (included by quade_kernels.cu)
#pragma once

namespace quade {
// ------------------------------------------------------------------------------------------------------------
// launchers
// ------------------------------------------------------------------------------------------------------------
static int g_sms = 0, g_occ_frame = 0, g_occ_validate = 0, g_occ_compact = 0, g_occ_compact3 = 0, g_compact_variant = 3;
void set_compact_variant(int v) { g_compact_variant = v; }
static int g_frame_variant = 10, g_occ_frame2 = 0, g_occ_frame3 = 0;  // 10 = k_frame6 (default); 2..9: earlier single-pass variants
void set_frame_variant(int v) { g_frame_variant = v; }
int frame_variant() { return g_frame_variant; }
int frame_launches() { return g_frame_variant == 10 ? 4 : 1; }   // k_frame6, k_tilescan, k_densify, k_frame6<redo>

static int g_f6_ctas = 0, g_c3_ctas = 0;    // CTAs per SM when k_frame6 and k_compact3 share the SMs (0: all that fit)

static void init_launch_config()
{
    if (g_sms) return;
    if (const char *v = getenv("QUADE_F6_CTAS")) g_f6_ctas = atoi(v);
    if (const char *v = getenv("QUADE_C3_CTAS")) g_c3_ctas = atoi(v);
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_sms, cudaDevAttrMultiProcessorCount, dev);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame, k_frame, FT, 0);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_validate, k_validate, 256, 0);
    cudaFuncSetAttribute(k_compact, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kCompactSmem);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact, k_compact, KT, kCompactSmem);
    if (g_occ_frame < 1) g_occ_frame = 1;
    cudaFuncSetAttribute(k_frame2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F2_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame2, k_frame2, F2T, F2_SMEM);
    if (g_occ_frame2 < 1) g_occ_frame2 = 1;
    cudaFuncSetAttribute(k_frame3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F3_SMEM);
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_frame3, k_frame3, F3T, F3_SMEM);
    if (g_occ_frame3 < 1) g_occ_frame3 = 1;
    if (g_occ_validate < 1) g_occ_validate = 1;
    if (g_occ_compact < 1) g_occ_compact = 1;
    cudaFuncSetAttribute(k_compact3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(C3_SMEM * C3_WPC));
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&g_occ_compact3, k_compact3, 32 * C3_WPC, C3_SMEM * C3_WPC);
    if (g_occ_compact3 < 1) g_occ_compact3 = 1;
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

void launch_frame2(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    const uint32_t total = b.frame_tile_start[b.n_streams];
    if (!total) return;
    if (g_frame_variant == 4 || g_frame_variant == 5) {      // 4: 16 KiB tiles, 5: 8 KiB tiles
        static int backoff = -1, occ16 = 0, occ8 = 0;
        if (backoff < 0) {
            const char *v = getenv("QUADE_F3_BACKOFF");
            backoff = v ? atoi(v) : 256;
            cudaFuncSetAttribute(k_frame4<16384>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F4<16384>::SMEM);
            cudaFuncSetAttribute(k_frame4<8192>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F4<8192>::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ16, k_frame4<16384>, F4<16384>::NT, F4<16384>::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ8, k_frame4<8192>, F4<8192>::NT, F4<8192>::SMEM);
            if (const char *o = getenv("QUADE_F4_OCC")) occ16 = occ8 = atoi(o);
        }
        const uint32_t tb = g_frame_variant == 4 ? 16384u : 8192u;
        uint32_t tiles = 0;
        for (int i = 0; i < b.n_streams; ++i) tiles += b.st[i].n ? b.st[i].n / tb + 1u : 0u;
        uint32_t g4 = (uint32_t)(g_sms * max(1, g_frame_variant == 4 ? occ16 : occ8));
        if (g4 > tiles) g4 = tiles;
        if (g_frame_variant == 4) k_frame4<16384><<<g4, F4<16384>::NT, F4<16384>::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        else k_frame4<8192><<<g4, F4<8192>::NT, F4<8192>::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        return;
    }
    if (g_frame_variant == 10) {                               // k_frame6: independent tiles, scan + densify after
        static int occ6 = 0;
        if (!occ6) {
            cudaFuncSetAttribute(k_frame6<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F6_SMEM);
            cudaFuncSetAttribute(k_frame6<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F6_SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ6, k_frame6<false>, F6T, F6_SMEM);
            if (occ6 < 1) occ6 = 1;
        }
        uint32_t g6 = (uint32_t)(g_sms * (g_f6_ctas > 0 && g_f6_ctas < occ6 ? g_f6_ctas : occ6));
        if (g6 > total) g6 = total;
        k_frame6<false><<<g6, F6T, F6_SMEM, s>>>(b, total);
        return;
    }
    if (g_frame_variant >= 7 && g_frame_variant <= 9) {        // k_frame5: scan / look-back / parse warps
        static int backoff = -1, occ[3] = {0, 0, 0};
        auto k7 = k_frame5<16384, 2, 4>;
        auto k8 = k_frame5<16384, 3, 3>;
        auto k9 = k_frame5<16384, 4, 3>;
        using G7 = F5<16384, 2>;
        using G8 = F5<16384, 3>;
        using G9 = F5<16384, 4>;
        if (backoff < 0) {
            const char *v = getenv("QUADE_F3_BACKOFF");
            backoff = v ? atoi(v) : 256;
            cudaFuncSetAttribute(k7, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G7::SMEM);
            cudaFuncSetAttribute(k8, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G8::SMEM);
            cudaFuncSetAttribute(k9, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G9::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[0], k7, G7::NT, G7::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[1], k8, G8::NT, G8::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[2], k9, G9::NT, G9::SMEM);
        }
        uint32_t tiles = 0;
        for (int i = 0; i < b.n_streams; ++i) tiles += b.st[i].n ? b.st[i].n / 16384u + 1u : 0u;
        const int v = g_frame_variant - 7;
        uint32_t g5 = (uint32_t)(g_sms * max(1, occ[v]));
        if (g5 > tiles) g5 = tiles;
        if (v == 0) k7<<<g5, G7::NT, G7::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        else if (v == 1) k8<<<g5, G8::NT, G8::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        else k9<<<g5, G9::NT, G9::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        return;
    }
    if (g_frame_variant == 6) {                                // 32 KiB tiles
        static int backoff = -1, occ32 = 0;
        if (backoff < 0) {
            const char *v = getenv("QUADE_F3_BACKOFF");
            backoff = v ? atoi(v) : 256;
            cudaFuncSetAttribute(k_frame4<32768>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)F4<32768>::SMEM);
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ32, k_frame4<32768>, F4<32768>::NT, F4<32768>::SMEM);
        }
        uint32_t tiles = 0;
        for (int i = 0; i < b.n_streams; ++i) tiles += b.st[i].n ? b.st[i].n / 32768u + 1u : 0u;
        uint32_t g4 = (uint32_t)(g_sms * max(1, occ32));
        if (g4 > tiles) g4 = tiles;
        k_frame4<32768><<<g4, F4<32768>::NT, F4<32768>::SMEM, s>>>(b, tiles, (uint32_t)backoff);
        return;
    }
    if (g_frame_variant == 3) {
        uint32_t grid3 = (uint32_t)(g_sms * g_occ_frame3);
        if (grid3 > total) grid3 = total;
        static int backoff = -1;
        if (backoff < 0) {
            const char *v = getenv("QUADE_F3_BACKOFF");
            backoff = v ? atoi(v) : 256;
        }
        k_frame3<<<grid3, F3T, F3_SMEM, s>>>(b, total, (uint32_t)backoff);
        return;
    }
    uint32_t grid = (uint32_t)(g_sms * g_occ_frame2);
    if (grid > total) grid = total;
    k_frame2<<<grid, F2T, F2_SMEM, s>>>(b, total);
}

void launch_fix(const BatchDev &b, cudaStream_t s)
{
    init_launch_config();
    if (g_frame_variant == 10) {                               // the second half of the k_frame6 pipeline
        const uint32_t total = b.frame_tile_start[b.n_streams];
        if (total) {
            uint32_t max_tiles = 0;
            for (int i = 0; i < b.n_streams; ++i) max_tiles = max(max_tiles, b.st[i].n ? b.st[i].n / (uint32_t)F2_TILE + 1u : 0u);
            k_tilescan<<<dim3((max_tiles + kTileChunk - 1) / kTileChunk, b.n_streams), kTileChunk, 0, s>>>(b);
            k_densify<<<(uint32_t)g_sms * (uint32_t)F6_DENSIFY_MINB, 256, 0, s>>>(b, total);
            k_frame6<true><<<(uint32_t)g_sms, F6T, F6_SMEM, s>>>(b, total);
        }
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
    init_launch_config();
    bool in_smem;
    size_t sm = classify_smem_bytes(b, &in_smem);
    cudaFuncSetAttribute(k_classify, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm);
    int occ = 1;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, k_classify, CT, sm);
    if (occ < 1) occ = 1;
    uint32_t grid = (uint32_t)(g_sms * occ);
    if (grid > b.cap_blocks) grid = b.cap_blocks;
    k_classify<<<grid, CT, sm, s>>>(b, in_smem ? 1 : 0);
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
    if (g_compact_variant == 3) {
        uint32_t grid3 = (uint32_t)(g_sms * (g_c3_ctas > 0 && g_c3_ctas < g_occ_compact3 ? g_c3_ctas : g_occ_compact3));
        if (grid3 > b.cap_pairs) grid3 = b.cap_pairs;
        if (grid3 < 1) grid3 = 1;
        k_compact3<<<grid3, 32 * C3_WPC, C3_SMEM * C3_WPC, s>>>(b);
        return;
    }
    uint32_t grid = (uint32_t)(g_sms * g_occ_compact);
    uint32_t need = (b.cap_pairs + 255) / 256;
    if (need < 1) need = 1;
    if (grid > need) grid = need;
    k_compact<<<grid, KT, kCompactSmem, s>>>(b);
}

#ifdef QUADE_F3_PROF
extern "C" void quade_debug_f3_prof(unsigned long long *out)
{
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(out, g_f3_prof, sizeof(unsigned long long) * 16);
    unsigned long long z[16] = {0};
    cudaMemcpyToSymbol(g_f3_prof, z, sizeof(z));
}
#endif

int kernel_launch_count(int stage)
{
    switch (stage) {
    case 0: case 1: case 2: case 3: case 5: case 6: return 1;
    case 4: return 2;
    default: return 0;
    }
}

} // namespace quade

