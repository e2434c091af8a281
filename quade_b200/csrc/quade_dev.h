// quade_dev.h -- structures shared by the kernels (quade_kernels.cu) and the C-ABI (quade_api.cu).
//
// HBM layout of one batch slot (all offsets are u32: a stream batch is < 3.75 GiB)
//   per stream s (R1, R2, I1[, I2]):
//     data[n]                     the raw FASTQ bytes (16-byte aligned base)
//     tile_tab[2 * tiles]         per frame tile: {line ends | k0, guess | ok, record slots, scratch base},
//                                 {records with len(seq) != len(qual), first slot with a blank header, -, -}
//     SCRATCH, indexed by the position v a tile reserved (tile order of arrival, records of a tile contiguous):
//       rec[v] (uint4)            insert reads: {name_off, seq_off, qual_off, seq_len | invalid << 31}
//                                 index reads:  {key lo, key hi, N mask, in | mn << 8 | lowq << 16 | bad << 24 | invalid << 31}
//       nv[v] (uint2)             {name_len, position just after the record's last line}
//       sfx[v * sfx_slot]         index reads: raw index bytes at [0, in), raw molecular bytes at [wi4, wi4 + mn)
//     loc[j]                      scratch position of the j-th VALID record of the file = its pairing ordinal
//                                 (FastqReader.py:48-50 skips a bad record in its own file only)
//     vraw[j]                     raw record index of the j-th valid record (only written for files that hold invalid records)
//   per pair p:
//     cls[p]   bin | suffix_len << 16 | written << 31
//     len1[p], len2[p]            output record bytes (0 when the bin's writer is off)
//     out1[p], out2[p]            byte offset of the pair's R1 / R2 record inside the output arena
//     suffix[p * suf_cap]         ':' index [':' molecular] '\n' as raw bytes
//   per block of pairs_per_block pairs: hist[block][2][nbins] output bytes per bin (R1, R2)
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace quade {

constexpr int kMaxStreams = 4;
constexpr uint64_t kMaxStreamBytes = 0xF0000000ull;
#ifndef QUADE_C3_TP
#define QUADE_C3_TP 8            /* 2 * TP records <= 16 record lanes of k_compact3 */
#endif
#ifndef QUADE_C3_CAP
#define QUADE_C3_CAP 3072
#endif
#ifndef QUADE_C3_MARGIN
#define QUADE_C3_MARGIN 8         /* the staged span of a tile is sized for mean record size * (1 + 1/MARGIN) */
#endif
constexpr int kCompactTilePairs = QUADE_C3_TP;   // k_compact3: pairs per tile (upper bound) ...
constexpr int kCompactStageBytes = QUADE_C3_CAP; // ... and staged bytes per read stream and tile
constexpr int kTileChunk = 256;    // k_frame / k_tilescan: tiles per chunk of the tile table
constexpr int kMaxChunks = 1024;   // >= kMaxStreamBytes / (16 KiB * kTileChunk)

// device-resident scalars of one batch; copied back to the host as a whole
struct BatchState {
    uint32_t n_lines[kMaxStreams];
    uint32_t n_valid[kMaxStreams];
    uint32_t n_invalid[kMaxStreams];   // complete records with len(seq) != len(qual)
    uint32_t first_blank[kMaxStreams]; // ~(lowest raw record index with an empty name token), 0 = none
    uint32_t n_pairs;
    uint32_t n_blocks;
    uint32_t consumed[kMaxStreams];
    uint32_t n_skipped[kMaxStreams];
    uint32_t err_capacity;             // some array would have overflowed
    uint32_t err_blank;                // blank header inside the parsed range
    uint32_t chunk_end;                // a file ran dry: its buffer ends at end-of-file and holds no further valid record
    uint32_t ticket_frame;             // dynamic tile counter of k_frame
    uint32_t compact_tp;               // pairs per k_compact tile, from the mean record size (k_pairing)
    uint32_t scratch_cursor[kMaxStreams]; // k_frame: records reserved in the per-stream scratch arrays
    uint32_t n_redo, ticket_redo;      // tiles the fast pass could not frame (wrong guess, end of buffer, long records)
    unsigned long long out_total[2];   // bytes of the R1 / R2 arenas
    uint32_t chunk_lines[kMaxStreams][kMaxChunks]; // k_frame: line ends per chunk of kTileChunk tiles (not copied to the host)
};

struct StreamDev {
    const uint8_t *data;
    uint32_t n;
    uint32_t cap_rec;                    // record capacity of the scratch arrays and of loc / vraw
    uint4 *rec;
    uint2 *nv;
    uint8_t *sfx;
    uint32_t *loc, *vraw;
    uint32_t sfx_slot, wi, wm;           // slot bytes (multiple of 16), widths of the index / molecular spans
    uint4 *tile_tab;
};

struct HashTable {
    const uint4 *entries;     // {key lo, key hi, nmask, len | (sample+1) << 8}
    const uint16_t *disp;     // per-bucket displacement
    uint32_t n_buckets_mask;  // buckets - 1 (power of two)
    uint32_t n_slots_mask;    // slots - 1 (power of two)
    uint64_t seed;
};

struct BatchDev {
    StreamDev st[kMaxStreams];
    int n_streams;
    int dual;
    int last;                  // bit s: the buffer of stream s ends at end-of-file
    int span[4][2];
    int min_qual;
    int wr_und, wr_pass, wr_fail;
    int n_bins;
    uint32_t cap_pairs;
    uint32_t pairs_per_block;
    uint32_t cap_blocks;
    uint32_t suf_cap;          // bytes per suffix slot (multiple of 16)
    unsigned long long out_cap[2];
    BatchState *state;
    HashTable table;
    uint32_t *cls, *len1, *len2, *out1, *out2;
    uint8_t *suffix;
    uint32_t *hist;            // [cap_blocks][2][n_bins]
    unsigned long long *bin_total; // [2][n_bins] scratch
    unsigned long long *bin_off;   // [2][n_bins+1]
    long long *counters;       // [n_bins] pair counts of this batch
    uint8_t *out[2];
    uint32_t frame_tile_start[kMaxStreams + 1];
    uint32_t *redo_list;       // stream << 28 | tile of the tiles to frame again
    uint32_t redo_cap;
};

// hash shared by the host builder and the device probe
__host__ __device__ inline uint64_t mix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline uint64_t key_hash(uint64_t key, uint32_t nmask, uint32_t len, uint64_t seed)
{
    return mix64(key ^ mix64(((uint64_t)nmask << 8 | len) ^ seed));
}
__host__ __device__ inline uint32_t key_slot(uint64_t h, uint32_t disp, uint32_t slots_mask)
{
    uint32_t stride = ((uint32_t)(h >> 12) & slots_mask) | 1u;
    return ((uint32_t)(h >> 36) + disp * stride) & slots_mask;
}

// launchers (quade_kernels.cu).  All enqueue on `s`; sizes unknown to the host are read from BatchState.
void launch_frame(const BatchDev &b, cudaStream_t s);      // k_frame<fast>
void launch_fix(const BatchDev &b, cudaStream_t s);        // k_tilescan, k_frame<redo>, k_locate_rare
void launch_pairing(const BatchDev &b, cudaStream_t s);
void launch_classify(const BatchDev &b, cudaStream_t s);
void launch_binscan(const BatchDev &b, cudaStream_t s);    // k_binscan_blocks, k_binscan_bins
void launch_offsets(const BatchDev &b, cudaStream_t s);
void launch_compact(const BatchDev &b, cudaStream_t s);
size_t classify_smem_bytes(const BatchDev &b, bool *table_in_smem);
void debug_check_tally(unsigned int out[2]);   // QUADE_CHECKS builds: {violations, first code}
void set_last_error(const char *text);         // quade_api.cu: the calling thread's quade_last_error() text (for quade_static.cu)

#ifndef QUADE_F_TILE
#define QUADE_F_TILE 32768
#endif
constexpr int kFrameTile = QUADE_F_TILE;   // bytes per frame tile (k_frame), 64 per thread

} // namespace quade
