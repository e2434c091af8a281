/*
 * quade_b200.h -- C-ABI of libquade_b200.so: the B200-native demultiplexing hot path of a-slide/Quade.
 *
 * The reference (pure Python 2.7) has no FFI; the boundary replaced here is the body of its hot loop.
 * Each entry point cites the reference interface it stands in for (paths relative to the reference root):
 *
 *   quade_ctx_create        <- Quade.__init__ deriving flags/spans/min_qual          src/Quade.py:96-129
 *                              + Sample.CLASS_INIT(write_*, min_qual)                 src/Sample.py:48-54
 *   quade_ctx_set_samples   <- Sample.__init__ filling INDEX_TO_SAMPLE                src/Sample.py:132-153
 *   quade_submit/peek/collect <- one slice of the `while True` loop of
 *                              double_index_parser / simple_index_parser              src/Quade.py:195-254
 *                              = FastqReader x3/4                                     src/pyFastq/FastqReader.py:25-58
 *                              + FastqSeq slicing/concat                              src/pyFastq/FastqSeq.py:91-101
 *                              + Sample.FINDER                                        src/Sample.py:56-91
 *                              + FastqWriter.__call__ / FastqSeq.fastqstr             src/FastqWriter.py:48-74,
 *                                                                                     src/pyFastq/FastqSeq.py:68-82
 *   quade_counters, quade_allreduce_counters
 *                           <- Sample.TOTAL/PASS_QUAL/FAIL_QUAL/UNDETERMINED and
 *                              per-sample pass_qual/fail_qual                         src/Sample.py:32,144
 *
 * Plain C, no exceptions cross the boundary.  Every call returns 0 (QUADE_OK) or a negative quade_err;
 * quade_last_error() gives a thread-local message.  A ctx is single-owner (one host thread per GPU);
 * different ctxs are independent.  There is NO CPU fallback: without a CUDA device every compute entry
 * point fails with QUADE_ERR_CUDA.
 */
#ifndef QUADE_B200_H
#define QUADE_B200_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define QUADE_ABI_VERSION 2

/* stream order = the order in which the reference pulls records inside one loop iteration
 * (src/Quade.py:211-214, :241-243): insert read 1, insert read 2, index read 1[, index read 2] */
enum { QUADE_R1 = 0, QUADE_R2 = 1, QUADE_I1 = 2, QUADE_I2 = 3 };
enum { QUADE_SPAN_IDX1 = 0, QUADE_SPAN_IDX2 = 1, QUADE_SPAN_MOL1 = 2, QUADE_SPAN_MOL2 = 3 };

typedef enum {
    QUADE_OK = 0,
    QUADE_ERR_BLANK_HEADER = -1, /* a parsed record has no name token: the reference dies with IndexError
                                    (src/pyFastq/FastqReader.py:42) */
    QUADE_ERR_NOMEM = -2,
    QUADE_ERR_BAD_SAMPLE = -3,   /* non-ATCGN / duplicate / empty sample index (src/Sample.py:140-141,171-175) */
    QUADE_ERR_CUDA = -4,         /* no device, or a CUDA call failed */
    QUADE_ERR_ARG = -5,
    QUADE_ERR_CAPACITY = -6,     /* batch exceeds the capacities given at ctx creation (bytes, records, pairs) */
    QUADE_ERR_UNSUPPORTED = -7,  /* fused sample index longer than 32 bases, suffix wider than 255 bytes */
    QUADE_ERR_STATE = -8         /* collect without submit, samples not set, ... */
} quade_err;

/* Values Quade.__init__ derives from the config file (src/Quade.py:96-129). */
typedef struct {
    int32_t dual;               /* self.idx2 */
    int32_t span[4][2];         /* python slices [start,end) of idx1, idx2, mol1, mol2; disabled = 0:0 */
    int32_t min_qual;           /* [quality] minimal_qual */
    int32_t write_undetermined; /* [output] flags: counters always count, bins with the flag off get no bytes */
    int32_t write_pass;
    int32_t write_fail;
    int32_t n_samples;
    /* capacities of one batch (0 = defaults: 256 MiB per stream, max_stream_bytes/64 pairs,
     * max_stream_bytes/24 records per file).  A batch that holds more pairs than max_pairs is not an error: the
     * pairs beyond are left unconsumed (quade_result.consumed) for the next batch.  More records in one file than
     * max_records is QUADE_ERR_CAPACITY. */
    uint64_t max_stream_bytes;
    uint64_t max_pairs;
    int32_t n_slots;            /* batches in flight (1..8, 0 = 2): submit N+1 while batch N is collected */
    uint32_t flags;             /* QUADE_CTX_* */
    uint64_t max_records;
} quade_params;

#define QUADE_CTX_DEVICE_ONLY 1u /* only QUADE_BATCH_DEVICE batches will be submitted: no pinned output arenas are
                                    allocated at creation (they are, lazily, if a host batch comes after all) */

#define QUADE_BATCH_LAST 1u   /* buffers end at end-of-file: a final line without '\n' still is a line and the
                                 leftovers are dropped (src/pyFastq/FastqReader.py:37,57-58) */
#define QUADE_BATCH_EOF(s) (1u << (8 + (s))) /* the same for stream s only: the host sets it per stream while the
                                 other files of the chunk still have data to stage */
#define QUADE_BATCH_DEVICE 2u /* stream[] are DEVICE pointers (16-byte aligned); outputs stay on the device */

typedef struct {
    const uint8_t *stream[4];
    uint64_t nbytes[4];
    uint32_t flags;
    uint32_t reserved;
    uint64_t ticket; /* caller's tag, returned by collect */
} quade_batch;

/* kernel stages timed per batch (ms_kernel[]).  FRAME: k_frame (newline scan, record framing, index digest);
 * VALIDATE: what turns its per-tile output into per-record tables -- k_tilescan, the redo pass for tiles whose text
 * defeated the record-start guess, and k_locate_rare for files with len(seq) != len(qual) records
 * (FastqReader.py:48-50); PAIRING .. COMPACT: one kernel each (BINSCAN: two). */
enum { QUADE_K_FRAME = 0, QUADE_K_VALIDATE, QUADE_K_PAIRING, QUADE_K_CLASSIFY, QUADE_K_BINSCAN, QUADE_K_OFFSETS,
       QUADE_K_COMPACT, QUADE_K_COUNT };

/* bins: 0 = Undetermined, 1+2s = <sample s>_pass, 2+2s = <sample s>_fail (src/Sample.py:44,147-148) */
typedef struct {
    uint64_t ticket;
    uint64_t n_pairs;          /* Sample.TOTAL increment of this batch */
    uint64_t consumed[4];      /* bytes of each stream covered by the paired records; the rest is carry-over */
    uint64_t n_skipped[4];     /* records with len(seq) != len(qual) inside consumed[] (FastqReader.py:48-50) */
    const uint8_t *out[2];     /* R1 / R2 output arenas, bins concatenated in bin order (pinned host memory, or
                                  device memory for QUADE_BATCH_DEVICE); valid until the slot is reused */
    const uint64_t *bin_off[2];/* nbins+1 offsets into out[k] (host memory) */
    const int64_t *counters;   /* nbins pair counts of this batch (host memory) */
    int32_t nbins;
    int32_t err;               /* QUADE_OK or QUADE_ERR_BLANK_HEADER / QUADE_ERR_CAPACITY for this batch */
    int32_t chunk_end;         /* 1: a file whose buffer ends at end-of-file has no further valid record -- the
                                  reference's StopIteration, the chunk ends with this batch (src/Quade.py:223,252) */
    int32_t reserved;
    float ms_h2d, ms_kernels, ms_d2h;     /* CUDA-event durations of the three legs */
    float ms_kernel[QUADE_K_COUNT];       /* per kernel stage, events on the compute stream */
    uint64_t in_bytes;         /* sum of consumed[] */
    uint64_t out_bytes;        /* bin_off[0][nbins] + bin_off[1][nbins] */
    uint64_t gpu_launches;     /* kernels launched for this batch */
    uint64_t reframed_tiles;   /* tiles framed a second time: end of a buffer, or text that defeated the record-start guess */
} quade_result;

int quade_version(void);
const char *quade_last_error(void);
int quade_device_count(int *count);

/* pinned, caller-visible, library-owned host memory for staging */
int quade_host_alloc(void **ptr, uint64_t nbytes);
int quade_host_free(void *ptr);

typedef struct quade_ctx quade_ctx;
int quade_ctx_create(int device, const quade_params *params, quade_ctx **out);
int quade_ctx_destroy(quade_ctx *ctx);

/* keys: concatenated fused sample indexes exactly as written in the config (index1_seq + index2_seq);
 * key_off: n_samples+1 offsets.  Builds the perfect-hash table and uploads it. */
int quade_ctx_set_samples(quade_ctx *ctx, const uint8_t *keys, const uint32_t *key_off);

/* asynchronous: H2D copy (copy stream) -> kernels (compute stream); returns once everything is enqueued */
int quade_submit(quade_ctx *ctx, const quade_batch *batch);
/* blocks until the kernels of the oldest submitted batch are done and describes it -- n_pairs, consumed[], n_skipped[],
 * bin_off, counters, errors; out[] stays NULL -- without releasing it and without waiting for its bins: their D2H
 * copy is only ENQUEUED here (host contexts).  The caller learns the carry-over (consumed[]) early, submits the next
 * batch, and only then collects this one, so that the D2H copy of batch N runs beside the H2D copy and the kernels of
 * batch N+1 and beside the caller's own work */
int quade_peek(quade_ctx *ctx, quade_result *res);
/* blocks until the oldest submitted batch is done and its bins are back (D2H stream; the copy is enqueued here unless
 * quade_peek already did) and describes them */
int quade_collect(quade_ctx *ctx, quade_result *res);

/* cumulative pair counts since ctx creation / reset: int64[1+2*n_samples] in bin order */
int quade_counters(quade_ctx *ctx, int64_t *out);
int quade_counters_reset(quade_ctx *ctx);
/* The path's only collective (src/Sample.py:32,144: the counters are class-level sums): replaces the cumulative
 * counters of each of the n contexts -- one per GPU of this process, no batch in flight -- by their sum over all of
 * them (single-process ncclCommInitAll + one ncclAllReduce(int64, sum) of 1+2*n_samples values per GPU; NCCL is
 * loaded at run time, QUADE_ERR_UNSUPPORTED if libnccl.so.2 cannot be found).  Multi-process callers (one rank per
 * GPU) reduce quade_counters() with their own communicator instead (quade_b200/parallel.py). */
int quade_allreduce_counters(quade_ctx *const *ctxs, int n);
/* CUDA stream handle (cudaStream_t) kernels are launched on, for callers that time with their own events */
int quade_compute_stream(quade_ctx *ctx, void **stream);

/* Has NO counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The bin layout stage of
 * the pipeline (per-bin pair counts, per-bin byte ranges of the two output arenas, per-pair output offsets in input
 * order inside a bin) as a call that allocates nothing: no cudaMalloc / cudaFree / allocating library routine, no
 * context.  Its scratch (per-block per-bin histograms, 6 MiB) is a file-scope static __device__ array that every GPU
 * loading the library holds until the process exits; sizes beyond the limits below are refused with
 * QUADE_ERR_CAPACITY.  Works on the CURRENT device; all pointers are device pointers of that device:
 *   d_bin[n_pairs]      bin of each pair, 0 .. nbins-1 (a pair with a bin >= nbins is left out of everything)
 *   d_len1 / d_len2     output bytes of the pair's R1 / R2 record
 *   d_bin_off           out: uint64[2][nbins + 1], start of every bin in the R1 arena, then in the R2 arena
 *   d_counters          out: int64[nbins] pairs per bin
 *   d_off1 / d_off2     out: uint64[n_pairs] byte offset of each pair's record in its arena; both NULL = not wanted
 * Enqueues on `stream` (a cudaStream_t, NULL = the default stream) and returns without synchronising.  Every call on a
 * device shares the scratch, which keeps the last call's contents (each call rewrites all it reads, nothing needs
 * clearing): calls that OVERLAP on one GPU -- two streams, no dependency between them -- race on it. */
#define QUADE_STATIC_MAX_PAIRS (1u << 20)
#define QUADE_STATIC_MAX_BINS 1024
int quade_bin_offsets_static(const uint32_t *d_bin, const uint32_t *d_len1, const uint32_t *d_len2, uint64_t n_pairs,
                             int32_t nbins, uint64_t *d_bin_off, int64_t *d_counters, uint64_t *d_off1, uint64_t *d_off2,
                             void *stream);

#ifdef __cplusplus
}
#endif
#endif
