/*
 * quade_io.h -- C-ABI of libquade_io.so: the host file side of the B200 build (output sink, gzip input reader).
 *
 * Replaces the file side of the reference's FastqWriter (src/FastqWriter.py:76-90): every flush there opens the
 * bin's `<name>_R1.fastq.gz` / `_R2.fastq.gz` with gzip "ab" and appends one gzip member.  Here a batch's bins
 * arrive as byte ranges of the two pinned output arenas (quade_result.out / bin_off); the sink cuts one gzip member
 * per non-empty bin and read, compresses the members on a pool of host threads and appends each to its file in
 * submission order (members of one file never overtake each other; concatenated members are one gzip stream).
 * Parity is on the decompressed stream: gzip members carry mtime and the reference's own member boundaries are an
 * artefact of its 20-record buffer.
 *
 * Plain C, no exceptions cross the boundary; calls return 0 or a negative error, text via qio_last_error().
 * Lazy creation of the files (an empty member on a bin's first record, src/FastqWriter.py:55-57,76-81) stays with
 * the caller: the sink only appends to paths that exist.
 */
#ifndef QUADE_IO_H
#define QUADE_IO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { QIO_OK = 0, QIO_ERR_ARG = -1, QIO_ERR_IO = -2, QIO_ERR_NOMEM = -3, QIO_ERR_ZLIB = -4 };

typedef struct qio_sink qio_sink;

/* n_threads compression threads (0 = one per host core, at most 64); level = zlib level 0..9 (0: stored blocks,
 * "uncompressed-equivalent") */
int qio_sink_create(int n_threads, int level, qio_sink **out);
/* waits for every job, then frees */
int qio_sink_destroy(qio_sink *s);
/* paths of the R1 / R2 file of every bin (copied); waits for the jobs in flight first */
int qio_sink_set_paths(qio_sink *s, int n_bins, const char *const r1[], const char *const r2[]);
/* for every bin b with off_k[b+1] > off_k[b]: one gzip member holding out_k[off_k[b] .. off_k[b+1]) appended to path k
 * of bin b.  Returns at once; the arenas must stay untouched until qio_sink_wait(ticket) returns. */
int qio_sink_append(qio_sink *s, const uint8_t *out1, const uint64_t *off1, const uint8_t *out2, const uint64_t *off2,
                    uint64_t *ticket);
/* until every job of the batches up to and including `ticket` is on disk; reports the first error of any job */
int qio_sink_wait(qio_sink *s, uint64_t ticket);
/* bytes handed to / written by the sink so far (decompressed, compressed) */
int qio_sink_stats(qio_sink *s, uint64_t *bytes_in, uint64_t *bytes_out);
const char *qio_last_error(void);

/* The reading side (src/pyFastq/FastqReader.py:30: a path ending in "gz" is opened with gzip): a host thread per
 * file inflates ahead into `depth` blocks of `block_bytes`, with no interpreter in the loop, while the caller drains
 * them into its (pinned) staging buffers.  Any number of concatenated gzip members is one stream, as for Python's
 * gzip module (the reference's own writer appends a member per flush); zero padding after the last member is ignored. */
typedef struct qio_reader qio_reader;
int qio_reader_open(const char *path, uint64_t block_bytes, int depth, qio_reader **out);
/* up to `cap` decompressed bytes into dst; waits for data; *got = 0 only at the end of the stream.  A corrupt or
 * truncated file gives QIO_ERR_ZLIB once the bytes in front of the damage have been handed out. */
int qio_reader_read(qio_reader *r, uint8_t *dst, uint64_t cap, uint64_t *got);
int qio_reader_close(qio_reader *r);

#ifdef __cplusplus
}
#endif
#endif
