/*
 * quade_oracle.h -- CPU restatement of the reference demultiplexing hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load this library; the product (quade_b200/) never does.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle_*.py) against
 *   - the reference's only golden vector: test/dataset + test/result (a copy of the data lives in
 *     tests/golden/quade_dataset, made by tests/golden/make_golden.py), and
 *   - outputs of the reference's own source run through oracle/ref_shim.py on adversarial inputs
 *     (tests/golden/adversarial.json, same generating script).
 *
 * Every function cites the reference lines it follows (paths relative to /root/reference).
 */
#ifndef QUADE_ORACLE_H
#define QUADE_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* stream order = the order in which the reference pulls records inside one loop iteration:
 * R1, R2, I1[, I2]  (src/Quade.py:211-214, :241-243) */
enum { QO_R1 = 0, QO_R2 = 1, QO_I1 = 2, QO_I2 = 3 };
enum { QO_SPAN_IDX1 = 0, QO_SPAN_IDX2 = 1, QO_SPAN_MOL1 = 2, QO_SPAN_MOL2 = 3 };

enum {
    QO_OK = 0,
    QO_ERR_BLANK_HEADER = -1, /* reference dies with IndexError, FastqReader.py:42 */
    QO_ERR_NOMEM = -2,
    QO_ERR_BAD_SAMPLE = -3    /* Sample.py:140-141 asserts (non-ATCGN, duplicate) or empty index */
};

typedef struct {
    int32_t dual;               /* src/Quade.py:98   self.idx2 */
    int32_t span[4][2];         /* src/Quade.py:105-116: python slice [start,end) of idx1, idx2, mol1, mol2;
                                   disabled sub-indexes are 0:0 */
    int32_t min_qual;           /* src/Quade.py:96 */
    int32_t write_undetermined; /* src/Sample.py:48-54 */
    int32_t write_pass;
    int32_t write_fail;
    int32_t n_samples;
    const uint8_t *keys;        /* concatenated fused sample indexes, already upper-cased (Sample.py:136) */
    const uint32_t *key_off;    /* n_samples+1 offsets into keys */
} qo_params;

/* bins: 0 = Undetermined, 1+2s = <sample s>_pass, 2+2s = <sample s>_fail (Sample.py:44,147-148) */
typedef struct {
    uint8_t *out[2];       /* R1 / R2 output arenas: bins concatenated in bin order */
    uint64_t *bin_off[2];  /* nbins+1 offsets into out[k] */
    int64_t *counters;     /* nbins pair counts (always counted, Sample.py:62-88) */
    uint64_t n_pairs;
    uint64_t consumed[4];  /* bytes of each stream covered by the n_pairs paired records (+ skipped ones before) */
    uint64_t n_skipped[4]; /* invalid records (len(seq)!=len(qual)) inside consumed[] (FastqReader.py:48-50) */
    int32_t nbins;
    int32_t err;
    int32_t chunk_end;     /* the loop hit StopIteration: a file whose buffer ends at end-of-file ran dry (Quade.py:223,252) */
    int32_t reserved;
} qo_result;

/* Demultiplex one batch of a chunk.  `last` != 0 means the buffers end at end-of-file: a final line
 * without '\n' counts as a line and the leftovers are dropped (FastqReader.py:37,57-58).  With last == 0
 * only '\n'-terminated lines are used and consumed[] tells the caller what to carry over.  Bit 8+s of `last`
 * marks end-of-file for stream s alone. */
int qo_demux_batch(const qo_params *p, const uint8_t *const stream[4], const uint64_t nbytes[4], int last,
                   qo_result *res);
void qo_result_free(qo_result *res);

/* Throughput helper for bench.py's CPU baseline: runs qo_demux_batch(last=1) over `n_chunks` independent
 * chunks on `n_threads` POSIX threads (chunks are independent in the reference, src/Quade.py:198-206) and
 * returns the total number of pairs; outputs are produced and freed per chunk. seconds_out = wall time. */
int64_t qo_demux_chunks_mt(const qo_params *p, int n_chunks, const uint8_t *const *streams /* [n_chunks*4] */,
                           const uint64_t *nbytes /* [n_chunks*4] */, int n_threads, double *seconds_out,
                           uint64_t *out_bytes_total);

#ifdef __cplusplus
}
#endif
#endif
