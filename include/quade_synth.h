/*
 * quade_synth.h -- C-ABI of libquade_synth.so: the counter-based synthetic FASTQ generator of SURVEY.md section 8(d)
 * (BASELINE.json configs 2-5).  TEST AND BENCHMARK INFRASTRUCTURE: it has no counterpart in the reference, it is not
 * part of the demultiplexing path, and libquade_b200.so does not contain it.  Every field of pair i is a pure
 * function of (seed, i): host buffers (quade_synth_fill) and HBM-resident buffers (quade_synth_fill_device) get
 * the same bytes.
 */
#ifndef QUADE_SYNTH_H
#define QUADE_SYNTH_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

enum { QUADE_SYNTH_OK = 0, QUADE_SYNTH_ERR_ARG = -5, QUADE_SYNTH_ERR_CUDA = -4 };

typedef struct {
    uint64_t seed;
    int32_t dual;         /* 0: R1,R2,I1   1: R1,R2,I1,I2 */
    int32_t read_len;     /* insert read length (150) */
    int32_t index_len;    /* sample barcode bases per index read */
    int32_t mol_len;      /* molecular bases following the barcode on each index read */
    int32_t n_samples;
    int32_t floor_mode;   /* 0: per-pair phred floor 30 w.p. 0.9 else 2;  1: floor uniform 2..40 */
    double undetermined;  /* probability that the barcode is random */
} quade_synth_params;

/* barcodes: n_samples * index_len * (dual?2:1) bytes, pairwise Hamming distance >= 3 per index */
int quade_synth_barcodes(const quade_synth_params *p, uint8_t *barcodes);
/* sizes of the 4 streams for pairs [first, first+count) */
int quade_synth_sizes(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                      uint64_t nbytes[4]);
/* fill the streams (host buffers sized by quade_synth_sizes) using n_threads host threads */
int quade_synth_fill(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                     uint8_t *const stream[4], int n_threads);
/* the same bytes written by CUDA kernels into DEVICE buffers (each sized by quade_synth_sizes, or generously:
 * the sizes come back in nbytes); barcodes is a HOST pointer; cuda_stream is a cudaStream_t (NULL = default stream);
 * returns after the kernels have finished */
int quade_synth_fill_device(const quade_synth_params *p, const uint8_t *barcodes, uint64_t first, uint64_t count,
                            uint8_t *const d_stream[4], uint64_t nbytes[4], void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif
