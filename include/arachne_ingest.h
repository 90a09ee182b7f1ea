/* libarachne_cuda -- barcode-grouped FASTQ ingest: the front of the hot path.
 *
 * Replaces src/fastqreader (reader.go:95-123 ParseHeader, :129-198 ReadOneLine, :208-291 ReadBarcodeSet; zipread.go:69-76) and the
 * per-barcode bookkeeping around it (worthRunningRFA, aligner.go:1018-1030; the md5 name seed, aligner.go:1413-1416) with a
 * multi-threaded reader whose batches ARE arachne_batch_in_t inputs: reads as codes 0..4, CSR offsets, work-unit offsets, the
 * per-unit unique / run_rfa flags and the per-pair PRNG seeds.  Host code only; gzip via zlib (a thread per file) or plain text.
 */
#ifndef ARACHNE_INGEST_H_
#define ARACHNE_INGEST_H_

#include "arachne_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

#define ARACHNE_INGEST_EOF 1

typedef struct arachne_ingest arachne_ingest_t;

typedef struct {
    int64_t n_pairs, n_units;
    const uint8_t *seq;              /* R1, R2 of pair 0, R1, R2 of pair 1, ...: codes 0..4 (nst_nt4_table; '-' is 5 as in the reference) */
    const int64_t *seq_off;          /* 2*n_pairs+1 */
    const char *qual;                /* ASCII qualities, same offsets as seq */
    const int64_t *unit_off;         /* n_units+1 pair indices: one entry per ReadBarcodeSet result = arachne_batch_in_t.barcode_off */
    const uint8_t *unique;           /* n_units: ReadBarcodeSet's full_barcode -> WorkUnit.unique_barcode = arachne_batch_in_t.attach_bx */
    const uint8_t *run_rfa;          /* n_units: worthRunningRFA = arachne_batch_in_t.run_rfa */
    const uint8_t *valid;            /* n_units: VX:i:1 on the unit's first record (ParseHeader's third result; unused by the aligner) */
    const int64_t *rand_seed;        /* n_pairs: md5(ReadInfo)[0:8] = arachne_batch_in_t.rand_seed */
    const char *names;               /* ReadInfo of every pair, concatenated */
    const int64_t *name_off;         /* n_pairs+1 */
    const char *barcodes;            /* barcode of every unit, concatenated */
    const int64_t *barcode_char_off; /* n_units+1 */
    const char *read_groups;         /* FastQRecord.ReadGroupId of every pair (reader.go:143-152: the header's last field), concatenated */
    const int64_t *read_group_off;   /* n_pairs+1 */
    void *priv;
} arachne_ingest_batch_t;

/* r1 / r2: FASTQ files, barcode-sorted, headers `@name/1 BX:Z:<barcode> VX:i:<0|1> [...]`; each file plain text, gzip (one zlib
 * stream per file: the slow case) or BGZF / bgzip (blocks inflated by up to n_threads threads per file); told apart by their first bytes.
 * n_threads: parser threads (<= 0: 4); two more threads read and split the files.  A read or format error in either file is returned
 * by the arachne_ingest_next call that reaches it.  ARACHNE_INGEST_STATS=1 prints the CPU seconds of every pipeline step at close. */
int arachne_ingest_open(const char *r1, const char *r2, int n_threads, arachne_ingest_t **out);
/* The next batch: whole work units, as many as it takes to reach target_pairs pairs (at least one unit).
 * ARACHNE_OK, ARACHNE_INGEST_EOF when the input is exhausted (*out NULL), or a negative ARACHNE_ERR_*. */
int arachne_ingest_next(arachne_ingest_t *in, int64_t target_pairs, arachne_ingest_batch_t **out);
void arachne_ingest_batch_free(arachne_ingest_batch_t *batch);
void arachne_ingest_close(arachne_ingest_t *in);

#ifdef __cplusplus
}
#endif
#endif
