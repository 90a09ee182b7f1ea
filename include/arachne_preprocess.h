/* libarachne_cuda -- FASTQ preprocessing: what `arachne preprocess` does in front of the reader.
 *
 * Replaces src/preprocess:
 *   preprocess.go:43-110   runSamtoolsPipelineWithPipes:  samtools import -T '*' -1 R1 -2 R2 | samtools sort -t <TAG> -O SAM |
 *                          samtools fastq -c 4 -T '*' -1 <TAG>_sort.R1 -2 <TAG>_sort.R2       (needs samtools in PATH)
 *   standardize.go:16-25   the format patterns (standard, haplotagging, stLFR, TELLseq)
 *   standardize.go:98-125  findFastqFormat: the first 200 records decide
 *   standardize.go:73-95   the standard header: <name>/1<TAB>BX:Z:<barcode><TAB>VX:i:<0|1>
 *   standardize.go:128-186 the three converters
 * with one in-process pass: read both files, bring every header to the standard form, order the pairs by barcode with a stable
 * device radix sort, write the two files the reader (arachne_ingest.h) takes.  No samtools, no temporary SAM.
 *
 * The reference's standardize.go does not compile (`:188`, `:74-84`) and samtools is not in this image, so this row has
 * no executable reference: the behaviour below follows the compilable half (preprocess.go), samtools' documented semantics and the
 * evident intent of the broken half; where the broken half contradicts itself the choice is listed in DESIGN.md.  Parity: unpinned.
 */
#ifndef ARACHNE_PREPROCESS_H_
#define ARACHNE_PREPROCESS_H_

#include "arachne_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

enum {
    ARACHNE_FQ_UNKNOWN = 0,
    ARACHNE_FQ_STANDARD = 1,      /* BX:Z:<x> and VX:i:[01] present (isStandardized, standardize.go:28-41) */
    ARACHNE_FQ_HAPLOTAGGING = 2,  /* BX:Z:AxxCxxBxxDxx (isHaplotagging, :44-51) */
    ARACHNE_FQ_STLFR = 3,         /* name ends in #<n>_<n>_<n> (isStlfr, :54-61) */
    ARACHNE_FQ_TELLSEQ = 4        /* name ends in :<ACGTN barcode> (isTellseq, :64-71) */
};

/* findFastqFormat (standardize.go:98-125): the format of the first of the first 200 R1 headers that matches any pattern, tried in the
 * order standard, haplotagging, stLFR, TELLseq; ARACHNE_FQ_UNKNOWN when none does.  r2 is only checked to be readable. */
int arachne_fastq_format(const char *r1, const char *r2, int *format);

/* The pattern test for one header line (without '@'; a trailing newline is appended if the line has none): the building block of the
 * call above, exported for the tests. */
int arachne_fastq_header_format(const char *header);

typedef struct {
    const char *tag;       /* SAM tag to sort by (preprocess.go:118 `-tag`, default "BX" when NULL) */
    int32_t standardize;   /* 1: rewrite non-standard headers to the standard form first (fastqStandardize); 0: headers are kept */
    int32_t sort;          /* 1: stable sort of the pairs by the tag's value, pairs without the tag first (samtools sort -t); 0: input order */
    int32_t device;        /* CUDA device of the sort */
    int32_t compress_level;/* gzip level of outputs whose name ends in .gz (preprocess.go:55 `-c 4`; <= 0: 4) */
} arachne_preprocess_opt_t;

typedef struct {
    int64_t n_pairs, n_barcodes /* distinct tag values */, n_no_tag, n_invalid /* VX:i:0 written */;
    int32_t format;        /* what arachne_fastq_format saw */
    float ms_read, ms_sort, ms_write;
} arachne_preprocess_stats_t;

/* r1 / r2: paired FASTQ, plain or gzip.  out_r1 / out_r2: the files to write (the reference names them <TAG>_sort.<basename>,
 * preprocess.go:44-45; the caller chooses here).  Mate suffixes /1 /2 are cut from the names on input and written back on output, header
 * fields that look like SAM tags (XX:T:value) are carried through tab-separated, as samtools import / fastq -T '*' do.  The whole input
 * is held in host memory.  Sort keys longer than 32 bytes are refused (ARACHNE_ERR_ARG). */
int arachne_preprocess(const char *r1, const char *r2, const char *out_r1, const char *out_r2, const arachne_preprocess_opt_t *opt,
                       arachne_preprocess_stats_t *stats);

#ifdef __cplusplus
}
#endif
#endif
