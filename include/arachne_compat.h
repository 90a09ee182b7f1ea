/* libarachne_cuda -- the gobwa-shaped layer: what src/gobwa binds, served from GPU batches.
 *
 * The reference's Go host calls BWA once per read pair and once per candidate (src/aligner/aligner.go:1638,1500,1519 through
 * src/gobwa/gobwa.go).  A GPU wants batches.  This layer keeps the per-pair / per-candidate CALLS -- so src/aligner stays
 * untouched -- and builds the batches underneath them: every worker goroutine of the host (aligner.go:335-337, `-threads`)
 * blocks in arachne_compat_mem_mate_sw(), the calls that are in flight at the same moment are collected into one
 * arachne_align_batch() (stages REGIONS | RESCUE | CIGAR), and each caller gets back exactly what the reference's sequence
 *     mem_align1_core x2  ->  mem_matesw loop  (gobwa.go:226-337)
 * would have left in its two mem_alnreg_v.  The CIGAR of every region of the pair was computed by the same batch, so the later
 *     mem_reg2aln(region)  (gobwa.go:400-415)
 * is a lookup by region address.  integration/gobwa/gobwa_cuda.go is the Go side (same exported API as gobwa.go);
 * INTEGRATION.md shows the build line.  All structs below have the layout of the BWA structs cgo dereferences today.
 */
#ifndef ARACHNE_COMPAT_H_
#define ARACHNE_COMPAT_H_

#include "arachne_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

/* mem_alnreg_v (bwa/bwamem.h:106): gobwa.go reads .n and .a (gobwa.go:262-283,326-335) */
typedef struct { size_t n, m; arachne_alnreg_t *a; } arachne_alnreg_v;

/* GO_mem_aln_t (src/gobwa/bwa_bridge.h:19-29) = mem_aln_t with its bit-field word exposed as flag2:
 * is_rev:1, is_alt:1, mapq:8, NM:22 (gobwa.go:452-484 reads flag2 & 1, flag2 & 2, flag2 >> 10) */
typedef struct {
    int64_t pos;
    int rid;
    int flag;
    uint32_t flag2;
    int n_cigar;
    uint32_t *cigar;   /* len<<4|op, MIDSH = 01234; owned by the pair's result block (valid until arachne_compat_release) */
    char *XA;          /* always NULL: the Go side never reads alternative mappings from here */
    int score, sub, alt_sc;
} arachne_go_aln_t;

typedef struct arachne_compat arachne_compat_t;

/* One batcher per loaded index.  max_batch_pairs: a batch is launched as soon as this many calls are waiting (<= 0: 4096);
 * max_wait_us: ... or when the oldest waiting call is this old (<= 0: 200 us).  opt is copied (mem_opt_init defaults when NULL). */
int arachne_compat_open(arachne_index_t *idx, const arachne_mem_opt_t *opt, int max_batch_pairs, int max_wait_us, arachne_compat_t **out);
void arachne_compat_close(arachne_compat_t *c);

/* GoBwaMemMateSW (gobwa.go:226-337) for one pair.  r1 / r2: the reads as codes 0..4 (SequenceConvert, gobwa.go:159-167), either
 * may be empty.  Blocks until the pair's batch has run.  out1 / out2: the region lists after mate rescue; the arrays live in
 * *owner, stay valid AND UNMOVED until arachne_compat_release(*owner) -- the lifetime gobwa's Arena gives the reference's
 * malloc'ed arrays (EasyAlignment.ChainedHit points into them and comes back through GoBwaSmithWaterman, gobwa.go:326-327,400). */
int arachne_compat_mem_mate_sw(arachne_compat_t *c, const uint8_t *r1, int l1, const uint8_t *r2, int l2, int score_delta,
                               arachne_alnreg_v *out1, arachne_alnreg_v *out2, void **owner);

/* GoBwaSmithWaterman -> mem_reg2aln (gobwa.go:400-415; bwa/bwamem.c:1086-1156) for a region handed out by the call above and not
 * yet released: pos, rid, strand, NM, CIGAR, score, sub as the reference computes them.  mapq (mem_approx_mapq_se) is left 0:
 * the Go side masks it with the wrong bits and never uses it (gobwa.go:476).  ARACHNE_ERR_ARG when `reg` is not a live region. */
int arachne_compat_reg2aln(arachne_compat_t *c, const arachne_alnreg_t *reg, arachne_go_aln_t *out);

/* arena.Free() (gobwa.go:113-118): drops one pair's result block. */
void arachne_compat_release(arachne_compat_t *c, void *owner);

/* GoBwaReference.GetSeq -> bns_fetch_seq (gobwa.go:50-80; bwa/bntseq.c:421-446): malloc'ed bases (codes 0..3) of the clamped
 * range, to be released with free() as gobwa.go:78 does.  NULL on error (arachne_last_error()). */
uint8_t *arachne_compat_fetch_seq(const arachne_index_t *idx, int64_t *beg, int64_t mid, int64_t *end, int *rid);

/* How the calls were batched so far (for the host's log and the tests). */
typedef struct { uint64_t n_calls, n_batches, max_batch, n_live_blocks; } arachne_compat_stats_t;
void arachne_compat_stats(arachne_compat_t *c, arachne_compat_stats_t *out);

#ifdef __cplusplus
}
#endif
#endif
