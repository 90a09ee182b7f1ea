/* libarachne_cuda -- C ABI of the B200-native candidate-alignment / read-placement hot path.
 *
 * This is the seam the reference crosses through cgo in src/gobwa/gobwa.go (it links `-lbwa` there,
 * gobwa.go:7).  Everything below is plain C: pointers and sizes, int status returns, no aborts.
 * A call made without a usable CUDA device fails with ARACHNE_ERR_CUDA -- there is no CPU fallback.
 *
 * Two layers:
 *   (1) batch API (new): one call per batch of barcodes instead of one cgo call per read pair / candidate;
 *   (2) stage entry points: the individual kernels on caller-supplied problems -- these are what the parity
 *       tests and the config-5 ksw_extend2 microbench drive.
 * The per-pair Go entry points (GoBwaMemMateSW / GoBwaSmithWaterman / GetSeq) are kept on the Go side as thin shims over
 * this batch ABI (INTEGRATION.md sections 2 and 4); BWA's own C symbol names are not re-exported; include/arachne_compat.h is the gobwa-shaped layer.
 */
#ifndef ARACHNE_CUDA_H_
#define ARACHNE_CUDA_H_

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

enum {
    ARACHNE_OK = 0,
    ARACHNE_ERR_ARG = -1,      /* bad argument (null pointer, read longer than ARACHNE_MAX_READ_LEN, ...) */
    ARACHNE_ERR_IO = -2,       /* index files missing / unreadable / inconsistent (replaces bwa_idx_load -> NULL, gobwa.go:130-135) */
    ARACHNE_ERR_CUDA = -3,     /* no device, launch failure, out of device memory */
    ARACHNE_ERR_OVERFLOW = -4, /* an internal arena stayed too small after the automatic retries */
    ARACHNE_ERR_OPTS = -5      /* option values the kernels do not implement (non bwa_fill_scmat matrix, ...) */
};

#define ARACHNE_MAX_READ_LEN 255

/* Options: field-for-field the reference's mem_opt_t (bwa/bwamem.h:27-66; defaults mem_opt_init, bwamem.c:48-84),
 * so GoBwaAllocSettings (gobwa.go:149-153) can keep handing an opaque pointer around. */
typedef struct {
    int a, b;
    int o_del, e_del;
    int o_ins, e_ins;
    int pen_unpaired;
    int pen_clip5, pen_clip3;
    int w;
    int zdrop;
    uint64_t max_mem_intv;
    int T;
    int flag;
    int min_seed_len;
    int min_chain_weight;
    int max_chain_extend;
    float split_factor;
    int split_width;
    int max_occ;
    int max_chain_gap;
    int n_threads;
    int chunk_size;
    float mask_level;
    float drop_ratio;
    float XA_drop_ratio;
    float mask_level_redun;
    float mapQ_coef_len;
    int mapQ_coef_fac;
    int max_ins;
    int max_matesw;
    int max_XA_hits, max_XA_hits_alt;
    int8_t mat[25];
} arachne_mem_opt_t;

/* Alignment region: byte-for-byte mem_alnreg_t (bwa/bwamem.h:86-104, 88 bytes); the fields gobwa.go dereferences
 * (rb re qb qe rid score secondary secondary_all, gobwa.go:352-369) sit at the same offsets. */
typedef struct {
    int64_t rb, re;
    int qb, qe;
    int rid;
    int score;
    int truesc;
    int sub;
    int alt_sc;
    int csub;
    int sub_n;
    int w;
    int seedcov;
    int secondary;
    int secondary_all;
    int seedlen0;
    int n_comp_is_alt; /* n_comp:30, is_alt:2 */
    float frac_rep;
    uint64_t hash;
} arachne_alnreg_t;

/* Per-region alignment record: what mem_reg2aln (bwamem.c:1086-1156) + InterpretSingleReadAlignment
 * (gobwa.go:449-488) + the CIGAR/reference walk of GetAlignments (aligner.go:1505-1569) produce. */
typedef struct {
    int64_t pos;          /* 0-based leftmost forward-strand position within the contig (mem_aln_t.pos) */
    int32_t rid;
    int32_t is_rev;
    int32_t nm;           /* edit distance (mem_aln_t.NM) */
    int32_t n_cigar;
    int64_t cigar_off;    /* into the batch CIGAR pool; words are len<<4|op, MIDSH = 01234 */
    int32_t score, sub;
    int32_t matches, mismatches, indels, soft_clipped, soft_clipped_length; /* aligner.go:1505-1569 */
    int32_t n_mm;         /* number of mismatch loci */
    int64_t mm_off;       /* into the batch mismatch pool: n_mm reference loci then n_mm read offsets */
} arachne_aln_t;

/* One entry of Go's `alignments [][]*Alignment` after DoRFAForOneBarcode (aligner.go:65-114,440-501): the candidates of a
 * read that survived `score >= best - delta`, or its single unmapped placeholder; CSR by read (cand_off). */
typedef struct {
    int32_t reg;          /* index into the batch region array; -1 = unmapped placeholder (aligner.go:1664-1676) */
    int32_t read;         /* batch-global read index (2*pair + mate) */
    int32_t id;           /* hit_id within the barcode (aligner.go:1645-1712) */
    int32_t active;       /* the alignment selected for this read */
    int32_t is_proper;
    int32_t bwa_pick;
    int32_t mapq;         /* int(min(method 1, method 2, 60)) (aligner.go:901-919) */
    int32_t molecule_id;  /* candidate molecule within the barcode after scrapMolecules; -1 = none */
    int32_t active_molecule;
    int32_t second_best;  /* mapq_data.second_best as a batch-global candidate index; -1 = nil (set on active alignments) */
    int32_t mate;         /* mate_alignment as a batch-global candidate index; -1 = nil */
    int32_t duplicate;    /* markDuplicates (aligner.go:611-641) */
    int64_t pos, aend;    /* Alignment.pos / aend (aligner.go:1582-1587) */
    double log_alignment_probability;
    double as_score;      /* mapq_data.score -> AS tag */
    double xs_score;      /* mapq_data.second_best_score -> XS tag */
    double molecule_difference;          /* DM tag */
    double sum_move_probability_change;
    double molecule_confidence;
    double raw_mapq;      /* the value before int() truncation */
} arachne_cand_t;

/* One SAM record as the reference's BAM writer assembles it (AppendBam, src/aligner/bamwriter.go:283-568; DoDumpToBam :635-658),
 * fields only: ARACHNE_STAGE_RECORDS runs CheckSplitReads (split.go:31-160) and the flag /
 * mate / TLEN / tag logic on the device; arachne_records_sam() turns the fields into SAM text on the host.  CSR by read
 * (rec_off): the record of the read's active alignment, then its split alignment if one was chosen. */
typedef struct {
    int32_t read;          /* batch-global read index: QNAME, SEQ, QUAL source */
    int32_t reg;           /* region of this record's alignment (CIGAR, AC:Z), -1 = unmapped placeholder */
    int32_t primary;       /* 1: the read's active alignment; 0: its split alignment (flag 0x100) */
    int32_t flag;          /* SAM FLAG */
    int32_t rid;           /* RNAME contig, -1 = '*' */
    int32_t mapq;
    int32_t mate_rid;      /* RNEXT, -1 = '*' */
    int32_t duplicate;
    int64_t pos;           /* POS, 0-based, -1 unmapped */
    int64_t mate_pos;      /* PNEXT, 0-based, -1 */
    int64_t tlen;
    int32_t reversed;      /* SEQ is reverse-complemented, QUAL reversed (bamwriter.go:371-374) */
    int32_t clip_start, clip_end; /* SEQ / QUAL = oriented read [clip_start, clip_end): HardClip of split records (:664-689) */
    int32_t hard;          /* leading / trailing S of the CIGAR are written as H */
    int32_t as, xs;        /* AS:i, XS:i = int() of mapq_data.score / second_best_score (:416-418) */
    int32_t has_xc, xc_reg;/* XC:Z lists the mismatch loci of region xc_reg (mapq_data.second_best); 0 / -1: empty string */
    int32_t xm, am, xt;    /* XM:Z, AM:Z, XT:i */
    int32_t sa_reg, sa_rid, sa_rev, sa_hard, sa_mapq, sa_nm; /* SA:Z (:459-492); sa_reg -1: no tag */
    int64_t sa_pos;
    int32_t attach_bx, has_dm; /* BX:Z + VX:i:1 (:556-565); DM:Z */
    double dm;
} arachne_record_t;

typedef struct arachne_index arachne_index_t;

typedef struct {
    int64_t n_pairs;
    const uint8_t *seq;        /* concatenated reads as codes 0..4 (SequenceConvert, gobwa.go:159-167): R1,R2 of pair 0, ... */
    const int64_t *seq_off;    /* 2*n_pairs+1 offsets into seq */
    int64_t n_barcodes;        /* work units (fastqreader.ReadBarcodeSet, reader.go:208-291); may be 0 => one unit */
    const int64_t *barcode_off;/* n_barcodes+1 pair indices */
    int32_t score_delta;       /* GetChains passes 25 (aligner.go:454) */
    int32_t stages;            /* bitmask of ARACHNE_STAGE_*; 0 = all */
    int32_t flags;             /* ARACHNE_IN_DEVICE | ARACHNE_OUT_DEVICE */
    int32_t max_read_len;      /* ARACHNE_IN_DEVICE only: upper bound of the read lengths (checked on the device) */
    int64_t n_bases;           /* ARACHNE_IN_DEVICE only: seq_off[2*n_pairs] */
    /* ---- ARACHNE_STAGE_PLACE inputs (device pointers under ARACHNE_IN_DEVICE, like seq / seq_off / barcode_off) ---- */
    const uint8_t *run_rfa;    /* n_barcodes flags = worthRunningRFA (aligner.go:1018-1030); NULL => no barcode runs the optimizer */
    const int64_t *rand_seed;  /* n_pairs: arachne_name_seed(read name) = first 8 bytes of md5(ReadInfo) (aligner.go:1413-1416); NULL => seed 1 */
    double improper_penalty;   /* -i / --improper-pair-penalty, default -4.0 (main.go:29); 0 in a zeroed struct means "use -4.0" only with ARACHNE_PLACE_DEFAULTS */
    int32_t delta;             /* GetAlignments keeps score >= best - delta; DoRFAForOneBarcode passes 17 (aligner.go:455); 0 => 17 */
    int32_t place_flags;       /* ARACHNE_PLACE_* */
    const int64_t *centro_start, *centro_end; /* per contig id, HOST pointers, or NULL (-c / --centromeres, aligner.go:906-916) */
    const uint8_t *attach_bx;  /* ARACHNE_STAGE_RECORDS: n_barcodes flags = WorkUnit.unique_barcode (aligner.go:474,498: BX/VX/DM are written);
                                  NULL => run_rfa.  Device pointer under ARACHNE_IN_DEVICE */
} arachne_batch_in_t;

enum {
    ARACHNE_PLACE_DEFAULTS = 1 /* ignore improper_penalty and use the reference's default -4.0 */
};

enum {
    ARACHNE_IN_DEVICE = 1,     /* seq / seq_off are device pointers on the index's device (inputs already resident in HBM) */
    ARACHNE_OUT_DEVICE = 2,    /* leave the results in the handle's device workspace: only counts, counters and stage times come back */
    ARACHNE_OUT_NO_REGIONS = 4 /* do not copy the mem_alnreg_t array to the host (out->regs = NULL; reg_off, alns, cigar, mm and the
                                  placement / record arrays come back as usual).  For callers that consume alignments, candidates or
                                  records: the regions are an intermediate of the path (gobwa.go:326-335 keeps them only to hand each
                                  one back to mem_reg2aln) and are 1/3 of the result bytes */
};

enum {
    ARACHNE_STAGE_REGIONS = 1, /* seeding, chaining, extension, dedup (mem_align1_core) */
    ARACHNE_STAGE_RESCUE = 2,  /* mem_matesw as driven by GoBwaMemMateSW */
    ARACHNE_STAGE_CIGAR = 4,   /* mem_reg2aln + GetAlignments walk */
    ARACHNE_STAGE_PLACE = 8,   /* tagBestAlignments .. estimateMapQualities, markDuplicates */
    ARACHNE_STAGE_RECORDS = 16 /* CheckSplitReads, the fields of every SAM record (implies PLACE) */
};

typedef struct {
    int64_t n_reads, n_regs, n_cigar, n_mm;
    int64_t *reg_off;          /* n_reads+1 */
    arachne_alnreg_t *regs;    /* n_regs */
    arachne_aln_t *alns;       /* n_regs (ARACHNE_STAGE_CIGAR) */
    uint32_t *cigar;           /* n_cigar */
    int32_t *mm;               /* mismatch pool */
    /* counters for roofline accounting: */
    uint64_t n_occ_lookups;    /* bwt_2occ4-equivalent calls in seeding */
    uint64_t n_occ_blocks;     /* distinct 64-byte occ blocks those calls touch (1 if k>>7 == l>>7 else 2; bwt.c:194-196) */
    uint64_t n_sa_lookups, n_lf_steps;
    uint64_t n_ext_calls, n_ext_cells;  /* ksw_extend2 calls and in-band cells exactly as ksw.c:424-448 visits them */
    uint64_t n_glb_cells;
    uint64_t n_kernel_launches; /* kernels this batch launched (bench.py reports it as gpu_launches) */
    float ms_seed, ms_sa, ms_chain, ms_extend, ms_rescue, ms_cigar, ms_place, ms_total; /* CUDA-event stage times */
    /* ---- ARACHNE_STAGE_PLACE outputs ---- */
    int64_t n_cands;
    int64_t *cand_off;         /* n_reads+1 */
    arachne_cand_t *cands;     /* n_cands */
    double *log_molecule_penalty; /* n_barcodes (calculateLogMoleculePenalty, aligner.go:722-755) */
    /* ---- ARACHNE_STAGE_RECORDS outputs ---- */
    int64_t n_records;
    int64_t *rec_off;          /* n_reads+1 */
    arachne_record_t *records; /* n_records */
    float ms_records;
    void *priv;
} arachne_batch_out_t;

/* ---- library ---- */
const char *arachne_last_error(void);
int arachne_device_count(void);

/* ---- index: replaces bwa_idx_load(path, BWA_IDX_ALL) (gobwa.go:130; bwa/bwa.c:262-294).  Reads
 * <prefix>.bwt .sa .pac .ann .amb as written by `bwa index` and keeps the FM-index, the sampled SA and the
 * packed reference resident in HBM of `device`. */
int arachne_index_load(const char *prefix, int device, arachne_index_t **out);
void arachne_index_free(arachne_index_t *idx);
int64_t arachne_index_l_pac(const arachne_index_t *idx);
int32_t arachne_index_n_seqs(const arachne_index_t *idx);
int arachne_index_contig(const arachne_index_t *idx, int32_t rid, const char **name, int64_t *offset, int32_t *len);
size_t arachne_index_device_bytes(const arachne_index_t *idx);
/* bns_fetch_seq (bwa/bntseq.c:421-446) as GoBwaReference.GetSeq calls it (gobwa.go:50-80): bases (codes 0..3) of the
 * forward-reverse coordinate range [*beg, *end) clamped to the contig (and strand) that holds `mid`; the clamped range is
 * written back, *rid receives the contig.  `out` must hold the unclamped end - beg bytes.  Host only: served from the
 * handle's host copy of the packed reference.  Returns the number of bases or a negative ARACHNE_ERR_*. */
int64_t arachne_index_fetch_seq(const arachne_index_t *idx, int64_t *beg, int64_t mid, int64_t *end, int32_t *rid, uint8_t *out);
/* ---- index construction: replaces the BWT/SA half of `bwa index` = bwa_idx_build steps 2, 3 and 5 (bwa/bwtindex.c:244-300:
 * bwt_pac2bwt :61-123, bwt_bwtupdate_core :151-173, bwt_cal_sa bwa/bwt.c:62-84) on the device.  Reads <prefix>.pac (forward
 * strand, as bns_fasta2bntseq(for_only=1) leaves it, bwa/bntseq.c:275-328) and <prefix>.ann, writes <prefix>.bwt and
 * <prefix>.sa byte-identical to the reference's (bwt_dump_bwt / bwt_dump_sa, bwa/bwt.c:385-409).  batch_cap = suffixes
 * sorted per radix-sort batch (0: 2^28).  info may be NULL. */
typedef struct {
    int64_t seq_len, primary;      /* bwt_t.seq_len (= 2 * l_pac), bwt_t.primary */
    int64_t n_batches, max_batch;  /* suffix-sort batches and the largest one */
    int64_t tied_suffixes;         /* suffixes still tied after the first 29 bases */
    int64_t max_rounds;            /* deepest refinement (29 bases per round) any batch needed */
    int64_t device_peak_bytes;
    double seconds_read, seconds_build, seconds_write;
} arachne_index_build_info_t;
int arachne_index_build(const char *prefix, int device, int64_t batch_cap, arachne_index_build_info_t *info);
/* The pack step in front of it: bns_fasta2bntseq(for_only = 1) + bns_dump (bwa/bntseq.c:229-328,66-96).  FASTA (plain or gzip) ->
 * <prefix>.pac / .ann / .amb byte-identical to what `bwa index` leaves, ambiguous bases replaced by lrand48() & 3 under srand48(11) as
 * the reference does (bntseq.c:261,290-291).  Host only.  arachne_index_pack + arachne_index_build = `bwa index`. */
int arachne_index_pack(const char *fasta, const char *prefix, int64_t *l_pac);

/* Run this handle's batches on the caller's CUDA stream (a cudaStream_t; use != 0) or on the handle's own stream (use == 0,
 * the default), so that a caller can bracket batches with its own CUDA events. */
int arachne_index_set_stream(arachne_index_t *idx, void *cuda_stream, int use);

/* Runtime knobs of a handle; results never depend on them.  "seed_kmer" 0|1 (+4): use the HBM k-mer tables built at load in
 * seeding -- LAST-like pass, forward windows and short backward extensions of the SMEM passes (default 1; 0 together with
 * "seed_direct" 0 makes every bwt_extend call of the reference, which is how bench.py counts the algorithmic occ-block
 * bytes; +4 completes a forward window's table copies in a loop trip of its own).  "seed_direct" 0|1: extend unique forward
 * matches by comparing read and text (needs the inverse SA the loader builds when HBM has room; default 1).
 * "seed_blocks_per_sm" 1..4: resident seeding blocks per SM (default 3).  "seed_shared_list" -1..16: interval-list entries
 * per lane kept in shared memory (-1: the default, 10 or as many as fit).  "seed_fwd_levels" 0..15: table levels a forward
 * window replays (0: the default, 8). */
int arachne_index_set_option(arachne_index_t *idx, const char *name, int64_t value);

/* int64(binary.LittleEndian.Uint64(md5(name)[0:8])): the per-read PRNG seed of tagBestAlignments (aligner.go:1413-1416). Host only. */
int64_t arachne_name_seed(const char *name, int32_t len);

/* ---- options: mem_opt_init defaults (gobwa.go:151) ---- */
int arachne_opts_default(arachne_mem_opt_t *opt);

/* ---- batch hot path: replaces the per-pair GoBwaMemMateSW loop of GetChains (aligner.go:1637-1638) and the
 * per-candidate GoBwaSmithWaterman / GetSeq calls of GetAlignments (aligner.go:1500,1519). */
int arachne_align_batch(arachne_index_t *idx, const arachne_mem_opt_t *opt, const arachne_batch_in_t *in, arachne_batch_out_t **out);
void arachne_batch_free(arachne_batch_out_t *out);

/* SAM text of every record of a batch that ran ARACHNE_STAGE_RECORDS with host outputs: one line per record, fields
 * QNAME FLAG RNAME POS MAPQ CIGAR RNEXT PNEXT TLEN SEQ QUAL and the tags in AppendBam's order (RG XS XC AC AS XM AM XT SA BX VX DM;
 * bamwriter.go:399-566).  `in` is the batch's input (host seq / seq_off / barcode_off).  names: n_pairs read names (FastQRecord.ReadInfo);
 * quals: 2*n_pairs ASCII quality strings or NULL ('*'); barcodes: n_barcodes strings or NULL; read_groups: n_pairs strings
 * (FastQRecord.ReadGroupId, reader.go:143-152; an empty string writes no RG tag) or NULL.
 * *sam is malloc'ed: release it with free().  BAM encoding itself (biogo/hts in the reference) stays with the host. */
int arachne_records_sam(const arachne_index_t *idx, const arachne_batch_in_t *in, const arachne_batch_out_t *out, const char *const *names,
                        const char *const *quals, const char *const *barcodes, const char *const *read_groups, char **sam, size_t *sam_len);

/* ---- stage entry points (device kernels on caller problems) ---- */
/* ksw_extend2 (bwa/ksw.c:380-479) on n independent problems; query i = q[q_off[i]..q_off[i+1]), same for t.
 * out: n x 6 ints {score, qle, tle, gtle, gscore, max_off}.  *cells receives the in-band cell count;
 * *kernel_ms the CUDA-event time of `repeat` back-to-back launches divided by repeat. */
int arachne_ksw_extend_batch(int device, const arachne_mem_opt_t *opt, int64_t n, const uint8_t *q, const int64_t *q_off, const uint8_t *t,
                             const int64_t *t_off, const int32_t *h0, const int32_t *w, int end_bonus, int zdrop, int32_t *out,
                             uint64_t *cells, float *kernel_ms, int repeat);
/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call cudaMalloc (code under
 * CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The same ksw_extend2 batch as arachne_ksw_extend_batch,
 * but the call allocates NOTHING, on the device or through any library: the problem queue, the per-problem results and the control
 * words live in one static device array of the library (8 KB + 72 B per problem = 4.5 MB, held by every GPU that loads the library
 * until the process exits), so a call takes at most ARACHNE_KSW_STATIC_MAX_PROBLEMS problems and is refused with ARACHNE_ERR_ARG
 * beyond that.  ALL pointers except opt are DEVICE pointers on `device`: q / t bases as codes 0..4, q_off / t_off n+1 offsets, h0 / w per
 * problem; max_qlen an upper bound of the query lengths (1..ARACHNE_MAX_READ_LEN).  d_out: n x 6 ints as above ({-1 x 6} for a problem the
 * kernels refuse: qlen outside 1..max_qlen, h0 <= 0, h0 + qlen * a > 8191); d_status: one int, 0 or 1 + the index of the first refused
 * problem; d_cells: in-band cell count (may be NULL).  The work is queued on cuda_stream (NULL: the default stream) and the call returns
 * without waiting.  Every call shares the static array: calls on one stream are safe, calls that OVERLAP on the GPU (other streams, other
 * host threads) race. */
#define ARACHNE_KSW_STATIC_MAX_PROBLEMS 65536
int arachne_ksw_extend_static(int device, const arachne_mem_opt_t *opt, int64_t n, const uint8_t *d_q, const int64_t *d_q_off,
                              const uint8_t *d_t, const int64_t *d_t_off, const int32_t *d_h0, const int32_t *d_w, int max_qlen,
                              int end_bonus, int zdrop, int32_t *d_out, int32_t *d_status, uint64_t *d_cells, void *cuda_stream);
/* Measured DPX issue peak of the device (independent 32-bit VIADDMNMX streams): the denominator of the extension kernel's roofline
 * (not in MEASURED_PEAKS.json).  gops = 1e9 viaddmax per second. */
int arachne_dpx_peak(int device, double *gops, float *kernel_ms);
/* ksw_global2 (bwa/ksw.c:504-606): out scores[n]; cigars as CSR (cigar_off n+1, pool of cap words) when cigar_off != NULL */
int arachne_ksw_global_batch(int device, const arachne_mem_opt_t *opt, int64_t n, const uint8_t *q, const int64_t *q_off, const uint8_t *t,
                             const int64_t *t_off, const int32_t *w, int32_t *scores, int64_t *cigar_off, uint32_t *cigar, int64_t cigar_cap);
/* bwt_2occ4 (bwa/bwt.c:189-220): cnt is n x 8 (cntk[4], cntl[4]) */
int arachne_occ_batch(arachne_index_t *idx, int64_t n, const uint64_t *k, const uint64_t *l, uint64_t *cnt);
/* bwt_sa (bwa/bwt.c:86-96) */
int arachne_sa_batch(arachne_index_t *idx, int64_t n, const uint64_t *k, uint64_t *sa);
/* mem_collect_intv (bwa/bwamem.c:114-162) per read: CSR of {x0,x1,x2,info} */
int arachne_seed_batch(arachne_index_t *idx, const arachne_mem_opt_t *opt, int64_t n_reads, const uint8_t *seq, const int64_t *seq_off,
                       int64_t *intv_off, uint64_t *intv, int64_t intv_cap);

#ifdef __cplusplus
}
#endif
#endif
