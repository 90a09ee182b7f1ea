/* TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's Go placement code.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use anything under oracle/.
 *
 * PARITY UNPINNED against Go execution: this box has no Go toolchain and the reference tree does not compile as shipped
 * (SURVEY.md "facts").  What is pinned: Go's math/rand source (gorand_table.h, checked against Go's published known
 * outputs) and MD5 (RFC 1321 vectors).  sort.Sort's pdqsort tie order on equal positions is NOT reproduced (stable order used).
 */
#ifndef ORACLE_RESTATE_PLACE_H_
#define ORACLE_RESTATE_PLACE_H_
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* one record per entry of Go's `alignments [][]*Alignment` (src/aligner/aligner.go:65-114), CSR by read */
typedef struct {
    int32_t reg;          /* index into the batch region array, -1 = unmapped placeholder (aligner.go:1664-1676) */
    int32_t read;         /* batch-global read index */
    int32_t id;           /* hit_id */
    int32_t active, is_proper, bwa_pick;
    int32_t mapq;
    int32_t molecule_id;  /* within the barcode, after scrapMolecules; -1 none */
    int32_t active_molecule;
    int32_t second_best;  /* mapq_data.second_best as a batch-global candidate index, -1 nil (active alignments only) */
    int32_t mate;         /* mate_alignment as a batch-global candidate index, -1 nil */
    int32_t duplicate;
    int64_t pos, aend;
    double log_alignment_probability;
    double as_score;      /* mapq_data.score */
    double xs_score;      /* mapq_data.second_best_score */
    double molecule_difference;
    double sum_move_probability_change;
    double molecule_confidence;
    double raw_mapq;      /* the float before int() truncation (tolerance checks) */
} orc_cand_t;

typedef struct {
    int64_t l_pac;
    int32_t n_seqs;
    const int64_t *ctg_off;   /* n_seqs */
    const int32_t *ctg_len;
    const uint8_t *pac;
    const int64_t *centro_start, *centro_end; /* per rid, or NULL: -1,-1 (aligner.go:906-916) */
} orc_ref_t;

/* regs: 88-byte mem_alnreg_t records; aln: 8 int64 per region {pos rid is_rev NM n_cigar cigar_off score sub} as oracle/ref_shim.c emits */
int orc_place(const orc_ref_t *ref, int64_t n_pairs, const uint8_t *seq, const int64_t *off, int64_t n_bc, const int64_t *bc_off,
              const uint8_t *run_rfa, const int64_t *rand_seed, const int64_t *reg_off, const void *regs, const int64_t *aln,
              const uint32_t *cigar, double improper_penalty, int delta, int64_t *cand_off /* 2*n_pairs+1 */, orc_cand_t *cands, int64_t cap,
              double *log_molecule_penalty /* per barcode, may be NULL */);

/* One SAM record as AppendBam assembles it (src/aligner/bamwriter.go:283-568), fields only: the strings (QNAME, SEQ, QUAL, CIGAR text,
 * XC / AC / SA tag bodies) are functions of these fields and of the batch arrays.  Same layout as arachne_record_t. */
typedef struct {
    int32_t read;          /* batch-global read index: QNAME, SEQ, QUAL source */
    int32_t reg;           /* region of this record's alignment, -1 = unmapped placeholder */
    int32_t primary;       /* 1: the read's active alignment, 0: its split alignment (flag 0x100) */
    int32_t flag;
    int32_t rid;           /* RNAME contig, -1 = '*' */
    int32_t mapq;
    int32_t mate_rid;      /* RNEXT, -1 = '*' */
    int32_t duplicate;
    int64_t pos;           /* POS, 0-based, -1 unmapped */
    int64_t mate_pos;      /* PNEXT, 0-based, -1 */
    int64_t tlen;
    int32_t reversed;      /* SEQ is reverse-complemented, QUAL reversed */
    int32_t clip_start, clip_end; /* SEQ / QUAL = oriented read [clip_start, clip_end) (HardClip of split records) */
    int32_t hard;          /* leading / trailing S of the CIGAR written as H */
    int32_t as, xs;        /* AS:i, XS:i */
    int32_t has_xc, xc_reg;/* XC:Z lists the mismatches of region xc_reg (second best alignment); has_xc 0 / xc_reg -1: empty string */
    int32_t xm, am, xt;    /* XM:Z, AM:Z, XT:i */
    int32_t sa_reg, sa_rid, sa_rev, sa_hard, sa_mapq, sa_nm; /* SA:Z (sa_reg -1: no tag) */
    int64_t sa_pos;
    int32_t attach_bx, has_dm; /* BX:Z + VX:i:1, DM:Z */
    double dm;
} orc_record_t;

/* orc_place + markDuplicates (aligner.go:611-641) + CheckSplitReads (split.go:31-160) + the record fields of DoDumpToBam / AppendBam.
 * attach_bx: per barcode (work.unique_barcode, aligner.go:474; forced true on the RFA path :498), NULL => run_rfa[b]. */
int orc_place_records(const orc_ref_t *ref, int64_t n_pairs, const uint8_t *seq, const int64_t *off, int64_t n_bc, const int64_t *bc_off,
                      const uint8_t *run_rfa, const uint8_t *attach_bx, const int64_t *rand_seed, const int64_t *reg_off, const void *regs,
                      const int64_t *aln, const uint32_t *cigar, double improper_penalty, int delta, int64_t *cand_off, orc_cand_t *cands, int64_t cap,
                      double *log_molecule_penalty, orc_record_t *recs, int64_t rec_cap, int64_t *n_recs);

/* GetAlignments walk for one region (aligner.go:1505-1569): out5 = matches mismatches indels soft_clipped soft_clipped_length;
 * mm receives n ref loci then n read offsets; returns n */
int orc_walk(const orc_ref_t *ref, int l_query, const uint8_t *query, const void *reg, const int64_t *aln8, const uint32_t *cigar, int *out5, int32_t *mm);

/* how equal positions are ordered by the position-list sort (sort.Sort(ByPosition), aligner.go:1477-1479): 0 insertion order (default, = the device),
 * 1 reversed, 2 a restatement of Go's pdqsort.  Test knob. */
void orc_set_tie_mode(int mode);

int64_t orc_name_seed(const char *name, int len);        /* int64(LittleEndian.Uint64(md5(name)[0:8])) */
void orc_md5(const uint8_t *msg, int64_t len, uint8_t out[16]);
void orc_gorand_float64(int64_t seed, int n, double *out); /* first n Float64() of rand.New(rand.NewSource(seed)) */

#ifdef __cplusplus
}
#endif
#endif
