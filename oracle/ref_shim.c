/* TEST INFRASTRUCTURE ONLY -- never linked or executed by the product path.
 *
 * Thin C driver around the reference's own vendored BWA C sources (compiled where they lie
 * under /root/reference by oracle/Makefile into oracle/_ref/libbwa_ref.so).  It exposes flat,
 * ctypes-friendly entry points so tests can (1) pin our CPU restatement (oracle/restate_*.c)
 * and (2) compare the CUDA path with "the reference itself run here".
 *
 * The orchestration below mirrors what the reference's Go<->C bridge does for one read pair,
 * src/gobwa/gobwa.go:226-337 (GoBwaMemMateSW) and :400-415 (GoBwaSmithWaterman); every BWA
 * call it makes is the unmodified reference function.
 */
#include <stdlib.h>
#include <string.h>
#include <stdio.h>
#include <pthread.h>
#include "bwa_bridge.h"   /* the reference's own re-declarations (src/gobwa/bwa_bridge.h) */
#include "bwa.h"
#include "ksw.h"
#include "kvec.h"

extern int bwa_verbose;

void *ref_idx_load(const char *prefix) { bwa_verbose = 1; return bwa_idx_load(prefix, BWA_IDX_ALL); }
void ref_idx_free(void *idx) { bwa_idx_destroy((bwaidx_t*)idx); }
int ref_idx_build(const char *fa, const char *prefix) { bwa_verbose = 1; return bwa_idx_build(fa, prefix, 0, -1); }
int64_t ref_idx_lpac(void *idx) { return ((bwaidx_t*)idx)->bns->l_pac; }
int ref_idx_nseqs(void *idx) { return ((bwaidx_t*)idx)->bns->n_seqs; }
int ref_sizeof_alnreg(void) { return (int)sizeof(mem_alnreg_t); }
int ref_sizeof_opt(void) { return (int)sizeof(mem_opt_t); }
void ref_opt_default(void *out) { mem_opt_t *o = mem_opt_init(); memcpy(out, o, sizeof(mem_opt_t)); free(o); }

/* ---- primitives, for known-answer vectors ---- */
int ref_ksw_extend2(int qlen, const uint8_t *q, int tlen, const uint8_t *t, int w, int end_bonus, int zdrop, int h0, int *out5)
{
	mem_opt_t *o = mem_opt_init(); int sc;
	sc = ksw_extend2(qlen, q, tlen, t, 5, o->mat, o->o_del, o->e_del, o->o_ins, o->e_ins, w, end_bonus, zdrop, h0,
					 &out5[0], &out5[1], &out5[2], &out5[3], &out5[4]); /* qle tle gtle gscore max_off */
	free(o); return sc;
}
int ref_ksw_global2(int qlen, const uint8_t *q, int tlen, const uint8_t *t, int w, int *n_cigar, uint32_t *cigar, int cap)
{
	mem_opt_t *o = mem_opt_init(); int sc, n = 0; uint32_t *c = 0;
	sc = ksw_global2(qlen, q, tlen, t, 5, o->mat, o->o_del, o->e_del, o->o_ins, o->e_ins, w, n_cigar? &n : 0, n_cigar? &c : 0);
	if (n_cigar) { *n_cigar = n; if (n <= cap) memcpy(cigar, c, 4 * n); free(c); }
	free(o); return sc;
}
/* out7: score te qe score2 te2 tb qb */
void ref_ksw_align2(int qlen, const uint8_t *q, int tlen, const uint8_t *t, int xtra, int *out7)
{
	mem_opt_t *o = mem_opt_init(); kswr_t r;
	uint8_t *qq = malloc(qlen + 1), *tt = malloc(tlen + 1);
	memcpy(qq, q, qlen); memcpy(tt, t, tlen);
	r = ksw_align2(qlen, qq, tlen, tt, 5, o->mat, o->o_del, o->e_del, o->o_ins, o->e_ins, xtra, 0);
	out7[0] = r.score; out7[1] = r.te; out7[2] = r.qe; out7[3] = r.score2; out7[4] = r.te2; out7[5] = r.tb; out7[6] = r.qb;
	free(qq); free(tt); free(o);
}
void ref_bwt_2occ4(void *idx, uint64_t k, uint64_t l, uint64_t *cntk, uint64_t *cntl) { bwt_2occ4(((bwaidx_t*)idx)->bwt, k, l, cntk, cntl); }
uint64_t ref_bwt_sa(void *idx, uint64_t k) { return bwt_sa(((bwaidx_t*)idx)->bwt, k); }
void ref_bwt_meta(void *idx, uint64_t *out7)
{
	bwt_t *b = ((bwaidx_t*)idx)->bwt;
	out7[0] = b->primary; memcpy(out7 + 1, b->L2, 40); out7[6] = b->seq_len;
}
/* SMEMs through x; out rows are {x0,x1,x2,info}; returns next x, *n rows */
int ref_smem1(void *idx, int len, const uint8_t *q, int x, int min_intv, uint64_t *out, int cap, int *n)
{
	bwtintv_v mem = {0,0,0}; int ret, i;
	ret = bwt_smem1(((bwaidx_t*)idx)->bwt, len, q, x, min_intv, &mem, 0);
	*n = (int)mem.n;
	for (i = 0; i < (int)mem.n && i < cap; ++i) memcpy(out + 4 * i, &mem.a[i], 32);
	free(mem.a); return ret;
}
int ref_seed_strategy1(void *idx, int len, const uint8_t *q, int x, int min_len, int max_intv, uint64_t *out4)
{
	bwtintv_t m; int ret = bwt_seed_strategy1(((bwaidx_t*)idx)->bwt, len, q, x, min_len, max_intv, &m);
	memcpy(out4, &m, 32); return ret;
}
/* mem_chain (+ optional mem_chain_flt): chain rows {pos, rid, n, w, kept, first_seed_row}, seed rows {rbeg,qbeg,len,score} */
int ref_mem_chain(void *idx_, int len, const uint8_t *seq, int do_flt, int64_t *chains, int cap_c, int64_t *seeds, int cap_s, int *n_seeds, float *frac_rep)
{
	bwaidx_t *idx = idx_; mem_opt_t *o = mem_opt_init(); mem_chain_v c; int i, j, ns = 0;
	uint8_t *s = malloc(len + 1); memcpy(s, seq, len);
	c = mem_chain(o, idx->bwt, idx->bns, len, s, 0);
	if (do_flt) { extern int mem_chain_flt(const mem_opt_t*, int, mem_chain_t*); c.n = mem_chain_flt(o, c.n, c.a); }
	for (i = 0; i < (int)c.n; ++i) {
		mem_chain_t *p = &c.a[i];
		if (i < cap_c) { int64_t *r = chains + 6 * i; r[0] = p->pos; r[1] = p->rid; r[2] = p->n; r[3] = p->w; r[4] = p->kept; r[5] = ns; }
		for (j = 0; j < p->n; ++j, ++ns)
			if (ns < cap_s) { int64_t *r = seeds + 4 * ns; r[0] = p->seeds[j].rbeg; r[1] = p->seeds[j].qbeg; r[2] = p->seeds[j].len; r[3] = p->seeds[j].score; }
		if (i == 0 && frac_rep) *frac_rep = p->frac_rep;
		free(p->seeds);
	}
	*n_seeds = ns; i = (int)c.n; free(c.a); free(s); free(o); return i;
}
int ref_fetch_seq(void *idx_, int64_t *beg, int64_t mid, int64_t *end, uint8_t *out, int cap)
{
	bwaidx_t *idx = idx_; int rid; uint8_t *s = bns_fetch_seq(idx->bns, idx->pac, beg, mid, end, &rid);
	if (*end - *beg <= cap) memcpy(out, s, *end - *beg);
	free(s); return rid;
}
int ref_contig(void *idx_, int rid, int64_t *offset, int32_t *len, char *name, int cap)
{
	bwaidx_t *idx = idx_; if (rid < 0 || rid >= idx->bns->n_seqs) return -1;
	*offset = idx->bns->anns[rid].offset; *len = idx->bns->anns[rid].len;
	strncpy(name, idx->bns->anns[rid].name, cap - 1); name[cap - 1] = 0; return 0;
}

/* ---- one read, no rescue: mem_align1_core (gobwa.go:244) ---- */
int ref_align1(void *idx_, int len, const uint8_t *seq, mem_alnreg_t *out, int cap)
{
	bwaidx_t *idx = idx_; mem_opt_t *o = mem_opt_init(); mem_alnreg_v r; int n;
	char *s = malloc(len + 1); memcpy(s, seq, len);
	r = mem_align1_core(o, idx->bwt, idx->bns, idx->pac, len, s, 0);
	n = (int)r.n; if (n <= cap) memcpy(out, r.a, n * sizeof(mem_alnreg_t));
	free(r.a); free(s); free(o); return n;
}

/* ---- one pair: gobwa.go:226-337 ---- */
static void pair_regions(const mem_opt_t *o, const bwaidx_t *idx, int l1, const uint8_t *s1_, int l2, const uint8_t *s2_, int score_delta, int do_rescue, mem_alnreg_v *r1, mem_alnreg_v *r2)
{
	mem_pestat_t pes[4]; int i, num, best1 = 0, best2 = 0, n2_0, n1_1;
	char *s1 = malloc(l1 + 1), *s2 = malloc(l2 + 1);
	int *sc;
	memset(pes, 0, sizeof(pes));
	pes[0].failed = 1; pes[2].failed = 1; pes[3].failed = 1;          /* gobwa.go:229-237 */
	pes[1].low = -35; pes[1].high = 500; pes[1].failed = 0; pes[1].avg = 200.0; pes[1].std = 100.0;
	memcpy(s1, s1_, l1); memcpy(s2, s2_, l2);
	kv_init(*r1); kv_init(*r2);
	if (l1 > 0) *r1 = mem_align1_core(o, idx->bwt, idx->bns, idx->pac, l1, s1, 0);   /* :243-251 */
	if (l2 > 0) *r2 = mem_align1_core(o, idx->bwt, idx->bns, idx->pac, l2, s2, 0);   /* :252-260 */
	if (do_rescue) {
		for (i = 0; i < (int)r1->n; ++i) if (r1->a[i].score > best1) best1 = r1->a[i].score;   /* :267-283 */
		for (i = 0; i < (int)r2->n; ++i) if (r2->a[i].score > best2) best2 = r2->a[i].score;
		/* :285-300; the Go loop walks the pre-rescue snapshot of read 2's list (scores and pointers into r2->a, which
		 * is not modified in this loop) */
		n2_0 = (int)r2->n;
		for (i = 0, num = 0; i < n2_0 && num < 50 && l1 > 0; ++i)
			if (r2->a[i].score >= best2 - score_delta) {
				++num;
				mem_matesw(o, idx->bns, idx->pac, pes, &r2->a[i], l1, (uint8_t*)s1, r1);
			}
		/* :302-324; the snapshot of read 1 is taken after its rescue; r1->a is not modified in this loop */
		n1_1 = (int)r1->n;
		sc = malloc(sizeof(int) * (n1_1 + 1));
		for (i = 0; i < n1_1; ++i) sc[i] = r1->a[i].score;
		for (i = 0, num = 0; i < n1_1 && num < 50 && l2 > 0; ++i)
			if (sc[i] >= best1 - score_delta) {
				++num;
				mem_matesw(o, idx->bns, idx->pac, pes, &r1->a[i], l2, (uint8_t*)s2, r2);
			}
		free(sc);
	}
	free(s1); free(s2);
}

int ref_pair(void *idx_, int l1, const uint8_t *s1, int l2, const uint8_t *s2, int score_delta, int do_rescue,
			 mem_alnreg_t *out1, int cap1, int *n1, mem_alnreg_t *out2, int cap2, int *n2)
{
	mem_opt_t *o = mem_opt_init(); mem_alnreg_v r1, r2;
	pair_regions(o, idx_, l1, s1, l2, s2, score_delta, do_rescue, &r1, &r2);
	*n1 = (int)r1.n; *n2 = (int)r2.n;
	if (*n1 <= cap1 && *n1) memcpy(out1, r1.a, *n1 * sizeof(mem_alnreg_t));
	if (*n2 <= cap2 && *n2) memcpy(out2, r2.a, *n2 * sizeof(mem_alnreg_t));
	free(r1.a); free(r2.a); free(o);
	return 0;
}

/* ---- GoBwaSmithWaterman (gobwa.go:400-415) + InterpretSingleReadAlignment (:449-488) ----
 * out: pos rid is_rev NM n_cigar score sub ; cigar words len<<4|op */
int ref_reg2aln(void *idx_, int len, const uint8_t *seq, const mem_alnreg_t *reg, int64_t *out7, uint32_t *cigar, int cap)
{
	bwaidx_t *idx = idx_; mem_opt_t *o = mem_opt_init(); mem_aln_t a;
	a = mem_reg2aln(o, idx->bns, idx->pac, len, (const char*)seq, reg);
	out7[0] = a.pos; out7[1] = a.rid; out7[2] = a.is_rev; out7[3] = a.NM; out7[4] = a.n_cigar; out7[5] = a.score; out7[6] = a.sub;
	if (a.n_cigar <= cap && a.n_cigar > 0) memcpy(cigar, a.cigar, 4 * a.n_cigar);
	free(a.cigar); free(o); return a.n_cigar;
}

/* ---- batch driver (threads over pairs), used for goldens and as the timed CPU reference ----
 * seq: concatenated 0..4 codes; off[2*n_pairs+1].  Outputs: CSR of regions per read plus per-region alignment
 * record and CIGAR pool, all malloc'ed here and released by ref_batch_free. */
typedef struct {
	int64_t n_reads, n_regs, n_cigar;
	int64_t *reg_off;      /* n_reads+1 */
	mem_alnreg_t *regs;    /* n_regs */
	int64_t *aln;          /* n_regs x 8: pos rid is_rev NM n_cigar cigar_off score sub */
	uint32_t *cigar;       /* n_cigar */
} ref_batch_t;

typedef struct {
	const bwaidx_t *idx; const mem_opt_t *opt; int64_t n_pairs; const uint8_t *seq; const int64_t *off;
	int score_delta, do_rescue, do_cigar, tid, n_threads;
	mem_alnreg_v *regs; mem_aln_t **alns;
} worker_t;

static void *batch_worker(void *data)
{
	worker_t *w = data; int64_t i; int k, j;
	for (i = w->tid; i < w->n_pairs; i += w->n_threads) {
		const uint8_t *s[2] = { w->seq + w->off[2*i], w->seq + w->off[2*i+1] };
		int l[2] = { (int)(w->off[2*i+1] - w->off[2*i]), (int)(w->off[2*i+2] - w->off[2*i+1]) };
		pair_regions(w->opt, w->idx, l[0], s[0], l[1], s[1], w->score_delta, w->do_rescue, &w->regs[2*i], &w->regs[2*i+1]);
		if (w->do_cigar)
			for (k = 0; k < 2; ++k) {
				mem_alnreg_v *r = &w->regs[2*i+k];
				mem_aln_t *a = w->alns[2*i+k] = calloc(r->n + 1, sizeof(mem_aln_t));
				for (j = 0; j < (int)r->n; ++j)
					a[j] = mem_reg2aln(w->opt, w->idx->bns, w->idx->pac, l[k], (const char*)s[k], &r->a[j]);
			}
	}
	return 0;
}

ref_batch_t *ref_batch(void *idx_, int64_t n_pairs, const uint8_t *seq, const int64_t *off, int score_delta, int do_rescue, int do_cigar, int n_threads)
{
	ref_batch_t *b = calloc(1, sizeof(ref_batch_t)); mem_opt_t *o = mem_opt_init();
	int64_t i, nr = 0, nc = 0; int t, j;
	mem_alnreg_v *regs = calloc(2 * n_pairs + 1, sizeof(mem_alnreg_v));
	mem_aln_t **alns = calloc(2 * n_pairs + 1, sizeof(mem_aln_t*));
	worker_t *w; pthread_t *th;
	if (n_threads < 1) n_threads = 1;
	w = calloc(n_threads, sizeof(worker_t)); th = calloc(n_threads, sizeof(pthread_t));
	for (t = 0; t < n_threads; ++t) {
		w[t].idx = idx_; w[t].opt = o; w[t].n_pairs = n_pairs; w[t].seq = seq; w[t].off = off; w[t].score_delta = score_delta;
		w[t].do_rescue = do_rescue; w[t].do_cigar = do_cigar; w[t].tid = t; w[t].n_threads = n_threads; w[t].regs = regs; w[t].alns = alns;
		pthread_create(&th[t], 0, batch_worker, &w[t]);
	}
	for (t = 0; t < n_threads; ++t) pthread_join(th[t], 0);
	b->n_reads = 2 * n_pairs;
	b->reg_off = calloc(b->n_reads + 1, 8);
	for (i = 0; i < b->n_reads; ++i) {
		b->reg_off[i] = nr; nr += regs[i].n;
		if (do_cigar) for (j = 0; j < (int)regs[i].n; ++j) nc += alns[i][j].n_cigar;
	}
	b->reg_off[b->n_reads] = nr; b->n_regs = nr; b->n_cigar = nc;
	b->regs = calloc(nr + 1, sizeof(mem_alnreg_t)); b->aln = calloc(8 * (nr + 1), 8); b->cigar = calloc(nc + 1, 4);
	for (i = 0, nc = 0; i < b->n_reads; ++i) {
		if (regs[i].n) memcpy(b->regs + b->reg_off[i], regs[i].a, regs[i].n * sizeof(mem_alnreg_t));
		if (do_cigar) {
			for (j = 0; j < (int)regs[i].n; ++j) {
				mem_aln_t *a = &alns[i][j]; int64_t *r = b->aln + 8 * (b->reg_off[i] + j);
				r[0] = a->pos; r[1] = a->rid; r[2] = a->is_rev; r[3] = a->NM; r[4] = a->n_cigar; r[5] = nc; r[6] = a->score; r[7] = a->sub;
				if (a->n_cigar) memcpy(b->cigar + nc, a->cigar, 4 * a->n_cigar);
				nc += a->n_cigar; free(a->cigar);
			}
			free(alns[i]);
		}
		free(regs[i].a);
	}
	free(regs); free(alns); free(w); free(th); free(o);
	return b;
}
void ref_batch_free(ref_batch_t *b) { if (!b) return; free(b->reg_off); free(b->regs); free(b->aln); free(b->cigar); free(b); }
