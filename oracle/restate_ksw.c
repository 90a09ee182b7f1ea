/* TEST INFRASTRUCTURE ONLY (CPU oracle) -- never linked, imported or executed by the product path.
 *
 * Plain-C restatement of the three dynamic-programming kernels on the hot path, written from the reference's
 * algorithm (not its source text):
 *   orc_extend   <- ksw_extend2   src/gobwa/bwa/ksw.c:380-479
 *   orc_global   <- ksw_global2   src/gobwa/bwa/ksw.c:504-606
 *   orc_local    <- ksw_align2 + ksw_qinit + ksw_u8 / ksw_i16   src/gobwa/bwa/ksw.c:63-365
 * Pinned by tests/test_oracle_ksw.py against oracle/_ref (the reference's own ksw.c compiled by oracle/Makefile)
 * and against the committed vectors in tests/golden/ksw_*.json that were generated from it.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct { int a, b, o_del, e_del, o_ins, e_ins; } orc_sc_t;

static inline int sc_of(const orc_sc_t *s, int t, int q) { return (t > 3 || q > 3) ? -1 : (t == q ? s->a : -s->b); } /* bwa.c:109-118 */
static inline int imax(int x, int y) { return x > y ? x : y; }
static inline int imin(int x, int y) { return x < y ? x : y; }

/* ------------------------------------------------------------------ banded, z-dropped seed extension ---- */
/* out5 = {qle, tle, gtle, gscore, max_off}; *cells = number of inner-loop iterations (in-band cells) */
int orc_extend(const orc_sc_t *s, int qlen, const uint8_t *q, int tlen, const uint8_t *t, int w, int end_bonus, int zdrop, int h0,
			   int *out5, uint64_t *cells)
{
	int *H = calloc(qlen + 2, sizeof(int)), *E = calloc(qlen + 2, sizeof(int)); /* H[j] = H(i-1,j-1), E[j] = E(i,j)   (:425) */
	int oe_del = s->o_del + s->e_del, oe_ins = s->o_ins + s->e_ins;
	int i, j, best = h0, best_i = -1, best_j = -1, best_ie = -1, gscore = -1, max_off = 0, beg = 0, end = qlen, lim;
	int top = imax(imax(s->a, -s->b), -1);
	uint64_t nc = 0;
	/* first row (:395-397) */
	H[0] = h0;
	if (qlen >= 1) H[1] = h0 > oe_ins ? h0 - oe_ins : 0;
	for (j = 2; j <= qlen && H[j-1] > s->e_ins; ++j) H[j] = H[j-1] - s->e_ins;
	/* band limit from the longest affordable gap (:402-407) */
	lim = (int)((double)(qlen * top + end_bonus - s->o_ins) / s->e_ins + 1.); w = imin(w, imax(lim, 1));
	lim = (int)((double)(qlen * top + end_bonus - s->o_del) / s->e_del + 1.); w = imin(w, imax(lim, 1));
	for (i = 0; i < tlen; ++i) {
		int f = 0, left, rowmax = 0, rowj = -1;
		beg = imax(beg, i - w); end = imin(imin(end, i + w + 1), qlen);
		left = beg == 0 ? imax(h0 - (s->o_del + s->e_del * (i + 1)), 0) : 0;
		for (j = beg; j < end; ++j) {
			int diag = H[j], e = E[j], M, h;
			H[j] = left;
			M = diag ? diag + sc_of(s, t[i], q[j]) : 0;            /* M and H kept apart (:433) */
			h = imax(imax(M, e), f);
			left = h;
			if (h >= rowmax) rowj = j;                                /* last index wins ties (:437) */
			rowmax = imax(rowmax, h);
			E[j] = imax(e - s->e_del, imax(M - oe_del, 0));
			f = imax(f - s->e_ins, imax(M - oe_ins, 0));
			++nc;
		}
		H[end] = left; E[end] = 0;
		if (j == qlen) { if (gscore <= left) best_ie = i; gscore = imax(gscore, left); }
		if (rowmax == 0) break;
		if (rowmax > best) {
			best = rowmax; best_i = i; best_j = rowj;
			max_off = imax(max_off, abs(rowj - i));
		} else if (zdrop > 0) {
			int di = i - best_i, dj = rowj - best_j;
			if (di > dj) { if (best - rowmax - (di - dj) * s->e_del > zdrop) break; }
			else if (best - rowmax - (dj - di) * s->e_ins > zdrop) break;
		}
		for (j = beg; j < end && H[j] == 0 && E[j] == 0; ++j) {}
		beg = j;
		for (j = end; j >= beg && H[j] == 0 && E[j] == 0; --j) {}
		end = imin(j + 2, qlen);
	}
	free(H); free(E);
	out5[0] = best_j + 1; out5[1] = best_i + 1; out5[2] = best_ie + 1; out5[3] = gscore; out5[4] = max_off;
	if (cells) *cells = nc;
	return best;
}

/* ------------------------------------------------------------------ banded global alignment + traceback ---- */
#define NEG_INF (-0x40000000)
static int cig_push(uint32_t *c, int n, int cap, int op, int len)
{
	if (n > 0 && (int)(c[n-1] & 0xf) == op) { c[n-1] += (uint32_t)len << 4; return n; }
	if (n < cap) c[n] = (uint32_t)len << 4 | op;
	return n + 1;
}
/* returns the score; when cigar != NULL also the CIGAR (len<<4|op, M=0 I=1 D=2), *n_cigar ops (<= cap written) */
int orc_global(const orc_sc_t *s, int qlen, const uint8_t *q, int tlen, const uint8_t *t, int w, uint32_t *cigar, int cap, int *n_cigar)
{
	int *H = malloc((qlen + 2) * sizeof(int)), *E = malloc((qlen + 2) * sizeof(int));
	int oe_del = s->o_del + s->e_del, oe_ins = s->o_ins + s->e_ins, i, j, score;
	int n_col = imin(qlen, 2 * w + 1);
	uint8_t *z = cigar ? malloc((size_t)n_col * (tlen > 0 ? tlen : 1)) : 0;
	H[0] = 0; E[0] = NEG_INF;
	for (j = 1; j <= qlen; ++j) { H[j] = j <= w ? -(s->o_ins + s->e_ins * j) : NEG_INF; E[j] = NEG_INF; }
	for (i = 0; i < tlen; ++i) {
		int beg = imax(i - w, 0), end = imin(i + w + 1, qlen), f = NEG_INF;
		int left = beg == 0 ? -(s->o_del + s->e_del * (i + 1)) : NEG_INF;
		for (j = beg; j < end; ++j) {
			int m = H[j] + sc_of(s, t[i], q[j]), e = E[j], h, d, x;
			H[j] = left;
			d = m >= e ? 0 : 1; h = imax(m, e);
			if (h < f) { d = 2; h = f; }
			left = h;
			x = m - oe_del; e -= s->e_del;
			if (e > x) d |= 1 << 2; else e = x;
			E[j] = e;
			x = m - oe_ins; f -= s->e_ins;
			if (f > x) d |= 2 << 4; else f = x;
			if (z) z[(size_t)i * n_col + (j - beg)] = (uint8_t)d;
		}
		H[end] = left; E[end] = NEG_INF;
	}
	score = H[qlen];
	if (cigar) {
		int n = 0, state = 0, k;
		i = tlen - 1; k = imin(i + w + 1, qlen) - 1;
		while (i >= 0 && k >= 0) {
			state = z[(size_t)i * n_col + (k - imax(i - w, 0))] >> (state << 1) & 3;
			if (state == 0) { n = cig_push(cigar, n, cap, 0, 1); --i; --k; }
			else if (state == 1) { n = cig_push(cigar, n, cap, 2, 1); --i; }
			else { n = cig_push(cigar, n, cap, 1, 1); --k; }
		}
		if (i >= 0) n = cig_push(cigar, n, cap, 2, i + 1);
		if (k >= 0) n = cig_push(cigar, n, cap, 1, k + 1);
		if (n <= cap) for (i = 0; i < n >> 1; ++i) { uint32_t x = cigar[i]; cigar[i] = cigar[n-1-i]; cigar[n-1-i] = x; }
		*n_cigar = n;
	}
	free(H); free(E); free(z);
	return score;
}

/* ------------------------------------------------------------------ local SW with the striped-SIMD semantics ----
 * The reference's local aligner is Farrar-striped SSE2 (16 x u8 or 8 x i16 lanes).  Two of its properties leak
 * into results and are modelled lane by lane here: E is derived from H before the lazy-F correction
 * (ksw.c:165-168 vs :177-188) and the first-pass F restarts from zero in every lane (:147,170-172), being carried
 * across lanes only by the lazy-F rounds, which stop as soon as F cannot raise any H (:185-186, :293).
 */
#define XBYTE 0x10000
#define XSTOP 0x20000
#define XSUBO 0x40000
#define XSTART 0x80000
typedef struct { int score, te, qe, score2, te2, tb, qb; } orc_loc_t;

static int sat(int v, int hi) { return v < 0 ? 0 : (v > hi ? hi : v); }

static orc_loc_t striped_pass(const orc_sc_t *s, int bytes, int qlen, const uint8_t *q, int tlen, const uint8_t *t, int xtra)
{
	const int P = bytes == 1 ? 16 : 8, HI = bytes == 1 ? 255 : 32767;
	const int slen = (qlen + P - 1) / P, n = slen * P;
	const int shift = bytes == 1 ? imax(imax(s->b, 1), -s->a) : 0;   /* 256 - min(mat) as uint8 (:80-85); min(mat) = min(a,-b,-1) */
	const int top = imax(imax(s->a, -s->b), -1);
	const int oe_del = s->o_del + s->e_del, oe_ins = s->o_ins + s->e_ins;
	int *H0 = calloc(n, sizeof(int)), *H1 = calloc(n, sizeof(int)), *E = calloc(n, sizeof(int)), *Hm = calloc(n, sizeof(int));
	int *F = malloc(P * sizeof(int)), *hv = malloc(P * sizeof(int)), *mx = malloc(P * sizeof(int));
	uint64_t *b = 0; int n_b = 0, m_b = 0;
	int minsc = (xtra & XSUBO) ? xtra & 0xffff : 0x10000, endsc = (xtra & XSTOP) ? xtra & 0xffff : 0x10000;
	int i, j, k, gmax = 0, te = -1;
	orc_loc_t r = { 0, -1, -1, -1, -1, -1, -1 };
	/* cell (vector j, lane k) holds query position j + k*slen (:87-88); index = j*P + k */
	for (i = 0; i < tlen; ++i) {
		int rowmax = 0, done = 0, *tmp;
		for (k = 0; k < P; ++k) { F[k] = 0; mx[k] = 0; }
		hv[0] = 0;
		for (k = 1; k < P; ++k) hv[k] = H0[(slen - 1) * P + k - 1];                      /* byte-shifted last vector (:148-149) */
		for (j = 0; j < slen; ++j) {
			for (k = 0; k < P; ++k) {
				int pos = j + k * slen, sc = pos < qlen ? sc_of(s, t[i], q[pos]) : 0, h, e, x;
				if (bytes == 1) h = sat(sat(hv[k] + sc + shift, HI) - shift, HI);         /* adds_epu8 then subs_epu8 */
				else { h = hv[k] + sc; h = h > 32767 ? 32767 : (h < -32768 ? -32768 : h); }
				e = E[j * P + k];
				h = imax(imax(h, e), F[k]);
				mx[k] = imax(mx[k], h);
				H1[j * P + k] = h;
				x = sat(h - oe_del, HI); e = imax(sat(e - s->e_del, HI), x);
				E[j * P + k] = e;
				x = sat(h - oe_ins, HI); F[k] = imax(sat(F[k] - s->e_ins, HI), x);
				hv[k] = H0[j * P + k];
			}
		}
		for (k = 0; k < 16 && !done; ++k) {                                             /* lazy F: 16 rounds in both widths (:177,285) */
			int l;
			for (l = P - 1; l > 0; --l) F[l] = F[l-1];
			F[0] = 0;
			for (j = 0; j < slen; ++j) {
				int any = 0;
				for (l = 0; l < P; ++l) {
					int h = imax(H1[j * P + l], F[l]);
					H1[j * P + l] = h;
					h = sat(h - oe_ins, HI);
					F[l] = sat(F[l] - s->e_ins, HI);
					if (F[l] > h) any = 1;
				}
				if (!any) { done = 1; break; }
			}
		}
		for (k = 0; k < P; ++k) rowmax = imax(rowmax, mx[k]);
		if (rowmax >= minsc) {                                                          /* row maxima for the 2nd-best score (:192-200) */
			if (n_b == 0 || (int32_t)b[n_b-1] + 1 != i) {
				if (n_b == m_b) { m_b = m_b ? m_b << 1 : 8; b = realloc(b, 8 * m_b); }
				b[n_b++] = (uint64_t)rowmax << 32 | i;
			} else if ((int)(b[n_b-1] >> 32) < rowmax) b[n_b-1] = (uint64_t)rowmax << 32 | i;
		}
		if (rowmax > gmax) {
			gmax = rowmax; te = i;
			memcpy(Hm, H1, n * sizeof(int));
			if ((bytes == 1 && gmax + shift >= 255) || gmax >= endsc) break;
		}
		tmp = H0; H0 = H1; H1 = tmp;
	}
	r.score = (bytes == 1 && gmax + shift >= 255) ? 255 : gmax;
	r.te = te;
	if (bytes == 2 || r.score != 255) {
		int best = -1, low, high;
		for (i = 0; i < n; ++i) {                                                       /* query end: smallest position among maxima (:212-216) */
			int pos = i / P + i % P * slen;
			if (Hm[i] > best) { best = Hm[i]; r.qe = pos; }
			else if (Hm[i] == best && pos < r.qe) r.qe = pos;
		}
		if (b) {
			k = (r.score + top - 1) / top;
			low = te - k; high = te + k;
			for (i = 0; i < n_b; ++i) {
				int e = (int32_t)b[i];
				if ((e < low || e > high) && (int)(b[i] >> 32) > r.score2) { r.score2 = (int)(b[i] >> 32); r.te2 = e; }
			}
		}
	}
	free(H0); free(H1); free(E); free(Hm); free(F); free(hv); free(mx); free(b);
	return r;
}

static void rev_bytes(uint8_t *a, int n) { int i; for (i = 0; i < n >> 1; ++i) { uint8_t x = a[i]; a[i] = a[n-1-i]; a[n-1-i] = x; } }

/* out7 = {score, te, qe, score2, te2, tb, qb} */
void orc_local(const orc_sc_t *s, int qlen, const uint8_t *q_, int tlen, const uint8_t *t_, int xtra, int *out7)
{
	int bytes = (xtra & XBYTE) ? 1 : 2;
	uint8_t *q = malloc(qlen + 1), *t = malloc(tlen + 1);
	orc_loc_t r, rr;
	memcpy(q, q_, qlen); memcpy(t, t_, tlen);
	r = striped_pass(s, bytes, qlen, q, tlen, t, xtra);
	if ((xtra & XSTART) && !((xtra & XSUBO) && r.score < (xtra & 0xffff))) {
		/* start: rerun on the reversed prefixes until the same score is reached (:357-364); the target keeps its full
		 * length with only the first te+1 symbols reversed, exactly as the reference passes it */
		rev_bytes(q, r.qe + 1); rev_bytes(t, r.te + 1);
		rr = striped_pass(s, bytes, r.qe + 1, q, tlen, t, XSTOP | r.score);
		if (r.score == rr.score) { r.tb = r.te - rr.te; r.qb = r.qe - rr.qe; }
	}
	out7[0] = r.score; out7[1] = r.te; out7[2] = r.qe; out7[3] = r.score2; out7[4] = r.te2; out7[5] = r.tb; out7[6] = r.qb;
	free(q); free(t);
}
