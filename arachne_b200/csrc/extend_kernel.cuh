// k_extend2: mem_chain2aln + mem_sort_dedup_patch (bwa/bwamem.c:632-786, 437-489) for a whole batch, one read per LANE.
//
// ksw_extend2 (bwa/ksw.c:380-479) trims its band after every row from the complete previous row, so the rows of one
// alignment are strictly sequential, and at 2x150 bp a row is only ~40-130 cells: a warp-per-alignment kernel spends
// most of its instructions on per-row overhead (scan, reductions, band bookkeeping).  Here a lane runs the reference's
// scalar recurrence on its own alignment, so a cell costs ~18 integer instructions and nothing else, and convergence
// is forced where the work is: every iteration of the outer loop is "each live lane computes up to EXT_CHUNK cells of
// its current row"; row ends, band retries and the chain2aln driver (seed order, containment tests, left/right
// extension, region records) are divergent bookkeeping between chunks.
//
//   * DP columns live in shared memory, one 32-bit word per column, interleaved across the block's lanes (column j of
//     lane t at word j*blockDim+t: bank = t, conflict-free whatever j each lane is at).  The word packs
//     h (13 bits) | e (13 bits) | query base (3 bits): one LDS gives H(i-1,j-1), E(i,j) and q[j].
//   * max-plus steps use DPX instructions (__viaddmax_s32_relu / __vimax3_s32 / __vibmax_s32: VIADDMNMX, VIMNMX3).
//   * the target base of a row is read straight from the 2-bit packed reference (no window copy).
//   * stale columns beyond `end` keep their first-row fill exactly as the reference leaves them.
#pragma once
#include "extend.cuh"

namespace arachne {

constexpr int EXT_CHUNK = 16;
#ifndef ARACHNE_EXT_GROUP
#define ARACHNE_EXT_GROUP 4
#endif
constexpr int EXT_GROUP = ARACHNE_EXT_GROUP;  // cells per rolled group of the DP block (2 / 4 / 8 measured: see profiles)
constexpr int EXT_BLOCK = 64;
constexpr int COL_HMASK = 0x1fff;

// Scalar context for the (rare) region patching path of mem_sort_dedup_patch: one lane does everything.
struct ThreadCtx {
    int* hbuf;      // 2*(maxlen+2) ints of private global scratch
    uint8_t* rbuf;  // private window buffer
    __device__ bool leader() const { return true; }
    __device__ void sync() const {}
    __device__ int bcast(int v) const { return v; }
    __device__ void fetch_ref(const RefView& ref, int64_t beg, int64_t end, uint8_t* dst) const {
        for (int64_t k = 0; k < end - beg; ++k) dst[k] = ref_base(ref, beg + k);
    }
    // ksw_global2 (bwa/ksw.c:504-606), score only
    __device__ __noinline__ int ksw_global(const MemOpt& o, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts, int w,
                                           uint32_t*, int, int*) const {
        const int NEG = -0x40000000;
        int* H = hbuf;
        int* E = hbuf + qlen + 2;
        const int oe_del = o.o_del + o.e_del, oe_ins = o.o_ins + o.e_ins;
        H[0] = 0; E[0] = NEG;
        for (int j = 1; j <= qlen; ++j) { H[j] = j <= w ? -(o.o_ins + o.e_ins * j) : NEG; E[j] = NEG; }
        for (int i = 0; i < tlen; ++i) {
            int beg = i > w ? i - w : 0, end = i + w + 1 < qlen ? i + w + 1 : qlen, f = NEG;
            int left = beg == 0 ? -(o.o_del + o.e_del * (i + 1)) : NEG;
            const int tb = t[(long)i * ts];
            for (int j = beg; j < end; ++j) {
                int qb = q[(long)j * qs];
                int m = H[j] + ((tb > 3 || qb > 3) ? -1 : (tb == qb ? o.a : -o.b)), e = E[j];
                H[j] = left;
                int h = m > e ? m : e;
                h = h > f ? h : f;
                left = h;
                int x = m - oe_del; e -= o.e_del; E[j] = e > x ? e : x;
                x = m - oe_ins; f -= o.e_ins; f = f > x ? f : x;
            }
            H[end] = left; E[end] = NEG;
        }
        return H[qlen];
    }
};

// mem_sort_dedup_patch once per read, kept out of line: the kernel's hot loop should not share instruction-cache lines with it
__device__ __noinline__ int sort_dedup_patch_cold(ThreadCtx tctx, const MemOpt& opt, const RefView& ref, const uint8_t* query, int n_av, AlnReg* av, int rcap) {
    return sort_dedup_patch(tctx, opt, ref, query, n_av, av, tctx.rbuf, rcap);
}

enum { XS_FETCH = 0, XS_CHAIN, XS_SEED, XS_SETUP, XS_INIT, XS_ROWSTART, XS_CELLS, XS_ROWEND, XS_POST, XS_REGION, XS_FINISH, XS_DONE };

struct ExtArgs {
    const uint8_t* seq; const int64_t* off; int n_reads;
    const unsigned long long* sa_off;
    ChainBufs b;
    uint64_t* srt; AlnReg* regs; int* n_regs;
    int* hscr; uint8_t* rscr; int rcap; int maxlen;
    int ncols;  // shared-memory column words per lane: (maxlen - min_seed_len) + EXT_CHUNK, a flank is never longer than len - seed
    unsigned int* next; Counters* cnt;
};

__global__ void __launch_bounds__(EXT_BLOCK, 6) k_extend2(MemOpt opt, RefView ref, ExtArgs A) {
    extern __shared__ uint32_t cols[];  // [ncols][EXT_BLOCK]
    const int tix = threadIdx.x;
    const long gtid = (long)blockIdx.x * blockDim.x + tix;
    uint32_t* col = cols + tix;
#define COL(j) col[(j) * EXT_BLOCK]
    const int e_del = opt.e_del, e_ins = opt.e_ins, oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    const int sa = opt.a, snb = -opt.b;
    ThreadCtx tctx; tctx.hbuf = A.hscr + gtid * (2 * (A.maxlen + 2)); tctx.rbuf = A.rscr + gtid * A.rcap;

    int st = XS_FETCH;
    // read / chain / seed level
    int r = 0, len = 0, nk = 0, ci = 0, n_av = 0, cap_av = 0, k = 0, cn = 0, crid = 0;
    float cfrac = 0.f;
    unsigned long long o = 0;
    const uint8_t* query = A.seq;
    const Seed* seeds = nullptr;
    AlnReg* av = nullptr;
    uint64_t* srt = nullptr;
    int64_t rmax0 = 0, rmax1 = 0;
    int64_t s_rbeg = 0; int s_qbeg = 0, s_len = 0;
    // region being built
    int64_t a_rb = 0, a_re = 0; int a_qb = 0, a_qe = 0, a_score = 0, a_truesc = 0, aw0 = 0, aw1 = 0;
    // extension level
    int dir = 0, tr = 0, prev_sc = 0, sc0 = 0;
    // DP level (ksw_extend2 locals)
    int qlen = 0, tlen = 0, w = 0, h0 = 0, end_bonus = 0, i = 0, j = 0, beg = 0, end = 0, h1 = 0, f = 0, m = 0, mj = -1;
    int mx = 0, max_i = -1, max_j = -1, max_ie = -1, gscore = -1, max_off = 0, tbv = 0;
    int64_t tpos = 0; int tstep = 1;  // target base of row i sits at forward-reverse coordinate tpos + i*tstep
    int64_t tfp = 0, tw_idx = -1; int tfstep = 1; bool tcomp = false; uint32_t tw = 0;
    unsigned long long cells = 0, calls = 0;
    int err = 0;
    int firstnz = 0x7fffffff, lastnz = -1;  // first / last column of the current row whose stored (h, e) is non-zero
    const uint8_t* iq = A.seq; int iqs = 1;  // first-row fill request (XS_INIT): query pointer and stride of this ksw_extend2 call

    for (;;) {
        // ------------------------------------------------------------ divergent driver: advance until this lane has cells to compute
        while (st != XS_DONE && st != XS_INIT && !(st == XS_CELLS && j < end)) {
            switch (st) {
            case XS_FETCH: {
                unsigned t = atomicAdd(A.next, 1u);
                if (t >= (unsigned)A.n_reads) { st = XS_DONE; break; }
                r = (int)t; nk = A.b.n_chain[r]; n_av = 0;
                if (nk == 0) { A.n_regs[r] = 0; break; }
                o = A.sa_off[r];
                len = (int)(A.off[r + 1] - A.off[r]);
                query = A.seq + A.off[r];
                av = A.regs + o; cap_av = A.b.n_kseed[r]; srt = A.srt + o;
                ci = 0; err = 0;
                st = XS_CHAIN;
            } break;
            case XS_CHAIN: {  // head of mem_chain2aln for the next kept chain (bwamem.c:640-667)
                if (ci == nk || err) { st = XS_FINISH; break; }
                const Chain c = A.b.chains[o + A.b.order[o + ci]];
                ++ci;
                cn = c.n; crid = c.rid; cfrac = c.frac_rep; seeds = A.b.kseeds + o + c.first_seed;
                if (cn == 0) break;
                const int64_t l_pac = ref.l_pac;
                rmax0 = l_pac << 1; rmax1 = 0;
                for (int x = 0; x < cn; ++x) {
                    const Seed t = seeds[x];
                    int64_t b = t.rbeg - (t.qbeg + cal_max_gap(opt, t.qbeg));
                    int64_t e = t.rbeg + t.len + ((len - t.qbeg - t.len) + cal_max_gap(opt, len - t.qbeg - t.len));
                    rmax0 = rmax0 < b ? rmax0 : b;
                    rmax1 = rmax1 > e ? rmax1 : e;
                }
                rmax0 = rmax0 > 0 ? rmax0 : 0;
                rmax1 = rmax1 < l_pac << 1 ? rmax1 : l_pac << 1;
                if (rmax0 < l_pac && l_pac < rmax1) {
                    if (seeds[0].rbeg < l_pac) rmax1 = l_pac;
                    else rmax0 = l_pac;
                }
                fetch_window(ref, &rmax0, seeds[0].rbeg, &rmax1);
                for (int x = 0; x < cn; ++x) srt[x] = (uint64_t)seeds[x].score << 32 | (uint32_t)x;
                ks_introsort((size_t)cn, srt, [](uint64_t a, uint64_t b) { return a < b; });
                k = cn - 1;
                st = XS_SEED;
            } break;
            case XS_SEED: {  // bwamem.c:669-713
                if (k < 0) { st = XS_CHAIN; break; }
                const Seed s = seeds[(uint32_t)srt[k]];
                s_rbeg = s.rbeg; s_qbeg = s.qbeg; s_len = s.len;
                int x;
                for (x = 0; x < n_av; ++x) {  // has this seed been covered by an earlier extension of this read?
                    const AlnReg& p = av[x];
                    int64_t rd;
                    int qd, ww, max_gap;
                    if (s.rbeg < p.rb || s.rbeg + s.len > p.re || s.qbeg < p.qb || s.qbeg + s.len > p.qe) continue;
                    if (s.len - p.seedlen0 > .1 * len) continue;
                    qd = s.qbeg - p.qb; rd = s.rbeg - p.rb;
                    max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                    ww = max_gap < p.w ? max_gap : p.w;
                    if (qd - rd < ww && rd - qd < ww) break;
                    qd = p.qe - (s.qbeg + s.len); rd = p.re - (s.rbeg + s.len);
                    max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
                    ww = max_gap < p.w ? max_gap : p.w;
                    if (qd - rd < ww && rd - qd < ww) break;
                }
                if (x < n_av) {  // (almost) contained: extend only if an overlapping seed hints at a different alignment
                    for (x = k + 1; x < cn; ++x) {
                        if (srt[x] == 0) continue;
                        const Seed t = seeds[(uint32_t)srt[x]];
                        if (t.len < s.len * .95) continue;
                        if (s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg) break;
                        if (t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg) break;
                    }
                    if (x == cn) { srt[k] = 0; --k; break; }
                }
                if (n_av >= cap_av) { err = -3; st = XS_FINISH; break; }
                aw0 = aw1 = opt.w;
                a_score = a_truesc = -1;
                if (s_qbeg) { dir = 0; tr = 0; st = XS_SETUP; break; }
                a_score = a_truesc = s_len * sa; a_qb = 0; a_rb = s_rbeg;
                if (s_qbeg + s_len != len) { dir = 1; tr = 0; sc0 = a_score; st = XS_SETUP; }
                else { a_qe = len; a_re = s_rbeg + s_len; st = XS_REGION; }
            } break;
            case XS_SETUP: {  // one ksw_extend2 call: first row (ksw.c:395-397), band limit (:402-407)
                prev_sc = a_score;
                const uint8_t* q; int qs;
                if (dir == 0) {
                    qlen = s_qbeg; tlen = (int)(s_rbeg - rmax0); w = opt.w << tr; aw0 = w; end_bonus = opt.pen_clip5; h0 = s_len * sa;
                    tpos = s_rbeg - 1; tstep = -1; q = query + s_qbeg - 1; qs = -1;
                } else {
                    const int qe = s_qbeg + s_len;
                    qlen = len - qe; tlen = (int)(rmax1 - (s_rbeg + s_len)); w = opt.w << tr; aw1 = w; end_bonus = opt.pen_clip3; h0 = sc0;
                    tpos = s_rbeg + s_len; tstep = 1; q = query + qe; qs = 1;
                }
                if (qlen + EXT_CHUNK > A.ncols) { err = -3; st = XS_FINISH; break; }  // cannot happen with seeds >= min_seed_len
                ++calls;
                {   // forward-strand coordinate of row 0 and its step: the window never crosses l_pac (fetch_window)
                    tcomp = tpos >= ref.l_pac;
                    tfp = tcomp ? (ref.l_pac << 1) - 1 - tpos : tpos;
                    tfstep = tcomp ? -tstep : tstep;
                    tw_idx = -1;
                }
                iq = q; iqs = qs;  // the first row itself (ksw.c:395-397) is filled by the whole warp below
                int top = sa > snb ? sa : snb; top = top > -1 ? top : -1;
                // (int)((double)x / e + 1.) == (x + e) / e in C integer division (both truncate the same real number toward zero)
                int lim = (qlen * top + end_bonus - opt.o_ins + e_ins) / e_ins;
                lim = lim > 1 ? lim : 1; w = w < lim ? w : lim;
                lim = (qlen * top + end_bonus - opt.o_del + e_del) / e_del;
                lim = lim > 1 ? lim : 1; w = w < lim ? w : lim;
                mx = h0; max_i = max_j = max_ie = -1; gscore = -1; max_off = 0;
                beg = 0; end = qlen; i = 0;
                st = XS_INIT;
            } break;
            case XS_ROWSTART: {
                if (i >= tlen) { st = XS_POST; break; }
                {   // target base of row i from a cached 16-base word of the packed reference (bns_get_seq, bntseq.c:398-419)
                    const int64_t fp = tfp + (int64_t)i * tfstep;
                    if ((fp >> 4) != tw_idx) { tw_idx = fp >> 4; tw = __ldg(reinterpret_cast<const uint32_t*>(ref.pac) + tw_idx); }
                    int tb = (int)(tw >> ((((unsigned)fp >> 2) & 3u) << 3) >> ((~(unsigned)fp & 3u) << 1)) & 3;
                    tbv = tcomp ? 3 - tb : tb;
                }
                if (beg < i - w) beg = i - w;
                if (end > i + w + 1) end = i + w + 1;
                if (end > qlen) end = qlen;
                h1 = 0;
                if (beg == 0) { h1 = h0 - (opt.o_del + e_del * (i + 1)); if (h1 < 0) h1 = 0; }
                f = 0; m = 0; mj = -1; j = beg;
                firstnz = 0x7fffffff; lastnz = -1;
                cells += end > beg ? end - beg : 0;
                st = XS_CELLS;
            } break;
            case XS_CELLS:     // only reached with j >= end: the row is complete
            case XS_ROWEND: {  // ksw.c:449-469
                COL(end) = (COL(end) & (7u << 26)) | (uint32_t)h1;
                if (j == qlen) {
                    max_ie = gscore > h1 ? max_ie : i;
                    gscore = gscore > h1 ? gscore : h1;
                }
                if (m == 0) { st = XS_POST; break; }
                if (m > mx) {
                    mx = m; max_i = i; max_j = mj;
                    int d = mj - i; d = d < 0 ? -d : d;
                    max_off = max_off > d ? max_off : d;
                } else if (opt.zdrop > 0) {
                    bool drop;
                    if (i - max_i > mj - max_j) drop = mx - m - ((i - max_i) - (mj - max_j)) * e_del > opt.zdrop;
                    else drop = mx - m - ((mj - max_j) - (i - max_i)) * e_ins > opt.zdrop;
                    if (drop) { st = XS_POST; break; }
                }
                // band trimming (ksw.c:465-468) without re-reading the row: the cell loop recorded the first and last column
                // it stored a non-zero (h, e) into, and column `end` holds (h1, 0)
                beg = firstnz < end ? firstnz : end;
                const int x = h1 ? end : (lastnz >= beg ? lastnz : beg - 1);
                end = x + 2 < qlen ? x + 2 : qlen;
                ++i;
                st = XS_ROWSTART;
            } break;
            case XS_POST: {  // band retry (bwamem.c:723-734 / 751-762), then local-vs-to-end choice
                a_score = mx;
                const int aw = dir ? aw1 : aw0;
                if (!(a_score == prev_sc || max_off < (aw >> 1) + (aw >> 2)) && tr + 1 < 2) { ++tr; st = XS_SETUP; break; }
                const int qle = max_j + 1, tle = max_i + 1, gtle = max_ie + 1;
                if (dir == 0) {
                    if (gscore <= 0 || gscore <= a_score - opt.pen_clip5) { a_qb = s_qbeg - qle; a_rb = s_rbeg - tle; a_truesc = a_score; }
                    else { a_qb = 0; a_rb = s_rbeg - gtle; a_truesc = gscore; }
                    if (s_qbeg + s_len != len) { dir = 1; tr = 0; sc0 = a_score; st = XS_SETUP; }
                    else { a_qe = len; a_re = s_rbeg + s_len; st = XS_REGION; }
                } else {
                    if (gscore <= 0 || gscore <= a_score - opt.pen_clip3) { a_qe = s_qbeg + s_len + qle; a_re = s_rbeg + s_len + tle; a_truesc += a_score - sc0; }
                    else { a_qe = len; a_re = s_rbeg + s_len + gtle; a_truesc += gscore - sc0; }
                    st = XS_REGION;
                }
            } break;
            case XS_REGION: {  // bwamem.c:773-783
                AlnReg a;
                a.rb = a_rb; a.re = a_re; a.qb = a_qb; a.qe = a_qe; a.rid = crid; a.score = a_score; a.truesc = a_truesc;
                a.sub = a.alt_sc = a.csub = a.sub_n = 0; a.secondary = a.secondary_all = 0; a.n_comp_is_alt = 0; a.hash = 0;
                a.seedcov = 0;
                for (int x = 0; x < cn; ++x) {
                    const Seed t = seeds[x];
                    if (t.qbeg >= a_qb && t.qbeg + t.len <= a_qe && t.rbeg >= a_rb && t.rbeg + t.len <= a_re) a.seedcov += t.len;
                }
                a.w = aw0 > aw1 ? aw0 : aw1;
                a.seedlen0 = s_len;
                a.frac_rep = cfrac;
                av[n_av++] = a;
                --k;
                st = XS_SEED;
            } break;
            case XS_FINISH: {
                int rc = err;
                if (rc == 0) {
                    int nn = sort_dedup_patch_cold(tctx, opt, ref, query, n_av, av, A.rcap);
                    if (nn < 0) rc = nn; else n_av = nn;
                }
                if (rc != 0) { raise(A.cnt, ARACHNE_ERR_OVERFLOW, r); n_av = 0; }
                for (int x = 0; x < n_av; ++x)  // mem_align1_core tail, bwamem.c:1078-1082
                    if (av[x].rid >= 0 && ref.contigs[av[x].rid].is_alt) av[x].n_comp_is_alt |= 1 << 30;
                A.n_regs[r] = n_av;
                st = XS_FETCH;
            } break;
            default: st = XS_DONE; break;
            }
        }
        __syncwarp();
        // ------------------------------------------------------------ convergent: first rows (ksw.c:395-397) of the lanes that asked for one
        // A lane filling its own qlen+1 columns runs a ~10-instruction loop body 100+ times alone (15 % of the kernel's warp
        // instructions at 1.1 active lanes); here the warp fills them together, 32 columns per trip: h(0) = h0,
        // h(x) = max(h0 - o_ins - x*e_ins, 0), e = 0, query base in the top bits.
        {
            unsigned want = __ballot_sync(0xffffffffu, st == XS_INIT);
            const int lane = tix & 31;
            while (want) {
                const int src = __ffs(want) - 1;
                want &= want - 1;
                const uint8_t* bq = reinterpret_cast<const uint8_t*>(__shfl_sync(0xffffffffu, (unsigned long long)iq, src));
                const int bqs = __shfl_sync(0xffffffffu, iqs, src), bqlen = __shfl_sync(0xffffffffu, qlen, src), bh0 = __shfl_sync(0xffffffffu, h0, src);
                uint32_t* bcol = cols + (tix & ~31) + src;
                for (int x = lane; x <= bqlen; x += 32) {
                    int hv = bh0 - opt.o_ins - x * e_ins;
                    hv = x == 0 ? bh0 : (hv > 0 ? hv : 0);
                    const uint32_t qb = x < bqlen ? (uint32_t)__ldg(bq + (long)x * bqs) : 0u;
                    bcol[x * EXT_BLOCK] = (uint32_t)hv | qb << 26;
                }
            }
            __syncwarp();
            if (st == XS_INIT) st = XS_ROWSTART;
        }
        // ------------------------------------------------------------ convergent: up to EXT_CHUNK cells of the current row (ksw.c:424-448)
        // In groups of EXT_GROUP cells: a group's column words are loaded up front (the column array is padded by EXT_CHUNK
        // words, so reading past `end` is harmless), every cell of the group is computed, and cells at or beyond `end` are
        // simply not committed.  The group loop is rolled (the fully unrolled 16-cell block was 5.6 KB of code and the
        // kernel's top stall was instruction fetch) and stops at the longest remaining run in the warp.
        {
            const int nc = st == XS_CELLS ? (end - j < EXT_CHUNK ? end - j : EXT_CHUNK) : 0;  // >= 1 for a lane in XS_CELLS
            const int nc_warp = __reduce_max_sync(0xffffffffu, nc);
            int mkey = m << 8 | (mj & 0xff);  // row maximum and its column in one key: ties go to the larger column (:437)
            unsigned nzm = 0;                  // bit u: the (h, e) stored into column j+u is non-zero
            uint32_t* p = &COL(j);
#pragma unroll 1
            for (int g = 0; g < nc_warp; g += EXT_GROUP, p += EXT_GROUP * EXT_BLOCK) {
                if (g >= nc) continue;
                uint32_t cw[EXT_GROUP];
#pragma unroll
                for (int u = 0; u < EXT_GROUP; ++u) cw[u] = p[u * EXT_BLOCK];
#pragma unroll
                for (int u = 0; u < EXT_GROUP; ++u) {
                    const bool on = g + u < nc;
                    const int hd = (int)(cw[u] & COL_HMASK), e = (int)((cw[u] >> 13) & COL_HMASK), qb = (int)(cw[u] >> 26);
                    const int sc = qb == tbv ? sa : ((qb | tbv) > 3 ? -1 : snb);
                    const int M = hd ? hd + sc : 0;
                    const int h = __vimax3_s32(M, e, f);
                    const int e2 = __viaddmax_s32_relu(M, -oe_del, e - e_del);
                    const int f2 = __viaddmax_s32_relu(M, -oe_ins, f - e_ins);
                    const uint32_t he = (uint32_t)h1 | (uint32_t)e2 << 13;
                    nzm |= (on && he) ? 1u << (g + u) : 0u;
                    if (on) p[u * EXT_BLOCK] = he | (cw[u] & (7u << 26));
                    const int key = on ? (h << 8 | (j + g + u)) : 0;
                    mkey = mkey > key ? mkey : key;
                    f = on ? f2 : f;
                    h1 = on ? h : h1;
                }
            }
            if (nc) {
                if (nzm) {
                    if (firstnz == 0x7fffffff) firstnz = j + __ffs(nzm) - 1;
                    lastnz = j + 31 - __clz(nzm);
                }
                m = mkey >> 8;
                if (m > 0) mj = mkey & 0xff;  // with m == 0 the column is never read (ksw.c:454)
                j += nc;
            }
        }
        if (__all_sync(0xffffffffu, st == XS_DONE)) break;
    }
#undef COL
    if (cells || calls) { atomicAdd(&A.cnt->ext_cells, cells); atomicAdd(&A.cnt->ext_calls, calls); }
}

}  // namespace arachne
