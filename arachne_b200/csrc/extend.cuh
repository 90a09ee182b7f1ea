// Chain -> alignment regions: the extension *driver* and the region de-duplication, written once against a
// cooperative context `Ctx`:
//   * on the device Ctx is a warp (csrc/warp_ctx.cuh): every lane runs this scalar control flow redundantly
//     (uniform branches, broadcast loads), global stores are done by the leader lane, and the DP primitives
//     (ksw_extend / ksw_global) are warp-cooperative;
//   * in tests/hostsim Ctx is a plain scalar implementation, which pins this control logic on the CPU tier.
//
//   mem_chain2aln        bwa/bwamem.c:632-786   cal_max_gap :621-628
//   mem_sort_dedup_patch bwa/bwamem.c:437-489   mem_patch_reg :406-435
//   bns_fetch_seq        bwa/bntseq.c:421-446   (window clamp to the contig / strand)
//   bwa_gen_cigar2 (score-only use)  bwa/bwa.c:121-207
#pragma once
#include "chain.cuh"

namespace arachne {

AR_HD int cal_max_gap(const MemOpt& opt, int qlen) {
    // bwamem.c:621-628 computes (int)((double)x / e + 1.); (x + e) / e in C integer division truncates the same real number
    // toward zero (no FP64 on the device path)
    int l_del = (qlen * opt.a - opt.o_del + opt.e_del) / opt.e_del;
    int l_ins = (qlen * opt.a - opt.o_ins + opt.e_ins) / opt.e_ins;
    int l = l_del > l_ins ? l_del : l_ins;
    l = l > 1 ? l : 1;
    return l < opt.w << 1 ? l : opt.w << 1;
}

// bns_fetch_seq's clamping: [*beg,*end) is cut to the contig that holds `mid` (on mid's strand); returns rid
AR_HD int fetch_window(const RefView& ref, int64_t* beg, int64_t mid, int64_t* end) {
    int is_rev;
    if (*end < *beg) { int64_t t = *beg; *beg = *end; *end = t; }
    int rid = bns_pos2rid(ref, bns_depos(ref, mid, &is_rev));
    int64_t far_beg = ref.contigs[rid].offset;
    int64_t far_end = far_beg + ref.contigs[rid].len;
    if (is_rev) {
        int64_t tmp = far_beg;
        far_beg = (ref.l_pac << 1) - far_end;
        far_end = (ref.l_pac << 1) - tmp;
    }
    *beg = *beg > far_beg ? *beg : far_beg;
    *end = *end < far_end ? *end : far_end;
    return rid;
}

// one base of the forward-reverse reference (bns_get_seq, bntseq.c:398-419)
AR_HD uint8_t ref_base(const RefView& ref, int64_t pos) {
    if (pos >= ref.l_pac) {
        int64_t f = (ref.l_pac << 1) - 1 - pos;
        return 3 - ((ref.pac[f >> 2] >> ((~f & 3) << 1)) & 3);
    }
    return (ref.pac[pos >> 2] >> ((~pos & 3) << 1)) & 3;
}

struct ExtScratch {
    uint8_t* rseq;      // reference window buffer, >= win_cap bytes
    int win_cap;
    uint64_t* srt;      // >= seeds-in-chain entries
};

// mem_chain2aln.  seeds: the chain's seeds, contiguous [0,c.n).  av/n_av: the read's region list so far (shared by
// all chains of the read, bwamem.c:1062-1068).  Returns 0, or a negative error (scratch exhausted).
template <class Ctx>
AR_HD int chain2aln(Ctx& ctx, const MemOpt& opt, const RefView& ref, int l_query, const uint8_t* query, const Chain& c,
                    const Seed* seeds, AlnReg* av, int* n_av, int cap_av, ExtScratch& sc) {
    int i, k, rid, max_off[2], aw[2];
    int64_t l_pac = ref.l_pac, rmax[2], tmp, max = 0;
    if (c.n == 0) return 0;
    rmax[0] = l_pac << 1; rmax[1] = 0;
    for (i = 0; i < c.n; ++i) {
        const Seed& t = seeds[i];
        int64_t b = t.rbeg - (t.qbeg + cal_max_gap(opt, t.qbeg));
        int64_t e = t.rbeg + t.len + ((l_query - t.qbeg - t.len) + cal_max_gap(opt, l_query - t.qbeg - t.len));
        rmax[0] = rmax[0] < b ? rmax[0] : b;
        rmax[1] = rmax[1] > e ? rmax[1] : e;
        if (t.len > max) max = t.len;
    }
    rmax[0] = rmax[0] > 0 ? rmax[0] : 0;
    rmax[1] = rmax[1] < l_pac << 1 ? rmax[1] : l_pac << 1;
    if (rmax[0] < l_pac && l_pac < rmax[1]) {
        if (seeds[0].rbeg < l_pac) rmax[1] = l_pac;
        else rmax[0] = l_pac;
    }
    rid = fetch_window(ref, &rmax[0], seeds[0].rbeg, &rmax[1]);
    (void)rid;
    if (rmax[1] - rmax[0] > sc.win_cap) return -2;
    ctx.fetch_ref(ref, rmax[0], rmax[1], sc.rseq);
    const uint8_t* rseq = sc.rseq;

    if (ctx.leader()) {
        for (i = 0; i < c.n; ++i) sc.srt[i] = (uint64_t)seeds[i].score << 32 | (uint32_t)i;
        ks_introsort((size_t)c.n, sc.srt, [](uint64_t a, uint64_t b) { return a < b; });
    }
    ctx.sync();

    for (k = c.n - 1; k >= 0; --k) {
        const Seed& s = seeds[(uint32_t)sc.srt[k]];
        int nav = *n_av;
        for (i = 0; i < nav; ++i) {  // has this seed been covered by an earlier extension of this read?
            const AlnReg& p = av[i];
            int64_t rd;
            int qd, w, max_gap;
            if (s.rbeg < p.rb || s.rbeg + s.len > p.re || s.qbeg < p.qb || s.qbeg + s.len > p.qe) continue;
            if (s.len - p.seedlen0 > .1 * l_query) continue;
            qd = s.qbeg - p.qb; rd = s.rbeg - p.rb;
            max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
            w = max_gap < p.w ? max_gap : p.w;
            if (qd - rd < w && rd - qd < w) break;
            qd = p.qe - (s.qbeg + s.len); rd = p.re - (s.rbeg + s.len);
            max_gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
            w = max_gap < p.w ? max_gap : p.w;
            if (qd - rd < w && rd - qd < w) break;
        }
        if (i < nav) {  // (almost) contained: extend only if an overlapping seed hints at a different alignment
            for (i = k + 1; i < c.n; ++i) {
                if (sc.srt[i] == 0) continue;
                const Seed& t = seeds[(uint32_t)sc.srt[i]];
                if (t.len < s.len * .95) continue;
                if (s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg) break;
                if (t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg) break;
            }
            if (i == c.n) {
                if (ctx.leader()) sc.srt[k] = 0;
                ctx.sync();
                continue;
            }
        }
        if (nav >= cap_av) return -3;
        AlnReg a;
        a.rb = a.re = 0; a.qb = a.qe = 0; a.sub = a.alt_sc = a.csub = a.sub_n = 0; a.seedcov = 0; a.secondary = a.secondary_all = 0;
        a.seedlen0 = 0; a.n_comp_is_alt = 0; a.frac_rep = 0.f; a.hash = 0;
        a.w = aw[0] = aw[1] = opt.w;
        a.score = a.truesc = -1;
        a.rid = c.rid;

        if (s.qbeg) {  // left extension: both sequences reversed, read backwards from the seed
            int qle = 0, tle = 0, gtle = 0, gscore = 0;
            tmp = s.rbeg - rmax[0];
            for (i = 0; i < 2; ++i) {  // MAX_BAND_TRY
                int prev = a.score;
                aw[0] = opt.w << i;
                a.score = ctx.ksw_extend(opt, s.qbeg, query + s.qbeg - 1, -1, (int)tmp, rseq + tmp - 1, -1, aw[0], opt.pen_clip5,
                                         opt.zdrop, s.len * opt.a, &qle, &tle, &gtle, &gscore, &max_off[0]);
                if (a.score == prev || max_off[0] < (aw[0] >> 1) + (aw[0] >> 2)) break;
            }
            if (gscore <= 0 || gscore <= a.score - opt.pen_clip5) {
                a.qb = s.qbeg - qle; a.rb = s.rbeg - tle;
                a.truesc = a.score;
            } else {
                a.qb = 0; a.rb = s.rbeg - gtle;
                a.truesc = gscore;
            }
        } else { a.score = a.truesc = s.len * opt.a; a.qb = 0; a.rb = s.rbeg; }

        if (s.qbeg + s.len != l_query) {  // right extension
            int qle = 0, tle = 0, gtle = 0, gscore = 0, sc0 = a.score;
            int qe = s.qbeg + s.len;
            int64_t re = s.rbeg + s.len - rmax[0];
            for (i = 0; i < 2; ++i) {
                int prev = a.score;
                aw[1] = opt.w << i;
                a.score = ctx.ksw_extend(opt, l_query - qe, query + qe, 1, (int)(rmax[1] - rmax[0] - re), rseq + re, 1, aw[1],
                                         opt.pen_clip3, opt.zdrop, sc0, &qle, &tle, &gtle, &gscore, &max_off[1]);
                if (a.score == prev || max_off[1] < (aw[1] >> 1) + (aw[1] >> 2)) break;
            }
            if (gscore <= 0 || gscore <= a.score - opt.pen_clip3) {
                a.qe = qe + qle; a.re = rmax[0] + re + tle;
                a.truesc += a.score - sc0;
            } else {
                a.qe = l_query; a.re = rmax[0] + re + gtle;
                a.truesc += gscore - sc0;
            }
        } else { a.qe = l_query; a.re = s.rbeg + s.len; }

        for (i = 0, a.seedcov = 0; i < c.n; ++i) {
            const Seed& t = seeds[i];
            if (t.qbeg >= a.qb && t.qbeg + t.len <= a.qe && t.rbeg >= a.rb && t.rbeg + t.len <= a.re) a.seedcov += t.len;
        }
        a.w = aw[0] > aw[1] ? aw[0] : aw[1];
        a.seedlen0 = s.len;
        a.frac_rep = c.frac_rep;
        if (ctx.leader()) av[nav] = a;
        *n_av = nav + 1;
        ctx.sync();
    }
    return 0;
}

// bwa_gen_cigar2 used score-only by mem_patch_reg (n_cigar == NM == NULL): returns 1 and *score on success, 0 when the
// reference rejects the request (bwa.c:133-135).  scratch: rseq buffer of >= re-rb bytes.
template <class Ctx>
AR_HD int gen_score(Ctx& ctx, const MemOpt& opt, const RefView& ref, int w_, int l_query, const uint8_t* query, int64_t rb, int64_t re,
                    int* score, uint8_t* rbuf, int rcap) {
    int64_t l_pac = ref.l_pac;
    if (l_query <= 0 || rb >= re || (rb < l_pac && re > l_pac)) return 0;
    // bns_get_seq clamps to [0, 2*l_pac); a clamp means rlen != re-rb and the reference bails out
    if (rb < 0 || re > (l_pac << 1)) return 0;
    int64_t rlen = re - rb;
    if (rlen > rcap) return -2;
    ctx.fetch_ref(ref, rb, re, rbuf);
    int rev = rb >= l_pac;  // both sequences reversed so that indels are left-aligned on the forward strand
    if (l_query == rlen && w_ == 0) {
        int s = 0;
        for (int i = 0; i < l_query; ++i) s += opt.mat[rbuf[rev ? rlen - 1 - i : i] * 5 + query[rev ? l_query - 1 - i : i]];
        *score = s;
        return 1;
    }
    int w, max_gap, max_ins, max_del, min_w;
    max_ins = (int)((double)(((l_query + 1) >> 1) * opt.mat[0] - opt.o_ins) / opt.e_ins + 1.);
    max_del = (int)((double)(((l_query + 1) >> 1) * opt.mat[0] - opt.o_del) / opt.e_del + 1.);
    max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    int dl = (int)rlen - l_query; dl = dl < 0 ? -dl : dl;
    w = (max_gap + dl + 1) >> 1;
    w = w < w_ ? w : w_;
    min_w = dl + 3;
    w = w > min_w ? w : min_w;
    *score = ctx.ksw_global(opt, l_query, rev ? query + l_query - 1 : query, rev ? -1 : 1, (int)rlen, rev ? rbuf + rlen - 1 : rbuf,
                            rev ? -1 : 1, w, nullptr, 0, nullptr);
    return 1;
}

// mem_patch_reg
template <class Ctx>
AR_HD int patch_reg(Ctx& ctx, const MemOpt& opt, const RefView& ref, const uint8_t* query, const AlnReg& a, const AlnReg& b, int* w_out,
                    uint8_t* rbuf, int rcap) {
    int w, score = 0, q_s, r_s;
    double r;
    if (a.rb < ref.l_pac && b.rb >= ref.l_pac) return 0;
    if (a.qb >= b.qb || a.qe >= b.qe || a.re >= b.re) return 0;
    w = (int)((a.re - b.rb) - (a.qe - b.qb));
    w = w > 0 ? w : -w;
    r = (double)(a.re - b.rb) / (b.re - a.rb) - (double)(a.qe - b.qb) / (b.qe - a.qb);
    r = r > 0. ? r : -r;
    if (a.re < b.rb || a.qe < b.qb) {
        if (w > opt.w << 1 || r >= 0.05f) return 0;
    } else if (w > opt.w << 2 || r >= 0.05f * 2) return 0;
    w += a.w + b.w;
    w = w < opt.w << 2 ? w : opt.w << 2;
    int ok = gen_score(ctx, opt, ref, w, b.qe - a.qb, query + a.qb, a.rb, b.re, &score, rbuf, rcap);
    if (ok < 0) return ok;
    // when bwa_gen_cigar2 bails out *score is left untouched by the reference (uninitialised there); unreachable
    // for colinear same-strand regions, treated as score 0
    q_s = (int)((double)(b.qe - a.qb) / ((b.qe - b.qb) + (a.qe - a.qb)) * (b.score + a.score) + .499);
    r_s = (int)((double)(b.re - a.rb) / ((b.re - b.rb) + (a.re - a.rb)) * (b.score + a.score) + .499);
    if ((double)score / (q_s > r_s ? q_s : r_s) < 0.90f) return 0;
    *w_out = w;
    return score;
}

// mem_sort_dedup_patch.  query == nullptr disables patching (the call from mem_matesw, bwamem_pair.c:175).
// Returns the new count, or a negative error.
template <class Ctx>
AR_HD int sort_dedup_patch(Ctx& ctx, const MemOpt& opt, const RefView& ref, const uint8_t* query, int n, AlnReg* a, uint8_t* rbuf, int rcap) {
    int m, i, j;
    if (n <= 1) return n;
    if (ctx.leader()) {
        ks_introsort((size_t)n, a, [](const AlnReg& x, const AlnReg& y) { return x.re < y.re; });
        for (i = 0; i < n; ++i) a[i].n_comp_is_alt = (a[i].n_comp_is_alt & ~0x3fffffff) | 1;  // n_comp = 1
    }
    ctx.sync();
    for (i = 1; i < n; ++i) {
        AlnReg* p = &a[i];
        if (p->rid != a[i - 1].rid || p->rb >= a[i - 1].re + opt.max_chain_gap) continue;
        for (j = i - 1; j >= 0 && p->rid == a[j].rid && p->rb < a[j].re + opt.max_chain_gap; --j) {
            AlnReg* q = &a[j];
            int64_t orr, oq, mr, mq;
            int score, w = 0;
            if (q->qe == q->qb) continue;
            orr = q->re - p->rb;
            oq = q->qb < p->qb ? q->qe - p->qb : p->qe - q->qb;
            mr = q->re - q->rb < p->re - p->rb ? q->re - q->rb : p->re - p->rb;
            mq = q->qe - q->qb < p->qe - p->qb ? q->qe - q->qb : p->qe - p->qb;
            if (orr > opt.mask_level_redun * mr && oq > opt.mask_level_redun * mq) {
                if (p->score < q->score) {
                    if (ctx.leader()) p->qe = p->qb;
                    ctx.sync();
                    break;
                } else {
                    if (ctx.leader()) q->qe = q->qb;
                    ctx.sync();
                }
            } else if (q->rb < p->rb && query != nullptr) {
                score = patch_reg(ctx, opt, ref, query, *q, *p, &w, rbuf, rcap);
                if (score < 0) return score;
                if (score > 0) {
                    if (ctx.leader()) {
                        int ncp = (p->n_comp_is_alt & 0x3fffffff) + (q->n_comp_is_alt & 0x3fffffff) + 1;
                        p->n_comp_is_alt = (p->n_comp_is_alt & ~0x3fffffff) | (ncp & 0x3fffffff);
                        p->seedcov = p->seedcov > q->seedcov ? p->seedcov : q->seedcov;
                        p->sub = p->sub > q->sub ? p->sub : q->sub;
                        p->csub = p->csub > q->csub ? p->csub : q->csub;
                        p->qb = q->qb; p->rb = q->rb;
                        p->truesc = p->score = score;
                        p->w = w;
                        q->qb = q->qe;
                    }
                    ctx.sync();
                }
            }
        }
    }
    if (ctx.leader()) {
        for (i = 0, m = 0; i < n; ++i)
            if (a[i].qe > a[i].qb) {
                if (m != i) a[m++] = a[i];
                else ++m;
            }
        n = m;
        ks_introsort((size_t)n, a, [](const AlnReg& x, const AlnReg& y) {
            return x.score > y.score || (x.score == y.score && (x.rb < y.rb || (x.rb == y.rb && x.qb < y.qb)));
        });
        for (i = 1; i < n; ++i)
            if (a[i].score == a[i - 1].score && a[i].rb == a[i - 1].rb && a[i].qb == a[i - 1].qb) a[i].qe = a[i].qb;
        for (i = 1, m = 1; i < n; ++i)
            if (a[i].qe > a[i].qb) {
                if (m != i) a[m++] = a[i];
                else ++m;
            }
    } else m = 0;
    m = ctx.bcast(m);
    ctx.sync();
    return m;
}

}  // namespace arachne
