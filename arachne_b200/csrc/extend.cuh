// Region-level helpers shared by the extension driver, mate rescue and the CIGAR stage, written once against a cooperative
// context `Ctx` (a warp on the device, csrc/warp_ctx.cuh; one thread in ext_stage.cuh's ThreadCtx; a scalar context in tests/hostsim):
//   cal_max_gap          bwa/bwamem.c:621-628
//   bns_fetch_seq        bwa/bntseq.c:421-446   (window clamp to the contig / strand)
//   mem_sort_dedup_patch bwa/bwamem.c:437-489   mem_patch_reg :406-435
//   bwa_gen_cigar2 (score-only use)  bwa/bwa.c:121-207
// The mem_chain2aln driver itself is ext_driver.cuh.
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
