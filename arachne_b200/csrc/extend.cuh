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

// Score of the global alignment of query[0, l_query) against reference [rb, re) within band w_ -- what mem_patch_reg asks
// bwa_gen_cigar2 for (bwa.c:121-207 with n_cigar == NM == NULL).  Returns 1 and *score, 0 when the reference declines the request
// (empty query, empty or strand-crossing or clamped reference range, bwa.c:133-135), a negative error when scratch is too small.
template <class Ctx>
AR_HD int gen_score(Ctx& ctx, const MemOpt& opt, const RefView& ref, int w_, int l_query, const uint8_t* query, int64_t rb, int64_t re,
                    int* score, uint8_t* rbuf, int rcap) {
    const int64_t l_pac = ref.l_pac;
    if (l_query <= 0 || rb >= re || (rb < l_pac && re > l_pac)) return 0;
    if (rb < 0 || re > (l_pac << 1)) return 0;  // bns_get_seq would clamp: rlen != re - rb and the reference bails out
    const int64_t rlen = re - rb;
    if (rlen > rcap) return -2;
    ctx.fetch_ref(ref, rb, re, rbuf);
    const bool rev = rb >= l_pac;  // both sequences reversed on the reverse strand so that indels are left-aligned on the forward strand
    if (l_query == rlen && w_ == 0) {  // no gap possible: sum the substitution scores
        int sum = 0;
        for (int i = 0; i < l_query; ++i) {
            const int k = rev ? l_query - 1 - i : i;
            sum += opt.mat[rbuf[rev ? rlen - 1 - i : i] * 5 + query[k]];
        }
        *score = sum;
        return 1;
    }
    // band: half of the largest gap the score could pay for plus the length difference, at least the difference + 3, at most w_
    const int half = ((l_query + 1) >> 1) * opt.mat[0];
    const int max_ins = (int)((double)(half - opt.o_ins) / opt.e_ins + 1.), max_del = (int)((double)(half - opt.o_del) / opt.e_del + 1.);
    int max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    const int dl = (int)rlen > l_query ? (int)rlen - l_query : l_query - (int)rlen;
    int w = (max_gap + dl + 1) >> 1;
    w = w < w_ ? w : w_;
    w = w > dl + 3 ? w : dl + 3;
    const int step = rev ? -1 : 1;
    *score = ctx.ksw_global(opt, l_query, rev ? query + l_query - 1 : query, step, (int)rlen, rev ? rbuf + rlen - 1 : rbuf, step, w, nullptr, 0, nullptr);
    return 1;
}

// Should regions a (first on the reference) and b be replaced by ONE region spanning both?  (mem_patch_reg, bwamem.c:406-435.)
// They must be on one strand, in query and reference order, on nearby diagonals and of similar slope; then the span is re-aligned
// globally and accepted when it scores at least 90 % of what the two would predict.  Returns the merged score (> 0) and the band used,
// 0 for "keep them apart", a negative error.
template <class Ctx>
AR_HD int patch_reg(Ctx& ctx, const MemOpt& opt, const RefView& ref, const uint8_t* query, const AlnReg& a, const AlnReg& b, int* w_out,
                    uint8_t* rbuf, int rcap) {
    if (a.rb < ref.l_pac && b.rb >= ref.l_pac) return 0;
    if (a.qb >= b.qb || a.qe >= b.qe || a.re >= b.re) return 0;
    const int64_t ref_gap = a.re - b.rb, qry_gap = a.qe - b.qb;  // overlaps when positive
    int w = (int)(ref_gap - qry_gap);
    w = w > 0 ? w : -w;
    double slope = (double)ref_gap / (b.re - a.rb) - (double)qry_gap / (b.qe - a.qb);
    slope = slope > 0. ? slope : -slope;
    const bool apart = a.re < b.rb || a.qe < b.qb;
    if (apart ? (w > opt.w << 1 || slope >= 0.05f) : (w > opt.w << 2 || slope >= 0.05f * 2)) return 0;
    w += a.w + b.w;
    w = w < opt.w << 2 ? w : opt.w << 2;
    int score = 0;  // (left as is when gen_score declines: unreachable for colinear same-strand regions)
    const int ok = gen_score(ctx, opt, ref, w, b.qe - a.qb, query + a.qb, a.rb, b.re, &score, rbuf, rcap);
    if (ok < 0) return ok;
    const int both = b.score + a.score;
    const int by_query = (int)((double)(b.qe - a.qb) / ((b.qe - b.qb) + (a.qe - a.qb)) * both + .499);
    const int by_ref = (int)((double)(b.re - a.rb) / ((b.re - b.rb) + (a.re - a.rb)) * both + .499);
    if ((double)score / (by_query > by_ref ? by_query : by_ref) < 0.90f) return 0;
    *w_out = w;
    return score;
}

// mem_sort_dedup_patch (bwamem.c:437-489): order the read's regions by reference end; for every region look back at the regions that end
// within max_chain_gap before it starts: a pair that overlaps by more than mask_level_redun of the shorter one on BOTH axes loses its
// lower-scoring member, a colinear pair may be merged (patch_reg); then order by (score desc, rb, qb) and drop exact duplicates.
// query == nullptr disables merging (the call from mem_matesw, bwamem_pair.c:175).  A dropped region is marked by qe = qb.
// Returns the new count, or a negative error.
template <class Ctx>
AR_HD int sort_dedup_patch(Ctx& ctx, const MemOpt& opt, const RefView& ref, const uint8_t* query, int n, AlnReg* a, uint8_t* rbuf, int rcap) {
    if (n <= 1) return n;
    const bool lead = ctx.leader();
    if (lead) {
        ks_introsort((size_t)n, a, [](const AlnReg& x, const AlnReg& y) { return x.re < y.re; });
        for (int i = 0; i < n; ++i) a[i].n_comp_is_alt = (a[i].n_comp_is_alt & ~0x3fffffff) | 1;  // n_comp = 1
    }
    ctx.sync();
    auto compact = [&](int count) {  // squeeze out the dropped regions, order kept
        int m = 0;
        for (int i = 0; i < count; ++i)
            if (a[i].qe > a[i].qb) { if (m != i) a[m] = a[i]; ++m; }
        return m;
    };
    for (int i = 1; i < n; ++i) {
        AlnReg& cur = a[i];
        for (int j = i - 1; j >= 0 && cur.rid == a[j].rid && cur.rb < a[j].re + opt.max_chain_gap; --j) {
            AlnReg& old = a[j];
            if (old.qe == old.qb) continue;  // already dropped
            const int64_t ov_ref = old.re - cur.rb;
            const int64_t ov_qry = old.qb < cur.qb ? old.qe - cur.qb : cur.qe - old.qb;
            const int64_t len_ref = ar_min(old.re - old.rb, cur.re - cur.rb), len_qry = (int64_t)ar_min(old.qe - old.qb, cur.qe - cur.qb);
            if (ov_ref > opt.mask_level_redun * len_ref && ov_qry > opt.mask_level_redun * len_qry) {  // one of them is redundant
                const bool drop_cur = cur.score < old.score;
                ctx.sync();  // every lane has read the pair before the leader marks one of them (a lane still reading would take another branch)
                if (lead) { if (drop_cur) cur.qe = cur.qb; else old.qe = old.qb; }
                ctx.sync();
                if (drop_cur) break;
            } else if (old.rb < cur.rb && query != nullptr) {
                int w = 0;
                const int merged = patch_reg(ctx, opt, ref, query, old, cur, &w, rbuf, rcap);
                if (merged < 0) return merged;
                if (merged > 0) {
                    ctx.sync();
                    if (lead) {
                        const int n_comp = (cur.n_comp_is_alt & 0x3fffffff) + (old.n_comp_is_alt & 0x3fffffff) + 1;
                        cur.n_comp_is_alt = (cur.n_comp_is_alt & ~0x3fffffff) | (n_comp & 0x3fffffff);
                        cur.seedcov = ar_max(cur.seedcov, old.seedcov); cur.sub = ar_max(cur.sub, old.sub); cur.csub = ar_max(cur.csub, old.csub);
                        cur.qb = old.qb; cur.rb = old.rb;
                        cur.truesc = cur.score = merged;
                        cur.w = w;
                        old.qb = old.qe;
                    }
                    ctx.sync();
                }
            }
        }
    }
    int m = 0;
    ctx.sync();  // the loop's last reads are done before the leader compacts and re-sorts the array
    if (lead) {
        m = compact(n);
        ks_introsort((size_t)m, a, [](const AlnReg& x, const AlnReg& y) {
            return x.score > y.score || (x.score == y.score && (x.rb < y.rb || (x.rb == y.rb && x.qb < y.qb)));
        });
        for (int i = 1; i < m; ++i)  // identical hits
            if (a[i].score == a[i - 1].score && a[i].rb == a[i - 1].rb && a[i].qb == a[i - 1].qb) a[i].qe = a[i].qb;
        m = compact(m);
    }
    m = ctx.bcast(m);
    ctx.sync();
    return m;
}

}  // namespace arachne
