// Region -> alignment record:
//   mem_reg2aln       bwa/bwamem.c:1086-1156   (band inference infer_bw :792-799, up to three band doublings, leading /
//                                               trailing deletion squeeze, soft clips, forward-strand position)
//   bwa_gen_cigar2    bwa/bwa.c:121-207        (gap-free fast path, banded NW + traceback, NM)
//   GetAlignments     src/aligner/aligner.go:1505-1569,1618  (CIGAR/reference walk: matches, mismatch loci, indels, soft clips)
//   InterpretAlign    src/gobwa/gobwa.go:339-371  (Offset / Alignment_end)
// Written against the cooperative context (extend.cuh).  The MD string the reference appends to the CIGAR buffer
// is not produced: nothing on the Go side reads it.
#pragma once
#include "extend.cuh"

namespace arachne {

// field-for-field arachne_aln_t (include/arachne_cuda.h)
struct AlnRec {
    int64_t pos;
    int32_t rid, is_rev, nm, n_cigar;
    int64_t cigar_off;
    int32_t score, sub;
    int32_t matches, mismatches, indels, soft_clipped, soft_clipped_length;
    int32_t n_mm;
    int64_t mm_off;
};
static_assert(sizeof(AlnRec) == 72, "AlnRec must match arachne_aln_t");

AR_HD int infer_bw(int l1, int l2, int score, int a, int q, int r) {
    int w;
    if (l1 == l2 && l1 * a - score < (q + r - a) << 1) return 0;
    w = (int)((double)((l1 < l2 ? l1 : l2) * a - score - q) / r + 2.);
    int d = l1 - l2; d = d < 0 ? -d : d;
    if (w < d) w = d;
    return w;
}

// bwa_gen_cigar2 with CIGAR + NM.  rbuf receives the reference bases of [rb,re) in forward-reverse coordinates (kept
// for the caller's mismatch walk).  cigar: scratch of cap words.  Returns n_cigar (>= 0), or -1 when the reference
// returns no CIGAR (bwa.c:133-135), or <= -2 on scratch exhaustion.
template <class Ctx>
AR_HD int gen_cigar(Ctx& ctx, const MemOpt& opt, const RefView& ref, int w_, int l_query, const uint8_t* query, int64_t rb, int64_t re,
                    int* score, uint32_t* cigar, int cap, int* NM, uint8_t* rbuf, int rcap) {
    const int64_t l_pac = ref.l_pac;
    *NM = -1;
    if (l_query <= 0 || rb >= re || (rb < l_pac && re > l_pac)) return -1;
    if (rb < 0 || re > (l_pac << 1)) return -1;  // bns_get_seq would clamp -> rlen mismatch -> no CIGAR
    const int64_t rlen = re - rb;
    if (rlen > rcap) return -2;
    ctx.fetch_ref(ref, rb, re, rbuf);
    const int rev = rb >= l_pac;
    int n_cigar = 0;
    if (l_query == rlen && w_ == 0) {
        int s = 0;
        for (int i = 0; i < l_query; ++i) s += opt.mat[rbuf[rev ? rlen - 1 - i : i] * 5 + query[rev ? l_query - 1 - i : i]];
        *score = s;
        if (cap < 1) return -3;
        if (ctx.leader()) cigar[0] = (uint32_t)l_query << 4;
        ctx.sync();
        n_cigar = 1;
    } else {
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
                                rev ? -1 : 1, w, cigar, cap, &n_cigar);
        if (n_cigar > cap) return -3;
    }
    // NM (bwa.c:171-198): mismatches over M runs + gap lengths, a deletion at either end of the CIGAR not counted
    int x = 0, y = 0, n_mm = 0, n_gap = 0;
    for (int k = 0; k < n_cigar; ++k) {
        int op = cigar[k] & 0xf, len = (int)(cigar[k] >> 4);
        if (op == 0) {
            for (int i = 0; i < len; ++i) {
                int qi = x + i, ri = y + i;
                if (query[rev ? l_query - 1 - qi : qi] != rbuf[rev ? rlen - 1 - ri : ri]) ++n_mm;
            }
            x += len; y += len;
        } else if (op == 2) {
            if (k > 0 && k < n_cigar - 1) n_gap += len;
            y += len;
        } else if (op == 1) { x += len; n_gap += len; }
    }
    *NM = n_mm + n_gap;
    return n_cigar;
}

// mem_reg2aln + the GetAlignments walk for one region.  cigar: scratch/out of cap words (final CIGAR left there);
// mm: out of 2*l_query ints (n_mm reference loci then n_mm read offsets).
template <class Ctx>
AR_HD int reg2aln(Ctx& ctx, const MemOpt& opt, const RefView& ref, int l_query, const uint8_t* query, const AlnReg& ar, AlnRec* out,
                  uint32_t* cigar, int cap, int32_t* mm, uint8_t* rbuf, int rcap) {
    AlnRec a;
    a.pos = -1; a.rid = -1; a.is_rev = 0; a.nm = 0; a.n_cigar = 0; a.cigar_off = 0; a.score = 0; a.sub = 0;
    a.matches = a.mismatches = a.indels = a.soft_clipped = a.soft_clipped_length = 0; a.n_mm = 0; a.mm_off = 0;
    if (ar.rb < 0 || ar.re < 0) { *out = a; return 0; }
    const int qb = ar.qb, qe = ar.qe;
    const int64_t rb = ar.rb, re = ar.re;
    int tmp = infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_del, opt.e_del);
    int w2 = infer_bw(qe - qb, (int)(re - rb), ar.truesc, opt.a, opt.o_ins, opt.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > opt.w) w2 = w2 < ar.w ? w2 : ar.w;
    int i = 0, score = 0, NM = -1, n_cigar = 0, last_sc = -(1 << 30);
    do {
        w2 = w2 < opt.w << 2 ? w2 : opt.w << 2;
        n_cigar = gen_cigar(ctx, opt, ref, w2, qe - qb, query + qb, rb, re, &score, cigar, cap - 2, &NM, rbuf, rcap);
        if (n_cigar <= -2) return n_cigar;
        if (n_cigar < 0) n_cigar = 0;  // the reference carries on with an empty CIGAR and NM = -1
        if (score == last_sc || w2 == opt.w << 2) break;
        last_sc = score;
        w2 <<= 1;
    } while (++i < 3 && score < ar.truesc - opt.a);
    a.nm = NM;
    int is_rev;
    int64_t pos = bns_depos(ref, rb < ref.l_pac ? rb : re - 1, &is_rev);
    a.is_rev = is_rev;
    // squeeze out a leading, else a trailing, deletion (:1124-1133)
    int c0 = 0;
    if (n_cigar > 0) {
        if ((cigar[0] & 0xf) == 2) { pos += cigar[0] >> 4; --n_cigar; c0 = 1; }
        else if ((cigar[n_cigar - 1] & 0xf) == 2) --n_cigar;
    }
    int clip5 = 0, clip3 = 0;
    if (qb != 0 || qe != l_query) {
        clip5 = is_rev ? l_query - qe : qb;
        clip3 = is_rev ? qb : l_query - qe;
    }
    // final layout in place: [clip5] ops [clip3].  Every lane has to be done reading cigar[0] / cigar[n_cigar - 1] above before the leader moves them.
    ctx.sync();
    if (ctx.leader()) {
        int shift = (clip5 ? 1 : 0) - c0;
        if (shift > 0) for (int k = n_cigar - 1; k >= 0; --k) cigar[k + shift] = cigar[k + c0];
        else if (shift < 0) for (int k = 0; k < n_cigar; ++k) cigar[k] = cigar[k + c0];
        if (clip5) cigar[0] = (uint32_t)clip5 << 4 | 3;
        if (clip3) cigar[n_cigar + (clip5 ? 1 : 0)] = (uint32_t)clip3 << 4 | 3;
    }
    n_cigar += (clip5 ? 1 : 0) + (clip3 ? 1 : 0);
    ctx.sync();
    a.n_cigar = n_cigar;
    a.rid = bns_pos2rid(ref, pos);
    a.pos = pos - ref.contigs[a.rid].offset;
    a.score = ar.score;
    a.sub = ar.sub > ar.csub ? ar.sub : ar.csub;

    // ---- GetAlignments walk (aligner.go:1505-1569).  rbuf holds the forward-reverse reference over [rb,re), which is what
    // GetSeq returns in read orientation when the region lies inside one contig. ----
    const int64_t coff = ref.contigs[ar.rid].offset;
    int64_t ch_pos, ch_aend;  // InterpretAlign: Offset / Alignment_end
    if (rb < ref.l_pac) ch_pos = rb - coff; else ch_pos = ref.l_pac * 2 - 1 - rb - coff;
    if (re < ref.l_pac) ch_aend = re - coff; else ch_aend = ref.l_pac * 2 - 1 - re - coff;
    int64_t refStart = ch_pos, refEnd = ch_aend;
    if (is_rev) { refStart = ch_aend + 1; refEnd = ch_pos + 1; }
    const int64_t ref_len = re - rb;  // len(refSeq) == refEnd - refStart
    int matches = 0, indels = 0, indel_len = 0, sclip = 0, sclip_len = 0, n_mm = 0;
    int refOff = 0, readOff = 0;
    for (int kk = 0; kk < n_cigar; ++kk) {
        int k = is_rev ? n_cigar - 1 - kk : kk;
        int op = cigar[k] & 0xf, len = (int)(cigar[k] >> 4);
        if (op == 0) {
            matches += len;
            for (int m = 0; m < len; ++m) {
                if (refOff + m >= ref_len) continue;
                if (readOff + m < l_query && rbuf[refOff + m] != query[readOff + m]) {
                    if (ctx.leader()) {
                        mm[n_mm] = is_rev ? (int32_t)(refEnd - (refOff + m)) : (int32_t)(refOff + refStart + m);
                        mm[l_query + n_mm] = readOff + m;
                    }
                    ++n_mm;
                }
            }
            refOff += len; readOff += len;
        } else if (op == 1) { ++indels; indel_len += len; readOff += len; }
        else if (op == 2) { ++indels; indel_len += len; refOff += len; }
        else if (op == 3) { ++sclip; sclip_len += len; readOff += len; }
    }
    int mismatches = NM - indel_len;
    matches -= mismatches;
    if (mismatches < 0) mismatches = 0;
    a.matches = matches; a.mismatches = mismatches; a.indels = indels; a.soft_clipped = sclip; a.soft_clipped_length = sclip_len;
    a.n_mm = n_mm;
    ctx.sync();
    *out = a;
    return 0;
}

}  // namespace arachne
