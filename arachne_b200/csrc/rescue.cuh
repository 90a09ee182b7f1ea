// Mate rescue as the reference drives it for one read pair:
//   GoBwaMemMateSW   src/gobwa/gobwa.go:226-337  (hard-coded insert model, rescue read 1 from read 2's hits, then 2 from 1)
//   mem_matesw       bwa/bwamem_pair.c:111-180
//   mem_infer_dir    bwa/bwamem_pair.c:23-30
// Written against the cooperative context (see extend.cuh): ctx.ksw_local is the striped-semantics local SW.
#pragma once
#include "extend.cuh"

namespace arachne {

struct PeStat {  // mem_pestat_t, bwa/bwamem.h:108-112
    int low, high, failed;
    double avg, std;
};

AR_HD void gobwa_pestat(PeStat pes[4]) {  // gobwa.go:229-237
    for (int i = 0; i < 4; ++i) { pes[i].low = pes[i].high = 0; pes[i].failed = 1; pes[i].avg = pes[i].std = 0.; }
    pes[1].low = -35; pes[1].high = 500; pes[1].failed = 0; pes[1].avg = 200.0; pes[1].std = 100.0;
}

// Relative placement of a mate hit against an anchor hit, both given by their start in forward-reverse coordinates: the mate is
// mirrored onto the anchor's strand when the two lie on different strands, then `cls` = bit 0: different strands, bit 1: mate start
// not beyond the anchor's -- XOR-folded the way mem_infer_dir numbers its four classes (bwamem_pair.c:23-30: FF / FR / RF / RR as
// 0..3) -- and `dist` the distance between the two starts.
struct PairGeom { int cls; int64_t dist; };
AR_HD PairGeom pair_geometry(int64_t l_pac, int64_t anchor_rb, int64_t mate_rb) {
    const bool same_strand = (anchor_rb >= l_pac) == (mate_rb >= l_pac);
    const int64_t m = same_strand ? mate_rb : 2 * l_pac - 1 - mate_rb;
    PairGeom g;
    g.cls = (same_strand ? 0 : 1) ^ (m > anchor_rb ? 0 : 3);
    g.dist = m > anchor_rb ? m - anchor_rb : anchor_rb - m;
    return g;
}

// Bit r set = orientation class r has a usable insert model and none of the mate's n_ma hits sits at a distance the model accepts
// from the anchor: the classes mem_matesw runs a Smith-Waterman for (bwamem_pair.c:116-124).  0 = nothing to rescue; k_rescue asks
// this first so that pairs with nothing to do (most) never get a private copy of their lists.
AR_HD unsigned rescue_open_classes(int64_t l_pac, const PeStat pes[4], const AlnReg& anchor, const AlnReg* ma, int n_ma) {
    unsigned open = 0;
    for (int r = 0; r < 4; ++r) open |= (pes[r].failed ? 0u : 1u) << r;
    for (int i = 0; i < n_ma && open; ++i) {
        const PairGeom g = pair_geometry(l_pac, anchor.rb, ma[i].rb);
        if (g.dist >= pes[g.cls].low && g.dist <= pes[g.cls].high) open &= ~(1u << g.cls);
    }
    return open;
}

constexpr int KSW_XBYTE = 0x10000, KSW_XSTOP = 0x20000, KSW_XSUBO = 0x40000, KSW_XSTART = 0x80000;

struct LocalAln { int score, te, qe, score2, te2, tb, qb; };

// Where a mate of l_ms bases in orientation class `cls` can lie around an anchor that starts at anchor_rb: its start is low..high
// away, ahead of the anchor for classes 0 / 1 and behind it for 2 / 3; a mate on the anchor's strand (classes 0, 3) extends the
// window by its length at the far end, a mate on the other strand (1, 2: it is searched reverse-complemented) at the near end.
// Clipped to the forward-reverse text; the caller clips to the contig.  (bwamem_pair.c:136-144)
struct RefWindow { int64_t beg, end; };
AR_HD RefWindow rescue_window(int cls, bool mate_rc, const PeStat& pe, int64_t anchor_rb, int l_ms, int64_t l_pac) {
    const bool ahead = cls < 2;
    RefWindow w;
    w.beg = ahead ? anchor_rb + pe.low : anchor_rb - pe.high;
    w.end = ahead ? anchor_rb + pe.high : anchor_rb - pe.low;
    if (mate_rc) w.beg -= l_ms; else w.end += l_ms;
    if (w.beg < 0) w.beg = 0;
    if (w.end > 2 * l_pac) w.end = 2 * l_pac;
    return w;
}

// The region a rescue hit becomes (bwamem_pair.c:151-163): everything zero except the spans, score, csub, the anchor's contig and ALT
// flag, secondary = -1 and seedcov = half the shorter span.  A reverse-complemented mate's spans are mirrored back.
AR_HD AlnReg rescued_region(const AlnReg& anchor, const LocalAln& aln, bool mate_rc, int64_t win_beg, int l_ms, int64_t l_pac) {
    AlnReg b;
    b.truesc = b.sub = b.alt_sc = b.sub_n = b.w = b.secondary_all = b.seedlen0 = 0; b.frac_rep = 0.f; b.hash = 0;
    b.rid = anchor.rid;
    b.n_comp_is_alt = anchor.n_comp_is_alt & ~0x3fffffff;  // is_alt copied, n_comp = 0
    const int q_lo = aln.qb, q_hi = aln.qe + 1;
    const int64_t t_lo = win_beg + aln.tb, t_hi = win_beg + aln.te + 1;
    b.qb = mate_rc ? l_ms - q_hi : q_lo;
    b.qe = mate_rc ? l_ms - q_lo : q_hi;
    b.rb = mate_rc ? 2 * l_pac - t_hi : t_lo;
    b.re = mate_rc ? 2 * l_pac - t_lo : t_hi;
    b.score = aln.score;
    b.csub = aln.score2;
    b.secondary = -1;
    const int64_t r_span = b.re - b.rb, q_span = b.qe - b.qb;
    b.seedcov = (int)((r_span < q_span ? r_span : q_span) >> 1);
    return b;
}

// mem_matesw (bwamem_pair.c:111-180): look for the mate `ms` near anchor `a` in every orientation class that is still open;
// ma / n_ma is the mate's region list (capacity cap, kept in descending score order).  rbuf: window buffer of >= rcap bytes.
// Returns the number of SW runs, or a negative error.
template <class Ctx>
AR_HD int matesw(Ctx& ctx, const MemOpt& opt, const RefView& ref, const PeStat pes[4], const AlnReg& a, int l_ms, const uint8_t* ms,
                 AlnReg* ma, int* n_ma, int cap, uint8_t* rbuf, int rcap) {
    const int64_t l_pac = ref.l_pac;
    const unsigned open = rescue_open_classes(l_pac, pes, a, ma, *n_ma);
    if (!open) return 0;
    int n_sw = 0;
    for (int cls = 0; cls < 4; ++cls) {
        if (!(open >> cls & 1)) continue;
        const bool mate_rc = cls == 1 || cls == 2;  // the mate lies on the other strand: it enters the SW reverse-complemented
        RefWindow w = rescue_window(cls, mate_rc, pes[cls], a.rb, l_ms, l_pac);
        const int rid = w.beg < w.end ? fetch_window(ref, &w.beg, (w.beg + w.end) >> 1, &w.end) : -1;
        if (rid == a.rid && w.end - w.beg >= opt.min_seed_len) {
            if (w.end - w.beg > rcap) return -2;
            ctx.fetch_ref(ref, w.beg, w.end, rbuf);
            const int xtra = KSW_XSUBO | KSW_XSTART | (l_ms * opt.a < 250 ? KSW_XBYTE : 0) | (opt.min_seed_len * opt.a);
            LocalAln aln;
            ctx.ksw_local(opt, l_ms, ms, mate_rc, (int)(w.end - w.beg), rbuf, xtra, &aln);
            ++n_sw;
            if (aln.score >= opt.min_seed_len && aln.qb >= 0) {
                if (*n_ma >= cap) return -3;
                const AlnReg b = rescued_region(a, aln, mate_rc, w.beg, l_ms, l_pac);
                int at = 0;  // in front of the first hit that scores less
                while (at < *n_ma && ma[at].score >= b.score) ++at;
                ctx.sync();  // every lane has read the list before the leader shifts it
                if (ctx.leader()) {
                    for (int k = *n_ma; k > at; --k) ma[k] = ma[k - 1];
                    ma[at] = b;
                }
                ++*n_ma;
                ctx.sync();
            }
        }
        if (n_sw) {  // once any SW has run, the list is de-duplicated after every class (bwamem_pair.c:175)
            const int m = sort_dedup_patch(ctx, opt, ref, (const uint8_t*)nullptr, *n_ma, ma, rbuf, rcap);
            if (m < 0) return m;
            *n_ma = m;
        }
    }
    return n_sw;
}

// One direction of GoBwaMemMateSW's rescue (gobwa.go:285-300 / 302-324): every anchor among the first n_anchor entries of `anchors`
// that scores within score_delta of `best` -- at most 50 of them -- gets a mem_matesw call for the mate.  Only the mate's list changes; the
// anchors' list is the other read's and stays as it is for the whole walk.
template <class Ctx>
AR_HD int rescue_from(Ctx& ctx, const MemOpt& opt, const RefView& ref, const PeStat pes[4], int score_delta, const AlnReg* anchors, int n_anchor,
                      int best, int l_mate, const uint8_t* mate, AlnReg* mate_regs, int* n_mate, int cap, uint8_t* rbuf, int rcap) {
    if (l_mate <= 0) return 0;
    for (int i = 0, tried = 0; i < n_anchor && tried < 50; ++i) {
        if (anchors[i].score < best - score_delta) continue;
        ++tried;
        const AlnReg anchor = anchors[i];
        const int rc = matesw(ctx, opt, ref, pes, anchor, l_mate, mate, mate_regs, n_mate, cap, rbuf, rcap);
        if (rc < 0) return rc;
    }
    return 0;
}

// GoBwaMemMateSW after the two mem_align1_core calls.  r1/r2 are the reads' region lists with room for cap1/cap2.  Both best scores
// are taken BEFORE any rescue (gobwa.go:264-283): read 2's rescue still filters read 1's grown list by read 1's old best.
template <class Ctx>
AR_HD int pair_rescue(Ctx& ctx, const MemOpt& opt, const RefView& ref, int score_delta, int l1, const uint8_t* s1, int l2, const uint8_t* s2,
                      AlnReg* r1, int* n1, int cap1, AlnReg* r2, int* n2, int cap2, uint8_t* rbuf, int rcap) {
    PeStat pes[4];
    gobwa_pestat(pes);
    int best1 = 0, best2 = 0;
    for (int i = 0; i < *n1; ++i) best1 = r1[i].score > best1 ? r1[i].score : best1;
    for (int i = 0; i < *n2; ++i) best2 = r2[i].score > best2 ? r2[i].score : best2;
    // read 1 is looked for around read 2's hits, then read 2 around read 1's hits INCLUDING the ones just rescued
    int rc = rescue_from(ctx, opt, ref, pes, score_delta, r2, *n2, best2, l1, s1, r1, n1, cap1, rbuf, rcap);
    if (rc < 0) return rc;
    return rescue_from(ctx, opt, ref, pes, score_delta, r1, *n1, best1, l2, s2, r2, n2, cap2, rbuf, rcap);
}

}  // namespace arachne
