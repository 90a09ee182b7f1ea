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

AR_HD int infer_dir(int64_t l_pac, int64_t b1, int64_t b2, int64_t* dist) {
    int r1 = (b1 >= l_pac), r2 = (b2 >= l_pac);
    int64_t p2 = r1 == r2 ? b2 : (l_pac << 1) - 1 - b2;
    *dist = p2 > b1 ? p2 - b1 : b1 - p2;
    return (r1 == r2 ? 0 : 1) ^ (p2 > b1 ? 0 : 3);
}

constexpr int KSW_XBYTE = 0x10000, KSW_XSTOP = 0x20000, KSW_XSUBO = 0x40000, KSW_XSTART = 0x80000;

struct LocalAln { int score, te, qe, score2, te2, tb, qb; };

// mem_matesw: try to find the mate `ms` near anchor `a`; ma/n_ma is the mate's region list (capacity cap).
// rbuf: window buffer of >= rcap bytes.  Returns the number of SW runs, or a negative error.
template <class Ctx>
AR_HD int matesw(Ctx& ctx, const MemOpt& opt, const RefView& ref, const PeStat pes[4], const AlnReg& a, int l_ms, const uint8_t* ms,
                 AlnReg* ma, int* n_ma, int cap, uint8_t* rbuf, int rcap) {
    const int64_t l_pac = ref.l_pac;
    int i, r, skip[4], n = 0;
    for (r = 0; r < 4; ++r) skip[r] = pes[r].failed ? 1 : 0;
    for (i = 0; i < *n_ma; ++i) {
        int64_t dist;
        r = infer_dir(l_pac, a.rb, ma[i].rb, &dist);
        if (dist >= pes[r].low && dist <= pes[r].high) skip[r] = 1;
    }
    if (skip[0] + skip[1] + skip[2] + skip[3] == 4) return 0;
    for (r = 0; r < 4; ++r) {
        if (skip[r]) continue;
        int is_rev = (r >> 1 != (r & 1));
        int is_larger = !(r >> 1);
        int64_t rb, re;
        int rid = -1;
        if (!is_rev) {
            rb = is_larger ? a.rb + pes[r].low : a.rb - pes[r].high;
            re = (is_larger ? a.rb + pes[r].high : a.rb - pes[r].low) + l_ms;
        } else {
            rb = (is_larger ? a.rb + pes[r].low : a.rb - pes[r].high) - l_ms;
            re = is_larger ? a.rb + pes[r].high : a.rb - pes[r].low;
        }
        if (rb < 0) rb = 0;
        if (re > l_pac << 1) re = l_pac << 1;
        if (rb < re) rid = fetch_window(ref, &rb, (rb + re) >> 1, &re);
        if (a.rid == rid && re - rb >= opt.min_seed_len) {
            if (re - rb > rcap) return -2;
            ctx.fetch_ref(ref, rb, re, rbuf);
            int xtra = KSW_XSUBO | KSW_XSTART | (l_ms * opt.a < 250 ? KSW_XBYTE : 0) | (opt.min_seed_len * opt.a);
            LocalAln aln;
            // the mate enters reverse-complemented when is_rev (bwamem_pair.c:131-135)
            ctx.ksw_local(opt, l_ms, ms, is_rev, (int)(re - rb), rbuf, xtra, &aln);
            if (aln.score >= opt.min_seed_len && aln.qb >= 0) {
                AlnReg b;
                b.truesc = b.sub = b.alt_sc = b.sub_n = b.w = b.secondary_all = b.seedlen0 = 0; b.frac_rep = 0.f; b.hash = 0;
                b.rid = a.rid;
                b.n_comp_is_alt = a.n_comp_is_alt & ~0x3fffffff;  // is_alt copied, n_comp = 0
                b.qb = is_rev ? l_ms - (aln.qe + 1) : aln.qb;
                b.qe = is_rev ? l_ms - aln.qb : aln.qe + 1;
                b.rb = is_rev ? (l_pac << 1) - (rb + aln.te + 1) : rb + aln.tb;
                b.re = is_rev ? (l_pac << 1) - (rb + aln.tb) : rb + aln.te + 1;
                b.score = aln.score;
                b.csub = aln.score2;
                b.secondary = -1;
                b.seedcov = (int)((b.re - b.rb < b.qe - b.qb ? b.re - b.rb : b.qe - b.qb) >> 1);
                if (*n_ma >= cap) return -3;
                int nn = *n_ma + 1, pos;
                for (pos = 0; pos < nn - 1; ++pos)
                    if (ma[pos].score < b.score) break;
                if (ctx.leader()) {
                    for (i = nn - 1; i > pos; --i) ma[i] = ma[i - 1];
                    ma[pos] = b;
                }
                *n_ma = nn;
                ctx.sync();
            }
            ++n;
        }
        if (n) {
            int m = sort_dedup_patch(ctx, opt, ref, (const uint8_t*)nullptr, *n_ma, ma, rbuf, rcap);
            if (m < 0) return m;
            *n_ma = m;
        }
    }
    return n;
}

// GoBwaMemMateSW after the two mem_align1_core calls.  r1/r2 are the reads' region lists with room for cap1/cap2.
template <class Ctx>
AR_HD int pair_rescue(Ctx& ctx, const MemOpt& opt, const RefView& ref, int score_delta, int l1, const uint8_t* s1, int l2, const uint8_t* s2,
                      AlnReg* r1, int* n1, int cap1, AlnReg* r2, int* n2, int cap2, uint8_t* rbuf, int rcap) {
    PeStat pes[4];
    gobwa_pestat(pes);
    int best1 = 0, best2 = 0, i, num, rc;
    for (i = 0; i < *n1; ++i) if (r1[i].score > best1) best1 = r1[i].score;
    for (i = 0; i < *n2; ++i) if (r2[i].score > best2) best2 = r2[i].score;
    const int n2_0 = *n2;  // the Go loop walks the pre-rescue list of read 2 (gobwa.go:285-300); r2 is not modified here
    for (i = 0, num = 0; i < n2_0 && num < 50 && l1 > 0; ++i)
        if (r2[i].score >= best2 - score_delta) {
            ++num;
            const AlnReg anchor = r2[i];
            rc = matesw(ctx, opt, ref, pes, anchor, l1, s1, r1, n1, cap1, rbuf, rcap);
            if (rc < 0) return rc;
        }
    const int n1_1 = *n1;  // read 1's list after its own rescue (gobwa.go:302-324); r1 is not modified here
    for (i = 0, num = 0; i < n1_1 && num < 50 && l2 > 0; ++i)
        if (r1[i].score >= best1 - score_delta) {
            ++num;
            const AlnReg anchor = r1[i];
            rc = matesw(ctx, opt, ref, pes, anchor, l2, s2, r2, n2, cap2, rbuf, rcap);
            if (rc < 0) return rc;
        }
    return 0;
}

}  // namespace arachne
