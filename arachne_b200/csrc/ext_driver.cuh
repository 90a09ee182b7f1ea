// Extension driver: what mem_chain2aln (bwa/bwamem.c:632-786) decides AROUND its ksw_extend2 calls, as a resumable
// per-read automaton, so that the dynamic programming itself can run elsewhere as a queue of uniform problems.
//
// A read's kept chains are walked in filter order, a chain's seeds from the highest-scoring down; for a seed that is not
// already explained by a region of this read the automaton asks for at most four DP problems, one at a time: the left flank
// (read and reference reversed, starting at the seed), its retry with the band doubled, the right flank, its retry.
// ext_advance() runs until it needs a DP result (returns 1 with the problem filled in) or the read has no seed left
// (returns 0).  The caller feeds the result of the pending problem into the next call.  Nothing here touches a DP cell.
//
// The same source is compiled for the device (k_ext_drive, ext_stage.cuh) and, TEST-ONLY, for the host simulation
// (tests/hostsim), where the problems are solved by the oracle's scalar ksw_extend2.
#pragma once
#include "extend.cuh"

namespace arachne {

struct ExtProb {      // one ksw_extend2 call (bwa/ksw.c:380): query = read bases q0, q0+qs, ...; target = reference tpos, tpos+tstep, ...
    int32_t read;
    int32_t q0;       // offset of the first query base inside the read
    int32_t qlen, tlen;
    int32_t w, h0, end_bonus;
    int32_t dirs;     // bit 0: query stride is -1; bit 1: target step is -1
    int64_t tpos;     // forward-reverse coordinate of the first target base
    int64_t qoff;     // position of the first query base in the batch's base array (read offset + q0): filled by whoever queues the problem
};
static_assert(sizeof(ExtProb) == 48, "ExtProb is copied as three 16-byte words");

struct ExtRes {       // what ksw_extend2 hands back: score, qle-1, tle-1, gtle-1, gscore, max_off
    int32_t mx, max_i, max_j, max_ie, gscore, max_off;
};

enum { EXT_PH_CHAIN = 0, EXT_PH_SEED, EXT_PH_LEFT, EXT_PH_RIGHT, EXT_PH_DONE };

struct ExtState {
    int32_t phase;
    int32_t chain_i;        // next kept chain of the read
    int32_t seed_k;         // position in the chain's score-sorted seed list (descending walk)
    int32_t n_seeds;        // seeds of the current chain
    int32_t seed0;          // first seed of the current chain inside the read's kept-seed array
    int32_t rid;
    float frac_rep;
    int32_t n_regions;      // regions emitted for this read so far
    int64_t win_beg, win_end;  // reference window of the chain after clamping (rmax[0], rmax[1])
    // the seed being extended
    int64_t s_rbeg; int32_t s_qbeg, s_len;
    // the region under construction
    int64_t r_beg, r_end; int32_t q_beg, q_end, score, truesc, band_l, band_r;
    int32_t attempt;        // band retries made on the current flank
    int32_t score_before;   // the score the current DP call started from (mem_chain2aln's `prev`)
    int32_t score_left;     // score after the left flank (sc0 of the right flank)
};

AR_HD void ext_state_init(ExtState& S) {
    S.phase = EXT_PH_CHAIN; S.chain_i = 0; S.seed_k = -1; S.n_seeds = 0; S.seed0 = 0; S.rid = -1; S.frac_rep = 0.f; S.n_regions = 0;
    S.win_beg = S.win_end = 0; S.s_rbeg = 0; S.s_qbeg = S.s_len = 0; S.r_beg = S.r_end = 0; S.q_beg = S.q_end = 0; S.score = S.truesc = -1;
    S.band_l = S.band_r = 0; S.attempt = 0; S.score_before = 0; S.score_left = 0;
}

// Is seed s already inside region p, close enough to p's diagonal that extending it again would reproduce p?  (bwamem.c:676-691)
AR_HD bool ext_seed_explained(const MemOpt& opt, int l_query, const Seed& s, const AlnReg& p) {
    if (s.rbeg < p.rb || s.rbeg + s.len > p.re || s.qbeg < p.qb || s.qbeg + s.len > p.qe) return false;  // not contained
    if (s.len - p.seedlen0 > .1 * l_query) return false;                                                  // much longer than the seed that made p
    for (int side = 0; side < 2; ++side) {
        const int qd = side == 0 ? s.qbeg - p.qb : p.qe - (s.qbeg + s.len);
        const int64_t rd = side == 0 ? s.rbeg - p.rb : p.re - (s.rbeg + s.len);
        const int gap = cal_max_gap(opt, qd < rd ? qd : (int)rd);
        const int band = gap < p.w ? gap : p.w;
        if (qd - rd < band && rd - qd < band) return true;
    }
    return false;
}

// Does a seed already extended (sorted position > k, not struck out) overlap s on the read but sit on another diagonal?  (bwamem.c:693-701)
AR_HD bool ext_overlap_disagrees(const Seed* seeds, const uint64_t* order, int n, int k, const Seed& s) {
    for (int x = k + 1; x < n; ++x) {
        if (order[x] == 0) continue;
        const Seed t = seeds[(uint32_t)order[x]];
        if (t.len < s.len * .95) continue;
        const bool t_after = s.qbeg <= t.qbeg && s.qbeg + s.len - t.qbeg >= s.len >> 2 && t.qbeg - s.qbeg != t.rbeg - s.rbeg;
        const bool t_before = t.qbeg <= s.qbeg && t.qbeg + t.len - s.qbeg >= s.len >> 2 && s.qbeg - t.qbeg != s.rbeg - t.rbeg;
        if (t_after || t_before) return true;
    }
    return false;
}

// Reference window a chain may be extended into (bwamem.c:640-660 + bns_fetch_seq's clamp to the contig and strand).
AR_HD void ext_chain_window(const MemOpt& opt, const RefView& ref, int l_query, const Seed* seeds, int n, int64_t* beg, int64_t* end) {
    const int64_t two_l = ref.l_pac << 1;
    int64_t lo = two_l, hi = 0;
    for (int x = 0; x < n; ++x) {
        const Seed t = seeds[x];
        const int tail = l_query - t.qbeg - t.len;
        const int64_t b = t.rbeg - (t.qbeg + cal_max_gap(opt, t.qbeg));
        const int64_t e = t.rbeg + t.len + (tail + cal_max_gap(opt, tail));
        lo = lo < b ? lo : b;
        hi = hi > e ? hi : e;
    }
    lo = lo > 0 ? lo : 0;
    hi = hi < two_l ? hi : two_l;
    if (lo < ref.l_pac && ref.l_pac < hi) {  // never across the forward / reverse boundary: keep the side the first seed is on
        if (seeds[0].rbeg < ref.l_pac) hi = ref.l_pac; else lo = ref.l_pac;
    }
    fetch_window(ref, &lo, seeds[0].rbeg, &hi);
    *beg = lo; *end = hi;
}

AR_HD void ext_emit(const MemOpt& opt, const ExtState& S, int read, int l_query, bool right, ExtProb* p) {
    p->read = read;
    if (!right) {
        p->q0 = S.s_qbeg - 1; p->qlen = S.s_qbeg; p->tlen = (int)(S.s_rbeg - S.win_beg); p->tpos = S.s_rbeg - 1; p->dirs = 3;
        p->w = opt.w << S.attempt; p->h0 = S.s_len * opt.a; p->end_bonus = opt.pen_clip5;
    } else {
        const int qe = S.s_qbeg + S.s_len;
        p->q0 = qe; p->qlen = l_query - qe; p->tlen = (int)(S.win_end - (S.s_rbeg + S.s_len)); p->tpos = S.s_rbeg + S.s_len; p->dirs = 0;
        p->w = opt.w << S.attempt; p->h0 = S.score_left; p->end_bonus = opt.pen_clip3;
    }
}

// chains / order / kseeds / av / srt: this read's slices of the batch arrays (k_chain's outputs, region and sort scratch).
// res: result of the problem emitted by the previous call (ignored in the first call).  Returns 1: *prob to be solved;
// 0: finished, S.n_regions regions in av; -3: more regions than the read's capacity (cannot happen: one per kept seed).
AR_HD int ext_advance(const MemOpt& opt, const RefView& ref, int read, int l_query, int n_chains, const Chain* chains, const int32_t* order,
                      const Seed* kseeds, AlnReg* av, int cap_av, uint64_t* srt, ExtState& S, const ExtRes* res, ExtProb* prob) {
    for (;;) {
        switch (S.phase) {
        case EXT_PH_CHAIN: {
            if (S.chain_i >= n_chains) { S.phase = EXT_PH_DONE; return 0; }
            const Chain c = chains[order[S.chain_i++]];
            if (c.n == 0) break;
            S.n_seeds = c.n; S.seed0 = c.first_seed; S.rid = c.rid; S.frac_rep = c.frac_rep;
            const Seed* seeds = kseeds + S.seed0;
            ext_chain_window(opt, ref, l_query, seeds, c.n, &S.win_beg, &S.win_end);
            for (int x = 0; x < c.n; ++x) srt[x] = (uint64_t)seeds[x].score << 32 | (uint32_t)x;  // ties fall as ks_introsort leaves them (bwamem.c:664-667)
            ks_introsort((size_t)c.n, srt, [](uint64_t a, uint64_t b) { return a < b; });
            S.seed_k = c.n - 1;
            S.phase = EXT_PH_SEED;
        } break;
        case EXT_PH_SEED: {
            if (S.seed_k < 0) { S.phase = EXT_PH_CHAIN; break; }
            const Seed* seeds = kseeds + S.seed0;
            const Seed s = seeds[(uint32_t)srt[S.seed_k]];
            bool explained = false;
            for (int x = 0; x < S.n_regions && !explained; ++x) explained = ext_seed_explained(opt, l_query, s, av[x]);
            if (explained && !ext_overlap_disagrees(seeds, srt, S.n_seeds, S.seed_k, s)) { srt[S.seed_k] = 0; --S.seed_k; break; }
            if (S.n_regions >= cap_av) return -3;
            S.s_rbeg = s.rbeg; S.s_qbeg = s.qbeg; S.s_len = s.len;
            S.band_l = S.band_r = opt.w; S.score = S.truesc = -1; S.attempt = 0;
            if (s.qbeg > 0) { S.score_before = S.score; S.phase = EXT_PH_LEFT; ext_emit(opt, S, read, l_query, false, prob); return 1; }
            S.score = S.truesc = s.len * opt.a; S.q_beg = 0; S.r_beg = s.rbeg;  // the seed starts the read: nothing on the left
            S.phase = EXT_PH_RIGHT; S.attempt = -1;                          // -1: enter the right flank without a result to consume
        } break;
        case EXT_PH_LEFT: {  // bwamem.c:715-744
            S.band_l = opt.w << S.attempt;
            S.score = res->mx;
            const bool settled = S.score == S.score_before || res->max_off < (S.band_l >> 1) + (S.band_l >> 2);
            if (!settled && S.attempt + 1 < 2) { ++S.attempt; S.score_before = S.score; ext_emit(opt, S, read, l_query, false, prob); return 1; }
            if (res->gscore <= 0 || res->gscore <= S.score - opt.pen_clip5) {  // local end beats reaching the read start
                S.q_beg = S.s_qbeg - (res->max_j + 1); S.r_beg = S.s_rbeg - (res->max_i + 1); S.truesc = S.score;
            } else {
                S.q_beg = 0; S.r_beg = S.s_rbeg - (res->max_ie + 1); S.truesc = res->gscore;
            }
            S.phase = EXT_PH_RIGHT; S.attempt = -1;
        } break;
        case EXT_PH_RIGHT: {  // bwamem.c:746-771
            const int seed_end = S.s_qbeg + S.s_len;
            if (S.attempt < 0) {  // first visit
                if (seed_end == l_query) { S.q_end = l_query; S.r_end = S.s_rbeg + S.s_len; }
                else { S.attempt = 0; S.score_left = S.score; S.score_before = S.score; ext_emit(opt, S, read, l_query, true, prob); return 1; }
            } else {
                S.band_r = opt.w << S.attempt;
                S.score = res->mx;
                const bool settled = S.score == S.score_before || res->max_off < (S.band_r >> 1) + (S.band_r >> 2);
                if (!settled && S.attempt + 1 < 2) { ++S.attempt; S.score_before = S.score; ext_emit(opt, S, read, l_query, true, prob); return 1; }
                if (res->gscore <= 0 || res->gscore <= S.score - opt.pen_clip3) {
                    S.q_end = seed_end + res->max_j + 1; S.r_end = S.s_rbeg + S.s_len + res->max_i + 1; S.truesc += S.score - S.score_left;
                } else {
                    S.q_end = l_query; S.r_end = S.s_rbeg + S.s_len + res->max_ie + 1; S.truesc += res->gscore - S.score_left;
                }
            }
            // the region is complete (bwamem.c:773-783)
            const Seed* seeds = kseeds + S.seed0;
            AlnReg a;
            a.rb = S.r_beg; a.re = S.r_end; a.qb = S.q_beg; a.qe = S.q_end; a.rid = S.rid; a.score = S.score; a.truesc = S.truesc;
            a.sub = a.alt_sc = a.csub = a.sub_n = 0; a.secondary = a.secondary_all = 0; a.n_comp_is_alt = 0; a.hash = 0;
            int cov = 0;
            for (int x = 0; x < S.n_seeds; ++x) {
                const Seed t = seeds[x];
                if (t.qbeg >= a.qb && t.qbeg + t.len <= a.qe && t.rbeg >= a.rb && t.rbeg + t.len <= a.re) cov += t.len;
            }
            a.seedcov = cov;
            a.w = S.band_l > S.band_r ? S.band_l : S.band_r;
            a.seedlen0 = S.s_len;
            a.frac_rep = S.frac_rep;
            av[S.n_regions++] = a;
            --S.seed_k;
            S.phase = EXT_PH_SEED;
        } break;
        default: return 0;
        }
    }
}

}  // namespace arachne
