// Seed chaining and chain filtering, thread-sequential per read (k_chain runs one read per lane).  The DECISIONS are the reference's --
// which chain a seed joins, where equal keys land in the B-tree, how ties sort -- because every one of them is observable in the
// regions that come out; the code that takes them is written for this library.
//   mem_chain        bwa/bwamem.c:251-315   (B-tree keyed by chain pos; kbtree.h semantics reproduced exactly,
//                                            including where equal keys land, because that decides `lower`)
//   test_and_merge   bwa/bwamem.c:190-211
//   mem_chain_weight bwa/bwamem.c:213-232
//   mem_chain_flt    bwa/bwamem.c:327-385   (klib introsort by weight: tie order is observable)
//   bns_pos2rid / bns_intv2rid / bns_depos  bwa/bntseq.c:349-373, bntseq.h:86-89
#pragma once
#include "common.cuh"

namespace arachne {

// forward-strand coordinate of a forward-reverse coordinate, and which strand it was on (bns_depos, bntseq.h:86-89)
AR_HD int64_t bns_depos(const RefView& r, int64_t pos, int* is_rev) {
    *is_rev = pos >= r.l_pac;
    return *is_rev ? 2 * r.l_pac - 1 - pos : pos;
}

// contig holding forward coordinate pos_f: the last contig whose offset is <= pos_f (bns_pos2rid, bntseq.c:346-363 returns the
// same index: its bisection stops exactly there); -1 beyond the reference
AR_HD int bns_pos2rid(const RefView& r, int64_t pos_f) {
    if (pos_f >= r.l_pac) return -1;
    int lo = 0, hi = r.n_seqs;          // invariant: offset[lo] <= pos_f (offset[0] == 0), answer in [lo, hi)
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (r.contigs[mid].offset <= pos_f) lo = mid; else hi = mid;
    }
    return lo;
}

// contig of the interval [rb, re) when it lies on one strand and in one contig; -2 across the strand boundary, -1 across contigs
// (bns_intv2rid, bntseq.c:365-373)
AR_HD int bns_intv2rid(const RefView& r, int64_t rb, int64_t re) {
    if (rb < r.l_pac && re > r.l_pac) return -2;
    int strand;
    const int first = bns_pos2rid(r, bns_depos(r, rb, &strand));
    const int last = rb < re ? bns_pos2rid(r, bns_depos(r, re - 1, &strand)) : first;
    return first == last ? first : -1;
}

// ---- kbtree with t = 5 (KB_DEFAULT_SIZE 512, 40-byte keys: kbtree.h:55-71), keys are chain ids ----
constexpr int KB_T = 5;
constexpr int KB_MAXKEYS = 2 * KB_T - 1;

struct BNode {
    int32_t n, internal;
    int32_t key[KB_MAXKEYS];
    int32_t child[KB_MAXKEYS + 1];
};

struct BTree {
    BNode* nodes;
    const Chain* chains;
    int n_nodes, cap_nodes, root, n_keys;

    AR_HD void init(BNode* mem, int cap, const Chain* ch) {
        nodes = mem; cap_nodes = cap; chains = ch; n_nodes = 1; root = 0; n_keys = 0;
        nodes[0].n = 0; nodes[0].internal = 0;
    }
    AR_HD int alloc_node() {
        int z = n_nodes++;
        nodes[z].n = 0; nodes[z].internal = 0;
        return z;
    }
    // __kb_getp_aux: index of the first key equal to pos (r = 0) else of the last key smaller (r != 0)
    AR_HD int getp_aux(const BNode& x, int64_t pos, int* r) const {
        int begin = 0, end = x.n;
        if (x.n == 0) return -1;
        while (begin < end) {
            int mid = (begin + end) >> 1;
            if (chains[x.key[mid]].pos < pos) begin = mid + 1;
            else end = mid;
        }
        if (begin == x.n) { *r = 1; return x.n - 1; }
        int64_t kp = chains[x.key[begin]].pos;
        *r = (kp < pos) - (pos < kp);
        if (*r < 0) --begin;
        return begin;
    }
    // kb_intervalp: the chain id playing `lower`, or -1
    AR_HD int lower(int64_t pos) const {
        int x = root, low = -1, r = 0;
        while (x >= 0) {
            const BNode& nd = nodes[x];
            int i = getp_aux(nd, pos, &r);
            if (i >= 0 && r == 0) return nd.key[i];
            if (i >= 0) low = nd.key[i];
            if (!nd.internal) return low;
            x = nd.child[i + 1];
        }
        return low;
    }
    AR_HD void split(int xi, int i, int yi) {
        int zi = alloc_node();
        BNode &x = nodes[xi], &y = nodes[yi], &z = nodes[zi];
        z.internal = y.internal;
        z.n = KB_T - 1;
        for (int k = 0; k < KB_T - 1; ++k) z.key[k] = y.key[KB_T + k];
        if (y.internal) for (int k = 0; k < KB_T; ++k) z.child[k] = y.child[KB_T + k];
        y.n = KB_T - 1;
        for (int k = x.n; k > i; --k) x.child[k + 1] = x.child[k];
        x.child[i + 1] = zi;
        for (int k = x.n - 1; k >= i; --k) x.key[k + 1] = x.key[k];
        x.key[i] = y.key[KB_T - 1];
        ++x.n;
    }
    AR_HD bool put(int cid) {
        if (n_nodes + 2 > cap_nodes) return false;
        int64_t pos = chains[cid].pos;
        int r = root, rr = 0;
        ++n_keys;
        if (nodes[r].n == KB_MAXKEYS) {
            int s = alloc_node();
            root = s; nodes[s].internal = 1; nodes[s].n = 0; nodes[s].child[0] = r;
            split(s, 0, r);
            r = s;
        }
        int x = r;
        for (;;) {
            BNode& nd = nodes[x];
            if (!nd.internal) {
                int i = getp_aux(nd, pos, &rr);
                for (int k = nd.n - 1; k > i; --k) nd.key[k + 1] = nd.key[k];
                nd.key[i + 1] = cid;
                ++nd.n;
                return true;
            }
            int i = getp_aux(nd, pos, &rr) + 1;
            if (nodes[nd.child[i]].n == KB_MAXKEYS) {
                if (n_nodes + 1 > cap_nodes) return false;
                split(x, i, nd.child[i]);
                int64_t kp = chains[nodes[x].key[i]].pos;
                if (pos > kp) ++i;
            }
            x = nodes[x].child[i];
        }
    }
    // in-order traversal (__kb_traverse) into out[]; returns the count
    AR_HD int traverse(int32_t* out) const {
        int sx[24], si[24], sp = 0, n = 0;
        sx[0] = root; si[0] = 0;
        for (;;) {
            while (sx[sp] >= 0 && si[sp] <= nodes[sx[sp]].n) {
                const BNode& nd = nodes[sx[sp]];
                sx[sp + 1] = nd.internal ? nd.child[si[sp]] : -1;
                si[sp + 1] = 0;
                ++sp;
            }
            --sp;
            if (sp >= 0) {
                if (sx[sp] >= 0 && si[sp] < nodes[sx[sp]].n) out[n++] = nodes[sx[sp]].key[si[sp]];
                ++si[sp];
            } else break;
        }
        return n;
    }
};

// Can seed `sid` join chain c?  (test_and_merge, bwamem.c:190-211.)  Three outcomes: it lies inside the chain's query and reference
// span (absorbed: nothing changes), it continues the chain's last seed on a nearby diagonal (appended), or it cannot (0).
AR_HD int test_and_merge(const MemOpt& opt, int64_t l_pac, Chain& c, Seed* seeds, int sid, int seed_rid) {
    if (seed_rid != c.rid) return 0;
    const Seed s = seeds[sid], head = seeds[c.first_seed], tail = seeds[c.last_seed];
    const int64_t chain_qend = tail.qbeg + tail.len, chain_rend = tail.rbeg + tail.len;
    const bool inside = s.qbeg >= head.qbeg && s.qbeg + s.len <= chain_qend && s.rbeg >= head.rbeg && s.rbeg + s.len <= chain_rend;
    if (inside) return 1;
    const bool chain_on_forward = tail.rbeg < l_pac || head.rbeg < l_pac;
    if (chain_on_forward && s.rbeg >= l_pac) return 0;  // a chain never continues onto the other strand
    const int64_t dq = s.qbeg - tail.qbeg, dr = s.rbeg - tail.rbeg;  // offsets from the last seed on the read / on the reference
    const bool colinear = dr >= 0 && dq - dr <= opt.w && dr - dq <= opt.w;
    const bool close = dq - tail.len < opt.max_chain_gap && dr - tail.len < opt.max_chain_gap;
    if (!(colinear && close)) return 0;
    seeds[c.last_seed].next = sid;
    c.last_seed = sid;
    ++c.n;
    return 1;
}

// Bases of the read (or of the reference) covered by the chain's seeds, overlaps counted once; the chain's weight is the smaller
// of the two coverages, capped at 2^30 - 1 (mem_chain_weight, bwamem.c:213-232)
AR_HD int chain_weight(const Chain& c, const Seed* seeds) {
    int64_t q_cov = 0, r_cov = 0, q_end = 0, r_end = 0;
    for (int k = 0, s = c.first_seed; k < c.n; ++k, s = seeds[s].next) {
        const Seed sd = seeds[s];
        const int64_t qe = (int64_t)sd.qbeg + sd.len, re = sd.rbeg + sd.len;
        if (sd.qbeg >= q_end) q_cov += sd.len; else if (qe > q_end) q_cov += qe - q_end;
        if (sd.rbeg >= r_end) r_cov += sd.len; else if (re > r_end) r_cov += re - r_end;
        q_end = q_end > qe ? q_end : qe;
        r_end = r_end > re ? r_end : re;
    }
    // the reference accumulates each coverage in an int; both stay far below 2^31 for reads of a few hundred bases
    const int64_t w = q_cov < r_cov ? q_cov : r_cov;
    return w < (1 << 30) ? (int)w : (1 << 30) - 1;
}

// mem_chain over precomputed SA values (sa_vals in the reference's visiting order: intervals in sorted order,
// occurrences k = 0, step, 2*step ...).  seeds/chains have room for n_sa entries; nodes for n_sa/4+4; order for n_sa.
// Returns the number of chains (in-order in order[]), or -1 if scratch ran out.
AR_HD int chain_seeds(const MemOpt& opt, const RefView& ref, int len, const Intv* intv, int n_intv, const uint64_t* sa_vals,
                      Seed* seeds, Chain* chains, BNode* nodes, int cap_nodes, int32_t* order, float* frac_rep_out) {
    int i, b, e, l_rep, n_seed = 0, n_chain = 0;
    if (len < opt.min_seed_len) { *frac_rep_out = 0.f; return 0; }
    BTree tree;
    tree.init(nodes, cap_nodes, chains);
    for (i = 0, b = e = l_rep = 0; i < n_intv; ++i) {  // frac_rep
        int sb = (int)(intv[i].info >> 32), se = (int)(uint32_t)intv[i].info;
        if (intv[i].x2 <= (uint64_t)opt.max_occ) continue;
        if (sb > e) { l_rep += e - b; b = sb; e = se; }
        else e = e > se ? e : se;
    }
    l_rep += e - b;
    int64_t vi = 0;
    for (i = 0; i < n_intv; ++i) {
        const Intv& p = intv[i];
        int slen = (int)((uint32_t)p.info - (uint32_t)(p.info >> 32));
        int64_t step = p.x2 > (uint64_t)opt.max_occ ? (int64_t)(p.x2 / opt.max_occ) : 1;
        int count = 0;
        for (int64_t k = 0; k < (int64_t)p.x2 && count < opt.max_occ; k += step, ++count) {
            int sid = n_seed;
            Seed& s = seeds[sid];
            s.rbeg = (int64_t)sa_vals[vi++];
            s.qbeg = (int32_t)(p.info >> 32);
            s.score = s.len = slen;
            s.next = -1;
            int rid = bns_intv2rid(ref, s.rbeg, s.rbeg + s.len);
            if (rid < 0) continue;  // bridging contigs or the forward-reverse boundary
            ++n_seed;
            int to_add = 0;
            if (tree.n_keys) {
                int lo = tree.lower(s.rbeg);
                if (lo < 0 || !test_and_merge(opt, ref.l_pac, chains[lo], seeds, sid, rid)) to_add = 1;
            } else to_add = 1;
            if (to_add) {
                Chain& c = chains[n_chain];
                c.pos = s.rbeg; c.rid = rid; c.n = 1; c.first_seed = c.last_seed = sid;
                c.w = 0; c.first = -1; c.kept = 0; c.is_alt = (int8_t)(ref.contigs[rid].is_alt != 0); c.pad = 0; c.frac_rep = 0.f;
                if (!tree.put(n_chain)) return -1;
                ++n_chain;
            }
        }
    }
    int n = tree.traverse(order);
    float fr = (float)l_rep / len;
    for (i = 0; i < n; ++i) chains[order[i]].frac_rep = fr;
    *frac_rep_out = fr;
    return n;
}

// mem_chain_flt (bwamem.c:327-385) on order[0..n_chn).  Chains are ranked by weight (klib introsort: the permutation of equal weights is
// part of the result), then taken greedily: a chain is KEPT unless an already kept chain overlaps it on the read by at least
// mask_level of the shorter one and outweighs it clearly (drop_ratio, 2 x min_seed_len).  A kept chain remembers the first chain
// it shadowed (`first`), which is kept too, as class 1; class 3 = kept without any significant overlap, 2 = kept despite one.
// Of the classes 1 and 2 only the first max_chain_extend survive.  kept_idx: scratch of n ints.  Leaves the surviving ids, in rank
// order, in order[0..ret).
struct QuerySpan { int beg, end; };
AR_HD QuerySpan chain_query_span(const Chain& c, const Seed* seeds) {
    QuerySpan q; q.beg = seeds[c.first_seed].qbeg; q.end = seeds[c.last_seed].qbeg + seeds[c.last_seed].len; return q;
}
AR_HD int chain_filter(const MemOpt& opt, int n_chn, int32_t* order, Chain* chains, const Seed* seeds, int32_t* kept_idx) {
    if (n_chn == 0) return 0;
    int n = 0;
    for (int i = 0; i < n_chn; ++i) {  // weigh, drop the light ones
        Chain& c = chains[order[i]];
        c.first = -1; c.kept = 0;
        c.w = chain_weight(c, seeds);
        if (c.w >= opt.min_chain_weight) order[n++] = order[i];
    }
    if (n == 0) return 0;
    const Chain* by_id = chains;
    ks_introsort((size_t)n, order, [by_id](int32_t a, int32_t b) { return by_id[a].w > by_id[b].w; });
    auto at = [&](int rank) -> Chain& { return chains[order[rank]]; };
    int n_kept = 0;
    at(0).kept = 3;
    kept_idx[n_kept++] = 0;
    for (int i = 1; i < n; ++i) {
        Chain& cand = at(i);
        const QuerySpan ci = chain_query_span(cand, seeds);
        bool overlapped = false, dropped = false;
        for (int k = 0; k < n_kept && !dropped; ++k) {
            Chain& keeper = at(kept_idx[k]);
            const QuerySpan cj = chain_query_span(keeper, seeds);
            const int ov_beg = cj.beg > ci.beg ? cj.beg : ci.beg, ov_end = cj.end < ci.end ? cj.end : ci.end;
            if (ov_end <= ov_beg || (keeper.is_alt && !cand.is_alt)) continue;  // an ALT chain never shadows a primary one
            const int len_i = ci.end - ci.beg, len_j = cj.end - cj.beg, shorter = len_i < len_j ? len_i : len_j;
            if (!(ov_end - ov_beg >= shorter * opt.mask_level && shorter < opt.max_chain_gap)) continue;
            overlapped = true;
            if (keeper.first < 0) keeper.first = i;
            dropped = cand.w < keeper.w * opt.drop_ratio && keeper.w - cand.w >= opt.min_seed_len << 1;
        }
        if (!dropped) { kept_idx[n_kept++] = i; cand.kept = overlapped ? 2 : 3; }
    }
    for (int k = 0; k < n_kept; ++k) {  // the first chain each kept chain shadows stays as class 1 (it may already be class 2 / 3: overwritten as in the reference)
        const int sh = at(kept_idx[k]).first;
        if (sh >= 0) at(sh).kept = 1;
    }
    int quota = 0, i = 0;
    for (; i < n; ++i) {  // classes 1 and 2 count against max_chain_extend ...
        const int kc = at(i).kept;
        if (kc == 0 || kc == 3) continue;
        if (++quota >= opt.max_chain_extend) break;
    }
    for (; i < n; ++i)     // ... everything of those classes from the chain that exhausts it onwards is dropped (that chain included, as at bwamem.c:376-379)
        if (at(i).kept < 3) at(i).kept = 0;
    int out = 0;
    for (i = 0; i < n; ++i)
        if (at(i).kept != 0) order[out++] = order[i];
    return out;
}

}  // namespace arachne
