// Seed chaining and chain filtering, thread-sequential per read.
//   mem_chain        bwa/bwamem.c:251-315   (B-tree keyed by chain pos; kbtree.h semantics reproduced exactly,
//                                            including where equal keys land, because that decides `lower`)
//   test_and_merge   bwa/bwamem.c:190-211
//   mem_chain_weight bwa/bwamem.c:213-232
//   mem_chain_flt    bwa/bwamem.c:327-385   (klib introsort by weight: tie order is observable)
//   bns_pos2rid / bns_intv2rid / bns_depos  bwa/bntseq.c:349-373, bntseq.h:86-89
#pragma once
#include "common.cuh"

namespace arachne {

AR_HD int64_t bns_depos(const RefView& r, int64_t pos, int* is_rev) {
    return (*is_rev = (pos >= r.l_pac)) ? (r.l_pac << 1) - 1 - pos : pos;
}

AR_HD int bns_pos2rid(const RefView& r, int64_t pos_f) {
    int left, mid, right;
    if (pos_f >= r.l_pac) return -1;
    left = 0; mid = 0; right = r.n_seqs;
    while (left < right) {
        mid = (left + right) >> 1;
        if (pos_f >= r.contigs[mid].offset) {
            if (mid == r.n_seqs - 1) break;
            if (pos_f < r.contigs[mid + 1].offset) break;
            left = mid + 1;
        } else right = mid;
    }
    return mid;
}

AR_HD int bns_intv2rid(const RefView& r, int64_t rb, int64_t re) {
    int is_rev, rid_b, rid_e;
    if (rb < r.l_pac && re > r.l_pac) return -2;
    rid_b = bns_pos2rid(r, bns_depos(r, rb, &is_rev));
    rid_e = rb < re ? bns_pos2rid(r, bns_depos(r, re - 1, &is_rev)) : rid_b;
    return rid_b == rid_e ? rid_b : -1;
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

// test_and_merge: 1 if the seed is absorbed by (or appended to) chain c
AR_HD int test_and_merge(const MemOpt& opt, int64_t l_pac, Chain& c, Seed* seeds, int sid, int seed_rid) {
    const Seed& p = seeds[sid];
    const Seed& last = seeds[c.last_seed];
    const Seed& first = seeds[c.first_seed];
    int64_t qend = last.qbeg + last.len, rend = last.rbeg + last.len, x, y;
    if (seed_rid != c.rid) return 0;
    if (p.qbeg >= first.qbeg && p.qbeg + p.len <= qend && p.rbeg >= first.rbeg && p.rbeg + p.len <= rend) return 1;
    if ((last.rbeg < l_pac || first.rbeg < l_pac) && p.rbeg >= l_pac) return 0;
    x = p.qbeg - last.qbeg;
    y = p.rbeg - last.rbeg;
    if (y >= 0 && x - y <= opt.w && y - x <= opt.w && x - last.len < opt.max_chain_gap && y - last.len < opt.max_chain_gap) {
        seeds[c.last_seed].next = sid;
        c.last_seed = sid;
        ++c.n;
        return 1;
    }
    return 0;
}

AR_HD int chain_weight(const Chain& c, const Seed* seeds) {
    int64_t end;
    int j, s, w = 0, tmp;
    for (j = 0, end = 0, s = c.first_seed; j < c.n; ++j, s = seeds[s].next) {
        const Seed& sd = seeds[s];
        if (sd.qbeg >= end) w += sd.len;
        else if (sd.qbeg + sd.len > end) w += (int)(sd.qbeg + sd.len - end);
        end = end > sd.qbeg + sd.len ? end : sd.qbeg + sd.len;
    }
    tmp = w; w = 0;
    for (j = 0, end = 0, s = c.first_seed; j < c.n; ++j, s = seeds[s].next) {
        const Seed& sd = seeds[s];
        if (sd.rbeg >= end) w += sd.len;
        else if (sd.rbeg + sd.len > end) w += (int)(sd.rbeg + sd.len - end);
        end = end > sd.rbeg + sd.len ? end : sd.rbeg + sd.len;
    }
    w = w < tmp ? w : tmp;
    return w < (1 << 30) ? w : (1 << 30) - 1;
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

// mem_chain_flt on order[0..n): weights, introsort, keep/shadow marking.  kept_idx is scratch of n ints.
// Leaves the surviving chain ids (in the reference's output order) in order[0..ret).
AR_HD int chain_filter(const MemOpt& opt, int n_chn, int32_t* order, Chain* chains, const Seed* seeds, int32_t* kept_idx) {
    int i, k, n_kept = 0;
    if (n_chn == 0) return 0;
    for (i = k = 0; i < n_chn; ++i) {
        Chain& c = chains[order[i]];
        c.first = -1; c.kept = 0;
        c.w = chain_weight(c, seeds);
        if (c.w < opt.min_chain_weight) continue;
        order[k++] = order[i];
    }
    n_chn = k;
    const Chain* cc = chains;
    ks_introsort((size_t)n_chn, order, [cc](int32_t a, int32_t b) { return cc[a].w > cc[b].w; });
#define CH(i) chains[order[i]]
#define CBEG(c) (seeds[(c).first_seed].qbeg)
#define CEND(c) (seeds[(c).last_seed].qbeg + seeds[(c).last_seed].len)
    if (n_chn == 0) return 0;
    CH(0).kept = 3;
    kept_idx[n_kept++] = 0;
    for (i = 1; i < n_chn; ++i) {
        int large_ovlp = 0;
        Chain& ci = CH(i);
        int bi = CBEG(ci), ei = CEND(ci);
        for (k = 0; k < n_kept; ++k) {
            int j = kept_idx[k];
            Chain& cj = CH(j);
            int bj = CBEG(cj), ej = CEND(cj);
            int b_max = bj > bi ? bj : bi;
            int e_min = ej < ei ? ej : ei;
            if (e_min > b_max && (!cj.is_alt || ci.is_alt)) {
                int li = ei - bi, lj = ej - bj;
                int min_l = li < lj ? li : lj;
                if (e_min - b_max >= min_l * opt.mask_level && min_l < opt.max_chain_gap) {
                    large_ovlp = 1;
                    if (cj.first < 0) cj.first = i;
                    if (ci.w < cj.w * opt.drop_ratio && cj.w - ci.w >= opt.min_seed_len << 1) break;
                }
            }
        }
        if (k == n_kept) {
            kept_idx[n_kept++] = i;
            ci.kept = large_ovlp ? 2 : 3;
        }
    }
    for (i = 0; i < n_kept; ++i) {
        Chain& c = CH(kept_idx[i]);
        if (c.first >= 0) CH(c.first).kept = 1;
    }
    for (i = k = 0; i < n_chn; ++i) {
        if (CH(i).kept == 0 || CH(i).kept == 3) continue;
        if (++k >= opt.max_chain_extend) break;
    }
    for (; i < n_chn; ++i)
        if (CH(i).kept < 3) CH(i).kept = 0;
    for (i = k = 0; i < n_chn; ++i)
        if (CH(i).kept != 0) order[k++] = order[i];
#undef CH
#undef CBEG
#undef CEND
    return k;
}

}  // namespace arachne
