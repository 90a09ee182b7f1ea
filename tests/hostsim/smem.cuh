// TEST-ONLY (not part of libarachne_cuda): a thread-sequential restatement of bwt_smem1a (bwa/bwt.c:289-351),
// bwt_seed_strategy1 (:358-379) and the three-pass mem_collect_intv (bwa/bwamem.c:114-162) over the product's FM-index
// primitives (csrc/fmindex.cuh).  The host simulation uses it only as the SOURCE of seed intervals for the chaining /
// extension-driver / dedup code it pins; the shipped seeding kernel is k_seed2 (csrc/seed_kernel.cuh), pinned on the GPU.
#pragma once
#include "../../arachne_b200/csrc/fmindex.cuh"

namespace arachne {

AR_HD void intv_reverse(Intv* a, int n) {
    for (int j = 0; j < n >> 1; ++j) {
        Intv t = a[n - 1 - j]; a[n - 1 - j] = a[j]; a[j] = t;
    }
}

// All SMEMs through position x.  mem receives them sorted by start; buf0/buf1 are scratch of >= len+1 entries.
// Returns the position to continue from.  max_intv is 0 on the path Arachne takes (bwt_smem1).
AR_HD int smem1(const FMIndex& fm, int len, const uint8_t* q, int x, int min_intv, Intv* mem, int* n_mem, Intv* buf0, Intv* buf1, OccStat* st = nullptr) {
    Intv ik, ok[4];
    Intv *prev = buf0, *curr = buf1, *swap;
    int i, j, c, ret, n_curr = 0, n_prev, nm = 0;
    *n_mem = 0;
    if (q[x] > 3) return x + 1;
    if (min_intv < 1) min_intv = 1;
    fm_set_intv(fm, q[x], ik);
    ik.info = x + 1;
    for (i = x + 1; i < len; ++i) {  // forward search
        if (q[i] < 4) {
            c = 3 - q[i];
            fm_extend(fm, ik, ok, 0, st);
            if (ok[c].x2 != ik.x2) {  // change of the interval size
                curr[n_curr++] = ik;
                if (ok[c].x2 < (uint64_t)min_intv) break;
            }
            ik = ok[c]; ik.info = i + 1;
        } else {
            curr[n_curr++] = ik;
            break;
        }
    }
    if (i == len) curr[n_curr++] = ik;
    intv_reverse(curr, n_curr);  // longer matches first
    ret = (int)curr[0].info;
    swap = curr; curr = prev; prev = swap; n_prev = n_curr;

    for (i = x - 1; i >= -1; --i) {  // backward search for MEMs
        c = i < 0 ? -1 : q[i] < 4 ? q[i] : -1;
        for (j = 0, n_curr = 0; j < n_prev; ++j) {
            const Intv p = prev[j];
            if (c >= 0) fm_extend(fm, p, ok, 1, st);
            if (c < 0 || ok[c].x2 < (uint64_t)min_intv) {
                if (n_curr == 0) {  // no longer match has survived this step
                    if (nm == 0 || (uint64_t)(i + 1) < (mem[nm - 1].info >> 32)) {  // skip contained matches
                        ik = p; ik.info |= (uint64_t)(i + 1) << 32;
                        mem[nm++] = ik;
                    }
                }
            } else if (n_curr == 0 || ok[c].x2 != curr[n_curr - 1].x2) {
                ok[c].info = p.info;
                curr[n_curr++] = ok[c];
            }
        }
        if (n_curr == 0) break;
        swap = curr; curr = prev; prev = swap; n_prev = n_curr;
    }
    intv_reverse(mem, nm);
    *n_mem = nm;
    return ret;
}

// LAST-like seeding: first extension whose interval drops below max_intv once >= min_len long
AR_HD int seed_strategy1(const FMIndex& fm, int len, const uint8_t* q, int x, int min_len, int max_intv, Intv* mem, OccStat* st = nullptr) {
    Intv ik, ok[4];
    mem->x0 = mem->x1 = mem->x2 = mem->info = 0;
    if (q[x] > 3) return x + 1;
    fm_set_intv(fm, q[x], ik);
    for (int i = x + 1; i < len; ++i) {
        if (q[i] < 4) {
            int c = 3 - q[i];
            fm_extend(fm, ik, ok, 0, st);
            if (ok[c].x2 < (uint64_t)max_intv && i - x >= min_len) {
                *mem = ok[c];
                mem->info = (uint64_t)x << 32 | (uint32_t)(i + 1);
                return i + 1;
            }
            ik = ok[c];
        } else return i + 1;
    }
    return len;
}

// mem_collect_intv.  out: capacity cap; scratch: 3*(len+1) Intv.  Returns the number of intervals, or -1 on overflow.
AR_HD int collect_intv(const MemOpt& opt, const FMIndex& fm, int len, const uint8_t* seq, Intv* out, int cap, Intv* scratch, OccStat* st = nullptr) {
    Intv* mem1 = scratch;
    Intv* b0 = scratch + (len + 1);
    Intv* b1 = scratch + 2 * (len + 1);
    int n = 0, n1, x = 0, old_n, i, k;
    int split_len = (int)(opt.min_seed_len * opt.split_factor + .499);
    while (x < len) {  // pass 1: all SMEMs
        if (seq[x] < 4) {
            x = smem1(fm, len, seq, x, 1, mem1, &n1, b0, b1, st);
            for (i = 0; i < n1; ++i) {
                int slen = (int)((uint32_t)mem1[i].info - (uint32_t)(mem1[i].info >> 32));
                if (slen >= opt.min_seed_len) {
                    if (n >= cap) return -1;
                    out[n++] = mem1[i];
                }
            }
        } else ++x;
    }
    old_n = n;
    for (k = 0; k < old_n; ++k) {  // pass 2: re-seed inside long, rare SMEMs
        int start = (int)(out[k].info >> 32), end = (int32_t)out[k].info;
        if (end - start < split_len || out[k].x2 > (uint64_t)opt.split_width) continue;
        smem1(fm, len, seq, (start + end) >> 1, (int)(out[k].x2 + 1), mem1, &n1, b0, b1, st);
        for (i = 0; i < n1; ++i)
            if ((int)((uint32_t)mem1[i].info - (uint32_t)(mem1[i].info >> 32)) >= opt.min_seed_len) {
                if (n >= cap) return -1;
                out[n++] = mem1[i];
            }
    }
    if (opt.max_mem_intv > 0) {  // pass 3: LAST-like
        x = 0;
        while (x < len) {
            if (seq[x] < 4) {
                Intv m;
                x = seed_strategy1(fm, len, seq, x, opt.min_seed_len, (int)opt.max_mem_intv, &m, st);
                if (m.x2 > 0) {
                    if (n >= cap) return -1;
                    out[n++] = m;
                }
            } else ++x;
        }
    }
    ks_introsort((size_t)n, out, [](const Intv& a, const Intv& b) { return a.info < b.info; });
    return n;
}

}  // namespace arachne
