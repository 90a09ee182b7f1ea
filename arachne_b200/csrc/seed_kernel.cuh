// k_seed2: mem_collect_intv (bwa/bwamem.c:114-162) for a whole batch as a convergent per-lane state machine.
//
// The reference nests four loops per read (bwt_smem1a forward pass, backward pass over a shrinking interval list,
// re-seeding, the LAST-like pass).  Run one-read-per-lane as written, a warp executes ~4 of 32 lanes at a time
// (profiles/r1a_seed_baseline.md).  Here every lane is a small automaton and EVERY iteration of the one outer loop is
// "each live lane performs one bwt_extend" (bwt.c:262-275): the sector loads, popcounts and interval arithmetic are
// executed convergently by the whole warp, only the cheap bookkeeping around them diverges.
//
//   * work item = (read, kind): kind 0 = passes 1+2 (SMEMs, re-seeding), kind 1 = pass 3 (bwt_seed_strategy1); the
//     three passes only meet in the final sort by `info`, and equal info means an identical interval, so the order in
//     which they emit is not observable.  Lanes are persistent and pull the next item from a global counter per kind
//     (next[0], next[1]); a warp stays on one kind while its queue lasts.
//   * interval lists: bwt_smem1a's prev/curr swap is an in-place compaction (curr never outruns prev), so one list per
//     lane suffices; entries are packed to 16 bytes (x0,x1,x2 40 bits each + the 8-bit query end) and stored
//     row-interleaved across lanes (entry j of every lane of a warp is one 512-byte line).
//   * occ blocks are the 32-byte sectors of fmindex.cuh: k and l cost one sector each (the same one when they share it).
#pragma once
#include "fmindex.cuh"

namespace arachne {

struct SeedOut {
    Intv* intv;            // [n_reads * cap]
    int* intv_n;           // [n_reads], zeroed before launch
    unsigned int* n_sa;    // [n_reads], zeroed before launch
    int cap;
};

__device__ __forceinline__ uint4 pk_pack(uint64_t x0, uint64_t x1, uint64_t x2, unsigned end) {
    uint64_t lo = x0 | (x1 << 40), hi = (x1 >> 24) | (x2 << 16) | ((uint64_t)end << 56);
    return make_uint4((unsigned)lo, (unsigned)(lo >> 32), (unsigned)hi, (unsigned)(hi >> 32));
}
__device__ __forceinline__ void pk_unpack(uint4 v, uint64_t& x0, uint64_t& x1, uint64_t& x2, unsigned& end) {
    uint64_t lo = (uint64_t)v.y << 32 | v.x, hi = (uint64_t)v.w << 32 | v.z;
    x0 = lo & 0xffffffffffull;
    x1 = (lo >> 40) | ((hi & 0xffffull) << 24);
    x2 = (hi >> 16) & 0xffffffffffull;
    end = (unsigned)(hi >> 56);
}

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(a), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory"); }

enum { SP_IDLE = 0, SP_P1_SCAN, SP_FWD, SP_BWD, SP_P2_NEXT, SP_P3_SCAN, SP_P3, SP_DONE };

template <int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS) k_seed2(FMIndex fm, MemOpt opt, const uint8_t* __restrict__ seq, const int64_t* __restrict__ off,
                                                  int n_reads, int maxlen, uint4* lists, int rows, SeedOut out, unsigned int* next, Counters* cnt,
                                                  int use_kmer, int slist_n) {
    const unsigned T = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint4* list = lists + tid;                       // entry j (j >= slist_n) at list[j * T]
    // the first slist_n entries of every lane's list live in shared memory ([entry][lane], 16 bytes each: the bank depends
    // on the lane only, so lanes at different entries never conflict): with ~17 entries per forward pass the lists of the
    // SM's lanes (0.3 MB) do not fit L1, and every backward step would otherwise start with an L2 round trip
    extern __shared__ uint4 slist[];
    uint4* sl = slist + threadIdx.x;
    auto lput = [&](int e, uint4 v) { if (e < slist_n) sl[e * 256] = v; else list[(size_t)e * T] = v; };
    auto lget = [&](int e) -> uint4 { return e < slist_n ? sl[e * 256] : list[(size_t)e * T]; };
    uint4* recs = lists + (size_t)(rows >> 1) * T + tid;  // re-seed records in the upper half
    const int split_len = (int)(opt.min_seed_len * opt.split_factor + .499f);
    // task kinds cost about 2:1 in bwt_extend calls (passes 1+2 vs pass 3): every third warp starts on the LAST-like queue
    // (1:5 when pass 3 starts from the k-mer table)
    const int kmer_k = (use_kmer && fm.kmer && fm.kmer_k <= opt.min_seed_len) ? fm.kmer_k : 0;
    // the SMEM passes' forward extension starts from the table too (all levels 1..K, fetched with K independent copies into
    // the lane's shared list slots) when those slots can hold them
    const bool kmer_fwd = kmer_k && fm.kmer_all && slist_n >= kmer_k;
    bool fresh = false;  // a forward pass was just set up at x
    const int pref_kind = (opt.max_mem_intv > 0 && (tid >> 5) % (kmer_k ? 6 : 3) == 2) ? 1 : 0;

    int phase = SP_IDLE, r = 0, len = 0, x = 0, i = 0, pass = 1;
    const uint8_t* q = seq;
    uint64_t kx0 = 0, kx1 = 0, kx2 = 0, min_intv = 1;  // ik
    unsigned kend = 0;                                  // ik.info (query end)
    int n = 0, lo = 0, hi = 0, j = 0, ncur = 0, last_start = 0x7fffffff, ret = 0, n_rec = 0, rec_i = 0;
    uint64_t lastcur_x2 = 0;
    uint64_t px0 = 0, px1 = 0, px2 = 0; unsigned pend = 0;  // prev[j] of the backward pass
    unsigned st_calls = 0, st_blocks = 0;

    auto emit = [&](uint64_t e0, uint64_t e1, uint64_t e2, int start, int end, bool filter) {
        const int slen = end - start;
        if (filter && slen < opt.min_seed_len) return;
        int slot = atomicAdd(&out.intv_n[r], 1);
        if (slot >= out.cap) { raise(cnt, ARACHNE_ERR_OVERFLOW, r); return; }
        Intv v; v.x0 = e0; v.x1 = e1; v.x2 = e2; v.info = (uint64_t)start << 32 | (unsigned)end;
        out.intv[(size_t)r * out.cap + slot] = v;
        atomicAdd(&out.n_sa[r], e2 > (uint64_t)opt.max_occ ? (unsigned)opt.max_occ : (unsigned)e2);
        if (pass == 1 && filter && slen >= split_len && e2 <= (uint64_t)opt.split_width)
            recs[(size_t)(n_rec++) * T] = make_uint4((unsigned)start, (unsigned)end, (unsigned)e2, 0u);
    };

    for (;;) {
        // ------------------------------------------------------------------ PRE: bookkeeping up to the next extend
        bool need = false;
        int c = 0, is_back = 0;
        uint64_t sx0 = 0, sx1 = 0, sx2 = 0;
        do {
            if (phase == SP_DONE) break;
            if (phase == SP_IDLE) {
                // one queue per task kind; a warp keeps to its own kind until that queue is empty, so its lanes walk the
                // same automaton states (a warp mixing SMEM and LAST-like tasks pays for both code paths every trip)
                int kind = pref_kind;
                unsigned t = atomicAdd(next + kind, 1u);
                if (t >= (unsigned)n_reads) {
                    kind ^= 1;
                    t = atomicAdd(next + kind, 1u);
                    if (t >= (unsigned)n_reads) { phase = SP_DONE; break; }
                }
                r = (int)t;
                len = (int)(off[r + 1] - off[r]);
                q = seq + off[r];
                if (len > maxlen) { if (kind == 0) raise(cnt, ARACHNE_ERR_ARG, r); break; }
                if (len < opt.min_seed_len) break;
                x = 0; pass = 1; n_rec = 0; rec_i = 0;
                phase = kind ? ((opt.max_mem_intv > 0) ? SP_P3_SCAN : SP_IDLE) : SP_P1_SCAN;
                if (phase == SP_IDLE) break;
            }
            if (phase == SP_P2_NEXT) {
                if (rec_i < n_rec) {
                    uint4 rc = recs[(size_t)(rec_i++) * T];
                    x = (int)((rc.x + rc.y) >> 1);
                    min_intv = (uint64_t)rc.z + 1;
                    pass = 2;
                    int b = __ldg(q + x);
                    if (b > 3) break;  // bwt_smem1a returns at once on an ambiguous base (cannot happen inside a match)
                    kx0 = fm_L2(fm, b) + 1; kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); kx1 = fm_L2(fm, 3 - b) + 1; kend = x + 1;
                    i = x + 1; n = 0; last_start = 0x7fffffff;
                    phase = SP_FWD; fresh = true;
                } else { phase = SP_IDLE; break; }
            }
            if (phase == SP_P1_SCAN) {
                while (x < len && __ldg(q + x) > 3) ++x;
                if (x >= len) { phase = SP_P2_NEXT; break; }
                int b = __ldg(q + x);
                kx0 = fm_L2(fm, b) + 1; kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); kx1 = fm_L2(fm, 3 - b) + 1; kend = x + 1;
                min_intv = 1;
                i = x + 1; n = 0; last_start = 0x7fffffff;
                phase = SP_FWD; fresh = true;
            }
            if (phase == SP_P3_SCAN) {
                // bwt_seed_strategy1 (bwt.c:358-379) from x: nothing is tested before the match is min_seed_len + 1 long, so
                // (a) a start with fewer bases left cannot seed, (b) the first kmer_k bases need no stepping: their bi-interval
                // is one table entry (an ambiguous base among them ends the call without a seed, as in the reference).
                for (;;) {
                    while (x < len && __ldg(q + x) > 3) ++x;
                    if (len - x <= opt.min_seed_len) { phase = SP_IDLE; break; }
                    int b = __ldg(q + x);
                    if (kmer_k) {
                        uint32_t km = (uint32_t)b;
                        int p = x + 1;
                        for (; p < x + kmer_k; ++p) { const int bb = __ldg(q + p); if (bb > 3) break; km = km << 2 | (uint32_t)bb; }
                        if (p < x + kmer_k) { x = p + 1; continue; }
                        unsigned spare;
                        pk_unpack(__ldg(reinterpret_cast<const uint4*>(fm.kmer) + km), kx0, kx1, kx2, spare);
                        i = x + kmer_k;
                        st_calls += kmer_k - 1; st_blocks += 2 * (kmer_k - 1);  // the skipped calls; intervals this large span two blocks
                    } else {
                        kx0 = fm_L2(fm, b) + 1; kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); kx1 = fm_L2(fm, 3 - b) + 1;
                        i = x + 1;
                    }
                    phase = SP_P3;
                    break;
                }
                if (phase == SP_IDLE) break;
            }
            if (fresh) {
                // bwt_smem1a's forward loop (bwt.c:300-316) over the first m <= K bases without stepping: the interval after
                // each base is a table entry; what the loop does with them (push on a size change, stop below min_intv) is
                // replayed from shared memory.  An ambiguous base or the read end inside the window ends the window there and
                // the ordinary stepping takes over (and closes the pass).
                fresh = false;
                if (kmer_fwd) {
                    uint32_t km = 0;
                    int m = 0;
                    const int lim = len - x < kmer_k ? len - x : kmer_k;
                    for (; m < lim; ++m) { const int bb = __ldg(q + x + m); if (bb > 3) break; km = km << 2 | (uint32_t)bb; }
                    if (m >= 2) {
                        const uint4* all = reinterpret_cast<const uint4*>(fm.kmer_all);
                        for (int L = 1; L <= m; ++L) cp_async16(&sl[(L - 1) * 256], all + ((1ull << (2 * L)) - 4) / 3 + (km >> (2 * (m - L))));
                        cp_async_wait_all();
                        unsigned spare;
                        pk_unpack(sl[0], kx0, kx1, kx2, spare);
                        bool stop = false;
                        int L = 1;
                        for (; L < m; ++L) {
                            uint64_t a0, a1, a2;
                            pk_unpack(sl[L * 256], a0, a1, a2, spare);
                            if (a2 != kx2) {
                                lput(n++, pk_pack(kx0, kx1, kx2, (unsigned)(x + L)));
                                if (a2 < min_intv) { stop = true; break; }
                            }
                            kx0 = a0; kx1 = a1; kx2 = a2;
                        }
                        st_calls += (unsigned)(stop ? L : L - 1); st_blocks += 2u * (unsigned)(stop ? L : L - 1);
                        if (stop) {
                            kend = (unsigned)(x + L); ret = (int)kend;
                            hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
                            phase = SP_BWD;
                        } else { kend = (unsigned)(x + m); i = x + m; }
                    }
                }
            }
            if (phase == SP_FWD) {
                int b = i < len ? (int)__ldg(q + i) : 4;
                if (b < 4) { need = true; is_back = 0; c = 3 - b; sx0 = kx0; sx1 = kx1; sx2 = kx2; }
                else {  // ambiguous base or end of read: close the forward pass (bwt.c:311-316)
                    lput(n++, pk_pack(kx0, kx1, kx2, kend));
                    ret = (int)kend;
                    hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
                    phase = SP_BWD;
                }
            }
            if (phase == SP_BWD && !need) {
                int b = i < 0 ? 4 : (int)__ldg(q + i);
                pk_unpack(lget(j), px0, px1, px2, pend);
                if (b > 3) {
                    // no extension possible: only the longest survivor can be a new SMEM (bwt.c:331-341), then the call ends
                    if (ncur == 0 && j == hi - 1 && i + 1 < last_start) emit(px0, px1, px2, i + 1, (int)pend, true);
                    if (pass == 1) { x = ret; phase = SP_P1_SCAN; } else phase = SP_P2_NEXT;
                    break;
                }
                need = true; is_back = 1; c = b; sx0 = px0; sx1 = px1; sx2 = px2;
            }
            if (phase == SP_P3) {
                int b = i < len ? (int)__ldg(q + i) : 5;
                if (b < 4) { need = true; is_back = 0; c = 3 - b; sx0 = kx0; sx1 = kx1; sx2 = kx2; }
                else if (b == 4) { x = i + 1; phase = SP_P3_SCAN; break; }
                else { phase = SP_IDLE; break; }
            }
        } while (0);
        __syncwarp();
        // ------------------------------------------------------------------ EXTEND (convergent): bwt_extend on (sx, is_back), base c
        uint64_t ox_a = 0, ox_b = 0, ox2 = 0;  // ok[c].x[!is_back], ok[c].x[is_back], ok[c].x[2]
        if (need) {
            const uint64_t xa = is_back ? sx0 : sx1, xb = is_back ? sx1 : sx0;
            const uint64_t k = xa - 1, l = xa - 1 + sx2;
            ++st_calls; st_blocks += fm_ref_blocks(fm, k, l);
            uint64_t tk[4], tl[4];
            if (k == (uint64_t)-1) {
                tk[0] = tk[1] = tk[2] = tk[3] = 0;
                fm_occ4(fm, l, tl);
            } else {
                const uint64_t k_ = k - (k >= fm.primary), l_ = l - (l >= fm.primary);
                const uint64_t kb = k_ >> 6, lb = l_ >> 6;
                Blk32 bk = fm_load_block(fm, kb);
                Blk32 bl = bk;
                if (lb != kb) bl = fm_load_block(fm, lb);
                fm_occ4_blk(bk, k_, tk);
                fm_occ4_blk(bl, l_, tl);
            }
            uint64_t ns3 = tl[3] - tk[3], ns2 = tl[2] - tk[2], ns1 = tl[1] - tk[1], ns0 = tl[0] - tk[0];
            uint64_t nb = xb + ((xa <= fm.primary && xa + sx2 - 1 >= fm.primary) ? 1 : 0);
            // nb[3] = that; nb[2] = nb[3]+ns3; nb[1] = nb[2]+ns2; nb[0] = nb[1]+ns1
            if (c < 3) nb += ns3;
            if (c < 2) nb += ns2;
            if (c < 1) nb += ns1;
            ox_b = nb;
            ox_a = fm_L2(fm, c) + 1 + (c == 0 ? tk[0] : c == 1 ? tk[1] : c == 2 ? tk[2] : tk[3]);
            ox2 = c == 0 ? ns0 : c == 1 ? ns1 : c == 2 ? ns2 : ns3;
        }
        // ------------------------------------------------------------------ POST
        if (!need) {
        } else if (phase == SP_FWD) {
            // forward extension: ok.x1 = looked-up coordinate, ok.x0 = the other
            bool stop = false;
            if (ox2 != kx2) {
                lput(n++, pk_pack(kx0, kx1, kx2, kend));
                if (ox2 < min_intv) stop = true;
            }
            if (stop) {
                ret = (int)kend;
                hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
                phase = SP_BWD;
            } else {
                kx0 = ox_b; kx1 = ox_a; kx2 = ox2; kend = (unsigned)(i + 1);
                ++i;
            }
        } else if (phase == SP_BWD) {
            // backward extension of prev[j]: ok.x0 = looked-up coordinate, ok.x1 = the other
            if (ox2 < min_intv) {
                if (ncur == 0 && i + 1 < last_start) { emit(px0, px1, px2, i + 1, (int)pend, true); last_start = i + 1; }
            } else if (ncur == 0 || ox2 != lastcur_x2) {
                lput(hi - 1 - ncur, pk_pack(ox_a, ox_b, ox2, pend));
                ++ncur; lastcur_x2 = ox2;
            }
            if (--j < lo) {
                if (ncur == 0) { if (pass == 1) { x = ret; phase = SP_P1_SCAN; } else phase = SP_P2_NEXT; }
                else { lo = hi - ncur; ncur = 0; j = hi - 1; --i; }
            }
        } else {  // SP_P3: bwt_seed_strategy1 (bwt.c:358-379)
            if (ox2 < opt.max_mem_intv && i - x >= opt.min_seed_len) {
                if (ox2 > 0) emit(ox_b, ox_a, ox2, x, i + 1, false);
                x = i + 1; phase = SP_P3_SCAN;
            } else { kx0 = ox_b; kx1 = ox_a; kx2 = ox2; ++i; }
        }
        if (__all_sync(0xffffffffu, phase == SP_DONE)) break;
    }
    if (st_calls) { atomicAdd(&cnt->occ, (unsigned long long)st_calls); atomicAdd(&cnt->occ_blocks, (unsigned long long)st_blocks); }
}

// Sort each read's intervals by info (bwamem.c:161; equal info => identical interval, so any sort gives the same
// array) and expand them into SA rows (bwamem.c:278-283).  Thread per read.
__global__ void k_sort_intv(int n_reads, Intv* intv, const int* __restrict__ intv_n, int cap) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += gridDim.x * blockDim.x) {
        Intv* a = intv + (size_t)r * cap;
        int n = intv_n[r];
        n = n < cap ? n : cap;
        for (int i = 1; i < n; ++i) {
            Intv v = a[i];
            int k = i - 1;
            while (k >= 0 && a[k].info > v.info) { a[k + 1] = a[k]; --k; }
            a[k + 1] = v;
        }
    }
}

}  // namespace arachne
