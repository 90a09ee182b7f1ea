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
//   * a trip is PRE (per-lane bookkeeping, shared memory only) -> ISSUE (every lane's loads of the trip, of whatever kind:
//     occ sectors, a k-mer table entry, an SA / inverse-SA entry, text words) -> warp barrier -> CONSUME -> POST.  The kernel
//     is bound by the dependent chain inside a trip (profiles/r1h_summary.md), so nothing global is read in PRE: the lane's
//     read is kept 2-bit packed with its N / end mask in shared memory (k_pack_reads packs the batch once).
//   * shortcuts that return the integers the stepping would (tests/test_gpu_parity.py::test_seed_intervals_every_shortcut):
//     k-mer tables (FMIndex::kmer_all) for the LAST-like pass, the first levels of every forward pass and every backward
//     extension whose result is at most K long; direct text comparison once a forward match is unique (FMIndex::isa / text).
#pragma once
#include "fmindex.cuh"

namespace arachne {

struct SeedOut {
    Intv* intv;            // [n_reads * cap]: the SMEM passes fill a read's slots from 0, the LAST-like pass from cap / 2
    int* cnt2;             // [n_reads * 2], zeroed before launch: intervals written by each of the two (one lane per read and kind)
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

#if defined(__CUDACC__)
__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
    const unsigned a = (unsigned)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(a), "l"(gmem));
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.commit_group;\n\tcp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void ld_sector(uint32_t* o, const void* p) {  // one 32-byte sector, LDG.E.256
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3]), "=r"(o[4]), "=r"(o[5]), "=r"(o[6]), "=r"(o[7]) : "l"(p));
}
__device__ __forceinline__ void ld_nc8(uint32_t* o, const void* p) {   // 8 bytes past L1 (an SA / inverse-SA entry, a text word)
    asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(o[0]), "=r"(o[1]) : "l"(p));
}
__device__ __forceinline__ void ld_nc16(uint32_t* o, uint64_t p) {     // 16 bytes past L1 (a k-mer table entry)
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(o[0]), "=r"(o[1]), "=r"(o[2]), "=r"(o[3]) : "l"(p));
}
#else  // TEST-ONLY host build (tests/hostsim/hostsim_seed.cpp runs this kernel one lane at a time): the same bytes by plain copies
inline void cp_async16(void* smem, const void* gmem) { memcpy(smem, gmem, 16); }
inline void cp_async_wait_all() {}
inline void ld_sector(uint32_t* o, const void* p) { memcpy(o, p, 32); }
inline void ld_nc8(uint32_t* o, const void* p) { memcpy(o, p, 8); }
inline void ld_nc16(uint32_t* o, uint64_t p) { memcpy(o, reinterpret_cast<const void*>(p), 16); }
#endif
// bits [s, s + 64) of the 128-bit string a.b (s even, 0..62)
__device__ __forceinline__ uint64_t fsh64(uint64_t a, uint64_t b, int s) { return (a << s) | ((b >> 1) >> (63 - s)); }

// first entry of table level L in FMIndex::kmer_all: (4^L - 4) / 3 = 0b0101..01 (L - 1 pairs) << 2, without the 64-bit division
__device__ __forceinline__ uint32_t klev(int L) { return (0x55555555u & ((1u << (2 * L - 2)) - 1u)) << 2; }

enum { SP_IDLE = 0, SP_P1_SCAN, SP_FWD, SP_BWD, SP_P2_NEXT, SP_P3_SCAN, SP_P3, SP_DONE, SP_FSA, SP_FTXT, SP_FISA, SP_FTAB };

// Reads as 2-bit words for the direct extension (below): per read `stride` pairs (bases, mask), 32 bases per u64 with base p
// at bits 62 - 2*(p & 31); the mask (one bit per base, base p at bit 31 - (p & 31)) is set at every ambiguous base and at every
// position from the read's end on, so a comparison stops there by itself.  One thread per (read, word).
constexpr int DIRECT_WORDS = 3;    // 64-bit words compared per trip (96 bases)
constexpr int DIRECT_MIN_LEFT = 8; // bases left in the read below which stepping is kept (the switch costs three trips)
__host__ __device__ inline int seed_qstride(int maxlen) { return (maxlen + 31) / 32 + DIRECT_WORDS + 1; }
// Dynamic shared memory of k_seed2 at `bpsm` resident blocks per SM: the read cache (nwq words + masks per lane) and as many
// interval-list entries per lane as fit beside it (at most 10, or `want` when >= 0).
inline size_t seed_smem_plan(int maxlen, int bpsm, int want, int* slist_n, int* nwq) {
    const int budget = (228 * 1024) / bpsm - 1024 - 3072;  // per-block reserve and the static counters
    *nwq = (maxlen > 0 ? (maxlen - 1) / 32 : 0) + 2;
    int sn = (budget - *nwq * 256 * 12) / (256 * 16);
    sn = sn > 16 ? 16 : sn < 0 ? 0 : sn;
    const int dflt = 10;  // more entries measured slower: the shared memory they take is L1 the list tails lose
    sn = want >= 0 ? (want < sn ? want : sn) : (dflt < sn ? dflt : sn);
    *slist_n = sn;
    return (size_t)sn * 256 * 16 + (size_t)*nwq * 256 * 12;
}
__global__ void __launch_bounds__(256) k_pack_reads(const uint8_t* __restrict__ seq, const int64_t* __restrict__ off, int n_reads, int stride, ulonglong2* out) {
    const size_t total = (size_t)n_reads * stride;
    for (size_t t = blockIdx.x * (size_t)blockDim.x + threadIdx.x; t < total; t += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(t / stride), w = (int)(t % stride);
        const int64_t o = off[r];
        const int len = (int)(off[r + 1] - o);
        uint64_t v = 0, m = 0;
        for (int b = 0; b < 32; ++b) {
            const int p = w * 32 + b;
            const unsigned c = p < len ? seq[o + p] : 4u;
            if (c > 3) m |= 1ull << (31 - b); else v |= (uint64_t)c << (62 - 2 * b);
        }
        out[t] = make_ulonglong2(v, m);
    }
}

template <int MIN_BLOCKS>
__global__ void __launch_bounds__(256, MIN_BLOCKS) k_seed2(FMIndex fm, MemOpt opt, const int64_t* __restrict__ off,
                                                  int n_reads, int maxlen, uint4* lists, int rows, SeedOut out, unsigned int* next, Counters* cnt,
                                                  int use_kmer, int slist_n, const ulonglong2* __restrict__ qpk, int qstride, int nwq, int use_direct) {
    const unsigned T = gridDim.x * blockDim.x, tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint4* list = lists + tid;                       // entry j (j >= slist_n) at list[j * T]
    // the first slist_n entries of every lane's list live in shared memory ([entry][lane], 16 bytes each: the bank depends
    // on the lane only, so lanes at different entries never conflict): with ~17 entries per forward pass the lists of the
    // SM's lanes (0.3 MB) do not fit L1, and every backward step would otherwise start with an L2 round trip
    extern __shared__ uint4 slist[];
    uint4* sl = slist + threadIdx.x;
    auto lput = [&](int e, uint4 v) { if (e < slist_n) sl[e * 256] = v; else list[(size_t)e * T] = v; };
    auto lget = [&](int e) -> uint4 { return e < slist_n ? sl[e * 256] : list[(size_t)e * T]; };
    // the lane's read, 2-bit packed with its mask (k_pack_reads), also in shared memory ([word][lane]): every trip starts by
    // looking at one base of the read, and the reads of the SM's lanes do not stay in L1 either
    uint64_t* qv = reinterpret_cast<uint64_t*>(slist + slist_n * 256) + threadIdx.x;
    uint32_t* qm = reinterpret_cast<uint32_t*>(reinterpret_cast<uint64_t*>(slist + slist_n * 256) + nwq * 256) + threadIdx.x;
    auto qb = [&](int p) -> int {  // base p of the read; 4 = ambiguous or at / past the end
        const int w = p >> 5, sft = p & 31;
        return ((qm[w * 256] >> (31 - sft)) & 1u) ? 4 : (int)((qv[w * 256] >> (62 - 2 * sft)) & 3);
    };
    auto qwin = [&](int p) -> uint64_t { const int w = p >> 5; return fsh64(qv[w * 256], qv[(w + 1) * 256], (p & 31) * 2); };  // 32 bases from p
    auto mwin = [&](int p) -> uint32_t { const int w = p >> 5; return __funnelshift_l(qm[(w + 1) * 256], qm[w * 256], p & 31); };  // their mask
    uint4* recs = lists + (size_t)(rows >> 1) * T + tid;  // re-seed records in the upper half
    const int split_len = (int)(opt.min_seed_len * opt.split_factor + .499f);
    // task kinds cost about 2:1 in bwt_extend calls (passes 1+2 vs pass 3): every third warp starts on the LAST-like queue
    // (1:5 when pass 3 starts from the k-mer table)
    const int kmer_k = ((use_kmer & 1) && fm.kmer && fm.kmer_k <= opt.min_seed_len) ? fm.kmer_k : 0;
    // the SMEM passes' forward extension starts from the table too (all levels 1..K, fetched with K independent copies into
    // the lane's shared list slots) when those slots can hold them
    // levels the forward window replays: at most the shared slots, and use_kmer bits 8..11 when set (the deep levels are HBM
    // round trips the warp waits for; the shallow ones stay in L2)
    const int kf_cap = (use_kmer >> 8) & 15;
    const int kf = min(min(kmer_k, slist_n), kf_cap ? kf_cap : 8);
    const bool kmer_fwd = kmer_k && fm.kmer_all && kf >= 4;
    // a backward extension whose result is at most kmer_k long is a table entry too: bwt_smem1a's backward pass extends every
    // interval of the forward list base by base (bwt.c:320-347), and most of those patterns are short
    const bool kmer_bwd = kmer_k && fm.kmer_all;
    bool fresh = false;  // a forward pass was just set up at x
    // direct extension: a forward pass whose interval has shrunk to ONE row can only go on matching the text at SA[row], so the
    // rest of the pass is SP_FSA (pos = SA[x0]) -> SP_FTXT (compare read and text, 96 bases per trip) -> SP_FISA (x1 =
    // ISA[n - pos - len]; x0 does not move) instead of one bwt_extend per base; the interval pushed at the end is the one the
    // stepping would have pushed (bwt.c:300-316).  Each of the three is one trip whose load is issued with the other lanes'.
    const bool direct = use_direct && fm.isa != nullptr && fm.text != nullptr && fm.sa_intv == 1;
    const int pref_kind = (opt.max_mem_intv > 0 && (tid >> 5) % (kmer_k ? 6 : 3) == 2) ? 1 : 0;

    int phase = SP_IDLE, r = 0, len = 0, x = 0, i = 0;  // passes 1 / 2 of a read: rec_i == 0 / > 0; during a forward pass ik.info's end is i
    uint64_t kx0 = 0, kx1 = 0, kx2 = 0;  // ik
    uint32_t min_intv = 1;
    int n = 0, lo = 0, hi = 0, j = 0, ncur = 0, last_start = 0x7fffffff, ret = 0, n_rec = 0, rec_i = 0;
    uint64_t lastcur_x2 = 0;
    uint4 praw = make_uint4(0, 0, 0, 0);  // prev[j] of the backward pass, packed (query end in the top byte)
    uint64_t& dpos = lastcur_x2;  // direct extension: text position of the match (the backward pass's register, idle until then)
    // bwt_extend calls / reference occ blocks (made or replaced), per lane in shared memory: two registers less in the loop
    __shared__ uint2 s_stat[256];
    __shared__ int s_emit[256];  // intervals the lane's current work item has written
    s_stat[threadIdx.x] = make_uint2(0u, 0u);
    auto stat = [&](unsigned calls, unsigned blocks) { uint2 t = s_stat[threadIdx.x]; t.x += calls; t.y += blocks; s_stat[threadIdx.x] = t; };

    auto emit = [&](uint64_t e0, uint64_t e1, uint64_t e2, int start, int end, bool filter) {
        const int slen = end - start;
        if (filter && slen < opt.min_seed_len) return;
        // no atomics: a read's two work items write disjoint halves, counted by their own lanes (k_sort_intv joins the halves)
        const int kind = filter ? 0 : 1, slot = s_emit[threadIdx.x];
        if (slot >= (out.cap >> 1)) { raise(cnt, ARACHNE_ERR_OVERFLOW, r); return; }
        s_emit[threadIdx.x] = slot + 1;
        out.cnt2[2 * r + kind] = slot + 1;
        Intv v; v.x0 = e0; v.x1 = e1; v.x2 = e2; v.info = (uint64_t)start << 32 | (unsigned)end;
        out.intv[(size_t)r * out.cap + kind * (out.cap >> 1) + slot] = v;
        if (rec_i == 0 && filter && slen >= split_len && e2 <= (uint64_t)opt.split_width)
            recs[(size_t)(n_rec++) * T] = make_uint4((unsigned)start, (unsigned)end, (unsigned)e2, 0u);
    };

    // the table window of a forward pass: intervals of q[x, x + L), L = 1..m, are in the lane's first list slots
    auto replay = [&]() {
        int m = __clz((int)mwin(x));
        m = m < kf ? m : kf;
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
        stat((unsigned)(stop ? L : L - 1), 2u * (unsigned)(stop ? L : L - 1));
        if (stop) {
            ret = x + L;
            hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
            phase = SP_BWD;
        } else { i = x + m; phase = SP_FWD; }
    };

    for (;;) {
        // ------------------------------------------------------------------ PRE: bookkeeping up to the next extend
        int need = 0;  // 0 nothing, 1 bwt_extend, 2 one SA / ISA entry, 3 text comparison, 4 one k-mer table entry
        int c = 0;  // base of the extension; it is a backward one (is_back) exactly in SP_BWD
        uint64_t tabp = 0;  // need == 4: address of the table entry
        do {
            if (phase >= SP_DONE) { need = phase == SP_DONE ? 0 : phase == SP_FTXT ? 3 : 2; break; }  // (SP_FTAB never lasts into a trip)
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
                if (len > maxlen) { if (kind == 0) raise(cnt, ARACHNE_ERR_ARG, r); break; }
                if (len < opt.min_seed_len) break;
                for (int w = 0; w < nwq; ++w) {
                    const ulonglong2 pw = __ldg(qpk + (size_t)r * qstride + w);
                    qv[w * 256] = pw.x; qm[w * 256] = (uint32_t)pw.y;
                }
                x = 0; n_rec = 0; rec_i = 0; s_emit[threadIdx.x] = 0;
                phase = kind ? ((opt.max_mem_intv > 0) ? SP_P3_SCAN : SP_IDLE) : SP_P1_SCAN;
                if (phase == SP_IDLE) break;
            }
            if (phase == SP_P2_NEXT) {
                if (rec_i < n_rec) {
                    uint4 rc = recs[(size_t)(rec_i++) * T];
                    x = (int)((rc.x + rc.y) >> 1);
                    min_intv = rc.z + 1;
                    int b = qb(x);
                    if (b > 3) break;  // bwt_smem1a returns at once on an ambiguous base (cannot happen inside a match)
                    kx0 = fm_L2(fm, b) + 1; kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); kx1 = fm_L2(fm, 3 - b) + 1;
                    i = x + 1; n = 0; last_start = 0x7fffffff;
                    phase = SP_FWD; fresh = true;
                } else { phase = SP_IDLE; break; }
            }
            if (phase == SP_P1_SCAN) {
                while (x < len && qb(x) > 3) ++x;
                if (x >= len) { phase = SP_P2_NEXT; break; }
                int b = qb(x);
                kx0 = fm_L2(fm, b) + 1; kx2 = fm_L2(fm, b + 1) - fm_L2(fm, b); kx1 = fm_L2(fm, 3 - b) + 1;
                min_intv = 1;
                i = x + 1; n = 0; last_start = 0x7fffffff;
                phase = SP_FWD; fresh = true;
            }
            if (phase == SP_P3_SCAN) {
                // bwt_seed_strategy1 (bwt.c:358-379) from x: nothing is tested before the match is min_seed_len + 1 long, so
                // (a) a start with fewer bases left cannot seed, (b) the first kmer_k bases need no stepping: their bi-interval
                // is one table entry (an ambiguous base among them ends the call without a seed, as in the reference).
                for (;;) {
                    while (x < len && qb(x) > 3) ++x;
                    if (len - x <= opt.min_seed_len) { phase = SP_IDLE; break; }
                    int b = qb(x);
                    if (kmer_k) {
                        const int zn = __clz((int)mwin(x));  // first ambiguous base of the window (the read is long enough)
                        if (zn < kmer_k) { x += zn + 1; continue; }
                        const uint32_t km = (uint32_t)(qwin(x) >> (64 - 2 * kmer_k));
                        unsigned spare;
                        pk_unpack(__ldg(reinterpret_cast<const uint4*>(fm.kmer) + km), kx0, kx1, kx2, spare);
                        i = x + kmer_k;
                        stat(kmer_k - 1, 2 * (kmer_k - 1));  // the skipped calls; intervals this large span two blocks
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
                    int m = __clz((int)mwin(x));  // unambiguous bases from x on (the mask also covers the read's end)
                    m = m < kf ? m : kf;
                    if (m >= 2) {
                        const uint32_t km = (uint32_t)(qwin(x) >> (64 - 2 * m));
                        const uint4* all = reinterpret_cast<const uint4*>(fm.kmer_all);
                        for (int L = 1; L <= m; ++L) cp_async16(&sl[(L - 1) * 256], all + klev(L) + (km >> (2 * (m - L))));
                        // the copies land while the other lanes' loads of this trip are in flight; replayed in POST
                        // waiting here costs the warp one table round trip; handing the window to the next trip (SP_FTAB, use_kmer
                        // bit 2) costs every forward pass a trip of its own and measured slower (57-60 vs 55 ms per 1 M pairs)
                        if (use_kmer & 4) { phase = SP_FTAB; need = 8; break; }
                        cp_async_wait_all(); replay();
                    }
                }
            }
            if (phase == SP_FWD) {
                int b = qb(i);
                if (b < 4) { need = 1; c = 3 - b; }
                else {  // ambiguous base or end of read: close the forward pass (bwt.c:311-316)
                    lput(n++, pk_pack(kx0, kx1, kx2, (unsigned)i));
                    ret = i;
                    hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
                    phase = SP_BWD;
                }
            }
            if (phase == SP_BWD && !need) {
                int b = i < 0 ? 4 : qb(i);
                praw = lget(j);
                if (b > 3) {
                    // no extension possible: only the longest survivor can be a new SMEM (bwt.c:331-341), then the call ends
                    if (ncur == 0 && j == hi - 1 && i + 1 < last_start) {
                        uint64_t p0, p1, p2; unsigned pe;
                        pk_unpack(praw, p0, p1, p2, pe);
                        emit(p0, p1, p2, i + 1, (int)(praw.w >> 24), true);
                    }
                    if (rec_i == 0) { x = ret; phase = SP_P1_SCAN; } else phase = SP_P2_NEXT;
                    break;
                }
                const int L = (int)(praw.w >> 24) - i;  // length of the extended match q[i, pend)
                if (kmer_bwd && L <= kmer_k) {
                    const uint64_t code = qwin(i) >> (64 - 2 * L);
                    tabp = (uint64_t)(reinterpret_cast<const uint4*>(fm.kmer_all) + klev(L) + code);
                    need = 4;
                } else { need = 1; c = b; }
            }
            if (phase == SP_P3) {
                int b = i < len ? qb(i) : 5;
                if (b < 4) { need = 1; c = 3 - b; }
                else if (b == 4) { x = i + 1; phase = SP_P3_SCAN; break; }
                else { phase = SP_IDLE; break; }
            }
        } while (0);
        __syncwarp();
        // ------------------------------------------------------------------ ISSUE: every lane's loads for this trip, nothing waits
        // (the barrier below keeps the compiler from pairing each kind's loads with its own consumer, which would serialise
        // the kinds' memory latencies in a mixed warp)
        uint32_t R[16];
        uint64_t xb = 0;          // x[is_back] of the interval being extended
        int kbit = 0, lbit = 0;   // k, l within their blocks
        bool kneg = false, same = true, hasp = false;
        if (need == 1) {
            uint64_t xa, x2;      // x[!is_back] is the looked-up coordinate
            if (phase == SP_BWD) { unsigned pe; pk_unpack(praw, xa, xb, x2, pe); } else { xa = kx1; xb = kx0; x2 = kx2; }
            const uint64_t k = xa - 1, l = xa - 1 + x2;
            stat(1u, fm_ref_blocks(fm, k, l));
            kneg = k == (uint64_t)-1;
            hasp = xa <= fm.primary && xa + x2 - 1 >= fm.primary;
            const uint64_t l_ = l - (l >= fm.primary), k_ = kneg ? l_ : k - (k >= fm.primary);
            const uint64_t kb = k_ >> 6, lb = l_ >> 6;
            kbit = (int)(k_ & 63); lbit = (int)(l_ & 63);
            same = kb == lb;
            ld_sector(R, fm.blocks + kb);
            if (!same) ld_sector(R + 8, fm.blocks + lb);
        } else if (need == 2) {
            const uint64_t* a = phase == SP_FSA ? fm.sa + kx0 : fm.isa + (fm.seq_len - dpos - (uint64_t)(i - x));
            ld_nc8(R, a);
        } else if (need == 3) {
            const uint64_t* tp = fm.text + ((dpos + (uint64_t)(i - x)) >> 5);
#pragma unroll
            for (int d = 0; d <= DIRECT_WORDS; ++d)
                ld_nc8(R + 2 * d, tp + d);
        } else if (need == 4) {
            stat(1u, 2u);
            ld_nc16(R, tabp);
        }
        __syncwarp();
        // ------------------------------------------------------------------ CONSUME: bwt_extend on (sx, is_back), base c / the loaded entry
        uint64_t ox_a = 0, ox_b = 0, ox2 = 0;  // ok[c].x[!is_back], ok[c].x[is_back], ok[c].x[2]
        uint64_t got = 0;                      // need == 2: the entry
        int dt = 32 * DIRECT_WORDS;            // need == 3: bases matched, why it stopped
        bool dstop = false, dnstop = false;
        if (need == 3) {
            if (len - i < dt) { dt = len - i; dnstop = true; }  // the read ends inside the window (a window starting past it is not read)
            // q[i, ..) against the text behind the match: first difference, ambiguous base or read end (mask) within 96 bases
            // (done before the extension below so that the words die early)
            const uint64_t ts = dpos + (uint64_t)(i - x), left = fm.seq_len - ts;
            const int tsh = (int)(ts & 31) * 2;
#pragma unroll
            for (int d = DIRECT_WORDS - 1; d >= 0; --d) {
                if (i + 32 * d >= len) continue;
                const uint64_t td = fsh64((uint64_t)R[2 * d + 1] << 32 | R[2 * d], (uint64_t)R[2 * d + 3] << 32 | R[2 * d + 2], tsh);
                const int zx = __clzll((long long)(td ^ qwin(i + 32 * d))) >> 1, zn = __clz((int)mwin(i + 32 * d));
                if (zx < 32 || zn < 32) { dt = 32 * d + (zx < zn ? zx : zn); dnstop = zn <= zx; }
            }
            dstop = dt < 32 * DIRECT_WORDS;
            if (left < (uint64_t)dt || (left == (uint64_t)dt && !dstop)) { dt = (int)left; dnstop = false; dstop = true; }  // the text ends first
        }
        if (need == 1) {
            uint64_t tk[4], tl[4];
            Blk32 bk, bl;
#pragma unroll
            for (int u = 0; u < 4; ++u) { bk.cnt[u] = R[u]; bk.w[u] = R[4 + u]; bl.cnt[u] = same ? R[u] : R[8 + u]; bl.w[u] = same ? R[4 + u] : R[12 + u]; }
            fm_occ4_blk(bk, (uint64_t)kbit, tk);
            fm_occ4_blk(bl, (uint64_t)lbit, tl);
            if (kneg) tk[0] = tk[1] = tk[2] = tk[3] = 0;
            uint64_t ns3 = tl[3] - tk[3], ns2 = tl[2] - tk[2], ns1 = tl[1] - tk[1], ns0 = tl[0] - tk[0];
            uint64_t nb = xb + (hasp ? 1 : 0);
            // nb[3] = that; nb[2] = nb[3]+ns3; nb[1] = nb[2]+ns2; nb[0] = nb[1]+ns1
            if (c < 3) nb += ns3;
            if (c < 2) nb += ns2;
            if (c < 1) nb += ns1;
            ox_b = nb;
            ox_a = fm_L2(fm, c) + 1 + (c == 0 ? tk[0] : c == 1 ? tk[1] : c == 2 ? tk[2] : tk[3]);
            ox2 = c == 0 ? ns0 : c == 1 ? ns1 : c == 2 ? ns2 : ns3;
        } else if (need == 2) {
            got = (uint64_t)R[1] << 32 | R[0];
        } else if (need == 8) {
            cp_async_wait_all();
        } else if (need == 4) {
            ox2 = ((uint64_t)R[3] << 32 | R[2]) >> 16 & 0xffffffffffull;
        }
        // ------------------------------------------------------------------ POST
        if (!need) {
        } else if (phase == SP_FWD) {
            // forward extension: ok.x1 = looked-up coordinate, ok.x0 = the other
            bool stop = false;
            if (ox2 != kx2) {
                lput(n++, pk_pack(kx0, kx1, kx2, (unsigned)i));
                if (ox2 < min_intv) stop = true;
            }
            if (stop) {
                ret = i;
                hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
                phase = SP_BWD;
            } else {
                kx0 = ox_b; kx1 = ox_a; kx2 = ox2;
                ++i;
                if (direct && kx2 == 1 && min_intv <= 1 && len - i >= DIRECT_MIN_LEFT) phase = SP_FSA;
            }
        } else if (phase == SP_BWD) {
            // backward extension of prev[j]: ok.x0 = looked-up coordinate, ok.x1 = the other
            if (ox2 < min_intv) {
                if (ncur == 0 && i + 1 < last_start) {
                    uint64_t p0, p1, p2; unsigned pe;
                    pk_unpack(praw, p0, p1, p2, pe);
                    emit(p0, p1, p2, i + 1, (int)(praw.w >> 24), true); last_start = i + 1;
                }
            } else if (ncur == 0 || ox2 != lastcur_x2) {
                // a table entry is already packed like a list entry: only the query end is added
                lput(hi - 1 - ncur, need == 4 ? make_uint4(R[0], R[1], R[2], (R[3] & 0xffffffu) | (praw.w & 0xff000000u)) : pk_pack(ox_a, ox_b, ox2, praw.w >> 24));
                ++ncur; lastcur_x2 = ox2;
            }
            if (--j < lo) {
                if (ncur == 0) { if (rec_i == 0) { x = ret; phase = SP_P1_SCAN; } else phase = SP_P2_NEXT; }
                else { lo = hi - ncur; ncur = 0; j = hi - 1; --i; }
            }
        } else if (phase == SP_P3) {  // bwt_seed_strategy1 (bwt.c:358-379)
            if (ox2 < opt.max_mem_intv && i - x >= opt.min_seed_len) {
                if (ox2 > 0) emit(ox_b, ox_a, ox2, x, i + 1, false);
                x = i + 1; phase = SP_P3_SCAN;
            } else { kx0 = ox_b; kx1 = ox_a; kx2 = ox2; ++i; }
        } else if (phase == SP_FTAB) {
            replay();
        } else if (phase == SP_FSA) {
            dpos = got;  // text position of the unique match q[x, i)
            phase = SP_FTXT;
        } else if (phase == SP_FTXT) {
            // the stepping would have made dt successful calls, and a failing one unless the read (or its unambiguous run) ended
            const unsigned calls = (unsigned)dt + ((dstop && !dnstop) ? 1u : 0u);
            stat(calls, calls);
            i += dt;
            if (dstop) phase = SP_FISA;
        } else if (phase == SP_FISA) {
            kx1 = got;
            lput(n++, pk_pack(kx0, kx1, kx2, (unsigned)i));
            ret = i;
            hi = n; lo = 0; j = hi - 1; ncur = 0; i = x - 1;
            phase = SP_BWD;
        }
        if (__all_sync(0xffffffffu, phase == SP_DONE)) break;
    }
    const uint2 st = s_stat[threadIdx.x];
    if (st.x) { atomicAdd(&cnt->occ, (unsigned long long)st.x); atomicAdd(&cnt->occ_blocks, (unsigned long long)st.y); }
}

// Join the two halves of each read's intervals, sort them by info (bwamem.c:161; equal info => identical interval, so any
// sort gives the same array), and count the SA rows they expand into (bwamem.c:278-283).  Thread per read.
__global__ void k_sort_intv(int n_reads, Intv* intv, const int* __restrict__ cnt2, int cap, int max_occ, int* intv_n, unsigned int* n_sa) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += gridDim.x * blockDim.x) {
        Intv* a = intv + (size_t)r * cap;
        const int h = cap >> 1;
        int n0 = cnt2[2 * r], n1 = cnt2[2 * r + 1];
        n0 = n0 < h ? n0 : h; n1 = n1 < h ? n1 : h;
        if (n0 < h) for (int t = 0; t < n1; ++t) a[n0 + t] = a[h + t];
        const int n = n0 + n1;
        unsigned rows = 0;
        for (int i = 0; i < n; ++i) {
            Intv v = a[i];
            rows += v.x2 > (uint64_t)max_occ ? (unsigned)max_occ : (unsigned)v.x2;
            int k = i - 1;
            while (k >= 0 && a[k].info > v.info) { a[k + 1] = a[k]; --k; }
            a[k + 1] = v;
        }
        intv_n[r] = n;
        n_sa[r] = rows;
    }
}

}  // namespace arachne
