// Extension stage: mem_chain2aln + mem_sort_dedup_patch (bwa/bwamem.c:632-786, 437-489) for a whole batch, in ROUNDS.
//
//   k_ext_drive    thread per read: ext_advance() (ext_driver.cuh) consumes the result of the read's pending DP problem and
//                  runs the driver until the next problem is known or the read is finished (then: region de-duplication).
//   k_ext_binbase  one block: counting sort of the round's problems by query length (longest first).
//   k_ext_scatter  thread per read of the round: its slot in the sorted queue.
//   k_ext_dp       ksw_extend2 (bwa/ksw.c:380-479), one problem per LANE, lanes persistent over the sorted queue.
//
// Why this shape: ksw_extend2 trims its band after every row from the complete previous row, so the rows of one problem are
// strictly sequential and a row is only ~40-130 cells; the parallelism worth having is ACROSS problems (3.5 M per 1 M pairs).
// The round-1 kernel kept the whole mem_chain2aln automaton inside the DP loop: 48 thread-instructions per cell (3 of them
// DPX), 18 of 32 lanes active, 168 registers.  Here the DP kernel does nothing but rows:
//   * row-synchronous lanes: one trip of the warp's loop = one row of every lane's problem; the queue is sorted by query
//     length, so the lanes of a warp run rows of nearly the same width; row start / row end execute convergently;
//   * a column is one 32-bit shared-memory word [column][lane] (bank = lane): E << 24 | H << 16 | selector.  The low half is
//     a ready-made PRMT selector (q * 0x1111 + 0x8880: "byte q, sign-replicated") into the row's 5-entry substitution-score
//     byte table, so a cell's score is ONE instruction; H and E are bytes (scores never exceed a * read length <= 255,
//     checked by the callers) extracted and re-packed by PRMT as well.  The integer ALU pipe is what bounds this kernel
//     (16 lanes per scheduler per clock, ncu: 70 % busy in the first version), so the cell is written for few ALU-pipe
//     instructions (~9: 3 PRMT, 1 shift, 4 DPX, the row-maximum key) and the rest as IMADs, which issue to the FMA pipe;
//   * max-plus steps on DPX (VIADDMNMX / VIMNMX3): M = max(min(H + s, H << 8), 0) reproduces `M ? M + s : 0` (a negative M and
//     a zero M are indistinguishable downstream: every consumer clamps at zero), E and F are one add + one VIADDMNMX.RELU;
//   * the band for the next row (ksw.c:465-468) is found by re-reading the few zero words at the row's two edges;
//   * columns beyond `end` keep what earlier rows left there, exactly as the reference's eh[] does.
#pragma once
#include "ext_driver.cuh"

namespace arachne {

// Scalar context for the (rare) region patching path of mem_sort_dedup_patch: one lane does everything.
struct ThreadCtx {
    int* hbuf;      // 2*(maxlen+2) ints of private global scratch
    uint8_t* rbuf;  // private window buffer
    __device__ bool leader() const { return true; }
    __device__ void sync() const {}
    __device__ int bcast(int v) const { return v; }
    __device__ void fetch_ref(const RefView& ref, int64_t beg, int64_t end, uint8_t* dst) const {
        for (int64_t k = 0; k < end - beg; ++k) dst[k] = ref_base(ref, beg + k);
    }
    // ksw_global2 (bwa/ksw.c:504-606), score only
    __device__ __noinline__ int ksw_global(const MemOpt& o, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts, int w,
                                           uint32_t*, int, int*) const {
        const int NEG = -0x40000000;
        int* H = hbuf;
        int* E = hbuf + qlen + 2;
        const int oe_del = o.o_del + o.e_del, oe_ins = o.o_ins + o.e_ins;
        H[0] = 0; E[0] = NEG;
        for (int j = 1; j <= qlen; ++j) { H[j] = j <= w ? -(o.o_ins + o.e_ins * j) : NEG; E[j] = NEG; }
        for (int i = 0; i < tlen; ++i) {
            int beg = i > w ? i - w : 0, end = i + w + 1 < qlen ? i + w + 1 : qlen, f = NEG;
            int left = beg == 0 ? -(o.o_del + o.e_del * (i + 1)) : NEG;
            const int tb = t[(long)i * ts];
            for (int j = beg; j < end; ++j) {
                int qb = q[(long)j * qs];
                int m = H[j] + ((tb > 3 || qb > 3) ? -1 : (tb == qb ? o.a : -o.b)), e = E[j];
                H[j] = left;
                int h = m > e ? m : e;
                h = h > f ? h : f;
                left = h;
                int x = m - oe_del; e -= o.e_del; E[j] = e > x ? e : x;
                x = m - oe_ins; f -= o.e_ins; f = f > x ? f : x;
            }
            H[end] = left; E[end] = NEG;
        }
        return H[qlen];
    }
};

struct ExtCtl {  // device-side control words of the round loop
    unsigned n_cur;        // reads the next k_ext_drive visits (through perm_cur); round 0 visits every read
    unsigned n_next;       // problems emitted by the round
    unsigned head;         // k_ext_dp's queue head
    unsigned pad;
    unsigned bin[256];     // problems per query length
    unsigned base[256];    // first slot of a query length in the sorted queue (longest first)
};

struct ExtBufs {
    const uint8_t* seq; const int64_t* off; int n_reads;
    const unsigned long long* sa_off;
    ChainBufs b;
    uint64_t* srt; AlnReg* regs; int* n_regs;
    int* hscr; uint8_t* rscr; int rcap; int maxlen;
    ExtState* state; ExtProb* prob; ExtRes* res; unsigned* rank;
    ExtProb* queue_cur; ExtProb* queue_next;  // the sorted problem queue of the previous round (= the reads this round visits) / of this round
    ExtCtl* ctl; Counters* cnt;
};

__global__ void __launch_bounds__(128) k_ext_drive(MemOpt opt, RefView ref, ExtBufs A, int round) {
    const unsigned n = round == 0 ? (unsigned)A.n_reads : A.ctl->n_cur;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const int r = round == 0 ? (int)t : A.queue_cur[t].read;
        const int nk = A.b.n_chain[r];
        ExtState S;
        if (round == 0) {
            if (nk == 0) { A.state[r].phase = EXT_PH_DONE; A.state[r].n_regions = 0; continue; }
            ext_state_init(S);
        } else S = A.state[r];
        const unsigned long long o = A.sa_off[r];
        const int64_t roff = A.off[r];
        const int len = (int)(A.off[r + 1] - roff);
        ExtProb P;
        const ExtRes R = A.res[r];
        int rc = ext_advance(opt, ref, r, len, nk, A.b.chains + o, A.b.order + o, A.b.kseeds + o, A.regs + o, A.b.n_kseed[r], A.srt + o, S, &R, &P);
        if (rc == 1 && (P.qlen > 255 || P.qlen <= 0)) rc = -3;  // cannot happen: a flank is shorter than the read
        if (rc == 1) {
            P.qoff = roff + P.q0;
            A.prob[r] = P;
            A.rank[r] = atomicAdd(&A.ctl->bin[P.qlen], 1u);
        } else {
            if (rc != 0) { raise(A.cnt, ARACHNE_ERR_OVERFLOW, r); S.n_regions = 0; }
            S.phase = EXT_PH_DONE;
        }
        A.state[r] = S;
    }
}

// mem_align1_core's tail (bwamem.c:1069-1082) once every read is through its rounds: region de-duplication / patching and the
// ALT flag.  Kept out of k_ext_drive: only this kernel needs per-thread scratch (the rare patch re-alignment).
__global__ void __launch_bounds__(128) k_ext_finish(MemOpt opt, RefView ref, ExtBufs A) {
    const long gtid = (long)blockIdx.x * blockDim.x + threadIdx.x;
    ThreadCtx tctx; tctx.hbuf = A.hscr + gtid * (2 * (A.maxlen + 2)); tctx.rbuf = A.rscr + gtid * A.rcap;
    for (int r = (int)gtid; r < A.n_reads; r += gridDim.x * blockDim.x) {
        int n_av = A.state[r].n_regions;
        if (n_av > 0) {
            AlnReg* av = A.regs + A.sa_off[r];
            if (n_av > 1) {
                const int nn = sort_dedup_patch(tctx, opt, ref, A.seq + A.off[r], n_av, av, tctx.rbuf, A.rcap);
                if (nn < 0) { raise(A.cnt, ARACHNE_ERR_OVERFLOW, r); n_av = 0; } else n_av = nn;
            }
            for (int x = 0; x < n_av; ++x)
                if (av[x].rid >= 0 && ref.contigs[av[x].rid].is_alt) av[x].n_comp_is_alt |= 1 << 30;
        }
        A.n_regs[r] = n_av;
    }
}

__global__ void __launch_bounds__(256) k_ext_binbase(ExtCtl* c) {
    __shared__ unsigned sh[256];
    const int t = threadIdx.x;
    const unsigned v = c->bin[255 - t];  // thread t owns query length 255 - t: longest first
    sh[t] = v;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
        const unsigned x = t >= d ? sh[t - d] : 0u;
        __syncthreads();
        sh[t] += x;
        __syncthreads();
    }
    c->base[255 - t] = sh[t] - v;
    c->bin[255 - t] = 0;
    if (t == 255) { c->n_next = sh[255]; c->head = 0; }
}

__global__ void k_ext_scatter(ExtBufs A, int round) {
    const unsigned n = round == 0 ? (unsigned)A.n_reads : A.ctl->n_cur;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const int r = round == 0 ? (int)t : A.queue_cur[t].read;
        if (A.state[r].phase == EXT_PH_DONE) continue;
        const uint4* src = reinterpret_cast<const uint4*>(A.prob + r);
        const uint4 a = src[0], b = src[1], c = src[2];
        uint4* dst = reinterpret_cast<uint4*>(A.queue_next + (A.ctl->base[a.z /* qlen */] + A.rank[r]));
        dst[0] = a; dst[1] = b; dst[2] = c;
    }
}

__global__ void k_ext_roll(ExtCtl* c) { c->n_cur = c->n_next; }

// ------------------------------------------------------------------------------------------------ k_ext_dp
constexpr int XDP_BLOCK = 128;
constexpr int XDP_REFILL = 8;  // idle lanes of a warp that trigger a refill

struct ExtDpArgs {
    const uint8_t* seq;                      // queries: problem p reads seq[p.qoff + x * (+1|-1)]
    const uint8_t* tbytes;                   // kPackedRef == false: targets as bytes, problem p reads tbytes[p.tpos + i]
    const ExtProb* queue;                    // the problems, longest query first
    ExtRes* res;                             // indexed by ExtProb::read
    const unsigned* n_items; unsigned* head;
    int ncols;                               // column words per lane, >= longest query + 2
    int zdrop;
    int one;                                 // 1 (see k_ext_dp)
    Counters* cnt;
};

// PRMT, default mode (see cig_prmt in cigar_kernel.cuh); the host definition serves the TEST-ONLY warp emulation (tests/hostsim/hostsim_extdp.cpp)
#if defined(__CUDACC__)
__device__ __forceinline__ int xdp_prmt(uint32_t lo, uint32_t hi, uint32_t sel) { int s; asm("prmt.b32 %0, %1, %2, %3;" : "=r"(s) : "r"(lo), "r"(hi), "r"(sel)); return s; }
#else
inline int xdp_prmt(uint32_t lo, uint32_t hi, uint32_t sel) { return (int)__byte_perm(lo, hi, sel); }
#endif

// Column-word codecs.  Narrow (the pipeline's): E << 24 | H << 16 | PRMT selector, scores <= 255.  Wide: E << 17 | H << 4 | q,
// 13-bit scores, plain mask-and-shift code -- only for callers of the stage entry point whose h0 + a * qlen exceeds a byte.
template <bool kWide> struct XCol;
template <> struct XCol<false> {
    static __device__ __forceinline__ uint32_t init(int h, uint32_t q) { return (uint32_t)h << 16 | (q * 0x1111u + 0x8880u); }
    static __device__ __forceinline__ int H(uint32_t w) { return (int)__byte_perm(w, 0u, 0x4442); }
    static __device__ __forceinline__ int E(uint32_t w) { return (int)(w >> 24); }
    static __device__ __forceinline__ int score(uint32_t lo, uint32_t hi, uint32_t w) { return xdp_prmt(lo, hi, w); }
    static __device__ __forceinline__ uint32_t pack(uint32_t w, int h, int e, int c256) { return __byte_perm(w, (uint32_t)(e * c256 + h), 0x5410); }
    static __device__ __forceinline__ uint32_t with_h(uint32_t w, int h) { return (w & 0xffffu) | (uint32_t)h << 16; }
    static __device__ __forceinline__ bool zero(uint32_t w) { return (w >> 16) == 0; }
    static constexpr int kMaxScore = 255;
};
template <> struct XCol<true> {
    static __device__ __forceinline__ uint32_t init(int h, uint32_t q) { return (uint32_t)h << 4 | q; }
    static __device__ __forceinline__ int H(uint32_t w) { return (int)((w >> 4) & 0x1fffu); }
    static __device__ __forceinline__ int E(uint32_t w) { return (int)(w >> 17); }
    static __device__ __forceinline__ int score(uint32_t lo, uint32_t hi, uint32_t w) { return (int)(int8_t)__byte_perm(lo, hi, w & 7u); }
    static __device__ __forceinline__ uint32_t pack(uint32_t w, int h, int e, int) { return (w & 7u) | (uint32_t)h << 4 | (uint32_t)e << 17; }
    static __device__ __forceinline__ uint32_t with_h(uint32_t w, int h) { return (w & 7u) | (uint32_t)h << 4; }
    static __device__ __forceinline__ bool zero(uint32_t w) { return (w >> 4) == 0; }
    static constexpr int kMaxScore = 8191;
};

template <bool kPackedRef, bool kWide>
__global__ void __launch_bounds__(XDP_BLOCK) k_ext_dp(MemOpt opt, RefView ref, ExtDpArgs A) {
    extern __shared__ uint32_t xcols[];  // [ncols][XDP_BLOCK]
    uint32_t* col = xcols + threadIdx.x;
#define XCOL(j) col[(j) * XDP_BLOCK]
    const unsigned lane = threadIdx.x & 31u;
    const unsigned n = *A.n_items;
    const int e_del = opt.e_del, e_ins = opt.e_ins, o_del = opt.o_del, o_ins = opt.o_ins;
    const int oe_del = o_del + e_del, oe_ins = o_ins + e_ins, zdrop = A.zdrop;
    const uint32_t mis4 = (uint32_t)(-opt.b & 0xff) * 0x01010101u, mat1 = (uint32_t)(opt.a & 0xff);
    const uint32_t lut_hi = 0xffffffffu;  // query N scores -1
    // multipliers the compiler cannot fold: `x * one + c` and `x * c256 + y` stay IMADs, which issue to the FMA pipe instead of the
    // saturated integer ALU pipe (as plain adds / LEAs they were 3 of a cell's 12 ALU-pipe instructions)
    const int one = A.one, c256 = A.one << 8;
    bool live = false, exhausted = false, need_init = false;
    uint4 pool0 = make_uint4(0, 0, 0, 0), pool1 = pool0, pool2 = pool0;
    int pool_n = 0, pool_pos = 0;
    int slot = 0, qlen = 0, tlen = 0, w = 0, h0 = 0, i = 0, beg = 0, end = 0;
    int mx = 0, max_i = -1, max_j = -1, max_ie = -1, gscore = -1, max_off = 0;
    int64_t tfp = 0; int tfstep = 1; bool tcomp = false; int64_t tw_idx = -1; uint32_t tw = 0;
    const uint8_t* qptr = A.seq; int qstep = 1;
    unsigned long long cells = 0; unsigned calls = 0;

    int tb_next = 0;
    // target base of row x: from a cached 16-base word of the packed reference (bns_get_seq, bntseq.c:398-419) or the caller's bytes.
    // Always asked for one row ahead, so the (rare) word load is in flight while a row's cells are computed.
    auto target_base = [&](int x) -> int {
        if (kPackedRef) {
            const int64_t fp = tfp + (int64_t)x * tfstep;
            if ((fp >> 4) != tw_idx) { tw_idx = fp >> 4; tw = __ldg(reinterpret_cast<const uint32_t*>(ref.pac) + tw_idx); }
            const int b = (int)(tw >> ((((unsigned)fp >> 2) & 3u) << 3) >> ((~(unsigned)fp & 3u) << 1)) & 3;
            return tcomp ? 3 - b : b;
        }
        return (int)__ldg(A.tbytes + tfp + x);
    };
    for (;;) {
        // ---- idle lanes take the next problems ----
        // The warp keeps a POOL of up to 32 descriptors in registers (lane k holds queue entry pool_base + k), loaded with one
        // coalesced burst right after the previous pool ran out, i.e. long before anybody needs them: taking a problem is a
        // handful of shuffles, not a dependent global load that stalls the 31 lanes still computing.  Lanes are refilled in
        // batches (XDP_REFILL idle lanes, or nobody live): every refill event costs the warp one first-row fill per problem.
        const unsigned idle = __ballot_sync(0xffffffffu, !live);
        const int n_idle = __popc(idle);
        if (pool_pos >= pool_n && !exhausted) {  // (re)load the pool; the registers are first read at the next take
            unsigned base = 0;
            if (lane == 0) base = atomicAdd(A.head, 32u);
            base = __shfl_sync(0xffffffffu, base, 0);
            if (base >= n) { exhausted = true; pool_n = 0; }
            else {
                pool_n = n - base < 32u ? (int)(n - base) : 32;
                if ((int)lane < pool_n) {
                    const uint4* src = reinterpret_cast<const uint4*>(A.queue + base + lane);
                    pool0 = __ldg(src); pool1 = __ldg(src + 1); pool2 = __ldg(src + 2);
                }
            }
            pool_pos = 0;
        }
        if (n_idle && pool_pos < pool_n && (n_idle >= XDP_REFILL || n_idle == 32)) {
            const int avail = pool_n - pool_pos, take = n_idle < avail ? n_idle : avail;
            const int rank = __popc(idle & ((1u << lane) - 1u));
            const bool mine = !live && rank < take;
            const int from = mine ? pool_pos + rank : (int)lane;
            uint4 d0, d1, d2;
            d0.x = __shfl_sync(0xffffffffu, pool0.x, from); d0.y = __shfl_sync(0xffffffffu, pool0.y, from);
            d0.z = __shfl_sync(0xffffffffu, pool0.z, from); d0.w = __shfl_sync(0xffffffffu, pool0.w, from);
            d1.x = __shfl_sync(0xffffffffu, pool1.x, from); d1.y = __shfl_sync(0xffffffffu, pool1.y, from);
            d1.z = __shfl_sync(0xffffffffu, pool1.z, from); d1.w = __shfl_sync(0xffffffffu, pool1.w, from);
            d2.x = __shfl_sync(0xffffffffu, pool2.x, from); d2.y = __shfl_sync(0xffffffffu, pool2.y, from);
            d2.z = __shfl_sync(0xffffffffu, pool2.z, from); d2.w = __shfl_sync(0xffffffffu, pool2.w, from);
            pool_pos += take;
            if (mine) {
                ExtProb P;
                {
                    uint4* dst = reinterpret_cast<uint4*>(&P);
                    dst[0] = d0; dst[1] = d1; dst[2] = d2;
                }
                slot = P.read;
                qlen = P.qlen; tlen = P.tlen; h0 = P.h0;
                qptr = A.seq + P.qoff; qstep = (P.dirs & 1) ? -1 : 1;
                ++calls;
                {   // band limit (ksw.c:402-407): (int)((double)x / e + 1.) == (x + e) / e in C integer division
                    int top = opt.a > -opt.b ? opt.a : -opt.b; top = top > -1 ? top : -1;
                    int lim = (qlen * top + P.end_bonus - opt.o_ins + opt.e_ins) / opt.e_ins;
                    lim = lim > 1 ? lim : 1; w = P.w < lim ? P.w : lim;
                    lim = (qlen * top + P.end_bonus - opt.o_del + opt.e_del) / opt.e_del;
                    lim = lim > 1 ? lim : 1; w = w < lim ? w : lim;
                }
                if (kPackedRef) {  // forward-strand coordinate of row 0 and its step: a window never crosses l_pac
                    tcomp = P.tpos >= ref.l_pac;
                    tfp = tcomp ? (ref.l_pac << 1) - 1 - P.tpos : P.tpos;
                    const int tstep = (P.dirs & 2) ? -1 : 1;
                    tfstep = tcomp ? -tstep : tstep;
                    tw_idx = -1;
                } else tfp = P.tpos;
                mx = h0; max_i = max_j = max_ie = -1; gscore = -1; max_off = 0; beg = 0; end = qlen; i = 0;
                if (tlen <= 0 || qlen + 2 > A.ncols || P.h0 + qlen * opt.a > XCol<kWide>::kMaxScore) {  // no row at all (ksw.c:409 never enters its loop) / caller sized the columns too small
                    ExtRes R; R.mx = tlen <= 0 ? P.h0 : -0x7fffffff; R.max_i = R.max_j = R.max_ie = -1; R.gscore = -1; R.max_off = 0;
                    A.res[slot] = R;
                } else { live = true; need_init = true; tb_next = target_base(0); }
            }
        }
        // ---- first row (ksw.c:395-397) of the lanes that just took a problem, filled by the whole warp, 32 columns per trip ----
        {
            unsigned want = __ballot_sync(0xffffffffu, need_init);
            while (want) {
                const int src = __ffs(want) - 1;
                want &= want - 1;
                const uint8_t* bq = reinterpret_cast<const uint8_t*>(__shfl_sync(0xffffffffu, (unsigned long long)qptr, src));
                const int bqs = __shfl_sync(0xffffffffu, qstep, src), bqlen = __shfl_sync(0xffffffffu, qlen, src), bh0 = __shfl_sync(0xffffffffu, h0, src);
                uint32_t* bcol = xcols + (threadIdx.x & ~31u) + src;
                for (int x0 = (int)lane; x0 <= bqlen; x0 += 128) {  // four independent loads in flight per trip
                    uint32_t qb[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) { const int x = x0 + 32 * k; qb[k] = x < bqlen ? (uint32_t)__ldg(bq + (long)x * bqs) : 0u; }
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const int x = x0 + 32 * k;
                        if (x <= bqlen) {
                            int hv = bh0 - o_ins - x * e_ins;
                            hv = x == 0 ? bh0 : (hv > 0 ? hv : 0);
                            bcol[x * XDP_BLOCK] = XCol<kWide>::init(hv, qb[k] > 4u ? 4u : qb[k]);
                        }
                    }
                }
            }
            __syncwarp();
            need_init = false;
        }
        // ---- row start (ksw.c:410-423) ----
        int nrow = 0, h1 = 0, f = 0;
        uint32_t lut_lo = lut_hi;
        if (live) {
            const int tb = tb_next;  // fetched while the previous row was computed (row 0: at problem start)
            if (tb < 4) lut_lo = (mis4 & ~(0xffu << (tb << 3))) | (mat1 << (tb << 3));
            if (beg < i - w) beg = i - w;
            if (end > i + w + 1) end = i + w + 1;
            if (end > qlen) end = qlen;
            if (beg == 0) { h1 = h0 - (o_del + e_del * (i + 1)); h1 = h1 > 0 ? h1 : 0; }
            nrow = end > beg ? end - beg : 0;
            cells += nrow;
            if (i + 1 < tlen) tb_next = target_base(i + 1);
        }
        // ---- the row's cells (ksw.c:424-448): groups of 4 to the warp's longest row, then up to 3 single cells ----
        int key = 0;  // 256 * (row maximum of H) + its column; ties go to the larger column (:437)
        {
            uint32_t* p = &XCOL(beg);
            const int nfull = nrow >> 2, nrem = nrow & 3;
            const int wfull = __reduce_max_sync(0xffffffffu, nfull);
            int jb = beg;
#define XDP_CELL(cw_, u_)                                                              \
    {                                                                                  \
        const int hd = XCol<kWide>::H(cw_), e = XCol<kWide>::E(cw_);                   \
        const int sc = XCol<kWide>::score(lut_lo, lut_hi, (cw_));                      \
        const int M = __viaddmin_s32_relu(hd, sc, hd << 8);                            \
        const int h = __vimax3_s32(M, e, f);                                           \
        const int e2 = __viaddmax_s32_relu(M, -oe_del, e * one - e_del);               \
        f = __viaddmax_s32(f, -e_ins, M * one - oe_ins);                               \
        p[(u_) * XDP_BLOCK] = XCol<kWide>::pack((cw_), h1, e2, c256);                  \
        k##u_ = h * c256 + (u_);                                                       \
        h1 = h;                                                                        \
    }
#pragma unroll 1
            for (int g = 0; g < wfull; ++g) {
                if (g < nfull) {
                    const uint32_t c0 = p[0], c1 = p[XDP_BLOCK], c2 = p[2 * XDP_BLOCK], c3 = p[3 * XDP_BLOCK];
                    int k0, k1, k2, k3;
                    XDP_CELL(c0, 0) XDP_CELL(c1, 1) XDP_CELL(c2, 2) XDP_CELL(c3, 3)
                    const int gk = __vimax3_s32(__vimax3_s32(k0, k1, k2), k3, -1);
                    key = __viaddmax_s32(gk, jb, key);
                    p += 4 * XDP_BLOCK; jb += 4;
                }
            }
            const int wrem = __reduce_max_sync(0xffffffffu, nrem);
#pragma unroll 1
            for (int u = 0; u < wrem; ++u) {
                if (u < nrem) {
                    const uint32_t c0 = p[0];
                    int k0;
                    XDP_CELL(c0, 0)
                    key = __viaddmax_s32(k0, jb, key);
                    p += XDP_BLOCK; ++jb;
                }
            }
#undef XDP_CELL
        }
        // ---- row end (ksw.c:449-469) ----
        if (live) {
            bool done = false;
            const int jfin = end > beg ? end : beg;
            XCOL(end) = XCol<kWide>::with_h(XCOL(end), h1);
            if (jfin == qlen) { max_ie = gscore > h1 ? max_ie : i; gscore = gscore > h1 ? gscore : h1; }
            const int m = key >> 8, mj = key & 0xff;
            if (m == 0) done = true;
            else if (m > mx) {
                mx = m; max_i = i; max_j = mj;
                int d = mj - i; d = d < 0 ? -d : d;
                max_off = max_off > d ? max_off : d;
            } else if (zdrop > 0) {
                if (i - max_i > mj - max_j) done = mx - m - ((i - max_i) - (mj - max_j)) * e_del > zdrop;
                else done = mx - m - ((mj - max_j) - (i - max_i)) * e_ins > zdrop;
            }
            if (!done) {  // band for the next row: first and last column holding a non-zero H or E (ksw.c:465-468)
                int j = beg;
                while (j < end && XCol<kWide>::zero(XCOL(j))) ++j;
                beg = j;
                j = end;
                while (j >= beg && XCol<kWide>::zero(XCOL(j))) --j;
                end = j + 2 < qlen ? j + 2 : qlen;
                if (++i >= tlen) done = true;
            }
            if (done) {
                ExtRes R;
                R.mx = mx; R.max_i = max_i; R.max_j = max_j; R.max_ie = max_ie; R.gscore = gscore; R.max_off = max_off;
                A.res[slot] = R;
                live = false;
            }
        }
        if (exhausted && pool_pos >= pool_n && __ballot_sync(0xffffffffu, live) == 0) break;
    }
#undef XCOL
    if (cells || calls) { atomicAdd(&A.cnt->ext_cells, cells); atomicAdd(&A.cnt->ext_calls, (unsigned long long)calls); }
}

}  // namespace arachne
