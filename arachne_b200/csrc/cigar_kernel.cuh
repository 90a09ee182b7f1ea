// Region -> alignment record, thread-per-region kernels (the warp-cooperative version in cigar.cuh/k_cigar stays as the
// wide-band tier).  Same reference functions as cigar.cuh:
//   mem_reg2aln bwa/bwamem.c:1086-1156, bwa_gen_cigar2 bwa/bwa.c:121-207, ksw_global2 bwa/ksw.c:504-606,
//   GetAlignments walk src/aligner/aligner.go:1505-1569, InterpretAlign src/gobwa/gobwa.go:339-371.
//
// Why three tiers.  ncu (profiles/r1b): the one-warp-per-region kernel spent 13.6 k warp instructions per region because
// (a) 32 lanes ran the gap-free path's 150-base loops redundantly and (b) the banded NW filled all qlen columns of every
// row although the inferred band is a few cells wide.  On the bench workload 86 % of regions are gap-free
// (bwa.c:141-149) and nearly all others need w <= 8.
//   tier 1  k_cigar_fast   one lane per region: band inference; the gap-free case is finished here (score, NM, clips, walk)
//   tier 2  k_cigar_band   one lane per queued region: exact ksw_global2 with the band in a 64-slot shared-memory ring
//                          (w <= 30); the traceback matrix is 4 bits per cell, accumulated in a register and stored 16 cells at
//                          a time, interleaved across lanes so that a warp's stores coalesce (byte-per-cell stores into
//                          per-lane slices made this kernel L2-transaction bound: 4.8 ms, profiles/r1d); the up-to-3 band
//                          doublings of mem_reg2aln
//   tier 3  k_cigar        (cigar.cuh) one warp per region for whatever needs a wider band or a long CIGAR
// Sequences are read where they lie: the read from the batch buffer, the reference straight from the 2-bit .pac.
#pragma once
#include "cigar.cuh"

namespace arachne {

constexpr int CIG_LOCAL = 24;     // CIGAR ops a lane keeps locally; longer ones go to tier 3
constexpr int BAND_RING = 64;     // ring slots per lane: covers band half-width w <= 30 (2w+3 <= 63)
constexpr int BAND_WMAX = 30;
constexpr int BAND_ZCHUNK = 4;    // 64 bits x 4 of 4-bit direction codes per row (stored as 8 32-bit words of 8 cells), 2w+1 <= 61 cells
constexpr int CIG_BLOCK = 128;

struct CigArgs {
    const uint8_t* seq; const int64_t* off;
    long n_regs; const AlnReg* regs; const int* reg_read; AlnRec* alns;
    uint32_t* cig_pool; unsigned long long cig_cap; int32_t* mm_pool; unsigned long long mm_cap;
    unsigned long long* used;          // [0] CIGAR words, [1] mismatch ints
    unsigned* list2; unsigned* n2;     // regions for tier 2 (k_cigar_fast appends, the host sorts by key2, k_cigar_band reads)
    unsigned* key2;
    unsigned* list3; unsigned* n3;     // regions for tier 3
    unsigned long long* z; int z_rows; // tier-2 direction codes: word (row, chunk) of lane t at z[(row*BAND_ZCHUNK + chunk)*T + t]
    Counters* cnt;
};

// view of one region's two sequences in the orientation bwa_gen_cigar2 aligns them in (both reversed on the reverse
// strand so that indels are left-aligned on the forward strand, bwa.c:135-140)
struct RegView {
    const uint8_t* q; const uint8_t* pac;
    int64_t tpos;   // forward-strand coordinate of target base 0
    int lq, rlen, qb, qe, rev;
    __device__ __forceinline__ int Q(int i) const { return rev ? q[qe - 1 - i] : q[qb + i]; }
    // 16 target bases per 32-bit load of the packed reference, kept across calls: the DP reads one target base per row and
    // the walks move along the region, so a load (a DRAM round trip on a 3.1 Gb reference) is needed every 16th base only
    mutable uint32_t tw = 0;
    mutable int64_t tw_at = -1;
    __device__ __forceinline__ int T(int i) const {
        const int64_t p = tpos + i;
        if ((p >> 4) != tw_at) { tw_at = p >> 4; tw = __ldg(reinterpret_cast<const uint32_t*>(pac) + tw_at); }
        const int b = (int)(tw >> ((((unsigned)p >> 2) & 3u) << 3) >> ((~(unsigned)p & 3u) << 1)) & 3;
        return rev ? 3 - b : b;
    }
};

__device__ __forceinline__ RegView make_view(const RefView& ref, const uint8_t* query, const AlnReg& ar) {
    RegView v;
    v.q = query; v.pac = ref.pac; v.qb = ar.qb; v.qe = ar.qe; v.lq = ar.qe - ar.qb; v.rlen = (int)(ar.re - ar.rb);
    v.rev = ar.rb >= ref.l_pac;
    v.tpos = v.rev ? (ref.l_pac << 1) - ar.re : ar.rb;
    return v;
}

// NM as bwa_gen_cigar2 computes it (bwa.c:171-198)
__device__ __forceinline__ int cig_nm(const RegView& v, const uint32_t* c, int n) {
    int x = 0, y = 0, n_mm = 0, n_gap = 0;
    for (int k = 0; k < n; ++k) {
        const int op = c[k] & 0xf, len = (int)(c[k] >> 4);
        if (op == 0) { for (int i = 0; i < len; ++i) n_mm += v.Q(x + i) != v.T(y + i); x += len; y += len; }
        else if (op == 2) { if (k > 0 && k < n - 1) n_gap += len; y += len; }
        else if (op == 1) { x += len; n_gap += len; }
    }
    return n_mm + n_gap;
}

// Everything mem_reg2aln does after the CIGAR loop (bwamem.c:1120-1155) + the GetAlignments walk + pool output.
// c[0..n) is the CIGAR of the aligned part in forward-strand order; c must have room for two more words.
__device__ __forceinline__ bool cig_finish(const CigArgs& A, const RefView& ref, long g, int r, int l_query, const uint8_t* query, const AlnReg& ar,
                                           const RegView& v, uint32_t* c, int n, int score, int NM) {
    AlnRec a;
    int is_rev;
    int64_t pos = bns_depos(ref, ar.rb < ref.l_pac ? ar.rb : ar.re - 1, &is_rev);
    a.is_rev = is_rev; a.nm = NM;
    int c0 = 0;
    if (n > 0) {  // squeeze out a leading, else a trailing, deletion (:1124-1133)
        if ((c[0] & 0xf) == 2) { pos += c[0] >> 4; --n; c0 = 1; }
        else if ((c[n - 1] & 0xf) == 2) --n;
    }
    int clip5 = 0, clip3 = 0;
    if (ar.qb != 0 || ar.qe != l_query) { clip5 = is_rev ? l_query - ar.qe : ar.qb; clip3 = is_rev ? ar.qb : l_query - ar.qe; }
    const int n_out = n + (clip5 ? 1 : 0) + (clip3 ? 1 : 0);
    a.n_cigar = n_out;
    a.rid = bns_pos2rid(ref, pos);
    a.pos = pos - ref.contigs[a.rid].offset;
    a.score = ar.score; a.sub = ar.sub > ar.csub ? ar.sub : ar.csub;
    // ---- GetAlignments walk (aligner.go:1505-1569): final CIGAR against the reference in read orientation ----
    const int64_t coff = ref.contigs[ar.rid].offset;
    const int64_t ch_pos = ar.rb < ref.l_pac ? ar.rb - coff : ref.l_pac * 2 - 1 - ar.rb - coff;
    const int64_t ch_aend = ar.re < ref.l_pac ? ar.re - coff : ref.l_pac * 2 - 1 - ar.re - coff;
    int64_t refStart = ch_pos, refEnd = ch_aend;
    if (is_rev) { refStart = ch_aend + 1; refEnd = ch_pos + 1; }
    const int ref_len = v.rlen;
    int matches = 0, indels = 0, indel_len = 0, sclip = 0, sclip_len = 0, n_mm = 0;
    unsigned long long mo = 0;
    // op k of the final CIGAR: [clip5] c[c0 .. c0+n) [clip3]
    auto op_at = [&](int k) -> uint32_t {
        if (clip5) { if (k == 0) return (uint32_t)clip5 << 4 | 3; --k; }
        return k < n ? c[c0 + k] : ((uint32_t)clip3 << 4 | 3);
    };
    for (int pass = 0; pass < 2; ++pass) {  // pass 0 counts the mismatch loci, pass 1 writes them
        int refOff = 0, readOff = 0, w = 0;
        for (int kk = 0; kk < n_out; ++kk) {
            const uint32_t cw = op_at(is_rev ? n_out - 1 - kk : kk);
            const int op = cw & 0xf, len = (int)(cw >> 4);
            if (op == 0) {
                for (int m = 0; m < len; ++m) {
                    if (refOff + m >= ref_len) continue;
                    // reference base k of the region in read orientation == target base (rev ? rlen-1-k : k) of the view
                    const int k = refOff + m;
                    if (readOff + m < l_query && v.T(v.rev ? v.rlen - 1 - k : k) != query[readOff + m]) {
                        if (pass) {
                            A.mm_pool[mo + w] = is_rev ? (int32_t)(refEnd - k) : (int32_t)(k + refStart);
                            A.mm_pool[mo + n_mm + w] = readOff + m;
                        }
                        ++w;
                    }
                }
                if (!pass) matches += len;
                refOff += len; readOff += len;
            } else if (op == 1) { if (!pass) { ++indels; indel_len += len; } readOff += len; }
            else if (op == 2) { if (!pass) { ++indels; indel_len += len; } refOff += len; }
            else if (op == 3) { if (!pass) { ++sclip; sclip_len += len; } readOff += len; }
        }
        if (!pass) {
            n_mm = w;
            if (n_mm == 0) break;
            mo = atomicAdd(&A.used[1], (unsigned long long)(2 * n_mm));
            if (mo + 2 * n_mm > A.mm_cap) { raise(A.cnt, -101, r); return false; }
        }
    }
    int mismatches = NM - indel_len;
    matches -= mismatches;
    if (mismatches < 0) mismatches = 0;
    a.matches = matches; a.mismatches = mismatches; a.indels = indels; a.soft_clipped = sclip; a.soft_clipped_length = sclip_len;
    a.n_mm = n_mm; a.mm_off = (int64_t)mo;
    const unsigned long long co = atomicAdd(&A.used[0], (unsigned long long)n_out);
    if (co + n_out > A.cig_cap) { raise(A.cnt, -101, r); return false; }
    for (int k = 0; k < n_out; ++k) A.cig_pool[co + k] = op_at(k);
    a.cigar_off = (int64_t)co;
    A.alns[g] = a;
    return true;
}

// band half-width mem_reg2aln starts from (bwamem.c:1105-1110)
__device__ __forceinline__ int cig_first_w(const MemOpt& opt, const AlnReg& ar) {
    const int l1 = ar.qe - ar.qb, l2 = (int)(ar.re - ar.rb);
    int tmp = infer_bw(l1, l2, ar.truesc, opt.a, opt.o_del, opt.e_del);
    int w2 = infer_bw(l1, l2, ar.truesc, opt.a, opt.o_ins, opt.e_ins);
    w2 = w2 > tmp ? w2 : tmp;
    if (w2 > opt.w) w2 = w2 < ar.w ? w2 : ar.w;
    return w2;
}

// the band half-width ksw_global2 is really called with (bwa.c:151-163)
__device__ __forceinline__ int cig_eff_w(const MemOpt& opt, int lq, int rlen, int w_) {
    int max_ins = (int)((double)(((lq + 1) >> 1) * opt.mat[0] - opt.o_ins) / opt.e_ins + 1.);
    int max_del = (int)((double)(((lq + 1) >> 1) * opt.mat[0] - opt.o_del) / opt.e_del + 1.);
    int max_gap = max_ins > max_del ? max_ins : max_del;
    max_gap = max_gap > 1 ? max_gap : 1;
    int dl = rlen - lq; dl = dl < 0 ? -dl : dl;
    int w = (max_gap + dl + 1) >> 1;
    w = w < w_ ? w : w_;
    return w > dl + 3 ? w : dl + 3;
}

// ------------------------------------------------------------------------------------------------ tier 1
__global__ void __launch_bounds__(CIG_BLOCK) k_cigar_fast(MemOpt opt, RefView ref, CigArgs A) {
    for (long g = blockIdx.x * (long)blockDim.x + threadIdx.x; g < A.n_regs; g += (long)gridDim.x * blockDim.x) {
        const AlnReg ar = A.regs[g];
        const int r = A.reg_read[g];
        const int l_query = (int)(A.off[r + 1] - A.off[r]);
        const uint8_t* query = A.seq + A.off[r];
        const int lq = ar.qe - ar.qb;
        // anything bwa_gen_cigar2 answers without an alignment (bwa.c:133-135), or that looks unusual, is the wide tier's
        const bool odd = ar.rb < 0 || ar.re < 0 || lq <= 0 || ar.rb >= ar.re || (ar.rb < ref.l_pac && ar.re > ref.l_pac) || ar.re > (ref.l_pac << 1);
        if (odd) { A.list3[atomicAdd(A.n3, 1u)] = (unsigned)g; continue; }
        int w2 = cig_first_w(opt, ar);
        w2 = w2 < opt.w << 2 ? w2 : opt.w << 2;
        const RegView v = make_view(ref, query, ar);
        if (!(lq == v.rlen && w2 == 0)) {
            // banded: tier 2 takes it when the first band fits its ring; the key lets the host sort tier 2's work so that
            // the lanes of a warp get regions with the same band width and similar length
            const int w = cig_eff_w(opt, lq, v.rlen, w2);
            if (w > BAND_WMAX) { A.list3[atomicAdd(A.n3, 1u)] = (unsigned)g; continue; }
            const unsigned t = atomicAdd(A.n2, 1u);
            A.list2[t] = (unsigned)g; A.key2[t] = (unsigned)w << 10 | (unsigned)(v.rlen < 1023 ? v.rlen : 1023);
            continue;
        }
        // gap-free (bwa.c:141-149).  The band-doubling loop of mem_reg2aln re-runs this with w2 = 0 << 1 = 0 and stops on
        // the unchanged score, so this is the final answer.
        int score = 0, nm = 0;
        for (int i = 0; i < lq; ++i) {
            const int qb = v.Q(i), tb = v.T(i);
            score += qb > 3 ? -1 : (qb == tb ? opt.a : -opt.b);
            nm += qb != tb;
        }
        uint32_t c[3];
        c[0] = (uint32_t)lq << 4;
        cig_finish(A, ref, g, r, l_query, query, ar, v, c, 1, score, nm);
    }
}

// PRMT (default mode): byte i of the result is byte (sel >> 4i) & 7 of the 8 bytes {hi:lo}, or that byte's sign replicated when bit 3 of
// the selector nibble is set.  The host definition (TEST-ONLY build, tests/hostsim/hostsim_cigar.cpp) states the same thing in C.
#if defined(__CUDACC__)
__device__ __forceinline__ int cig_prmt(unsigned lo, unsigned hi, int sel) {
    int r;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(lo), "r"(hi), "r"(sel));
    return r;
}
#else
inline int cig_prmt(unsigned lo, unsigned hi, int sel) {
    const unsigned long long src = (unsigned long long)hi << 32 | lo;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) {
        const unsigned n = ((unsigned)sel >> (4 * i)) & 0xfu, b = (unsigned)(src >> (8 * (n & 7u))) & 0xffu;
        r |= ((n & 8u) ? ((b & 0x80u) ? 0xffu : 0u) : b) << (8 * i);
    }
    return (int)r;
}
#endif

// ------------------------------------------------------------------------------------------------ tier 2
// One bwa_gen_cigar2 call for band parameter w_ (bwa.c:121-207) on the lane's ring.  Returns n_cigar >= 0 with the CIGAR
// in c (forward-strand order), or -1 when this lane cannot do it (band or CIGAR too large -> tier 3).
// Direction code per cell (ksw.c:551-563 keeps a byte d = h | e<<2 | f<<4 with e in {0,1}, f in {0,2}):
// nibble = h | e<<2 | (f>>1)<<3.
__device__ __forceinline__ int band_gen_cigar(const MemOpt& opt, const RegView& v, int w_, int2 (*ring)[CIG_BLOCK], unsigned* z32, long T,
                                              int z_rows, int* score, uint32_t* c, unsigned long long* cells) {
    const int lq = v.lq, rlen = v.rlen, tid = threadIdx.x;
    if (lq == rlen && w_ == 0) {
        int s = 0;
        for (int i = 0; i < lq; ++i) { const int qb = v.Q(i), tb = v.T(i); s += qb > 3 ? -1 : (qb == tb ? opt.a : -opt.b); }
        *score = s; c[0] = (uint32_t)lq << 4;
        return 1;
    }
    const int w = cig_eff_w(opt, lq, rlen, w_);
    if (w > BAND_WMAX || rlen > z_rows) return -1;
    // ksw_global2 (ksw.c:504-606): eh[j] = {H(i-1,j-1), E(i,j)} lives in ring slot j & 63: .x = H, .y = E << 16 | selector.
    // The selector (q * 0x1111 + 0x8880: "byte q, sign-replicated") is a ready-made PRMT selector into the row's substitution
    // scores {s(tb,0..3) in one register, s(N) = -1 in another}, so a cell's score is one instruction and the cell loop has no
    // global load: a column's base is fetched once, when the sliding band first reaches it.
    // "-infinity": -2^26 for H and F, -2^14 for the 16-bit E (the reference: -2^30).  Only two things about sentinels ever
    // matter -- that anything derived from one stays below every real score, and that a real cell's choice between a real and
    // a sentinel candidate goes to the real one; comparisons between two sentinel-derived values only set direction bits the
    // backtrack never reads (it reaches a cell's E / F bit only through a real E / F).  Real scores are above
    // -(len * (b + 2e) + 2o) and a sentinel drifts down by at most e per row and column; the guard below keeps both well
    // inside 16 bits, anything else is tier 3's.
    constexpr int NEG = -(1 << 26), NEGE = -(1 << 14);
    const int e_del = opt.e_del, e_ins = opt.e_ins, oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    {
        const int lmax = lq > rlen ? lq : rlen, emax = e_del > e_ins ? e_del : e_ins, omax = opt.o_del > opt.o_ins ? opt.o_del : opt.o_ins;
        if (lmax * (opt.b + 2 * emax + opt.a) + 2 * omax + (2 * w + 2) * emax > 6000) return -1;
    }
    const uint8_t* qp = v.rev ? v.q + v.qe - 1 : v.q + v.qb;
    const int qs = v.rev ? -1 : 1;
    auto slot_y = [](int e, int q) -> int { return (int)(((unsigned)e << 16) | (unsigned)((q > 4 ? 4 : q) * 0x1111 + 0x8880)); };
    for (int j = 0; j <= lq && j <= w + 1; ++j) {
        const int qb = j < lq ? qp[j * qs] : 0;
        ring[j & (BAND_RING - 1)][tid] = make_int2(j == 0 ? 0 : (j <= w ? -(opt.o_ins + e_ins * j) : NEG), slot_y(NEGE, qb));
    }
    const unsigned mis4 = (unsigned)(-opt.b & 0xff) * 0x01010101u, lut_hi = 0xffu;  // s(N, .) = -1 (bwa_fill_scmat)
    // direction codes: 8 cells per 32-bit word, lane-interleaved: word k of row i of lane t at z32[(i*8 + k)*T] (z32 already offset by t)
    for (int i = 0; i < rlen; ++i) {
        const int tb = v.T(i) & 3;
        const unsigned lut_lo = (mis4 & ~(0xffu << (tb << 3))) | ((unsigned)(opt.a & 0xff) << (tb << 3));
        const int beg = i > w ? i - w : 0, end = i + w + 1 < lq ? i + w + 1 : lq;
        const int qnext = end < lq ? qp[end * qs] : 0;  // base of the column the band reaches in the next row
        int f = NEG, left = beg == 0 ? -(opt.o_del + e_del * (i + 1)) : NEG;
        unsigned* zi = z32 + (long)i * (2 * BAND_ZCHUNK) * T;
#define BAND_CELL(slot_, u_)                                                                    \
    {                                                                                           \
        const int2 he = (slot_);                                                                \
        const int m = he.x + cig_prmt(lut_lo, lut_hi, he.y);                                    \
        int e = he.y >> 16, h, x;                                                               \
        unsigned d = m >= e ? 0u : 1u; h = m >= e ? m : e;                                      \
        if (h < f) { d = 2u; h = f; }                                                           \
        x = m - oe_del; e -= e_del;                                                             \
        if (e > x) d |= 4u; else e = x;                                                         \
        (slot_) = make_int2(left, (int)__byte_perm((unsigned)he.y, (unsigned)e, 0x5410));       \
        left = h;                                                                               \
        x = m - oe_ins; f -= e_ins;                                                             \
        if (f > x) d |= 8u; else f = x;                                                         \
        acc |= d << (4 * (u_));                                                                 \
    }
        for (int j0 = beg; j0 < end; j0 += 8) {
            unsigned acc = 0;
            const int s0 = j0 & (BAND_RING - 1);
            if (j0 + 8 <= end && s0 <= BAND_RING - 8) {  // a full group that does not wrap: constant slot offsets, no per-cell test
                int2* rp = &ring[s0][tid];
#pragma unroll
                for (int u = 0; u < 8; ++u) BAND_CELL(rp[u * CIG_BLOCK], u)
            } else {
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = j0 + u;
                    if (j < end) BAND_CELL(ring[j & (BAND_RING - 1)][tid], u)
                }
            }
            zi[(long)((j0 - beg) >> 3) * T] = acc;
        }
#undef BAND_CELL
        ring[end & (BAND_RING - 1)][tid] = make_int2(left, slot_y(NEGE, qnext));
        *cells += end > beg ? end - beg : 0;
    }
    *score = ring[lq & (BAND_RING - 1)][tid].x;
    // backtrack (ksw.c:585-603): ops are produced last-to-first
    int n = 0, state = 0, i = rlen - 1, k = (i + w + 1 < lq ? i + w + 1 : lq) - 1;
    auto push = [&](int op, int len) -> bool {
        if (n > 0 && (int)(c[n - 1] & 0xf) == op) { c[n - 1] += (uint32_t)len << 4; return true; }
        if (n >= CIG_LOCAL - 2) return false;
        c[n++] = (uint32_t)len << 4 | op;
        return true;
    };
    bool ok = true;
    while (i >= 0 && k >= 0 && ok) {
        const int ci = k - (i > w ? i - w : 0);
        if (ci < 0 || k >= (i + w + 1 < lq ? i + w + 1 : lq)) return -1;  // the path left the band: the reference reads a neighbouring row's byte here
        const unsigned nib = (z32[((long)i * (2 * BAND_ZCHUNK) + (ci >> 3)) * T] >> ((ci & 7) << 2)) & 15u;
        state = state == 0 ? (int)(nib & 3u) : (state == 1 ? (int)((nib >> 2) & 1u) : (int)((nib >> 3) & 1u) << 1);
        if (state == 0) { ok = push(0, 1); --i; --k; }
        else if (state == 1) { ok = push(2, 1); --i; }
        else { ok = push(1, 1); --k; }
    }
    if (ok && i >= 0) ok = push(2, i + 1);
    if (ok && k >= 0) ok = push(1, k + 1);
    if (!ok) return -1;
    for (int a = 0; a < n >> 1; ++a) { const uint32_t t = c[a]; c[a] = c[n - 1 - a]; c[n - 1 - a] = t; }
    return n;
}

__global__ void __launch_bounds__(CIG_BLOCK) k_cigar_band(MemOpt opt, RefView ref, CigArgs A) {
    extern __shared__ int2 ring_raw[];
    int2 (*ring)[CIG_BLOCK] = reinterpret_cast<int2 (*)[CIG_BLOCK]>(ring_raw);
    const unsigned n2 = *A.n2;
    const long T = (long)gridDim.x * blockDim.x;
    unsigned* z = reinterpret_cast<unsigned*>(A.z) + (blockIdx.x * (long)blockDim.x + threadIdx.x);
    unsigned long long cells = 0;
    for (unsigned t = blockIdx.x * blockDim.x + threadIdx.x; t < n2; t += gridDim.x * blockDim.x) {
        const long g = A.list2[t];
        const AlnReg ar = A.regs[g];
        const int r = A.reg_read[g];
        const int l_query = (int)(A.off[r + 1] - A.off[r]);
        const uint8_t* query = A.seq + A.off[r];
        const RegView v = make_view(ref, query, ar);
        uint32_t c[CIG_LOCAL];
        // the band-doubling loop of mem_reg2aln (bwamem.c:1111-1119)
        int w2 = cig_first_w(opt, ar), i = 0, score = 0, n = 0, last_sc = -(1 << 30);
        bool bail = false;
        do {
            w2 = w2 < opt.w << 2 ? w2 : opt.w << 2;
            n = band_gen_cigar(opt, v, w2, ring, z, T, A.z_rows, &score, c, &cells);
            if (n < 0) { bail = true; break; }
            if (score == last_sc || w2 == opt.w << 2) break;
            last_sc = score;
            w2 <<= 1;
        } while (++i < 3 && score < ar.truesc - opt.a);
        if (bail) { A.list3[atomicAdd(A.n3, 1u)] = (unsigned)g; continue; }
        cig_finish(A, ref, g, r, l_query, query, ar, v, c, n, score, cig_nm(v, c, n));
    }
    if (cells) atomicAdd(&A.cnt->glb_cells, cells);
}

}  // namespace arachne
