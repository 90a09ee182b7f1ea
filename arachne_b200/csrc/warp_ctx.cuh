// Warp-cooperative DP primitives for sm_100a used by mate rescue and the wide-band CIGAR tier: one warp per problem, the
// DP row striped over the lanes in registers (column j lives in lane j / CPL, slot j % CPL).
//
// ksw_global2 (bwa/ksw.c:504-606): E and F derive from M (the diagonal term), not from H, so the horizontal dependency
// F(i,j+1) = max(F(i,j) - e_ins, M(i,j) - oe_ins) is a max-plus prefix scan: CPL steps inside a lane plus a 5-step shuffle
// scan across lanes -- no lazy-F iteration; fixed band, -inf borders, the reference's direction byte per cell for the
// traceback.  All max-plus steps use the DPX instructions (__viaddmax_s32 / __vimax3_s32 -> VIADDMNMX, VIMNMX3).
// (ksw_extend2 does not live here: it runs one problem per lane in ext_stage.cuh.)
#pragma once
#include "extend.cuh"
#include "rescue.cuh"

namespace arachne {

constexpr int KSW_MINUS_INF = -0x40000000;
constexpr unsigned FULL = 0xffffffffu;

AR_D int sw_score(int a, int nb, int tb, int qb) {  // bwa_fill_scmat, bwa/bwa.c:109-118
    return (tb > 3 || qb > 3) ? -1 : (tb == qb ? a : nb);
}

// ksw_global2.  z != nullptr records the direction bytes (row stride n_col) and the caller backtracks.
template <int CPL>
__device__ __noinline__ int ksw_global_warp(const MemOpt& opt, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts,
                                            int w, uint8_t* z, int n_col) {
    const int lane = threadIdx.x & 31;
    const int e_del = opt.e_del, e_ins = opt.e_ins, oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    const int sa = opt.a, snb = -opt.b;
    const int j0 = lane * CPL;
    int hp[CPL], ep[CPL], qb[CPL];
#pragma unroll
    for (int c = 0; c < CPL; ++c) {
        int j = j0 + c;
        qb[c] = j < qlen ? q[(long)j * qs] : 4;
        hp[c] = j == 0 ? 0 : (j <= qlen && j <= w ? -(opt.o_ins + e_ins * j) : KSW_MINUS_INF);
        ep[c] = KSW_MINUS_INF;
    }
    for (int i = 0; i < tlen; ++i) {
        const int tb = t[(long)i * ts];
        int beg = i > w ? i - w : 0;
        int end = i + w + 1 < qlen ? i + w + 1 : qlen;
        int h1i = beg == 0 ? -(opt.o_del + e_del * (i + 1)) : KSW_MINUS_INF;
        int M[CPL], g[CPL + 1];
        g[0] = KSW_MINUS_INF;
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            int j = j0 + c;
            bool act = j >= beg && j < end;
            int m = hp[c] + sw_score(sa, snb, tb, qb[c]);
            M[c] = m;
            int tf = act ? m - oe_ins : KSW_MINUS_INF;
            g[c + 1] = __viaddmax_s32(g[c], -e_ins, tf);
        }
        int V = g[CPL];
        const int D = CPL * e_ins;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int up = __shfl_up_sync(FULL, V, d);
            if (lane >= d) V = __viaddmax_s32(up, -d * D, V);
        }
        int fin = __shfl_up_sync(FULL, V, 1);
        if (lane == 0) fin = KSW_MINUS_INF;
        int h[CPL];
#pragma unroll
        for (int c = 0; c < CPL; ++c) {
            int j = j0 + c;
            bool act = j >= beg && j < end;
            int f = __viaddmax_s32(fin, -c * e_ins, g[c]);   // F(i,j)
            int m = M[c], e = ep[c];
            unsigned d = m >= e ? 0u : 1u;
            int hh = m >= e ? m : e;
            d = hh >= f ? d : 2u;
            hh = hh >= f ? hh : f;
            h[c] = hh;
            int tt = m - oe_del;
            int e2 = e - e_del;
            d |= e2 > tt ? 1u << 2 : 0u;
            e2 = e2 > tt ? e2 : tt;
            tt = m - oe_ins;
            int f2 = f - e_ins;
            d |= f2 > tt ? 2u << 4 : 0u;
            if (act) {
                ep[c] = e2;
                if (z) z[(long)i * n_col + (j - beg)] = (uint8_t)d;
            }
        }
        int hup = __shfl_up_sync(FULL, h[CPL - 1], 1);
#pragma unroll
        for (int c = CPL - 1; c >= 0; --c) {
            int j = j0 + c;
            int left = c > 0 ? h[c - 1] : hup;
            if (end > beg) {
                if (j > beg && j <= end) hp[c] = left;
                else if (j == beg) hp[c] = h1i;
            } else if (j == end) hp[c] = h1i;
            if (j == end) ep[c] = KSW_MINUS_INF;
        }
    }
    int mine = 0;
#pragma unroll
    for (int c = 0; c < CPL; ++c) if (j0 + c == qlen) mine = hp[c];
    return __shfl_sync(FULL, mine, qlen / CPL);
}


// ---- local SW with the reference's striped-SIMD semantics (ksw_u8 / ksw_i16 / ksw_align2, bwa/ksw.c:111-365) ----
// Lane k of the first P lanes plays element k of the SSE2 vector (P = 16 for u8, 8 for i16); vector j, element k holds
// query position j + k*slen.  The recurrence, the saturation, the lazy-F rounds (with their early exit), the row-maximum
// list `b` and the query-end tie-break are the reference's, so scores, ends, score2 and starts are bit-identical.
constexpr int LSW_MAXSEG = 32;  // slen <= 32 for qlen <= 255 at P = 8

struct LswOut { int score, te, qe, score2, te2; };

template <class QF, class TF>
__device__ __noinline__ LswOut striped_pass_warp(const MemOpt& opt, int bytes, int qlen, QF qf, int tlen, TF tf, int xtra, uint64_t* blist,
                                                 int bcap, int* err) {
    const int lane = threadIdx.x & 31;
    const int P = bytes == 1 ? 16 : 8, HI = bytes == 1 ? 255 : 32767;
    const unsigned pmask = bytes == 1 ? 0xffffu : 0xffu;
    const bool on = lane < P;
    const int slen = (qlen + P - 1) / P;
    const int mb = opt.b > 1 ? opt.b : 1;
    const int shift = bytes == 1 ? (mb > -opt.a ? mb : -opt.a) : 0;
    int top = opt.a > -opt.b ? opt.a : -opt.b; top = top > -1 ? top : -1;
    const int e_del = opt.e_del, e_ins = opt.e_ins, oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    const int sa = opt.a, snb = -opt.b;
    short H[2][LSW_MAXSEG], E[LSW_MAXSEG], Hm[LSW_MAXSEG];
    unsigned char qb[LSW_MAXSEG];
    for (int j = 0; j < slen; ++j) {
        int pos = j + lane * slen;
        H[0][j] = H[1][j] = 0; E[j] = 0; Hm[j] = 0;
        qb[j] = (on && pos < qlen) ? (unsigned char)qf(pos) : (unsigned char)255;
    }
    const int minsc = (xtra & KSW_XSUBO) ? xtra & 0xffff : 0x10000, endsc = (xtra & KSW_XSTOP) ? xtra & 0xffff : 0x10000;
    int gmax = 0, te = -1, n_b = 0, cur = 0;
    auto sat = [HI](int v) { return v < 0 ? 0 : (v > HI ? HI : v); };
    for (int i = 0; i < tlen; ++i) {
        const int tb = tf(i);
        short* H0 = H[cur];
        short* H1 = H[cur ^ 1];
        int F = 0, mx = 0;
        int hv = __shfl_up_sync(FULL, (int)H0[slen - 1], 1);
        if (lane == 0) hv = 0;
        for (int j = 0; j < slen; ++j) {
            int q = qb[j];
            int sc = q == 255 ? 0 : sw_score(sa, snb, tb, q);
            int h;
            if (bytes == 1) h = sat(sat(hv + sc + shift) - shift);
            else { h = hv + sc; h = h > 32767 ? 32767 : (h < -32768 ? -32768 : h); }
            int e = E[j];
            h = __vimax3_s32(h, e, F);
            mx = mx > h ? mx : h;
            H1[j] = (short)h;
            int x = sat(h - oe_del);
            e = sat(e - e_del); e = e > x ? e : x;
            E[j] = (short)e;
            x = sat(h - oe_ins);
            F = sat(F - e_ins); F = F > x ? F : x;
            hv = H0[j];
        }
        bool done = false;
        for (int kk = 0; kk < 16 && !done; ++kk) {  // lazy F
            F = __shfl_up_sync(FULL, F, 1);
            if (lane == 0) F = 0;
            for (int j = 0; j < slen; ++j) {
                int h = H1[j];
                h = h > F ? h : F;
                H1[j] = (short)h;
                h = sat(h - oe_ins);
                F = sat(F - e_ins);
                unsigned any = __ballot_sync(FULL, on && F > h) & pmask;
                if (!any) { done = true; break; }
            }
        }
        int rowmax = __reduce_max_sync(FULL, on ? mx : 0);
        if (rowmax >= minsc) {
            if (lane == 0) {
                if (n_b == 0 || (int32_t)blist[n_b - 1] + 1 != i) {
                    if (n_b < bcap) blist[n_b] = (uint64_t)rowmax << 32 | (uint32_t)i;
                    else *err = 1;
                    ++n_b;
                } else if ((int)(blist[n_b - 1] >> 32) < rowmax) blist[n_b - 1] = (uint64_t)rowmax << 32 | (uint32_t)i;
            }
            n_b = __shfl_sync(FULL, n_b, 0);
        }
        bool brk = false;
        if (rowmax > gmax) {
            gmax = rowmax; te = i;
            for (int j = 0; j < slen; ++j) Hm[j] = H1[j];
            if ((bytes == 1 && gmax + shift >= 255) || gmax >= endsc) brk = true;
        }
        if (brk) break;
        cur ^= 1;
    }
    LswOut r;
    r.score = (bytes == 1 && gmax + shift >= 255) ? 255 : gmax;
    r.te = te; r.qe = -1; r.score2 = -1; r.te2 = -1;
    if (bytes == 2 || r.score != 255) {
        int best = -1, bpos = 0x7fffffff;
        if (on)
            for (int j = 0; j < slen; ++j) {
                int v = Hm[j], pos = j + lane * slen;
                if (v > best || (v == best && pos < bpos)) { best = v; bpos = pos; }
            }
        int gbest = __reduce_max_sync(FULL, best);
        int gpos = __reduce_min_sync(FULL, best == gbest ? bpos : 0x7fffffff);
        r.qe = gpos;
        __syncwarp();
        if (n_b > 0) {
            int k = (r.score + top - 1) / top, low = te - k, high = te + k, s2 = -1, t2 = -1;
            if (lane == 0)
                for (int i = 0; i < n_b && i < bcap; ++i) {
                    int e = (int32_t)blist[i];
                    if ((e < low || e > high) && (int)(blist[i] >> 32) > s2) { s2 = (int)(blist[i] >> 32); t2 = e; }
                }
            r.score2 = __shfl_sync(FULL, s2, 0);
            r.te2 = __shfl_sync(FULL, t2, 0);
        }
    }
    __syncwarp();
    return r;
}

#define AR_CPL_DISPATCH(ncols, CALL)                 \
    switch (((ncols) + 31) >> 5) {                   \
        case 0: case 1: { constexpr int CPL_ = 1; CALL; } break; \
        case 2: { constexpr int CPL_ = 2; CALL; } break;         \
        case 3: { constexpr int CPL_ = 3; CALL; } break;         \
        case 4: { constexpr int CPL_ = 4; CALL; } break;         \
        case 5: { constexpr int CPL_ = 5; CALL; } break;         \
        case 6: { constexpr int CPL_ = 6; CALL; } break;         \
        case 7: { constexpr int CPL_ = 7; CALL; } break;         \
        default: { constexpr int CPL_ = 8; CALL; } break;        \
    }

constexpr int AR_MAX_QLEN = 255;  // (qlen + 1) columns must fit 32 lanes x 8

struct WarpCtx {
    int lane;
    unsigned long long* cells;  // per-warp DP cell counter (roofline accounting), may be null
    uint8_t* z;                 // traceback scratch for ksw_global (n_col * tlen bytes), may be null
    long z_cap;
    uint64_t* blist = nullptr;  // row-maximum list scratch for ksw_local
    int bcap = 0;
    int lsw_err = 0;

    AR_D bool leader() const { return lane == 0; }
    AR_D void sync() const { __syncwarp(); }
    AR_D int bcast(int v) const { return __shfl_sync(FULL, v, 0); }

    // cooperative 2-bit -> byte unpack of [beg,end) of the forward-reverse reference into dst
    AR_D void fetch_ref(const RefView& ref, int64_t beg, int64_t end, uint8_t* dst) const {
        for (int64_t k = lane; k < end - beg; k += 32) dst[k] = ref_base(ref, beg + k);
        __syncwarp();
    }

    // ksw_align2 with the mate optionally reverse-complemented on the fly (bwamem_pair.c:131-135)
    AR_D void ksw_local(const MemOpt& opt, int qlen, const uint8_t* ms, int is_rev, int tlen, const uint8_t* t, int xtra, LocalAln* out) {
        const int bytes = (xtra & KSW_XBYTE) ? 1 : 2;
        auto qf = [=](int p) -> int { if (!is_rev) return ms[p]; int c = ms[qlen - 1 - p]; return c < 4 ? 3 - c : 4; };
        auto tf = [=](int i) -> int { return t[i]; };
        LswOut r = striped_pass_warp(opt, bytes, qlen, qf, tlen, tf, xtra, blist, bcap, &lsw_err);
        out->score = r.score; out->te = r.te; out->qe = r.qe; out->score2 = r.score2; out->te2 = r.te2; out->tb = -1; out->qb = -1;
        if ((xtra & KSW_XSTART) == 0 || ((xtra & KSW_XSUBO) && r.score < (xtra & 0xffff))) return;
        // start coordinates: same recurrence over the reversed prefixes; the target keeps its full length with only
        // its first te+1 symbols reversed (ksw.c:357-360)
        const int qe = r.qe, te = r.te;
        auto qf2 = [=](int p) -> int { return qf(qe - p); };
        auto tf2 = [=](int i) -> int { return i <= te ? t[te - i] : t[i]; };
        LswOut rr = striped_pass_warp(opt, bytes, qe + 1, qf2, tlen, tf2, KSW_XSTOP | r.score, blist, bcap, &lsw_err);
        if (r.score == rr.score) { out->tb = r.te - rr.te; out->qb = r.qe - rr.qe; }
    }

    // score only when cigar == nullptr; otherwise also backtracks into cigar (len<<4|op, already in forward order)
    AR_D int ksw_global(const MemOpt& opt, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts, int w, uint32_t* cigar,
                        int cigar_cap, int* n_cigar) const {
        int n_col = qlen < 2 * w + 1 ? qlen : 2 * w + 1;
        uint8_t* zz = cigar ? z : nullptr;
        int score = 0;
        AR_CPL_DISPATCH(qlen + 1, score = ksw_global_warp<CPL_>(opt, qlen, q, qs, tlen, t, ts, w, zz, n_col));
        if (cigar) {
            __syncwarp();
            int n = 0;
            if (lane == 0) {  // backtrack (bwa/ksw.c:588-603); ops are produced last-to-first
                int which = 0, i = tlen - 1, k = (i + w + 1 < qlen ? i + w + 1 : qlen) - 1;
                int last_op = -1;
                auto push = [&](int op, int len) {
                    if (op != last_op) { if (n < cigar_cap) cigar[n] = (uint32_t)len << 4 | op; ++n; last_op = op; }
                    else if (n <= cigar_cap) cigar[n - 1] += (uint32_t)len << 4;
                };
                while (i >= 0 && k >= 0) {
                    which = zz[(long)i * n_col + (k - (i > w ? i - w : 0))] >> (which << 1) & 3;
                    if (which == 0) { push(0, 1); --i; --k; }
                    else if (which == 1) { push(2, 1); --i; }
                    else { push(1, 1); --k; }
                }
                if (i >= 0) push(2, i + 1);
                if (k >= 0) push(1, k + 1);
                if (n <= cigar_cap)
                    for (int x = 0; x < n >> 1; ++x) { uint32_t tmp = cigar[x]; cigar[x] = cigar[n - 1 - x]; cigar[n - 1 - x] = tmp; }
            }
            n = __shfl_sync(FULL, n, 0);
            __syncwarp();
            if (n_cigar) *n_cigar = n;
        }
        return score;
    }
};

}  // namespace arachne
