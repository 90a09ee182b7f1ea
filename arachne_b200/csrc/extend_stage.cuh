// k_ksw_extend_lane: ksw_extend2 (bwa/ksw.c:380-479) on caller-supplied problems, one problem per LANE -- the DP core of
// k_extend2 (extend_kernel.cuh) without the mem_chain2aln driver around it, so that the config-5 microbench
// (arachne_ksw_extend_batch) measures the kernel design the pipeline really runs.  Same structure: a per-lane automaton,
// every trip of the outer loop is "each live lane computes up to EXT_CHUNK cells of its current row" as straight-line DPX
// code over packed shared-memory columns (h 13 bits | e 13 bits | query base 3 bits); row ends and the adaptive band are
// divergent bookkeeping between chunks.
#pragma once
#include "extend_kernel.cuh"

namespace arachne {

struct KswStageArgs {
    long n;
    const uint8_t* q; const int64_t* q_off; const uint8_t* t; const int64_t* t_off;
    const int* h0; const int* w; int end_bonus, zdrop;
    int* out;                       // n x 6: score qle tle gtle gscore max_off
    unsigned long long* cells_out;
    unsigned* next;
    int maxq;                       // columns available per lane: maxq + 1 + EXT_CHUNK
};

__global__ void __launch_bounds__(EXT_BLOCK) k_ksw_extend_lane(MemOpt opt, KswStageArgs A) {
    extern __shared__ uint32_t cols[];
    uint32_t* col = cols + threadIdx.x;
#define COL(j) col[(j) * EXT_BLOCK]
    const int e_del = opt.e_del, e_ins = opt.e_ins, oe_del = opt.o_del + opt.e_del, oe_ins = opt.o_ins + opt.e_ins;
    const int sa = opt.a, snb = -opt.b;
    enum { S_FETCH, S_INIT, S_ROWSTART, S_CELLS, S_ROWEND, S_OUT, S_DONE };
    int st = S_FETCH;
    long p = 0;
    const uint8_t* tq = A.t;
    int qlen = 0, tlen = 0, w = 0, h0 = 0, i = 0, j = 0, beg = 0, end = 0, h1 = 0, f = 0, m = 0, mj = -1;
    int mx = 0, max_i = -1, max_j = -1, max_ie = -1, gscore = -1, max_off = 0, tbv = 0;
    unsigned long long cells = 0;
    int firstnz = 0x7fffffff, lastnz = -1;
    const uint8_t* iq = A.q;
    for (;;) {
        while (st != S_DONE && st != S_INIT && !(st == S_CELLS && j < end)) {
            switch (st) {
            case S_FETCH: {
                const unsigned g = atomicAdd(A.next, 1u);
                if ((long)g >= A.n) { st = S_DONE; break; }
                p = (long)g;
                const uint8_t* q = A.q + A.q_off[p];
                qlen = (int)(A.q_off[p + 1] - A.q_off[p]); tlen = (int)(A.t_off[p + 1] - A.t_off[p]); tq = A.t + A.t_off[p];
                h0 = A.h0[p]; w = A.w[p];
                if (qlen > A.maxq) { int* o = A.out + 6 * p; o[0] = o[1] = o[2] = o[3] = o[4] = o[5] = -0x7fffffff; break; }  // caller sized the columns from the longest query
                iq = q;  // the first row (ksw.c:395-397) is filled by the whole warp below, as in k_extend2
                int top = sa > snb ? sa : snb; top = top > -1 ? top : -1;  // band limit (:402-407), integer form of (int)((double)x / e + 1.)
                int lim = (qlen * top + A.end_bonus - opt.o_ins + e_ins) / e_ins;
                lim = lim > 1 ? lim : 1; w = w < lim ? w : lim;
                lim = (qlen * top + A.end_bonus - opt.o_del + e_del) / e_del;
                lim = lim > 1 ? lim : 1; w = w < lim ? w : lim;
                mx = h0; max_i = max_j = max_ie = -1; gscore = -1; max_off = 0; beg = 0; end = qlen; i = 0;
                st = S_INIT;
            } break;
            case S_ROWSTART: {
                if (i >= tlen) { st = S_OUT; break; }
                tbv = __ldg(tq + i);
                if (beg < i - w) beg = i - w;
                if (end > i + w + 1) end = i + w + 1;
                if (end > qlen) end = qlen;
                h1 = 0;
                if (beg == 0) { h1 = h0 - (opt.o_del + e_del * (i + 1)); if (h1 < 0) h1 = 0; }
                f = 0; m = 0; mj = -1; j = beg;
                firstnz = 0x7fffffff; lastnz = -1;
                cells += end > beg ? end - beg : 0;
                st = S_CELLS;
            } break;
            case S_CELLS:
            case S_ROWEND: {  // ksw.c:449-469
                COL(end) = (COL(end) & (7u << 26)) | (uint32_t)h1;
                if (j == qlen) { max_ie = gscore > h1 ? max_ie : i; gscore = gscore > h1 ? gscore : h1; }
                if (m == 0) { st = S_OUT; break; }
                if (m > mx) {
                    mx = m; max_i = i; max_j = mj;
                    int d = mj - i; d = d < 0 ? -d : d;
                    max_off = max_off > d ? max_off : d;
                } else if (A.zdrop > 0) {
                    bool drop;
                    if (i - max_i > mj - max_j) drop = mx - m - ((i - max_i) - (mj - max_j)) * e_del > A.zdrop;
                    else drop = mx - m - ((mj - max_j) - (i - max_i)) * e_ins > A.zdrop;
                    if (drop) { st = S_OUT; break; }
                }
                beg = firstnz < end ? firstnz : end;  // band trimming (ksw.c:465-468) from the columns the cell loop recorded
                const int x = h1 ? end : (lastnz >= beg ? lastnz : beg - 1);
                end = x + 2 < qlen ? x + 2 : qlen;
                ++i;
                st = S_ROWSTART;
            } break;
            case S_OUT: {
                int* o = A.out + 6 * p;
                o[0] = mx; o[1] = max_j + 1; o[2] = max_i + 1; o[3] = max_ie + 1; o[4] = gscore; o[5] = max_off;
                st = S_FETCH;
            } break;
            default: st = S_DONE; break;
            }
        }
        __syncwarp();
        {   // first rows of the lanes that asked for one, 32 columns per trip (see k_extend2)
            unsigned want = __ballot_sync(0xffffffffu, st == S_INIT);
            const int lane = threadIdx.x & 31;
            while (want) {
                const int src = __ffs(want) - 1;
                want &= want - 1;
                const uint8_t* bq = reinterpret_cast<const uint8_t*>(__shfl_sync(0xffffffffu, (unsigned long long)iq, src));
                const int bqlen = __shfl_sync(0xffffffffu, qlen, src), bh0 = __shfl_sync(0xffffffffu, h0, src);
                uint32_t* bcol = cols + (threadIdx.x & ~31) + src;
                for (int x = lane; x <= bqlen; x += 32) {
                    int hv = bh0 - opt.o_ins - x * e_ins;
                    hv = x == 0 ? bh0 : (hv > 0 ? hv : 0);
                    bcol[x * EXT_BLOCK] = (uint32_t)hv | (x < bqlen ? (uint32_t)__ldg(bq + x) : 0u) << 26;
                }
            }
            __syncwarp();
            if (st == S_INIT) st = S_ROWSTART;
        }
        {   // up to EXT_CHUNK cells of the current row (ksw.c:424-448) in rolled groups of EXT_GROUP, as in k_extend2
            const int nc = st == S_CELLS ? (end - j < EXT_CHUNK ? end - j : EXT_CHUNK) : 0;
            const int nc_warp = __reduce_max_sync(0xffffffffu, nc);
            int mkey = m << 9 | (mj & 0x1ff);
            unsigned nzm = 0;
            uint32_t* pc = &COL(j);
#pragma unroll 1
            for (int g = 0; g < nc_warp; g += EXT_GROUP, pc += EXT_GROUP * EXT_BLOCK) {
                if (g >= nc) continue;
                uint32_t cw[EXT_GROUP];
#pragma unroll
                for (int u = 0; u < EXT_GROUP; ++u) cw[u] = pc[u * EXT_BLOCK];
#pragma unroll
                for (int u = 0; u < EXT_GROUP; ++u) {
                    const bool on = g + u < nc;
                    const int hd = (int)(cw[u] & COL_HMASK), e = (int)((cw[u] >> 13) & COL_HMASK), qb = (int)(cw[u] >> 26);
                    const int sc = (qb | tbv) > 3 ? -1 : (qb == tbv ? sa : snb);  // a caller's target may hold N (the packed reference never does)
                    const int M = hd ? hd + sc : 0;
                    const int h = __vimax3_s32(M, e, f);
                    const int e2 = __viaddmax_s32_relu(M, -oe_del, e - e_del);
                    const int f2 = __viaddmax_s32_relu(M, -oe_ins, f - e_ins);
                    const uint32_t he = (uint32_t)h1 | (uint32_t)e2 << 13;
                    nzm |= (on && he) ? 1u << (g + u) : 0u;
                    if (on) pc[u * EXT_BLOCK] = he | (cw[u] & (7u << 26));
                    const int key = on ? (h << 9 | (j + g + u)) : 0;
                    mkey = mkey > key ? mkey : key;
                    f = on ? f2 : f;
                    h1 = on ? h : h1;
                }
            }
            if (nc) {
                if (nzm) {
                    if (firstnz == 0x7fffffff) firstnz = j + __ffs(nzm) - 1;
                    lastnz = j + 31 - __clz(nzm);
                }
                m = mkey >> 9;
                if (m > 0) mj = mkey & 0x1ff;
                j += nc;
            }
        }
        if (__all_sync(0xffffffffu, st == S_DONE)) break;
    }
#undef COL
    if (cells) atomicAdd(A.cells_out, cells);
}

}  // namespace arachne
