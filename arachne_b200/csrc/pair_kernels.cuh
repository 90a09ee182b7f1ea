// Warp-per-item kernels after extension: mate rescue (k_rescue), the gather of final region lists, and the wide-band CIGAR tier
// (k_cigar).  Included by pipeline.cu; tests/hostsim compiles the same text for the host.
#pragma once
#include "counters.cuh"
#include "warp_ctx.cuh"
#include "cigar.cuh"
namespace arachne {

constexpr int EXT_WARPS = 4;

// ------------------------------------------------------------------------------------------------ k_rescue
// Warp per pair.  Most pairs already hold a consistent hit for every anchor, in which case mem_matesw returns before
// touching anything (bwamem_pair.c:118-124) and the lists stay where k_extend left them.  Only pairs that really run
// a rescue SW get private copies with room to grow (each mem_matesw call adds at most one region).
constexpr unsigned long long RESC_FLAG = 1ull << 63;

__global__ void __launch_bounds__(EXT_WARPS * 32) k_rescue(MemOpt opt, RefView ref, int score_delta, const uint8_t* __restrict__ seq,
                                                           const int64_t* __restrict__ off, int n_pairs, const unsigned long long* __restrict__ sa_off,
                                                           AlnReg* regs, int* n_regs, unsigned long long* reg_loc, AlnReg* resc,
                                                           unsigned long long resc_cap, unsigned long long* resc_used, uint8_t* win, int win_cap,
                                                           uint64_t* blists, int bcap, unsigned int* next, Counters* cnt) {
    const int lane = threadIdx.x & 31;
    const int gw = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    WarpCtx ctx;
    ctx.lane = lane; ctx.cells = nullptr; ctx.z = nullptr; ctx.z_cap = 0; ctx.blist = blists + (size_t)gw * bcap; ctx.bcap = bcap;
    uint8_t* rbuf = win + (size_t)gw * win_cap;
    PeStat pes[4];
    gobwa_pestat(pes);
    for (;;) {
        int p = 0;
        if (lane == 0) p = (int)atomicAdd(next, 1u);
        p = __shfl_sync(FULL, p, 0);
        if (p >= n_pairs) break;
        const int ra = 2 * p, rb = 2 * p + 1;
        const int l1 = (int)(off[ra + 1] - off[ra]), l2 = (int)(off[rb + 1] - off[rb]);
        AlnReg* r1 = regs + sa_off[ra];
        AlnReg* r2 = regs + sa_off[rb];
        int n1 = n_regs[ra], n2 = n_regs[rb];
        if (lane == 0) { reg_loc[ra] = sa_off[ra]; reg_loc[rb] = sa_off[rb]; }
        // would any mem_matesw call get past its early return?  (lanes split the anchors)
        int best1 = 0, best2 = 0;
        for (int i = 0; i < n1; ++i) best1 = best1 > r1[i].score ? best1 : r1[i].score;
        for (int i = 0; i < n2; ++i) best2 = best2 > r2[i].score ? best2 : r2[i].score;
        bool need = false;
        if (l1 > 0)
            for (int i = 0, num = 0; i < n2 && num < 50 && !need; ++i)
                if (r2[i].score >= best2 - score_delta) { ++num; need = rescue_open_classes(ref.l_pac, pes, r2[i], r1, n1) != 0; }
        if (l2 > 0)
            for (int i = 0, num = 0; i < n1 && num < 50 && !need; ++i)
                if (r1[i].score >= best1 - score_delta) { ++num; need = rescue_open_classes(ref.l_pac, pes, r1[i], r2, n2) != 0; }
        if (!need) continue;
        const int cap1 = n1 + 50, cap2 = n2 + 50;
        unsigned long long pos = 0;
        if (lane == 0) pos = atomicAdd(resc_used, (unsigned long long)(cap1 + cap2));
        pos = __shfl_sync(FULL, pos, 0);
        if (pos + cap1 + cap2 > resc_cap) { if (lane == 0) raise(cnt, -100, p); continue; }
        AlnReg* c1 = resc + pos;
        AlnReg* c2 = c1 + cap1;
        for (int i = lane; i < n1; i += 32) c1[i] = r1[i];
        for (int i = lane; i < n2; i += 32) c2[i] = r2[i];
        __syncwarp();
        ctx.lsw_err = 0;
        int rc = pair_rescue(ctx, opt, ref, score_delta, l1, seq + off[ra], l2, seq + off[rb], c1, &n1, cap1, c2, &n2, cap2, rbuf, win_cap);
        if (rc < 0 || ctx.lsw_err) { if (lane == 0) raise(cnt, ARACHNE_ERR_OVERFLOW, p); continue; }
        if (lane == 0) {
            n_regs[ra] = n1; n_regs[rb] = n2;
            reg_loc[ra] = RESC_FLAG | pos; reg_loc[rb] = RESC_FLAG | (pos + cap1);
        }
    }
}

__global__ void k_gather_regs2(int n_reads, const unsigned long long* __restrict__ reg_loc, const unsigned long long* __restrict__ reg_off,
                               const AlnReg* __restrict__ regs, const AlnReg* __restrict__ resc, AlnReg* out, int* reg_read) {
    const int lane = threadIdx.x & 31;
    for (int r = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < n_reads; r += (gridDim.x * blockDim.x) >> 5) {
        unsigned long long loc = reg_loc[r];
        const AlnReg* sp = (loc & RESC_FLAG) ? resc + (loc & ~RESC_FLAG) : regs + loc;
        const uint32_t* src = reinterpret_cast<const uint32_t*>(sp);
        uint32_t* dst = reinterpret_cast<uint32_t*>(out + reg_off[r]);
        int nr = (int)(reg_off[r + 1] - reg_off[r]);
        int nw = nr * (int)(sizeof(AlnReg) / 4);
        for (int k = lane; k < nw; k += 32) dst[k] = src[k];
        for (int k = lane; k < nr; k += 32) reg_read[reg_off[r] + k] = r;
    }
}

__global__ void k_init_loc(int n_reads, const unsigned long long* __restrict__ sa_off, unsigned long long* reg_loc) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += gridDim.x * blockDim.x) reg_loc[r] = sa_off[r];
}

// ------------------------------------------------------------------------------------------------ k_cigar
__global__ void __launch_bounds__(EXT_WARPS * 32) k_cigar(MemOpt opt, RefView ref, const uint8_t* __restrict__ seq, const int64_t* __restrict__ off,
                                                          long n_regs, const AlnReg* __restrict__ regs, const int* __restrict__ reg_read, AlnRec* alns,
                                                          uint32_t* cig_pool, unsigned long long cig_cap, int32_t* mm_pool, unsigned long long mm_cap,
                                                          unsigned long long* used /*[0]=cigar [1]=mm*/, uint8_t* win, int win_cap, uint8_t* zbuf,
                                                          long z_cap, uint32_t* cig_scr, int cig_scr_cap, int32_t* mm_scr, int mm_scr_cap,
                                                          unsigned int* next, Counters* cnt, const unsigned* list, const unsigned* n_list) {
    if (list) n_regs = (long)*n_list;  // wide-band tier: only the regions the thread-per-region tiers handed over
    const int lane = threadIdx.x & 31;
    const long gw = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
    WarpCtx ctx;
    ctx.lane = lane; ctx.cells = nullptr; ctx.z = zbuf + gw * z_cap; ctx.z_cap = z_cap;
    uint8_t* rbuf = win + gw * (long)win_cap;
    uint32_t* cig = cig_scr + gw * (long)cig_scr_cap;
    int32_t* mm = mm_scr + gw * (long)mm_scr_cap;
    for (;;) {
        long g = 0;
        if (lane == 0) g = (long)atomicAdd(next, 1u);
        g = __shfl_sync(FULL, g, 0);
        if (g >= n_regs) break;
        if (list) g = (long)list[g];
        const int r = reg_read[g];
        const int len = (int)(off[r + 1] - off[r]);
        const AlnReg ar = regs[g];
        AlnRec a;
        // traceback matrix must fit: n_col * tlen with n_col <= qlen
        if ((long)(ar.qe - ar.qb) * (ar.re - ar.rb) > z_cap || 2 * len > mm_scr_cap) { if (lane == 0) raise(cnt, ARACHNE_ERR_OVERFLOW, r); continue; }
        int rc = reg2aln(ctx, opt, ref, len, seq + off[r], ar, &a, cig, cig_scr_cap, mm, rbuf, win_cap);
        if (rc < 0) { if (lane == 0) raise(cnt, ARACHNE_ERR_OVERFLOW, r); continue; }
        unsigned long long co = 0, mo = 0;
        if (lane == 0) {
            co = atomicAdd(&used[0], (unsigned long long)a.n_cigar);
            mo = atomicAdd(&used[1], (unsigned long long)(2 * a.n_mm));
        }
        co = __shfl_sync(FULL, co, 0); mo = __shfl_sync(FULL, mo, 0);
        if (co + a.n_cigar > cig_cap || mo + 2 * a.n_mm > mm_cap) { if (lane == 0) raise(cnt, -101, r); continue; }
        for (int k = lane; k < a.n_cigar; k += 32) cig_pool[co + k] = cig[k];
        for (int k = lane; k < a.n_mm; k += 32) { mm_pool[mo + k] = mm[k]; mm_pool[mo + a.n_mm + k] = mm[len + k]; }
        a.cigar_off = (int64_t)co; a.mm_off = (int64_t)mo;
        if (lane == 0) alns[g] = a;
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------ k_ksw_global (stage entry point arachne_ksw_global_batch)
__global__ void __launch_bounds__(EXT_WARPS * 32) k_ksw_global(MemOpt opt, long n, const uint8_t* __restrict__ q, const int64_t* __restrict__ q_off,
                                                               const uint8_t* __restrict__ t, const int64_t* __restrict__ t_off,
                                                               const int* __restrict__ w, int* scores, uint8_t* zbuf, long z_cap,
                                                               uint32_t* cig, int cig_cap, int* n_cig) {
    const int lane = threadIdx.x & 31;
    const long gw = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
    WarpCtx ctx; ctx.lane = lane; ctx.cells = nullptr; ctx.z = zbuf ? zbuf + gw * z_cap : nullptr; ctx.z_cap = z_cap;
    for (long i = gw; i < n; i += ((long)gridDim.x * blockDim.x) >> 5) {
        int qlen = (int)(q_off[i + 1] - q_off[i]), tlen = (int)(t_off[i + 1] - t_off[i]);
        int nc = 0;
        int sc = ctx.ksw_global(opt, qlen, q + q_off[i], 1, tlen, t + t_off[i], 1, w[i], cig ? cig + i * (long)cig_cap : nullptr, cig_cap, &nc);
        if (lane == 0) { scores[i] = sc; if (n_cig) n_cig[i] = nc; }
    }
}

}  // namespace arachne
