// TEST-ONLY: the batch's stage kernels from SA intervals to final region lists, compiled by g++ from the shipped headers and launched in the
// order csrc/batch.inl launches them:
//   k_expand -> k_sa -> k_chain -> rounds of { k_ext_drive, k_ext_binbase, k_ext_scatter, k_ext_dp<true, *>, k_ext_roll } -> k_ext_finish
//   -> k_init_loc -> k_rescue -> k_gather_regs2                       (csrc/stage_kernels.cuh, ext_stage.cuh, pair_kernels.cuh)
// Every kernel runs as one block of fibers (cuda_warp_shim.h): thread-per-item kernels as a single warp whose grid-stride loop covers the batch,
// k_sa / k_ext_dp / k_rescue as the warp they cooperate in, k_ext_binbase as its full block of 256.  What stays host code here is what is host
// code in batch.inl: the two prefix sums (cub there), buffer sizing, the round loop's stop test.  k_ext_dp runs in the pipeline's instantiation
// (targets fetched from the 2-bit packed reference), which hostsim_extdp.cpp's byte-target entry does not reach.
#include <algorithm>
#include <vector>
#define __shared__ thread_local   /* one host thread runs all fibers of a block: block-scope __shared__ arrays get one copy per block, and
                                     `extern __shared__ T a[]` names the array defined below */
#include "cuda_warp_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/common.cuh"
#ifndef AR_D
#define AR_D inline
#endif
#define xcols xcols_stages  /* hostsim_extdp.cpp defines its own arachne::xcols */
namespace arachne { thread_local uint32_t xcols[(ARACHNE_MAX_READ_LEN + 8) * 128]; }  // k_ext_dp's dynamic shared memory: [ncols][XDP_BLOCK]
#include "../../arachne_b200/csrc/stage_kernels.cuh"
#include "../../arachne_b200/csrc/ext_stage.cuh"
#include "../../arachne_b200/csrc/pair_kernels.cuh"
#include "../../arachne_b200/csrc/xs_kernels.cuh"

using namespace arachne;
extern "C" const void* hs_ref(void* h);
extern "C" const void* hs_fm(void* h);

namespace {
template <class F> void warp(unsigned first_tid, unsigned block_dim, F&& f) {
    blockIdx.x = 0; blockDim.x = block_dim; gridDim.x = 1;
    warpsim::run_warp(first_tid, f);
}
// a warp-cooperative kernel either as ONE of its block's warps (slot = warp_id & 3; for a 256-thread block, one of the first four) or, with
// bit 3 of warp_id set, as the WHOLE block: all its warps as fibers, interleaving at every collective and sharing the queue counters
template <class F> void coop(int warp_id, unsigned block_dim, F&& f) {
    blockIdx.x = 0; blockDim.x = block_dim; gridDim.x = 1;
    if (warp_id & 8) warpsim::run_block((int)block_dim, f, 1 << 20);
    else warpsim::run_warp(32u * (unsigned)(warp_id & 3), f);
}
}  // namespace

extern "C" {

void hs_warp_order(int order) { warpsim::lane_order() = order; }  // one setting for the whole library (an inline function's static)

// intv: 4 u64 per SA interval {x0, x1, x2, info}, read r owns intv_off[r] .. intv_off[r + 1] (mem_collect_intv's output, sorted).
// stages: 1 = regions of mem_align1_core, 3 = + mate rescue (reads 2p / 2p + 1 are a pair).  warp_id: bits 0-1 which of a block's four warps runs the
// warp kernels, bit 2 forces k_ext_dp's 13-bit column codec (the pipeline takes it when a * max read length exceeds a byte), bit 3 runs
// the warp kernels as whole blocks (see coop above).  regs_out / reg_off (n_reads + 1): the final lists.  info: {rounds, total_sa, dp problems, pairs rescued, err, err_read}.
int hs_stages(void* h, const arachne_mem_opt_t* opt_, int64_t n_reads_, const uint8_t* seq, const int64_t* off, const uint64_t* intv, const int64_t* intv_off_in,
              int stages, int score_delta, int warp_id, arachne_alnreg_t* regs_out, int64_t regs_cap, int64_t* reg_off_out, int64_t* info) {
    static_assert(XDP_BLOCK == 128, "xcols is sized for 128 lanes");
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    const FMIndex fm = *static_cast<const FMIndex*>(hs_fm(h));
    const int n_reads = (int)n_reads_;
    int maxlen = 1, cap = 1;
    for (int r = 0; r < n_reads; ++r) {
        maxlen = std::max<int>(maxlen, (int)(off[r + 1] - off[r]));
        cap = std::max<int>(cap, (int)(intv_off_in[r + 1] - intv_off_in[r]));
    }
    if (maxlen > ARACHNE_MAX_READ_LEN || (long)opt.a * maxlen > 8191) return ARACHNE_ERR_ARG;
    // the interval arena as k_sort_intv leaves it: `cap` slots per read; seeds per read (bwamem.c:278-283) and their prefix sum
    std::vector<Intv> arena((size_t)n_reads * cap + 1);
    std::vector<int> intv_n((size_t)n_reads + 1);
    std::vector<unsigned long long> sa_off((size_t)n_reads + 2), intv_off((size_t)n_reads + 1);
    unsigned long long total_sa = 0;
    for (int r = 0; r < n_reads; ++r) {
        const int n = (int)(intv_off_in[r + 1] - intv_off_in[r]);
        intv_n[(size_t)r] = n;
        memcpy(arena.data() + (size_t)r * cap, intv + 4 * intv_off_in[r], (size_t)n * sizeof(Intv));
        sa_off[(size_t)r] = total_sa;
        for (int i = 0; i < n; ++i) {
            const uint64_t x2 = intv[4 * (intv_off_in[r] + i) + 2];
            total_sa += std::min<uint64_t>(x2, (uint64_t)opt.max_occ);
        }
    }
    sa_off[(size_t)n_reads] = sa_off[(size_t)n_reads + 1] = total_sa;
    const size_t tsa = (size_t)total_sa + 8;
    std::vector<uint64_t> rows(tsa), srt(tsa);
    std::vector<Seed> seeds(tsa), kseeds(tsa);
    std::vector<Chain> chains(tsa);
    std::vector<BNode> nodes(tsa / 4 + 4 * (size_t)n_reads + 8);
    std::vector<int32_t> order(tsa), kept(tsa);
    std::vector<int> n_chain((size_t)n_reads + 1), n_kseed((size_t)n_reads + 1), n_regs((size_t)n_reads + 1);
    std::vector<AlnReg> regs(tsa);
    Counters cnt; memset(&cnt, 0, sizeof(cnt));
    unsigned long long misc[8] = {0};

    warp(0, 32, [&] { k_expand(opt, n_reads, arena.data(), intv_off.data(), cap, intv_n.data(), sa_off.data(), rows.data()); });
    coop(warp_id, 256, [&] { k_sa(fm, total_sa, rows.data(), &misc[0], &cnt); });
    ChainBufs cb;
    cb.seeds = seeds.data(); cb.kseeds = kseeds.data(); cb.chains = chains.data(); cb.nodes = nodes.data();
    cb.order = order.data(); cb.kept = kept.data(); cb.n_chain = n_chain.data(); cb.n_kseed = n_kseed.data();
    warp(0, 32, [&] { k_chain(opt, ref, off, n_reads, arena.data(), intv_off.data(), intv_n.data(), sa_off.data(), rows.data(), cb, &cnt); });

    // ---- extension rounds ----
    const int win_cap = 2 * maxlen + 1024 + 4 * opt.w;
    const int drv_T = 32;
    std::vector<int> hscr((size_t)drv_T * 2 * (maxlen + 2));
    std::vector<uint8_t> rscr((size_t)drv_T * win_cap);
    std::vector<ExtState> state((size_t)n_reads + 1);
    std::vector<ExtProb> prob((size_t)n_reads + 1), queue(2 * (size_t)n_reads + 2);
    std::vector<ExtRes> res((size_t)n_reads + 1);
    std::vector<unsigned> rank((size_t)n_reads + 1);
    ExtCtl ctl; memset(&ctl, 0, sizeof(ctl));
    ExtBufs XB;
    XB.seq = seq; XB.off = off; XB.n_reads = n_reads; XB.sa_off = sa_off.data(); XB.b = cb;
    XB.srt = srt.data(); XB.regs = regs.data(); XB.n_regs = n_regs.data();
    XB.hscr = hscr.data(); XB.rscr = rscr.data(); XB.rcap = win_cap; XB.maxlen = maxlen;
    XB.state = state.data(); XB.prob = prob.data(); XB.res = res.data(); XB.rank = rank.data();
    XB.queue_cur = queue.data(); XB.queue_next = queue.data() + n_reads + 1;
    XB.ctl = &ctl; XB.cnt = &cnt;
    ExtDpArgs DA;
    DA.seq = seq; DA.tbytes = nullptr; DA.res = XB.res; DA.n_items = &ctl.n_next; DA.head = &ctl.head;
    DA.ncols = std::max(1, maxlen - opt.min_seed_len) + 2;
    DA.zdrop = opt.zdrop; DA.one = 1; DA.cnt = &cnt;
    const bool narrow = (long)opt.a * maxlen <= 255 && !(warp_id & 4);
    memset(arachne::xcols, 0xcd, sizeof(arachne::xcols));
    int round = 0; long n_prob = 0;
    for (;;) {
        warp(0, 32, [&] { k_ext_drive(opt, ref, XB, round); });
        blockIdx.x = 0; blockDim.x = 256; gridDim.x = 1;
        warpsim::run_block(256, [&] { k_ext_binbase(&ctl); });
        warp(0, 32, [&] { k_ext_scatter(XB, round); });
        DA.queue = XB.queue_next;
        n_prob += ctl.n_next;
        if (narrow) coop(warp_id, XDP_BLOCK, [&] { k_ext_dp<true, false>(opt, ref, DA); });
        else coop(warp_id, XDP_BLOCK, [&] { k_ext_dp<true, true>(opt, ref, DA); });
        warp(0, 1, [&] { if (threadIdx.x == 0) k_ext_roll(&ctl); });
        std::swap(XB.queue_cur, XB.queue_next);
        ++round;
        if (ctl.n_next == 0) break;
        if (round > 4096) return -20;
    }
    warp(0, 32, [&] { k_ext_finish(opt, ref, XB); });

    // ---- mate rescue ----
    std::vector<unsigned long long> reg_loc((size_t)n_reads + 1), reg_off((size_t)n_reads + 2);
    warp(0, 32, [&] { k_init_loc(n_reads, sa_off.data(), reg_loc.data()); });
    const unsigned long long resc_cap = (unsigned long long)n_reads * 64 + 4096;
    std::vector<AlnReg> resc((size_t)resc_cap);
    long n_rescued = 0;
    if (stages & 2) {
        const int bcap = win_cap, n_pairs = n_reads / 2;
        std::vector<uint8_t> win((size_t)EXT_WARPS * win_cap);
        std::vector<uint64_t> blists((size_t)EXT_WARPS * bcap);
        unsigned long long resc_used = 0; unsigned next = 0;
        coop(warp_id, EXT_WARPS * 32, [&] {
            k_rescue(opt, ref, score_delta, seq, off, n_pairs, sa_off.data(), regs.data(), n_regs.data(), reg_loc.data(), resc.data(), resc_cap, &resc_used,
                     win.data(), win_cap, blists.data(), bcap, &next, &cnt);
        });
        for (int r = 0; r < n_reads; r += 2) n_rescued += (reg_loc[(size_t)r] & RESC_FLAG) != 0;
    }
    unsigned long long total = 0;
    for (int r = 0; r < n_reads; ++r) { reg_off[(size_t)r] = total; total += (unsigned long long)n_regs[(size_t)r]; }
    reg_off[(size_t)n_reads] = total;
    if ((int64_t)total > regs_cap) return ARACHNE_ERR_OVERFLOW;
    std::vector<int> reg_read((size_t)total + 1);
    warp(0, 32, [&] { k_gather_regs2(n_reads, reg_loc.data(), reg_off.data(), regs.data(), resc.data(), reinterpret_cast<AlnReg*>(regs_out), reg_read.data()); });
    for (int r = 0; r <= n_reads; ++r) reg_off_out[r] = (int64_t)reg_off[(size_t)r];
    for (int r = 0; r < n_reads; ++r)
        for (unsigned long long g = reg_off[(size_t)r]; g < reg_off[(size_t)r + 1]; ++g)
            if (reg_read[(size_t)g] != r) return -21;
    if (info) { info[0] = round; info[1] = (int64_t)total_sa; info[2] = n_prob; info[3] = n_rescued; info[4] = cnt.err; info[5] = cnt.err_read; }
    return cnt.err;
}

// The wide-band CIGAR tier as shipped: k_cigar (csrc/pair_kernels.cuh) over the regions in `list` (what k_cigar_fast / k_cigar_band handed over;
// list == nullptr: every region), one warp of fibers.  alns: n_regs records, filled at the listed indices; pools as in batch.inl.
int hs_k_cigar(void* h, const arachne_mem_opt_t* opt_, int64_t n_reads, const uint8_t* seq, const int64_t* off, int64_t n_regs, const arachne_alnreg_t* regs,
               const int32_t* reg_read, const uint32_t* list, uint32_t n_list, int warp_id, arachne_aln_t* alns, uint32_t* cig_pool, uint64_t cig_cap,
               int32_t* mm_pool, uint64_t mm_cap, uint64_t* used2) {
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    int maxlen = 1;
    for (int64_t r = 0; r < n_reads; ++r) maxlen = std::max<int>(maxlen, (int)(off[r + 1] - off[r]));
    const int win_cap = 2 * maxlen + 1024 + 4 * opt.w;
    const long z_cap = (long)maxlen * (maxlen + 1024);  // batch.inl's sizing of the traceback scratch
    const int cig_scr_cap = 2 * maxlen + 64, mm_scr_cap = 2 * maxlen + 8;
    std::vector<uint8_t> win((size_t)EXT_WARPS * win_cap), zbuf((size_t)EXT_WARPS * z_cap);
    std::vector<uint32_t> cig_scr((size_t)EXT_WARPS * cig_scr_cap);
    std::vector<int32_t> mm_scr((size_t)EXT_WARPS * mm_scr_cap);
    Counters cnt; memset(&cnt, 0, sizeof(cnt));
    unsigned long long used[2] = {0, 0}; unsigned next = 0;
    const unsigned nl = n_list;
    coop(warp_id, EXT_WARPS * 32, [&] {
        k_cigar(opt, ref, seq, off, (long)n_regs, reinterpret_cast<const AlnReg*>(regs), reg_read, reinterpret_cast<AlnRec*>(alns), cig_pool, cig_cap, mm_pool, mm_cap,
                used, win.data(), win_cap, zbuf.data(), z_cap, cig_scr.data(), cig_scr_cap, mm_scr.data(), mm_scr_cap, &next, &cnt, list, list ? &nl : nullptr);
    });
    used2[0] = used[0]; used2[1] = used[1];
    return cnt.err;
}

// arachne_ksw_extend_static()'s launch sequence (csrc/noalloc.inl) on the host: control block cleared, k_xs_count, k_xs_base (its block of 256),
// k_xs_queue, k_ext_dp<false, true> over the sorted queue, k_xs_finish.  The queue / result arrays are NOT cleared between calls, as the static
// scratch is not: a call must write every slot it reads.  out: n x 6, {-1 x 6} for refused problems; status: 0 or 1 + lowest refused index.
int hs_xs_static(const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* q, const int64_t* q_off, const uint8_t* t, const int64_t* t_off, const int32_t* h0,
                 const int32_t* w, int max_qlen, int end_bonus, int zdrop, int warp_id, int32_t* out, int32_t* status, uint64_t* cells) {
    static std::vector<ExtProb> queue(65536);   // stand-ins for the static device scratch: they keep the last call's contents
    static std::vector<ExtRes> res(65536);
    static XsCtl ctl;
    if (n > 65536 || max_qlen < 1 || max_qlen > ARACHNE_MAX_READ_LEN) return ARACHNE_ERR_ARG;
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    XsIn I; I.n = (long)n; I.q_off = q_off; I.t_off = t_off; I.h0 = h0; I.w = w; I.end_bonus = end_bonus; I.max_qlen = max_qlen; I.a = opt.a;
    memset(&ctl, 0, sizeof(ctl));
    warp(0, 32, [&] { k_xs_count(I, &ctl, out); });
    blockIdx.x = 0; blockDim.x = 256; gridDim.x = 1;
    warpsim::run_block(256, [&] { k_xs_base(&ctl); });
    warp(0, 32, [&] { k_xs_queue(I, &ctl, queue.data()); });
    ExtDpArgs DA;
    DA.seq = q; DA.tbytes = t; DA.queue = queue.data(); DA.res = res.data(); DA.n_items = &ctl.n_items; DA.head = &ctl.head;
    DA.ncols = max_qlen + 2; DA.zdrop = zdrop; DA.one = 1; DA.cnt = &ctl.cnt;
    RefView no_ref; memset(&no_ref, 0, sizeof(no_ref));
    memset(arachne::xcols, 0xcd, sizeof(arachne::xcols));
    coop(warp_id, XDP_BLOCK, [&] { k_ext_dp<false, true>(opt, no_ref, DA); });
    warp(0, 32, [&] { k_xs_finish(I, &ctl, res.data(), out, status, reinterpret_cast<unsigned long long*>(cells)); });
    return ctl.cnt.err;
}

// The stage entry point arachne_ksw_global_batch's kernel (k_ksw_global: ksw_global2 with CIGAR on the warp context, a warp per problem), sized as
// the entry sizes it: traceback slice = min(qlen, 2w + 1) x tlen, CIGAR slots = qlen + tlen + 2.  cig: n x cig_cap words, n_cig: n.
int hs_k_ksw_global(const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* q, const int64_t* q_off, const uint8_t* t, const int64_t* t_off, const int32_t* w,
                    int warp_id, int32_t* scores, uint32_t* cig, int cig_cap, int32_t* n_cig) {
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    long zmax = 1;
    for (int64_t i = 0; i < n; ++i) {
        const long ql = q_off[i + 1] - q_off[i], tl = t_off[i + 1] - t_off[i];
        if (ql > ARACHNE_MAX_READ_LEN || ql < 1 || cig_cap < ql + tl + 2) return ARACHNE_ERR_ARG;
        zmax = std::max(zmax, std::min<long>(ql, 2L * w[i] + 1) * tl);
    }
    std::vector<uint8_t> zbuf((size_t)EXT_WARPS * zmax);
    // the kernel deals problems out statically (problem i to warp i mod 4): either the whole block, or its four warps one after the other
    for (int k = 0; k < ((warp_id & 8) ? 1 : EXT_WARPS); ++k)
        coop((warp_id & 8) | ((warp_id + k) & 3), EXT_WARPS * 32, [&] { k_ksw_global(opt, (long)n, q, q_off, t, t_off, w, scores, zbuf.data(), zmax, cig, cig_cap, n_cig); });
    return 0;
}

// arachne_occ_batch's kernel: bwt_2occ4 of n row pairs, 8 counts each
void hs_k_occ(void* h, int64_t n, const uint64_t* k, const uint64_t* l, uint64_t* cnt) {
    const FMIndex fm = *static_cast<const FMIndex*>(hs_fm(h));
    warp(0, 32, [&] { k_occ(fm, (long)n, k, l, cnt); });
}

}  // extern "C"
