// TEST-ONLY: the SHIPPED extension DP kernel (csrc/ext_stage.cuh: k_ext_dp -- ksw_extend2 with one problem per lane, row-synchronous, the
// register pool of problem descriptors dealt out by shuffles, first rows filled by the whole warp, DPX cells on byte-packed column words)
// compiled by g++ and run as ONE WARP of 32 fibers (cuda_warp_shim.h).  A warp is the kernel's unit of cooperation: its lanes share the
// descriptor pool and write each other's first rows; warps share only the queue head.  One warp draining the queue therefore computes what
// the grid computes.  Set-up as in arachne_ksw_extend_batch (csrc/pipeline.cu): the queue sorted by query length, longest first.
#include <algorithm>
#include <vector>
#include "cuda_warp_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/common.cuh"
#ifndef AR_D
#define AR_D inline
#endif
#include "../../arachne_b200/csrc/stage_kernels.cuh"  // Counters, ChainBufs (named by the round kernels of ext_stage.cuh, which hostsim_stages.cpp runs)

namespace arachne {
uint32_t xcols[(ARACHNE_MAX_READ_LEN + 8) * 128];  // k_ext_dp's dynamic shared memory: [ncols][XDP_BLOCK]
}  // namespace arachne
#include "../../arachne_b200/csrc/ext_stage.cuh"

using namespace arachne;

extern "C" {

// n ksw_extend2 problems (byte targets, as the stage entry point takes them).  codec: 0 = chosen as the entry point chooses it, 1 = force the
// 13-bit column codec.  warp: which of the block's four warps runs; + 8: all four.  out: n x 6 {score, qle, tle, gtle, gscore, max_off}.
int hs_ext_dp(const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* q, const int64_t* q_off, const uint8_t* t, const int64_t* t_off, const int32_t* h0,
              const int32_t* w, int end_bonus, int zdrop, int codec, int warp, int32_t* out, uint64_t* cells) {
    static_assert(XDP_BLOCK == 128, "xcols is sized for 128 lanes");
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    int maxq = 1; bool wide = codec == 1;
    std::vector<ExtProb> probs((size_t)n); std::vector<int> order((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const int64_t ql = q_off[i + 1] - q_off[i], tl = t_off[i + 1] - t_off[i];
        if (ql > ARACHNE_MAX_READ_LEN || ql < 1 || h0[i] <= 0 || tl < 0) return ARACHNE_ERR_ARG;
        if ((long)h0[i] + ql * opt.a > 8191) return ARACHNE_ERR_ARG;
        if ((long)h0[i] + ql * opt.a > 255) wide = true;
        maxq = std::max<int>(maxq, (int)ql);
        ExtProb& P = probs[(size_t)i];
        P.read = (int32_t)i; P.q0 = 0; P.qlen = (int32_t)ql; P.tlen = (int32_t)tl; P.w = w[i]; P.h0 = h0[i]; P.end_bonus = end_bonus; P.dirs = 0; P.tpos = t_off[i]; P.qoff = q_off[i];
        order[(size_t)i] = (int)i;
    }
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return probs[(size_t)x].qlen > probs[(size_t)y].qlen; });
    std::vector<ExtProb> queue((size_t)n + 1);
    for (int64_t i = 0; i < n; ++i) queue[(size_t)i] = probs[(size_t)order[(size_t)i]];
    std::vector<ExtRes> res((size_t)n + 1);
    Counters cnt; memset(&cnt, 0, sizeof(cnt));
    unsigned ctl[2] = {(unsigned)n, 0u};
    ExtDpArgs DA;
    DA.seq = q; DA.tbytes = t; DA.queue = queue.data(); DA.res = res.data(); DA.n_items = &ctl[0]; DA.head = &ctl[1];
    DA.ncols = maxq + 2; DA.zdrop = zdrop; DA.one = 1; DA.cnt = &cnt;
    RefView no_ref; memset(&no_ref, 0, sizeof(no_ref));
    memset(arachne::xcols, 0xcd, sizeof(arachne::xcols));
    blockIdx.x = 0; blockDim.x = XDP_BLOCK; gridDim.x = 1;
    auto launch = [&](const std::function<void()>& f) {  // one of the block's four warps, or (bit 3 of warp) the whole block of four sharing the queue head
        if (warp & 8) warpsim::run_block(XDP_BLOCK, f, 1 << 20); else warpsim::run_warp(32u * (unsigned)(warp & 3), f);
    };
    if (wide) launch([&] { k_ext_dp<false, true>(opt, no_ref, DA); });
    else launch([&] { k_ext_dp<false, false>(opt, no_ref, DA); });
    for (int64_t i = 0; i < n; ++i) {
        const ExtRes& R = res[(size_t)i];
        int32_t* o = out + 6 * i;
        o[0] = R.mx; o[1] = R.max_j + 1; o[2] = R.max_i + 1; o[3] = R.max_ie + 1; o[4] = R.gscore; o[5] = R.max_off;
    }
    if (cells) *cells = cnt.ext_cells;
    return cnt.err;
}

}  // extern "C"
