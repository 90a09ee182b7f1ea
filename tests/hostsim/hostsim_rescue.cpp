// TEST-ONLY: mate rescue as k_rescue runs it -- GoBwaMemMateSW's rescue loops and mem_matesw (csrc/rescue.cuh) on the warp-cooperative
// context of csrc/warp_ctx.cuh (the striped-semantics local Smith-Waterman with lanes as SSE2 vector elements, the cooperative reference
// fetch, leader writes between warp barriers) -- compiled by g++ and run as one warp of 32 fibers (cuda_warp_shim.h).  The host simulation's
// other rescue path (hostsim.cpp: HostCtx) swaps the warp primitives for the oracle's scalar kernels; this one runs the device's own.
#include <vector>
#include "cuda_warp_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/common.cuh"
#ifndef AR_D
#define AR_D inline
#endif
#include "../../arachne_b200/csrc/warp_ctx.cuh"
#include "../../arachne_b200/csrc/cigar.cuh"

using namespace arachne;
extern "C" const void* hs_ref(void* h);

extern "C" {

// r1 / r2: the reads' region lists after mem_align1_core (n1 / n2 entries, room for cap each); on return the lists after rescue.
int hs_pair_rescue_warp(void* h, const arachne_mem_opt_t* opt_, int l1, const uint8_t* s1, int l2, const uint8_t* s2, int score_delta,
                        arachne_alnreg_t* r1_, int* n1_, arachne_alnreg_t* r2_, int* n2_, int cap) {
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    AlnReg* r1 = reinterpret_cast<AlnReg*>(r1_); AlnReg* r2 = reinterpret_cast<AlnReg*>(r2_);
    const int win_cap = 2 * (l1 > l2 ? l1 : l2) + 1024 + 4 * opt.w, bcap = 4096;
    std::vector<uint8_t> win((size_t)win_cap + 64);
    std::vector<uint64_t> blist((size_t)bcap);
    int rc_out = 0, n1_out = *n1_, n2_out = *n2_, err_out = 0, lane_n[32][2];
    const int n1_in = *n1_, n2_in = *n2_;
    blockIdx.x = 0; blockDim.x = 128; gridDim.x = 1;
    warpsim::run_warp(64u, [&] {
        WarpCtx ctx;
        ctx.lane = (int)(threadIdx.x & 31u); ctx.cells = nullptr; ctx.z = nullptr; ctx.z_cap = 0; ctx.blist = blist.data(); ctx.bcap = bcap; ctx.lsw_err = 0;
        int n1 = n1_in, n2 = n2_in;  // per lane, as in k_rescue
        const int rc = pair_rescue(ctx, opt, ref, score_delta, l1, s1, l2, s2, r1, &n1, cap, r2, &n2, cap, win.data(), win_cap);
        lane_n[ctx.lane][0] = n1; lane_n[ctx.lane][1] = n2;
        if (ctx.lane == 0) { rc_out = rc; n1_out = n1; n2_out = n2; err_out = ctx.lsw_err; }
    });
    for (int l = 1; l < 32; ++l) if (lane_n[l][0] != n1_out || lane_n[l][1] != n2_out) return -8;  // the lanes' private counts must agree
    *n1_ = n1_out; *n2_ = n2_out;
    return rc_out < 0 ? rc_out : (err_out ? -9 : 0);
}


// mem_reg2aln + the GetAlignments walk of one region on the device's warp tier (csrc/cigar.cuh::reg2aln with WarpCtx: ksw_global2 striped over
// the lanes, byte traceback in scratch, leader backtrack) as k_cigar runs it.  mm: n_mm reference loci then n_mm read offsets.
int hs_reg2aln_warp(void* h, const arachne_mem_opt_t* opt_, int len, const uint8_t* seq, const arachne_alnreg_t* reg, arachne_aln_t* out, uint32_t* cigar, int cap, int32_t* mm) {
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    const AlnReg ar = *reinterpret_cast<const AlnReg*>(reg);
    const int win_cap = 2 * len + 1024 + 4 * opt.w;
    const long z_cap = (long)(len + 8) * (len + 4 * opt.w + 1024);
    std::vector<uint8_t> win((size_t)win_cap + 64), z((size_t)z_cap + 64);
    std::vector<int32_t> m((size_t)2 * len + 16);
    AlnRec res; int rc_out = 0;
    blockIdx.x = 0; blockDim.x = 128; gridDim.x = 1;
    warpsim::run_warp(96u, [&] {
        WarpCtx ctx;
        ctx.lane = (int)(threadIdx.x & 31u); ctx.cells = nullptr; ctx.z = z.data(); ctx.z_cap = z_cap;
        AlnRec a;
        const int rc = reg2aln(ctx, opt, ref, len, seq, ar, &a, cigar, cap, m.data(), win.data(), win_cap);
        if (ctx.lane == 0) { rc_out = rc; res = a; }
    });
    if (rc_out < 0) return rc_out;
    memcpy(out, &res, sizeof(res));
    for (int i = 0; i < res.n_mm; ++i) { mm[i] = m[i]; mm[res.n_mm + i] = m[len + i]; }
    return 0;
}

}  // extern "C"
