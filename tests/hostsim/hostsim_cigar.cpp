// TEST-ONLY: the SHIPPED lane-per-region CIGAR tiers (csrc/cigar_kernel.cuh: k_cigar_fast = band inference + the gap-free case, k_cigar_band
// = ksw_global2 on the shared-memory ring with 4-bit traceback, both ending in cig_finish = mem_reg2aln's tail + the GetAlignments walk)
// compiled by g++ through cuda_lane_shim.h and run one lane at a time: a lane owns its regions, its ring column and its traceback slice,
// and shares only pool / list counters with the others.  The staging follows csrc/batch.inl (fast tier -> tier-2 list sorted by key ->
// band tier; what neither takes lands in list3 for the warp tier, whose logic the caller runs through hs_reg2aln).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "cuda_lane_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/common.cuh"
#include "../../arachne_b200/csrc/counters.cuh"

namespace arachne {
}  // namespace arachne
#include "../../arachne_b200/csrc/cigar_kernel.cuh"
namespace arachne { int2 ring_raw[BAND_RING * CIG_BLOCK]; }  // k_cigar_band's dynamic shared memory

using namespace arachne;
extern "C" const void* hs_ref(void* h);   // hostsim.cpp: the loaded index's RefView
extern "C" int arachne_opts_default_host(arachne_mem_opt_t* o);

extern "C" {

// regs: n_regs mem_alnreg_t; reg_read[g]: read of region g.  alns: n_regs arachne_aln_t (those of list3 are left zeroed).
// cig / mm: pools of cig_cap / mm_cap words; used[2]: words used.  list3: regions neither tier took (capacity n_regs), counts[3] =
// {tier-2 regions, list3 length, NW cells of the band tier}.  Returns 0, or the kernels' error code.
int hs_cigar_tiers(void* h, const arachne_mem_opt_t* opt_, int64_t n_reads, const uint8_t* seq, const int64_t* off, int64_t n_regs, const arachne_alnreg_t* regs,
                   const int32_t* reg_read, arachne_aln_t* alns, uint32_t* cig, uint64_t cig_cap, int32_t* mm, uint64_t mm_cap, uint64_t* used,
                   uint32_t* list3, uint64_t* counts) {
    static_assert(sizeof(AlnRec) == sizeof(arachne_aln_t), "AlnRec layout");
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    int maxlen = 0;
    for (int64_t i = 0; i < n_reads; ++i) maxlen = std::max<int>(maxlen, (int)(off[i + 1] - off[i]));
    const int zband_rows = maxlen + 2 * BAND_WMAX + 4;
    const unsigned T = CIG_BLOCK;
    std::vector<unsigned> l2((size_t)n_regs + 1), key2((size_t)n_regs + 1, 0xffffffffu), l2s((size_t)n_regs + 1);
    std::vector<unsigned long long> zband((size_t)T * zband_rows * BAND_ZCHUNK + 8);
    unsigned n2 = 0, n3 = 0; unsigned long long used_[2] = {0, 0};
    Counters cnt; memset(&cnt, 0, sizeof(cnt));
    CigArgs A;
    A.seq = seq; A.off = off; A.n_regs = (long)n_regs; A.regs = reinterpret_cast<const AlnReg*>(regs); A.reg_read = reg_read;
    A.alns = reinterpret_cast<AlnRec*>(alns); A.cig_pool = cig; A.cig_cap = cig_cap; A.mm_pool = mm; A.mm_cap = mm_cap; A.used = used_;
    A.list2 = l2.data(); A.n2 = &n2; A.key2 = key2.data(); A.list3 = list3; A.n3 = &n3; A.z = zband.data(); A.z_rows = zband_rows; A.cnt = &cnt;
    blockIdx.x = 0; blockDim.x = T; gridDim.x = 1;
    for (unsigned t = 0; t < T; ++t) { threadIdx.x = t; k_cigar_fast(opt, ref, A); }
    // the host sorts tier 2's work by key (csrc/batch.inl: a stable radix sort on the key's low 16 bits)
    std::vector<unsigned> order(n2);
    for (unsigned i = 0; i < n2; ++i) order[i] = i;
    std::stable_sort(order.begin(), order.end(), [&](unsigned a, unsigned b) { return (key2[a] & 0xffffu) < (key2[b] & 0xffffu); });
    for (unsigned i = 0; i < n2; ++i) l2s[i] = l2[order[i]];
    A.list2 = l2s.data();
    memset(ring_raw, 0x5a, sizeof(ring_raw));
    for (unsigned t = 0; t < T; ++t) { threadIdx.x = t; k_cigar_band(opt, ref, A); }
    threadIdx.x = 0; blockDim.x = 1;
    used[0] = used_[0]; used[1] = used_[1];
    counts[0] = n2; counts[1] = n3; counts[2] = cnt.glb_cells;
    return cnt.err;
}

}  // extern "C"
