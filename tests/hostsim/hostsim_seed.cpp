// TEST-ONLY: the SHIPPED seeding kernel (csrc/seed_kernel.cuh: k_pack_reads, k_seed2, k_sort_intv) and the index builders it relies on
// (csrc/index_kernels.cuh: dense SA, k-mer bi-interval tables, inverse SA, both-strand text) compiled by g++ through cuda_lane_shim.h and run
// one lane at a time.  k_seed2's lanes are independent automata -- they share nothing but two work-queue counters -- so a single lane that
// drains both queues computes exactly what the grid computes, interval for interval; what the host run cannot show is timing.  The set-up
// below follows arachne_seed_batch (csrc/pipeline.cu).  Lets the CPU tier pin rows a3-a5 on the code the GPU runs, with every shortcut
// (tables, direct extension) on or off, and run it under ASan / UBSan and the differential fuzzer.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "cuda_lane_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/fmindex.cuh"
#include "../../arachne_b200/csrc/counters.cuh"

namespace arachne {
uint4 slist[(228 * 1024) / 16];  // the block's dynamic shared memory (k_seed2 declares it `extern __shared__`)
}  // namespace arachne
#include "../../arachne_b200/csrc/index_kernels.cuh"
#include "../../arachne_b200/csrc/seed_kernel.cuh"

using namespace arachne;

extern "C" const void* hs_fm(void* h);   // hostsim.cpp: the loaded index's FMIndex (file SA sampling, no tables)

namespace {
struct SeedIndex {  // what index.cuh builds at load time, built here by the same kernels
    FMIndex fm;
    std::vector<uint64_t> dense, isa, text;
    std::vector<uint4> kmer_all;
};
void launch_lane(unsigned tid, unsigned block, unsigned grid) { threadIdx.x = tid; blockIdx.x = 0; blockDim.x = block; gridDim.x = grid; }
}  // namespace

extern "C" {

// kmer_k: table depth (0 = none); direct: build the dense SA, inverse SA and text
void* hs_seed_index(void* h, const uint8_t* pac, int64_t l_pac, int kmer_k, int direct) {
    auto* S = new SeedIndex();
    S->fm = *static_cast<const FMIndex*>(hs_fm(h));
    FMIndex& fm = S->fm;
    fm.kmer = fm.kmer_all = nullptr; fm.kmer_k = 0; fm.isa = nullptr; fm.text = nullptr;
    launch_lane(0, 1, 1);
    if (direct) {
        S->dense.assign(fm.seq_len + 1, 0);
        k_sa_densify(fm, S->dense.data(), 1);
        fm.sa = S->dense.data(); fm.sa_intv = 1; fm.n_sa = fm.seq_len + 1;
        S->isa.assign(fm.seq_len + 1, 0);
        k_isa_fill(fm.sa, fm.seq_len + 1, S->isa.data());
        const uint64_t n_words = (fm.seq_len + 31) / 32 + DIRECT_WORDS + 2;
        S->text.assign(n_words, 0);
        k_text_both(pac, (uint64_t)l_pac, n_words, S->text.data());
        fm.isa = S->isa.data(); fm.text = S->text.data();
    }
    if (kmer_k > 0) {
        S->kmer_all.assign(kmer_level_off(kmer_k + 1) + 4, uint4{0, 0, 0, 0});
        for (int L = 0; L < kmer_k; ++L) { launch_lane(0, 4, 1); if (L == 0) { for (unsigned t = 0; t < 4; ++t) { threadIdx.x = t; k_kmer_level(fm, S->kmer_all.data(), 0); } } else { launch_lane(0, 1, 1); k_kmer_level(fm, S->kmer_all.data(), L); } }
        fm.kmer_all = S->kmer_all.data(); fm.kmer = S->kmer_all.data() + kmer_level_off(kmer_k); fm.kmer_k = kmer_k;
    }
    launch_lane(0, 1, 1);
    return S;
}
void hs_seed_index_free(void* s) { delete static_cast<SeedIndex*>(s); }

// mem_collect_intv for n_reads reads through k_pack_reads -> k_seed2 -> k_sort_intv.  intv: 4 u64 per interval, intv_off: n_reads + 1.
// lane: which of the block's 256 lanes runs (exercises the lane-strided shared and global layouts).  stats2: bwt_extend calls, occ blocks.
int hs_seed2(void* s, const arachne_mem_opt_t* opt_, int64_t n_reads, const uint8_t* seq, const int64_t* off, int use_kmer, int use_direct, int slist_want,
             int lane, uint64_t* intv, int64_t intv_cap, int64_t* intv_off, uint64_t* stats2) {
    SeedIndex* S = static_cast<SeedIndex*>(s);
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    int maxlen = 0;
    for (int64_t i = 0; i < n_reads; ++i) maxlen = std::max<int>(maxlen, (int)(off[i + 1] - off[i]));
    if (maxlen > ARACHNE_MAX_READ_LEN) return ARACHNE_ERR_ARG;
    const int L1 = maxlen + 1, rows = 2 * L1;
    const unsigned T = 256;
    std::vector<uint4> scr((size_t)T * rows);
    const int qstride = seed_qstride(maxlen);
    int slist_n = 0, nwq = 0;
    const size_t smem = seed_smem_plan(maxlen, 3, slist_want, &slist_n, &nwq);
    if (smem > sizeof(arachne::slist)) return ARACHNE_ERR_ARG;
    std::vector<ulonglong2> qpk((size_t)n_reads * qstride + 1);
    launch_lane(0, 1, 1);
    k_pack_reads(seq, off, (int)n_reads, qstride, qpk.data());
    int cap = 32;
    std::vector<Intv> arena; std::vector<int> c2, in_((size_t)n_reads + 1); std::vector<unsigned> nsa((size_t)n_reads + 1);
    Counters hc;
    for (int attempt = 0;; ++attempt) {
        arena.assign((size_t)n_reads * cap + 2, Intv{0, 0, 0, 0}); c2.assign((size_t)2 * n_reads + 2, 0);
        unsigned next[16] = {0}; memset(&hc, 0, sizeof(hc));
        memset(arachne::slist, 0xab, sizeof(arachne::slist));
        SeedOut so; so.intv = arena.data(); so.cnt2 = c2.data(); so.cap = cap;
        launch_lane((unsigned)lane, T, 1);
        k_seed2<3>(S->fm, opt, off, (int)n_reads, maxlen, scr.data(), rows, so, next, &hc, use_kmer, slist_n, qpk.data(), qstride, nwq,
                   use_direct && S->fm.isa && S->fm.text);
        if (!hc.err) break;
        if (hc.err != ARACHNE_ERR_OVERFLOW || attempt >= 3) return hc.err;
        cap *= 4;
    }
    launch_lane(0, 1, 1);
    k_sort_intv((int)n_reads, arena.data(), c2.data(), cap, opt.max_occ, in_.data(), nsa.data());
    int64_t o = 0;
    for (int64_t r = 0; r < n_reads; ++r) {
        intv_off[r] = o;
        if (o + in_[r] > intv_cap) return ARACHNE_ERR_OVERFLOW;
        memcpy(intv + 4 * o, arena.data() + (size_t)r * cap, (size_t)in_[r] * sizeof(Intv));
        o += in_[r];
    }
    intv_off[n_reads] = o;
    if (stats2) { stats2[0] = hc.occ; stats2[1] = hc.occ_blocks; }
    return ARACHNE_OK;
}

}  // extern "C"
