// TEST-ONLY: the SHIPPED placement stage (csrc/place.cuh: candidates, tagBestAlignments, molecules, the optimizer sweep of k_place_rfa,
// MAPQ; csrc/records.cuh: markDuplicates) compiled by g++ and run on the host.  A placement block is ONE warp (PLACE_T = 32): k_place_rfa
// runs as 32 fibers whose __syncthreads / shuffles are rendezvous (cuda_warp_shim.h), with its block-scope __shared__ variables as statics;
// the other kernels are independent grid-stride loops and run the same way without ever meeting.  The staging follows
// csrc/place_host.inl kernel for kernel, with std::partial_sum / std::stable_sort where that file calls CUB scans and radix sorts.
#include <algorithm>
#include <numeric>
#include <vector>
#include <cmath>
using std::isnan; using std::log10; using std::nan; using std::pow;   // the device's math functions by their unqualified names
#define ARACHNE_HOSTSIM 1   // gorand.cuh: compile the device's Go PRNG for the host
#define __shared__ static
#include "cuda_warp_shim.h"
#include "../../include/arachne_cuda.h"
#include "../../arachne_b200/csrc/common.cuh"
#include "../../arachne_b200/csrc/counters.cuh"

namespace arachne {
}  // namespace arachne
#include "../../arachne_b200/csrc/place.cuh"
#include "../../arachne_b200/csrc/records.cuh"

using namespace arachne;
extern "C" const void* hs_ref(void* h);

namespace {
// a kernel "launch": one block of 32 lanes (fibers); grid-stride loops cover everything, a kernel that indexes by block gets `blocks` launches
template <class F> void launch(int blocks, F body) {
    blockDim.x = 32; gridDim.x = (unsigned)blocks;
    for (int b = 0; b < blocks; ++b) { blockIdx.x = (unsigned)b; warpsim::run_warp(0u, body); }
    blockIdx.x = 0; gridDim.x = 1;
}
// cub::DeviceRadixSort::SortPairs on bits [0, end_bit): a stable sort by the masked key
void sort_pairs(const std::vector<unsigned long long>& kin, std::vector<unsigned long long>& kout, const std::vector<unsigned>& vin, std::vector<unsigned>& vout, long n, int end_bit) {
    const unsigned long long mask = end_bit >= 64 ? ~0ull : ((1ull << end_bit) - 1);
    std::vector<long> ord((size_t)n);
    std::iota(ord.begin(), ord.end(), 0L);
    std::stable_sort(ord.begin(), ord.end(), [&](long a, long b) { return (kin[(size_t)a] & mask) < (kin[(size_t)b] & mask); });
    for (long i = 0; i < n; ++i) { kout[(size_t)i] = kin[(size_t)ord[(size_t)i]]; vout[(size_t)i] = vin[(size_t)ord[(size_t)i]]; }
}
}  // namespace

extern "C" {

// ARACHNE_STAGE_PLACE on host arrays.  reg_off: n_reads + 1; regs / alns: per region (alns = the CIGAR stage's records); bc_off: n_bc + 1 pair
// indices; run_rfa: n_bc flags; rand_seed: n_pairs.  Outputs: cand_off (n_reads + 1), cands (capacity cand_cap), lmp (n_bc).  Returns the
// number of candidates, or a negative error.
// With recs_out != null also ARACHNE_STAGE_RECORDS (CheckSplitReads, the fields of every SAM record): attach_bx per barcode (null: run_rfa),
// cig_pool = the CIGAR words alns[].cigar_off points into; rec_off: n_reads + 1, recs_out: capacity 2 * n_reads; *n_recs receives the count.
long hs_place(void* h, int64_t n_reads, const int64_t* seq_off, const uint64_t* reg_off, const arachne_alnreg_t* regs, const arachne_aln_t* alns, int n_bc,
              const int64_t* bc_off, const uint8_t* run_rfa, const int64_t* rand_seed, double improper, int delta, int64_t* cand_off, arachne_cand_t* cands_out,
              long cand_cap, double* lmp_out, const uint8_t* attach_bx, const uint32_t* cig_pool, int64_t* rec_off_out, arachne_record_t* recs_out, long* n_recs) {
    static_assert(sizeof(AlnRec) == sizeof(arachne_aln_t) && sizeof(Cand) == sizeof(arachne_cand_t), "layouts");
    const RefView ref = *static_cast<const RefView*>(hs_ref(h));
    const int64_t n_pairs = n_reads / 2;
    const int n_ctg = ref.n_seqs;
    Counters cnt; memset(&cnt, 0, sizeof(cnt));
    std::vector<unsigned long long> pl_cnt((size_t)n_reads + 2, 0), pl_scan((size_t)n_reads + 2, 0), coff((size_t)n_reads + 2, 0), roff(reg_off, reg_off + n_reads + 1);
    std::vector<double> lmp((size_t)n_bc + 1, 0.0);
    launch(1, [&] { k_place_count((int)n_reads, roff.data(), reinterpret_cast<const AlnReg*>(regs), delta, pl_cnt.data()); });
    { unsigned long long run = 0; for (int64_t i = 0; i <= n_reads; ++i) { pl_scan[(size_t)i] = run; run += pl_cnt[(size_t)i]; } }
    const long n_cands = (long)(pl_scan[(size_t)n_reads] & 0xffffffffull);
    if (n_cands > cand_cap) return -2;
    std::vector<Cand> cands((size_t)n_cands + 1); std::vector<Hot> hot((size_t)n_cands + 1); std::vector<Aux> aux((size_t)n_cands + 1);
    PlaceArgs PA;
    PA.regs = reinterpret_cast<const AlnReg*>(regs); PA.alns = reinterpret_cast<const AlnRec*>(alns); PA.reg_off = roff.data(); PA.seq_off = seq_off;
    PA.contigs = ref.contigs; PA.l_pac = ref.l_pac; PA.bc_off = bc_off; PA.n_bc = n_bc; PA.n_reads = (int)n_reads;
    PA.run_rfa = run_rfa; PA.rand_seed = rand_seed; PA.centro_start = nullptr; PA.centro_end = nullptr; PA.improper = improper; PA.delta = delta;
    PA.scan = pl_scan.data(); PA.cand_off = coff.data(); PA.cands = cands.data(); PA.hot = hot.data(); PA.aux = aux.data(); PA.lmp = lmp.data(); PA.cnt = &cnt;
    launch(1, [&] { k_place_build(PA); });
    // ---- tagBestAlignments ----
    unsigned slow_cap = 4;   // small on purpose: the overflow-and-retry path of place_host.inl is exercised too
    std::vector<uint64_t> slow;
    for (int attempt = 0;; ++attempt) {
        slow.assign((size_t)slow_cap * GORAND_LEN, 0);
        unsigned slow_next = 0;
        launch(1, [&] { k_place_tag(PA, (int)n_pairs, slow.data(), slow_cap, &slow_next); });
        if (cnt.err == 0) break;
        if (cnt.err != -102 || attempt >= 6) return cnt.err;
        slow_cap *= 16; cnt.err = 0; cnt.err_read = 0;
    }
    // ---- molecules + optimizer for the barcodes that run RFA ----
    std::vector<int> slot_of_bc((size_t)n_bc, -1), bc_of_slot;
    for (int b = 0; b < n_bc; ++b) if (run_rfa && run_rfa[b]) { slot_of_bc[(size_t)b] = (int)bc_of_slot.size(); bc_of_slot.push_back(b); }
    const int n_slots = (int)bc_of_slot.size();
    if (n_slots > 0 && n_cands > 0) {
        const int n_ctg1 = n_ctg + 1;
        std::vector<unsigned> first((size_t)n_slots * n_ctg1, 0xffffffffu), vals_a((size_t)n_cands + 1), vals_b((size_t)n_cands + 1);
        std::vector<unsigned long long> keys_a((size_t)n_cands + 1), keys_b((size_t)n_cands + 1), segc((size_t)n_slots + 2, 0), seg_off((size_t)n_slots + 2, 0);
        launch(1, [&] { k_place_first(PA, n_cands, slot_of_bc.data(), n_ctg1, first.data()); });
        launch(1, [&] { k_place_keys1(PA, n_cands, keys_a.data(), vals_a.data()); });
        sort_pairs(keys_a, keys_b, vals_a, vals_b, n_cands, 40);
        launch(1, [&] { k_place_keys2(PA, n_cands, vals_b.data(), slot_of_bc.data(), n_ctg1, first.data(), keys_a.data()); });
        sort_pairs(keys_a, keys_b, vals_b, vals_a, n_cands, 64);
        launch((n_slots + 32) / 32, [&] { k_place_seg(PA, n_slots, bc_of_slot.data(), segc.data()); });
        { unsigned long long run = 0; for (int i = 0; i <= n_slots; ++i) { seg_off[(size_t)i] = run; run += segc[(size_t)i]; } }
        const size_t nc1 = (size_t)n_cands + 1;
        std::vector<int> mol(12 * nc1, 0);
        RfaArgs RA;
        RA.bc_of_slot = bc_of_slot.data(); RA.n_slots = n_slots; RA.ord = vals_a.data(); RA.seg_off = seg_off.data();
        int* mi = mol.data();
        RA.k_p = reinterpret_cast<double*>(mi); mi += 2 * nc1;
        RA.m_beg = mi; RA.m_end = mi + nc1; RA.m_nact = mi + 2 * nc1; RA.m_nreads = mi + 3 * nc1; RA.m_new = mi + 4 * nc1;
        RA.k_beg = mi + 5 * nc1; RA.k_end = mi + 6 * nc1; RA.k_nact = mi + 7 * nc1; RA.k_nreads = mi + 8 * nc1; RA.k_isact = mi + 9 * nc1;
        unsigned next = 0; RA.next = &next;
        launch(1, [&] { k_place_rfa(PA, RA); });   // one block = one warp pulls barcodes from the queue until it is empty
        if (cnt.err) return cnt.err;
    }
    launch(1, [&] { k_place_mapq(PA); });
    // ---- markDuplicates (closes DoRFAForOneBarcode) ----
    if (n_cands > 0) {
        std::vector<int> act((size_t)n_reads + 1, -1);
        std::vector<unsigned long long> ka((size_t)n_reads + 1), kb((size_t)n_reads + 1); std::vector<unsigned> va((size_t)n_reads + 1), vb((size_t)n_reads + 1);
        launch(1, [&] { k_rec_active(PA, act.data()); });
        launch(1, [&] { k_rec_dup_keys(PA, act.data(), ka.data(), va.data()); });
        sort_pairs(ka, kb, va, vb, n_reads, 64);
        launch(1, [&] { k_rec_dup_mark(PA, act.data(), kb.data(), vb.data()); });
        if (recs_out) {
            static_assert(sizeof(Record) == sizeof(arachne_record_t), "Record layout");
            const size_t nr1 = (size_t)n_reads + 1;
            std::vector<int> ri(4 * nr1, 0); std::vector<double> rd(2 * nr1, 0.0); std::vector<unsigned long long> rcnt((size_t)n_reads + 2, 0), roff2((size_t)n_reads + 2, 0);
            std::vector<Record> recs((size_t)2 * n_reads + 1);
            RecArgs RA2;
            RA2.P = PA; RA2.attach_bx = attach_bx;
            RA2.act = act.data(); RA2.split_reg = ri.data(); RA2.split_mapq = ri.data() + nr1; RA2.split_proper = ri.data() + 2 * nr1;
            RA2.split_as = rd.data(); RA2.split_xs = rd.data() + nr1; RA2.rec_cnt = rcnt.data(); RA2.rec_off = roff2.data(); RA2.recs = nullptr;
            launch(1, [&] { k_rec_split(RA2); });
            { unsigned long long run = 0; for (int64_t i = 0; i <= n_reads; ++i) { roff2[(size_t)i] = run; run += rcnt[(size_t)i]; } }
            RA2.recs = recs.data();
            launch(1, [&] { k_rec_emit(RA2, cig_pool); });
            const long nrec = (long)roff2[(size_t)n_reads];
            for (int64_t i = 0; i <= n_reads; ++i) rec_off_out[i] = (int64_t)roff2[(size_t)i];
            memcpy(recs_out, recs.data(), (size_t)nrec * sizeof(Record));
            *n_recs = nrec;
        }
    }
    if (cnt.err) return cnt.err;
    for (int64_t i = 0; i <= n_reads; ++i) cand_off[i] = (int64_t)coff[(size_t)i];
    if (n_cands) memcpy(cands_out, cands.data(), (size_t)n_cands * sizeof(Cand));
    for (int b = 0; b < n_bc; ++b) lmp_out[b] = lmp[(size_t)b];
    return n_cands;
}

}  // extern "C"
