// Placement stage (ARACHNE_STAGE_PLACE): the Go side of the reference's per-barcode work on the device.
//   GetChains / GetAlignments bookkeeping     src/aligner/aligner.go:1633-1715, 1484-1631
//   scoreAlignment / isPair                   :556-581, 1032-1063
//   tagBestAlignments                         :1397-1481
//   inferMolecules, markBestAlignmentForReadInMolecule, scrapMolecules, setMoleculeDifferences  :1300-1393, 991-1016, 527-545
//   optimizer.Optimize -> GenerateMove / fastScore / acceptMove   src/optimizer/optimizer.go:15-27, aligner.go:1065-1298
//   estimateMapQualities and helpers          :643-755, 797-922, 978-989
//
// Organisation (nothing like the Go pointer graph):
//   * candidates of the whole batch live in flat arrays, CSR by read; a 32-byte "hot" record per candidate carries what the
//     inner loops read (pos, log-probability, molecule, read, contig, strand, flags);
//   * a candidate molecule is a contiguous range of the batch-wide (barcode, contig first-appearance, pos) sort order, so
//     "alignments of molecule m" is a range scan and needs no per-molecule container;
//   * Go's ordered maps disappear: every read has exactly one active alignment, so molecule.active_alignments is "the
//     candidates of the range whose active flag is set" and best_alignment_for_read.Get(read) is a scan of that read's few
//     candidates for (molecule == m && best).  Iteration orders of the Go maps are not observable in any output
//     (all order-dependent sums are sums of dyadic rationals), which is what allows this;
//   * the optimizer runs one barcode per block: the M-way fastScore sweep of GenerateMove is spread over the block's
//     threads and reduced with the reference's tie rule; moves are applied in two phases (decide, then flip).
#pragma once
#include "common.cuh"
#include "gorand.cuh"
#include "cigar.cuh"
#include <cfloat>

namespace arachne {

struct Cand {  // == arachne_cand_t
    int32_t reg, read, id, active, is_proper, bwa_pick, mapq, molecule_id, active_molecule, second_best, mate, duplicate;
    int64_t pos, aend;
    double lap, as_score, xs_score, molecule_difference, sum_move, molecule_confidence, raw_mapq;
};
static_assert(sizeof(Cand) == 120, "Cand must match arachne_cand_t");

struct Hot {  // what the inner loops touch, one 32-byte sector per candidate
    int64_t pos;
    double lap;
    int32_t molid;  // kept-molecule id within the barcode, -1 none
    int32_t read;   // batch-global read index
    int32_t rid;    // contig, -1 for the unmapped placeholder ("" in Go)
    uint8_t reversed, best, active, sclip;
};
static_assert(sizeof(Hot) == 32, "Hot record");

struct Aux {  // cold per-candidate scalars
    double base;  // -2*mm -3*indels -5*#S -0.5*S_len: the single-end part of scoreAlignment
    double val;   // markBest scratch
    int32_t mism, molraw, rank, mv;
};

struct PlaceArgs {
    // inputs
    const AlnReg* regs; const AlnRec* alns; const unsigned long long* reg_off; const int64_t* seq_off;
    const Contig* contigs; int64_t l_pac;
    const int64_t* bc_off; int n_bc; int n_reads;
    const uint8_t* run_rfa; const int64_t* rand_seed;
    const int64_t* centro_start; const int64_t* centro_end;
    double improper; int delta;
    // candidate arrays
    const unsigned long long* scan;  // per read: low 32 = candidates before, high 32 = hits before (exclusive)
    unsigned long long* cand_off;    // n_reads+1
    Cand* cands; Hot* hot; Aux* aux;
    double* lmp;                     // per barcode
    Counters* cnt;
};

__device__ __forceinline__ int place_bc_of_read(const PlaceArgs& A, int r) {  // barcode holding read r
    const int64_t p = r >> 1;
    int lo = 0, hi = A.n_bc;  // bc_off[lo] <= p < bc_off[hi]
    while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (A.bc_off[mid] <= p) lo = mid; else hi = mid; }
    return lo;
}

__device__ __forceinline__ bool place_is_pair(const Hot& a, const Hot& b) {  // isPair, aligner.go:1032-1063
    if (a.reversed == b.reversed || a.rid != b.rid) return false;
    const int64_t dist = a.reversed ? a.pos - b.pos : b.pos - a.pos;  // reverse.pos - forward.pos
    return dist >= -35 && dist < 750;
}

// scoreAlignment(aln, mate, lmp) with both present (aligner.go:556-581); add_lmp = !aln.active_molecule
__device__ __forceinline__ double place_score(double base_a, double base_m, bool paired, double improper, bool add_lmp, double lmp) {
    double s = base_a + base_m;
    if (!paired) s += improper;
    if (add_lmp) s += lmp;
    return s;
}

// ------------------------------------------------------------------------------------------------ candidates
__global__ void k_place_count(int n_reads, const unsigned long long* __restrict__ reg_off, const AlnReg* __restrict__ regs, int delta,
                              unsigned long long* __restrict__ cnt) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= n_reads; r += gridDim.x * blockDim.x) {
        if (r == n_reads) { cnt[r] = 0; continue; }
        const long g0 = (long)reg_off[r], g1 = (long)reg_off[r + 1];
        int best = 0;
        for (long g = g0; g < g1; ++g) best = regs[g].score > best ? regs[g].score : best;
        unsigned kept = 0;
        for (long g = g0; g < g1; ++g) kept += regs[g].score >= best - delta;
        if (g1 == g0) kept = 1;
        const unsigned hits = g1 > g0 ? (unsigned)(g1 - g0) : 1u;
        cnt[r] = (unsigned long long)hits << 32 | kept;
    }
}

__global__ void k_place_build(PlaceArgs A) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r <= A.n_reads; r += gridDim.x * blockDim.x) {
        const unsigned long long sc = A.scan[r];
        A.cand_off[r] = sc & 0xffffffffull;
        if (r == A.n_reads) break;
        const int b = place_bc_of_read(A, r);
        const unsigned hit0 = (unsigned)(sc >> 32) - (unsigned)(A.scan[2 * A.bc_off[b]] >> 32);
        const long g0 = (long)A.reg_off[r], g1 = (long)A.reg_off[r + 1];
        long c = (long)(sc & 0xffffffffull);
        int best = 0;
        for (long g = g0; g < g1; ++g) best = A.regs[g].score > best ? A.regs[g].score : best;
        for (long g = g0 - (g1 == g0); g < g1; ++g) {
            Cand cd; Hot h; Aux x;
            int mism = 0, indels = 0, sclip = 0, sclen = 0;
            cd.read = r; cd.active = cd.is_proper = cd.bwa_pick = 0; cd.mapq = 0; cd.molecule_id = -1; cd.active_molecule = 0;
            cd.second_best = -1; cd.mate = -1; cd.duplicate = 0;
            cd.as_score = cd.xs_score = cd.molecule_difference = 0.0; cd.sum_move = 1.0; cd.molecule_confidence = 0.00075 * 0.025; cd.raw_mapq = 0.0;
            if (g >= g0) {
                const AlnReg rg = A.regs[g];
                if (rg.score < best - A.delta) continue;
                const AlnRec a = A.alns[g];
                const int64_t coff = A.contigs[rg.rid].offset;
                // InterpretAlign (gobwa.go:339-371), then GetAlignments' strand swap (aligner.go:1582-1587)
                const int64_t cpos = rg.rb < A.l_pac ? rg.rb - coff : A.l_pac * 2 - 1 - rg.rb - coff;
                const int64_t caend = rg.re < A.l_pac ? rg.re - coff : A.l_pac * 2 - 1 - rg.re - coff;
                cd.pos = cpos; cd.aend = caend;
                if (cpos != -1 && a.is_rev) { cd.pos = caend + 1; cd.aend = cpos + 1; }
                cd.reg = (int32_t)g; cd.id = (int32_t)(hit0 + (unsigned)(g - g0));
                h.rid = a.rid; h.reversed = (uint8_t)(a.is_rev != 0);
                mism = a.mismatches; indels = a.indels; sclip = a.soft_clipped; sclen = a.soft_clipped_length;
            } else {  // unmapped placeholder (aligner.go:1664-1676,1700-1712): the read-1 one omits hit_id
                cd.pos = -1; cd.aend = 0; cd.reg = -1; cd.id = (r & 1) ? (int32_t)hit0 : 0;
                h.rid = -1; h.reversed = 0;
            }
            double base = (double)(mism * -2 + indels * -3);
            if (sclip > 0) { base -= 5.0 * (double)sclip; base -= (double)sclen * 0.5; }
            cd.lap = (base + A.improper) - A.improper;  // scoreAlignment(a, nil, 0) - improper_pair_penalty (aligner.go:1618)
            h.pos = cd.pos; h.lap = cd.lap; h.molid = -1; h.read = r; h.best = 0; h.active = 0; h.sclip = (uint8_t)(sclip > 0);
            x.base = base; x.val = 0.0; x.mism = mism; x.molraw = -1; x.rank = 0; x.mv = -1;
            A.cands[c] = cd; A.hot[c] = h; A.aux[c] = x;
            ++c;
        }
    }
}

// ------------------------------------------------------------------------------------------------ tagBestAlignments
// One thread per pair.  Only read 1's generator ever decides: read 2 is always "touched" because every read has at least
// one candidate (aligner.go:1457-1471).  slow_pool: 607-word scratch slots for pairs that need more than 273 outputs.
__global__ void k_place_tag(PlaceArgs A, int n_pairs, uint64_t* slow_pool, unsigned slow_cap, unsigned* slow_next) {
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n_pairs; p += gridDim.x * blockDim.x) {
        const long a0 = (long)A.cand_off[2 * p], a1 = (long)A.cand_off[2 * p + 1], m1 = (long)A.cand_off[2 * p + 2];
        const int64_t seed = A.rand_seed ? A.rand_seed[p] : 1;
        bool stateful = (a1 - a0) * (m1 - a1) + 4 > GORAND_TAP;
        double best = -DBL_MAX; long ba = -1, bm = -1;
        for (int attempt = 0; attempt < 2; ++attempt) {
            GoRandDev rng; rng.init(seed);
            if (stateful) {
                const unsigned s = atomicAdd(slow_next, 1u);
                if (s >= slow_cap) { raise(A.cnt, -102, 2 * p); return; }
                rng.materialise(slow_pool + (size_t)s * GORAND_LEN);
            }
            bool need = false;
            best = -DBL_MAX; ba = bm = -1;
            for (long a = a0; a < a1 && !need; ++a) {
                const Hot ha = A.hot[a]; const double xa = A.aux[a].base;
                for (long m = a1; m < m1; ++m) {
                    const Hot hm = A.hot[m];
                    const double u = rng.float64(&need);
                    if (need) break;
                    const double total = place_score(xa, A.aux[m].base, place_is_pair(ha, hm), A.improper, true, 0.0) + (u / 2.0);
                    if (total > best) { best = total; ba = a; bm = m; }
                }
            }
            if (!need) break;
            stateful = true;  // Float64 retries ran past the stateless window: redo the pair on the stateful generator
        }
        if (ba < 0) continue;
        const bool proper = place_is_pair(A.hot[ba], A.hot[bm]);
        A.cands[ba].active = 1; A.cands[ba].bwa_pick = 1; A.hot[ba].active = 1;
        A.cands[bm].active = 1; A.cands[bm].bwa_pick = 1; A.hot[bm].active = 1;
        if (proper) { A.cands[ba].is_proper = 1; A.cands[bm].is_proper = 1; }
    }
}

// ------------------------------------------------------------------------------------------------ position lists
// tagBestAlignments appends every alignment to the list of its contig, lists ordered by first appearance of the contig
// within the barcode (aligner.go:1448-1455), then sorts each list by pos (:1477-1479).  Here: first[slot][rid+1] = smallest
// candidate index of that contig in the barcode; two stable radix sorts (pos, then slot<<32|first) give the same order for
// the whole batch at once.  sort.Sort is unstable on equal pos (pdqsort) -- unpinned; ties keep candidate order here.
__global__ void k_place_first(PlaceArgs A, long n_cands, const int* __restrict__ slot_of_bc, int n_ctg1, unsigned* first) {
    for (long c = blockIdx.x * (long)blockDim.x + threadIdx.x; c < n_cands; c += (long)gridDim.x * blockDim.x) {
        const Hot h = A.hot[c];
        const int slot = slot_of_bc[place_bc_of_read(A, h.read)];
        if (slot >= 0) atomicMin(&first[(size_t)slot * n_ctg1 + (h.rid + 1)], (unsigned)c);
    }
}

__global__ void k_place_keys1(PlaceArgs A, long n_cands, unsigned long long* keys, unsigned* vals) {
    for (long c = blockIdx.x * (long)blockDim.x + threadIdx.x; c < n_cands; c += (long)gridDim.x * blockDim.x) {
        keys[c] = (unsigned long long)(A.hot[c].pos + 1);  // pos >= -1
        vals[c] = (unsigned)c;
    }
}

__global__ void k_place_keys2(PlaceArgs A, long n_cands, const unsigned* __restrict__ vals, const int* __restrict__ slot_of_bc, int n_ctg1,
                              const unsigned* __restrict__ first, unsigned long long* keys) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n_cands; i += (long)gridDim.x * blockDim.x) {
        const Hot h = A.hot[vals[i]];
        const int slot = slot_of_bc[place_bc_of_read(A, h.read)];
        keys[i] = slot < 0 ? ~0ull : ((unsigned long long)(unsigned)slot << 32 | first[(size_t)slot * n_ctg1 + (h.rid + 1)]);
    }
}

// ------------------------------------------------------------------------------------------------ per-barcode optimizer
struct RfaArgs {
    const int* bc_of_slot; int n_slots;
    const unsigned* ord;               // candidates in (slot, contig first-appearance, pos) order
    const unsigned long long* seg_off; // per slot: range of ord (n_slots+1)
    // per-molecule scratch, indexed seg_off[slot] + molecule (a barcode has at most as many molecules as candidates)
    int *m_beg, *m_end, *m_nact, *m_nreads, *m_new;
    int *k_beg, *k_end, *k_nact, *k_nreads, *k_isact;
    double* k_p;
    unsigned* next;
};

constexpr int PLACE_T = 32;  // threads per barcode block: a 30-pair barcode has ~100 candidates; small blocks keep more barcodes in flight and barriers cheap

__device__ __forceinline__ int place_block_scan_excl(int v, int* total, int* sh /* PLACE_T/32 + 1 ints */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int x = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, d); if (lane >= d) x += y; }
    __syncthreads();
    if (lane == 31) sh[w] = x;
    __syncthreads();
    if (threadIdx.x == 0) { int s = 0; for (int i = 0; i < PLACE_T / 32; ++i) { int t = sh[i]; sh[i] = s; s += t; } sh[PLACE_T / 32] = s; }
    __syncthreads();
    *total = sh[PLACE_T / 32];
    return x - v + sh[w];
}

// best_alignment_for_read.Get(read) of kept molecule `mol`
__device__ __forceinline__ long place_lookup(const PlaceArgs& A, int mol, int read) {
    for (long k = (long)A.cand_off[read], e = (long)A.cand_off[read + 1]; k < e; ++k) { const Hot h = A.hot[k]; if (h.molid == mol && h.best) return k; }
    return -1;
}
__device__ __forceinline__ long place_active_of(const PlaceArgs& A, int read) {
    for (long k = (long)A.cand_off[read], e = (long)A.cand_off[read + 1]; k < e; ++k) if (A.hot[k].active) return k;
    return -1;
}
__device__ __forceinline__ bool place_is_active_molecule(int active, int potential) {  // isActiveMolecule, aligner.go:1239-1249
    if ((double)active <= 4) return false;
    if ((double)active / (double)potential < 0.1) return false;
    return true;
}
// the per-alignment test of fastScore (aligner.go:1133-1147): does source alignment c move to sink, and its score terms
__device__ __forceinline__ bool place_move_terms(const PlaceArgs& A, int src, int sink, long c, const Hot& hc, long* s_out, bool* moves, double* dchange) {
    const long s = place_lookup(A, sink, hc.read);
    if (s < 0) return false;
    const int mate_r = hc.read ^ 1;
    long sm = place_active_of(A, mate_r);
    if (sm >= 0 && A.hot[sm].molid != src) sm = -1;
    const bool has_mate = sm >= 0;
    const bool shmp = has_mate && place_is_pair(hc, A.hot[sm]);
    const Hot hs = A.hot[s];
    const long mt = place_lookup(A, sink, mate_r);
    const bool skmp = mt >= 0 && place_is_pair(hs, A.hot[mt]) && has_mate;
    *moves = !shmp || (has_mate && skmp);
    double d = hs.lap - hc.lap;
    *s_out = s;
    dchange[0] = d;
    dchange[1] = (shmp && !skmp) ? A.improper / 2.0 : ((!shmp && skmp) ? -(A.improper / 2.0) : 0.0);
    return true;
}

__device__ double place_fast_score(const PlaceArgs& A, const RfaArgs& R, long ob, int src, int sink, int* num_out) {  // fastScore, aligner.go:1109-1237
    double alignment_change = 0.0, change = 0.0;
    int num = 0;
    for (int j = R.k_beg[ob + src], e = R.k_end[ob + src]; j < e; ++j) {
        const long c = R.ord[ob + j];
        const Hot hc = A.hot[c];
        if (!hc.active) continue;
        long s; bool mv; double d[2];
        if (!place_move_terms(A, src, sink, c, hc, &s, &mv, d)) continue;
        alignment_change += d[0];
        if (d[1] != 0.0) alignment_change += d[1];
        ++num;
    }
    const int sa = R.k_nact[ob + src], sp = R.k_nreads[ob + src], ka = R.k_nact[ob + sink], kp = R.k_nreads[ob + sink];
    if (!place_is_active_molecule(sa - num, sp) && place_is_active_molecule(sa, sp)) change -= (double)sp * -0.5;
    if (place_is_active_molecule(ka + num, kp) && !place_is_active_molecule(ka, kp)) change += (double)kp * -0.5;
    if (sa - num == 0 && num > 0) change -= -3.0;
    if (ka == 0 && num > 0) change += -3.0;
    change += alignment_change;
    *num_out = num;
    return change;
}

// One barcode per block (slots pulled from a queue).  Steps follow DoRFAForOneBarcode (aligner.go:479-496).
__global__ void __launch_bounds__(PLACE_T) k_place_rfa(PlaceArgs A, RfaArgs R) {
    __shared__ int sh_scan[PLACE_T / 32 + 1];
    __shared__ int sh_slot, sh_best_idx, sh_accept, sh_ntouch;
    __shared__ double sh_wscore[PLACE_T / 32];
    __shared__ int sh_wlen[PLACE_T / 32], sh_widx[PLACE_T / 32];
    __shared__ long long sh_dna;
    const int tid = threadIdx.x;
    for (;;) {
        __syncthreads();
        if (tid == 0) sh_slot = (int)atomicAdd(R.next, 1u);
        __syncthreads();
        const int slot = sh_slot;
        if (slot >= R.n_slots) break;
        const int b = R.bc_of_slot[slot];
        const long ob = (long)R.seg_off[slot];
        const int n = (int)(R.seg_off[slot + 1] - R.seg_off[slot]);
        // ---- inferMolecules (aligner.go:1300-1338): a new molecule at every list head and at gaps > 50 kb ----
        int carry = 0;
        for (int j0 = 0; j0 < n; j0 += PLACE_T) {
            const int j = j0 + tid;
            int head = 0; long c = -1;
            if (j < n) {
                c = R.ord[ob + j];
                if (j == 0) head = 1;
                else { const Hot h = A.hot[c], hp = A.hot[R.ord[ob + j - 1]]; head = (h.rid != hp.rid) || (h.pos - hp.pos > 50000); }
            }
            int tot;
            const int ex = place_block_scan_excl(head, &tot, sh_scan);
            if (j < n) {
                const int mol = carry + ex + head - 1;
                A.aux[c].molraw = mol; A.aux[c].rank = j;
                if (head) { R.m_beg[ob + mol] = j; R.m_nact[ob + mol] = 0; R.m_nreads[ob + mol] = 0; if (mol > 0) R.m_end[ob + mol - 1] = j; }
            }
            carry += tot;
        }
        const int M0 = carry;
        if (tid == 0 && M0 > 0) R.m_end[ob + M0 - 1] = n;
        __syncthreads();
        // ---- markBestAlignmentForReadInMolecule (aligner.go:1340-1393) ----
        for (int j = tid; j < n; j += PLACE_T) {  // value of each alignment: best pair score with a mate alignment of the same molecule, else its own
            const long c = R.ord[ob + j];
            const Hot hc = A.hot[c]; const Aux xc = A.aux[c];
            double v = -DBL_MAX; bool has = false;
            for (long m = (long)A.cand_off[hc.read ^ 1], e = (long)A.cand_off[(hc.read ^ 1) + 1]; m < e; ++m) {
                if (A.aux[m].molraw != xc.molraw) continue;
                has = true;
                const double s = place_score(xc.base, A.aux[m].base, place_is_pair(hc, A.hot[m]), A.improper, true, 0.0);
                v = s > v ? s : v;
            }
            A.aux[c].val = has ? v : hc.lap;
        }
        __syncthreads();
        for (int j = tid; j < n; j += PLACE_T) {  // first alignment in position order with the largest value wins its (molecule, read)
            const long c = R.ord[ob + j];
            const Hot hc = A.hot[c]; const Aux xc = A.aux[c];
            bool best = true;
            for (long k = (long)A.cand_off[hc.read], e = (long)A.cand_off[hc.read + 1]; k < e; ++k) {
                if (k == c) continue;
                const Aux xk = A.aux[k];
                if (xk.molraw == xc.molraw && (xk.val > xc.val || (xk.val == xc.val && xk.rank < xc.rank))) { best = false; break; }
            }
            A.hot[c].best = (uint8_t)best;
            if (best) atomicAdd(&R.m_nreads[ob + xc.molraw], 1);
            if (hc.active) atomicAdd(&R.m_nact[ob + xc.molraw], 1);
        }
        __syncthreads();
        // ---- scrapMolecules (aligner.go:991-1016): drop molecules without an active alignment, renumber ----
        carry = 0;
        for (int m0 = 0; m0 < M0; m0 += PLACE_T) {
            const int m = m0 + tid;
            const int keep = m < M0 && R.m_nact[ob + m] > 0;
            int tot;
            const int ex = place_block_scan_excl(keep, &tot, sh_scan);
            if (m < M0) {
                const int id = keep ? carry + ex : -1;
                R.m_new[ob + m] = id;
                if (keep) { R.k_beg[ob + id] = R.m_beg[ob + m]; R.k_end[ob + id] = R.m_end[ob + m]; R.k_nact[ob + id] = R.m_nact[ob + m]; R.k_nreads[ob + id] = R.m_nreads[ob + m]; }
            }
            carry += tot;
        }
        const int K = carry;
        __syncthreads();
        for (int j = tid; j < n; j += PLACE_T) {
            const long c = R.ord[ob + j];
            const int id = R.m_new[ob + A.aux[c].molraw];
            A.hot[c].molid = id; A.cands[c].molecule_id = id;
        }
        __syncthreads();
        // ---- setMoleculeDifferences (aligner.go:527-545), first call ----
        for (int m = tid; m < K; m += PLACE_T) {
            int diff = 0;
            for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) { const long c = R.ord[ob + j]; if (A.hot[c].active) diff += A.aux[c].mism; }
            const double d = (double)diff / (double)R.k_nact[ob + m];
            for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) { const long c = R.ord[ob + j]; if (A.hot[c].active) A.cands[c].molecule_difference = d; }
        }
        __syncthreads();
        // ---- which molecules can ever be the source of a move: one of their reads has a best alignment in another kept molecule.
        // `best` and `molid` are fixed from here on (only `active` flips), so this is decided once; a GenerateMove step or a
        // moleculeMapqProbabilitySums row of any other molecule finds num_moved == 0 for every sink and changes nothing.
        // m_beg / m_end / m_nact are dead after scrapMolecules: reused as the sink stamps, the touched-sink list and this flag.
        int* k_mark = R.m_beg + ob; int* k_list = R.m_end + ob; int* k_alt = R.m_nact + ob;
        for (int m = tid; m < K; m += PLACE_T) { k_mark[m] = 0; k_alt[m] = 0; }
        __syncthreads();
        for (int j = tid; j < n; j += PLACE_T) {
            const long c = R.ord[ob + j];
            const Hot hc = A.hot[c];
            if (hc.molid < 0 || k_alt[hc.molid]) continue;
            for (long k = (long)A.cand_off[hc.read], e = (long)A.cand_off[hc.read + 1]; k < e; ++k) {
                const Hot hk = A.hot[k];
                if (hk.best && hk.molid >= 0 && hk.molid != hc.molid) { k_alt[hc.molid] = 1; break; }
            }
        }
        __syncthreads();
        // The sinks fastScore can report num_moved > 0 for = the molecules holding a best alignment of a read that is active in
        // src.  They are collected first (a handful), and only they are scored: O(A*C) per step instead of O(K*A*C) -- at 200+
        // pairs and 10+ molecules per barcode (stLFR) and 1500 molecules in a 30000-record work unit the dense sweep was 236 ms
        // of a 359 ms step.  Each collected sink is still scored by the same sequential fastScore, so sums keep their order.
        auto collect = [&](int src, int stamp) {
            if (tid == 0) sh_ntouch = 0;
            __syncthreads();
            for (int j = R.k_beg[ob + src] + tid; j < R.k_end[ob + src]; j += PLACE_T) {
                const Hot hc = A.hot[R.ord[ob + j]];
                if (!hc.active) continue;
                for (long k = (long)A.cand_off[hc.read], e = (long)A.cand_off[hc.read + 1]; k < e; ++k) {
                    const Hot hk = A.hot[k];
                    if (hk.best && hk.molid >= 0 && hk.molid != src && atomicExch(&k_mark[hk.molid], stamp) != stamp) k_list[atomicAdd(&sh_ntouch, 1)] = hk.molid;
                }
            }
            __syncthreads();
            return sh_ntouch;
        };
        // ---- optimizer.Optimize(model, 1, 2, 4*K): 8*K GenerateMove steps (optimizer.go:15-27, aligner.go:1065-1097) ----
        int cur = 0;
        for (long step = 0; step < 8L * K; ++step) {
            const int src = cur;
            cur = (cur + 1) % K;
            if (R.k_nact[ob + src] == 0 || !k_alt[src]) continue;  // uniform: every thread reads the same words after the barrier below
            const int ntouch = collect(src, (int)step + 1);
            double bs = -DBL_MAX; int bl = -1, bi = 0x7fffffff;
            for (int t = tid; t < ntouch; t += PLACE_T) {
                const int i = k_list[t];
                int num;
                const double sc = place_fast_score(A, R, ob, src, i, &num);
                const int len = R.k_nact[ob + i];
                if (num > 0 && (bi == 0x7fffffff || sc > bs || (sc == bs && (len > bl || (len == bl && i < bi))))) { bs = sc; bl = len; bi = i; }
            }
            // lexicographic max of (score, sink size), smallest index on ties == the sequential rule of GenerateMove
#pragma unroll
            for (int d = 16; d > 0; d >>= 1) {
                const double os = __shfl_down_sync(0xffffffffu, bs, d); const int ol = __shfl_down_sync(0xffffffffu, bl, d), oi = __shfl_down_sync(0xffffffffu, bi, d);
                if (oi != 0x7fffffff && (bi == 0x7fffffff || os > bs || (os == bs && (ol > bl || (ol == bl && oi < bi))))) { bs = os; bl = ol; bi = oi; }
            }
            if ((tid & 31) == 0) { sh_wscore[tid >> 5] = bs; sh_wlen[tid >> 5] = bl; sh_widx[tid >> 5] = bi; }
            __syncthreads();
            if (tid == 0) {
                for (int w = 1; w < PLACE_T / 32; ++w) {
                    const double os = sh_wscore[w]; const int ol = sh_wlen[w], oi = sh_widx[w];
                    if (oi != 0x7fffffff && (bi == 0x7fffffff || os > bs || (os == bs && (ol > bl || (ol == bl && oi < bi))))) { bs = os; bl = ol; bi = oi; }
                }
                sh_best_idx = bi;
                sh_accept = bi != 0x7fffffff && (bs > 0 || (bs == 0 && bl > R.k_nact[ob + src]));
            }
            __syncthreads();
            if (sh_accept) {  // acceptMove (aligner.go:1261-1298): decide on the pre-move state, then flip
                const int sink = sh_best_idx;
                for (int j = R.k_beg[ob + src] + tid; j < R.k_end[ob + src]; j += PLACE_T) {
                    const long c = R.ord[ob + j];
                    const Hot hc = A.hot[c];
                    long s = -1; bool mv = false; double d[2];
                    if (hc.active && place_move_terms(A, src, sink, c, hc, &s, &mv, d) && mv) A.aux[c].mv = (int32_t)s; else A.aux[c].mv = -1;
                }
                __syncthreads();
                for (int j = R.k_beg[ob + src] + tid; j < R.k_end[ob + src]; j += PLACE_T) {
                    const long c = R.ord[ob + j];
                    const int s = A.aux[c].mv;
                    if (s < 0) continue;
                    A.hot[c].active = 0; A.cands[c].active = 0; A.hot[s].active = 1; A.cands[s].active = 1;
                    atomicSub(&R.k_nact[ob + src], 1); atomicAdd(&R.k_nact[ob + sink], 1);
                }
            }
            __syncthreads();
        }
        // ---- moleculeMapqProbabilitySums (aligner.go:697-720): only sinks holding an alternative for an active read of src add anything ----
        for (int src = 0; src < K; ++src) {
            if (R.k_nact[ob + src] == 0 || !k_alt[src]) continue;
            const int ntouch = collect(src, 8 * K + 1 + src);
            for (int t = tid; t < ntouch; t += PLACE_T) {
                const int i = k_list[t];
                int num;
                R.k_p[ob + i] = pow(10.0, place_fast_score(A, R, ob, src, i, &num));
            }
            __syncthreads();
            for (int j = R.k_beg[ob + src] + tid; j < R.k_end[ob + src]; j += PLACE_T) {
                const long c = R.ord[ob + j];
                const Hot hc = A.hot[c];
                if (!hc.active) continue;
                double sum = A.cands[c].sum_move;
                int last = -1;
                for (;;) {  // sinks holding an alternative for this read, ascending (the order Go adds them in)
                    int nxt = 0x7fffffff;
                    for (long k = (long)A.cand_off[hc.read], e = (long)A.cand_off[hc.read + 1]; k < e; ++k) {
                        const Hot hk = A.hot[k];
                        if (hk.best && hk.molid > last && hk.molid != src && hk.molid < nxt) nxt = hk.molid;
                    }
                    if (nxt == 0x7fffffff) break;
                    sum += R.k_p[ob + nxt]; last = nxt;
                }
                A.cands[c].sum_move = sum;
            }
            __syncthreads();
        }
        // ---- updateAlignmentsMoleculeStatus (aligner.go:643-675): confidences, differences, active molecules ----
        for (int m = tid; m < K; m += PLACE_T) {
            const int nact = R.k_nact[ob + m];
            const double conf = (double)nact / (double)R.k_nreads[ob + m];
            int softc = 0, diff = 0;
            for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) { const long c = R.ord[ob + j]; const Hot h = A.hot[c]; if (h.active) { softc += h.sclip; diff += A.aux[c].mism; } }
            const double d = (double)diff / (double)nact;
            for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) {
                const long c = R.ord[ob + j];
                if (A.hot[c].active) { A.cands[c].molecule_confidence = conf; A.cands[c].molecule_difference = d; }
            }
            R.k_isact[ob + m] = (nact - softc > 4 && conf > 0.1) ? 1 : 0;
        }
        if (tid == 0) sh_dna = 1000;
        __syncthreads();
        for (int j = tid; j < n; j += PLACE_T) {
            const long c = R.ord[ob + j];
            const int id = A.hot[c].molid;
            if (id >= 0) A.cands[c].active_molecule = R.k_isact[ob + id];
        }
        // ---- calculateLogMoleculePenalty (aligner.go:722-755); all terms are integers, so the sum order is free ----
        for (int m = tid; m < K; m += PLACE_T) {
            long long add = 0;
            if (R.k_isact[ob + m]) {
                long long smallest = 0x7fffffffffffffffLL, biggest = -1;
                for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) {
                    const long c = R.ord[ob + j]; const Hot h = A.hot[c];
                    if (h.active) { if (h.pos > biggest) biggest = h.pos; if (h.pos < smallest) smallest = h.pos; }
                }
                if (biggest >= smallest) add = (biggest - smallest) + 1000;
            } else {
                for (int j = R.k_beg[ob + m]; j < R.k_end[ob + m]; ++j) { const long c = R.ord[ob + j]; if (A.hot[c].active) add += (A.cands[c].aend - A.cands[c].pos) * 2; }
            }
            if (add) atomicAdd((unsigned long long*)&sh_dna, (unsigned long long)add);
        }
        __syncthreads();
        if (tid == 0) A.lmp[b] = K > 0 ? log10((double)sh_dna / 3200000000.0 * 0.05) : 0.0;
    }
}

// ------------------------------------------------------------------------------------------------ estimateMapQualities, per read
// aligner.go:817-919.  One thread per read; it writes only its own read's candidates (the mate link of the active pair is
// derived locally, never read from what the mate's thread writes).
__global__ void k_place_mapq(PlaceArgs A) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < A.n_reads; r += gridDim.x * blockDim.x) {
        const int b = place_bc_of_read(A, r);
        const double lmp = A.lmp[b];
        const long a0 = (long)A.cand_off[r], a1 = (long)A.cand_off[r + 1];
        const long m0 = (long)A.cand_off[r ^ 1], m1 = (long)A.cand_off[(r ^ 1) + 1];
        if (a1 == a0) continue;
        double top[15]; int ntop = 0;  // the 15 largest pair scores, descending (aligner.go:895-898)
#pragma unroll
        for (int i = 0; i < 15; ++i) top[i] = -DBL_MAX;
        auto push = [&](double s) {
            int i;
            if (ntop < 15) i = ntop++;
            else if (s > top[14]) i = 14;
            else return;
            while (i > 0 && top[i - 1] < s) { top[i] = top[i - 1]; --i; }
            top[i] = s;
        };
        // pseudo-count alignment paired with the best single mate (appendPsuedocountAlignmentScore, :677-695)
        double bestSingle = -DBL_MAX; long am = -1;
        for (long m = m0; m < m1; ++m) {
            const double s = A.aux[m].base + A.improper;  // scoreAlignment(nil, mate, lmp)
            bestSingle = s > bestSingle ? s : bestSingle;
            if (A.hot[m].active) am = m;
        }
        double pseudo = 0.0;
        pseudo -= 10.0; pseudo -= ((double)(A.seq_off[r + 1] - A.seq_off[r]) - 25.0) * 0.5; pseudo += lmp;
        const double first = m1 > m0 ? bestSingle + pseudo : pseudo;
        push(first);
        // best pair score of every alignment; second-best bookkeeping (:829-873)
        double sb_lp = -1000.0, sb_raw = first; long sb = -1;
        for (long a = a0; a < a1; ++a) {
            const Hot ha = A.hot[a]; const double xa = A.aux[a].base; const bool add = !A.cands[a].active_molecule;
            double best = -DBL_MAX; long mate_of_a = -1;
            for (long m = m0; m < m1; ++m) {
                const bool paired = place_is_pair(ha, A.hot[m]);
                const double s = place_score(xa, A.aux[m].base, paired, A.improper, add, lmp);
                best = s > best ? s : best;
                if (!ha.active && s > sb_lp) { sb_lp = s; sb_raw = place_score(xa, A.aux[m].base, paired, A.improper, add, 0.0); sb = a; mate_of_a = m; }
            }
            if (m1 == m0) { best = xa + A.improper; if (add) best += lmp; }
            push(best);
            A.cands[a].mate = ha.active ? (int32_t)am : (int32_t)mate_of_a;
        }
        double total = 0.0;
        for (int i = 0; i < ntop; ++i) total += pow(10.0, top[i]);
        for (long a = a0; a < a1; ++a) {
            const Hot ha = A.hot[a]; const double xa = A.aux[a].base; const bool add = !A.cands[a].active_molecule;
            const long mt = A.cands[a].mate;
            if (ha.active) {
                A.cands[a].second_best = (int32_t)sb; A.cands[a].xs_score = sb_raw;
                A.cands[a].as_score = mt >= 0 ? place_score(xa, A.aux[mt].base, place_is_pair(ha, A.hot[mt]), A.improper, add, 0.0) : xa + A.improper + (add ? 0.0 : 0.0);
            }
            double score;
            if (mt >= 0) score = place_score(xa, A.aux[mt].base, place_is_pair(ha, A.hot[mt]), A.improper, add, lmp);
            else { score = xa + A.improper; if (add) score += lmp; }
            double mapq = -10.0 * log10(1.0 - pow(10.0, score) / total);                  // method 1
            const double mmq = -10.0 * log10(1.0 - (1.0 / A.cands[a].sum_move));          // method 2
            if (isnan(mapq) || isnan(mmq)) mapq = nan(""); else mapq = mapq < mmq ? mapq : mmq;  // math.Min propagates NaN
            if (!isnan(mapq)) mapq = 60.0 < mapq ? 60.0 : mapq;
            long long cs = -1, ce = -1;
            if (A.centro_start && ha.rid >= 0) { cs = A.centro_start[ha.rid]; ce = A.centro_end[ha.rid]; }
            if (ha.pos > cs && ha.pos <= ce) mapq = 0.0;
            A.cands[a].raw_mapq = mapq;
            A.cands[a].mapq = isnan(mapq) ? (int32_t)0x80000000 : (int32_t)mapq;
        }
    }
}

}  // namespace arachne

namespace arachne {
// candidates per RFA slot (range of the sort order)
__global__ void k_place_seg(PlaceArgs A, int n_slots, const int* __restrict__ bc_of_slot, unsigned long long* cnt) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s > n_slots) return;
    if (s == n_slots) { cnt[s] = 0; return; }
    const int b = bc_of_slot[s];
    cnt[s] = A.cand_off[2 * A.bc_off[b + 1]] - A.cand_off[2 * A.bc_off[b]];
}
}  // namespace arachne
