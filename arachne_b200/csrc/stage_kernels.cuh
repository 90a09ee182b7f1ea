// Stage kernels between seeding and extension: one SA row per seed occurrence (k_expand), bwt_sa (k_sa), chaining (k_chain).
// Included by pipeline.cu inside the build; tests/hostsim compiles the same text for the host.
#pragma once
#include "fmindex.cuh"
#include "counters.cuh"
#include "chain.cuh"
#include "warp_ctx.cuh"
namespace arachne {

// ------------------------------------------------------------------------------------------------ k_expand / k_sa
__global__ void k_expand(MemOpt opt, int n_reads, const Intv* __restrict__ arena, unsigned long long* __restrict__ intv_off, int cap,
                         const int* __restrict__ intv_n, const unsigned long long* __restrict__ sa_off, uint64_t* rows) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += gridDim.x * blockDim.x) {
        intv_off[r] = (unsigned long long)r * cap;
        const Intv* iv = arena + (size_t)r * cap;
        unsigned long long o = sa_off[r];
        for (int i = 0; i < intv_n[r]; ++i) {
            const Intv p = iv[i];
            int64_t step = p.x2 > (uint64_t)opt.max_occ ? (int64_t)(p.x2 / opt.max_occ) : 1;
            int count = 0;
            for (int64_t k = 0; k < (int64_t)p.x2 && count < opt.max_occ; k += step, ++count) rows[o++] = p.x0 + k;
        }
    }
}

// bwt_sa for every SA row of the batch.  Walk lengths are geometric (rows, not text positions, are sampled), so lanes
// are persistent: a lane that reaches a sampled row stores its result and takes the next row, and every loop iteration
// is one LF step (one 32-byte sector) for every live lane.  Rows are claimed from the global queue a warp-chunk at a
// time (one atomic per SA_CLAIM rows) and dealt to idle lanes by ballot rank: a per-lane atomic on the single queue
// counter serialises in L2 and was the kernel's bound on the 3.1 Gb index (15 M same-address atomics per 1 M pairs).
constexpr unsigned SA_CLAIM = 128;
__global__ void __launch_bounds__(256) k_sa(FMIndex fm, unsigned long long n, uint64_t* rows, unsigned long long* next, Counters* cnt) {
    const uint64_t mask = (uint64_t)fm.sa_intv - 1;
    const unsigned lane = threadIdx.x & 31;
    unsigned long long steps = 0, idx = 0;
    unsigned long long wbase = 0, wend = 0;  // warp-uniform: the claimed rows not yet dealt
    bool exhausted = false;                  // warp-uniform: the queue is empty
    uint64_t k = 0, sa = 0;
    bool live = false;
    for (;;) {
        const unsigned idle = __ballot_sync(FULL, !live);
        if (idle && !exhausted) {
            if (wbase == wend) {
                unsigned long long b = 0;
                if (lane == 0) b = atomicAdd(next, (unsigned long long)SA_CLAIM);
                b = __shfl_sync(FULL, b, 0);
                if (b >= n) exhausted = true;
                else { wbase = b; wend = b + SA_CLAIM < n ? b + SA_CLAIM : n; }
            }
            if (!exhausted) {
                const unsigned rank = __popc(idle & ((1u << lane) - 1u));
                const unsigned long long avail = wend - wbase;
                if (!live && rank < avail) { idx = wbase + rank; k = rows[idx]; sa = 0; live = true; }
                const unsigned want = __popc(idle);
                wbase += want < avail ? want : avail;
            }
        }
        if (live) {
            if (k & mask) { ++sa; k = fm_inv_psi(fm, k); }
            else { rows[idx] = sa + fm.sa[k / fm.sa_intv]; steps += sa; live = false; }
        }
        if (exhausted && __ballot_sync(FULL, live) == 0) break;
    }
    steps = __reduce_add_sync(FULL, (unsigned)steps);  // per-thread totals stay far below 2^32 / 32
    if (lane == 0 && steps) atomicAdd(&cnt->lf_steps, steps);
}

// ------------------------------------------------------------------------------------------------ k_chain
struct ChainBufs {
    Seed* seeds;     // [total_sa]   chaining order, linked
    Seed* kseeds;    // [total_sa]   seeds of kept chains, contiguous per chain
    Chain* chains;   // [total_sa]
    BNode* nodes;    // [total_sa/4 + 4*n_reads]
    int32_t* order;  // [total_sa]   chain ids: in-order, then sorted/filtered (kept chains first n_kept)
    int32_t* kept;   // [total_sa]   scratch
    int* n_chain;    // [n_reads] kept chains
    int* n_kseed;    // [n_reads] seeds in kept chains (region capacity)
};

__global__ void __launch_bounds__(128) k_chain(MemOpt opt, RefView ref, const int64_t* __restrict__ off, int n_reads, const Intv* __restrict__ arena,
                                               const unsigned long long* __restrict__ intv_off, const int* __restrict__ intv_n,
                                               const unsigned long long* __restrict__ sa_off, const uint64_t* __restrict__ sa_vals,
                                               ChainBufs b, Counters* cnt) {
    for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n_reads; r += gridDim.x * blockDim.x) {
        const int len = (int)(off[r + 1] - off[r]);
        const unsigned long long o = sa_off[r];
        const int nsa = (int)(sa_off[r + 1] - o);
        int nk = 0, nks = 0;
        if (nsa > 0) {
            Seed* seeds = b.seeds + o;
            Chain* chains = b.chains + o;
            int32_t* order = b.order + o;
            BNode* nodes = b.nodes + (o / 4 + 4ull * r);
            float fr;
            int n = chain_seeds(opt, ref, len, arena + intv_off[r], intv_n[r], sa_vals + o, seeds, chains, nodes, nsa / 4 + 4, order, &fr);
            if (n < 0) { raise(cnt, ARACHNE_ERR_OVERFLOW, r); n = 0; }
            nk = chain_filter(opt, n, order, chains, seeds, b.kept + o);
            Seed* ks = b.kseeds + o;
            for (int i = 0; i < nk; ++i) {  // materialise the kept chains' seeds contiguously
                Chain& c = chains[order[i]];
                int s = c.first_seed;
                c.first_seed = nks;
                for (int j = 0; j < c.n; ++j, s = seeds[s].next) ks[nks++] = seeds[s];
            }
        }
        b.n_chain[r] = nk; b.n_kseed[r] = nks;
    }
}

// ------------------------------------------------------------------------------------------------ k_occ (stage entry point arachne_occ_batch)
__global__ void k_occ(FMIndex fm, long n, const uint64_t* k, const uint64_t* l, uint64_t* cnt) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x)
        fm_2occ4(fm, k[i], l[i], cnt + 8 * i, cnt + 8 * i + 4);
}

}  // namespace arachne
