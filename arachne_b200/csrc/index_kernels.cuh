// The kernels that build the HBM-only index structures at load time (dense SA, k-mer bi-interval tables, inverse SA, both-strand text):
// independent grid-stride loops, kept in a header of their own so that the TEST-ONLY host emulation of the seeding kernel
// (tests/hostsim/hostsim_seed.cpp) builds ITS index structures with the same code.  Included by index.cuh.
#pragma once
#include "fmindex.cuh"

namespace arachne {

// HBM-resident SA at a denser sampling than the file's (bwtindex.c:309 samples every 32nd row; 180 GB of HBM hold every
// row of a 3.1 Gb index in 50 GB): bwt_sa (bwt.c:86-96) then needs (d-1)/2 LF steps on average instead of 15.5 -- none at
// d = 1 -- and returns the same integers.  The rows between two file samples form one LF chain (bwt_cal_sa, bwt.c:62-84,
// walks them in this order), so one lane per file sample fills its chain: seq_len LF steps in total.
__global__ void __launch_bounds__(256) k_sa_densify(FMIndex fm, uint64_t* dense, int d) {
    const uint64_t n_file = fm.n_sa, mask = (uint64_t)fm.sa_intv - 1, dmask = (uint64_t)d - 1;
    for (uint64_t j = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; j < n_file; j += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t k = j * (uint64_t)fm.sa_intv;
        uint64_t v = j == 0 ? fm.seq_len : fm.sa[j];  // row 0 is the empty suffix; its stored value stays the reference's -1 (bwt.c:83)
        if ((k & dmask) == 0) dense[k / d] = j == 0 ? (uint64_t)-1 : v;
        for (;;) {
            k = fm_inv_psi(fm, k);
            if ((k & mask) == 0) break;
            --v;
            if ((k & dmask) == 0) dense[k / d] = v;
        }
    }
}

// Bi-intervals of every L-mer, L = 1..K, level by level: level 1 is bwt_set_intv (bwt.h:80); an entry of level L+1 is one
// forward extension (bwt_extend(.., is_back = 0) with the complement base, bwt.c:304-307) of its level-L prefix -- the
// chain the reference's seeding would walk.  Entry = the seeding kernel's packed list entry (x0, x1, x2: 40 bits each).
__device__ __forceinline__ uint4 kmer_pack(const Intv& o) {
    const uint64_t lo = o.x0 | (o.x1 << 40), hi = (o.x1 >> 24) | (o.x2 << 16);
    return make_uint4((unsigned)lo, (unsigned)(lo >> 32), (unsigned)hi, (unsigned)(hi >> 32));
}
__host__ __device__ inline uint64_t kmer_level_off(int L) { return ((1ull << (2 * L)) - 4) / 3; }
__global__ void __launch_bounds__(256) k_kmer_level(FMIndex fm, uint4* all, int L) {  // fills level L+1 from level L (L = 0: level 1)
    if (L == 0) {
        if (blockIdx.x == 0 && threadIdx.x < 4) { Intv ik; fm_set_intv(fm, (int)threadIdx.x, ik); all[threadIdx.x] = kmer_pack(ik); }
        return;
    }
    const uint64_t n = 1ull << (2 * L);
    const uint4* src = all + kmer_level_off(L);
    uint4* dst = all + kmer_level_off(L + 1);
    for (uint64_t p = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; p < n; p += (uint64_t)gridDim.x * blockDim.x) {
        const uint4 v = src[p];
        const uint64_t lo = (uint64_t)v.y << 32 | v.x, hi = (uint64_t)v.w << 32 | v.z;
        Intv ik, ok[4];
        ik.x0 = lo & 0xffffffffffull; ik.x1 = (lo >> 40) | ((hi & 0xffffull) << 24); ik.x2 = (hi >> 16) & 0xffffffffffull; ik.info = 0;
        fm_extend(fm, ik, ok, 0);
        for (int b = 0; b < 4; ++b) dst[p * 4 + b] = kmer_pack(ok[3 - b]);
    }
}

// Inverse of the dense SA and the 2-bit text of both strands (FMIndex::isa / ::text): once a match is unique, its forward
// extension (bwt.c:300-316) is a comparison against the text at SA[row], and the reverse-strand row the bi-interval
// carries is ISA[n - pos - len] (the text equals its own reverse complement).
__global__ void __launch_bounds__(256) k_isa_fill(const uint64_t* __restrict__ sa, uint64_t n_rows, uint64_t* isa) {
    for (uint64_t k = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x + 1; k < n_rows; k += (uint64_t)gridDim.x * blockDim.x) isa[sa[k]] = k;
}
__global__ void __launch_bounds__(256) k_text_both(const uint8_t* __restrict__ pac, uint64_t l_pac, uint64_t n_words, uint64_t* text) {
    const uint64_t n = 2 * l_pac;
    for (uint64_t w = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; w < n_words; w += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t v = 0;
        for (int b = 0; b < 32; ++b) {
            const uint64_t p = w * 32 + b;
            if (p >= n) break;
            const uint64_t f = p < l_pac ? p : n - 1 - p;
            unsigned c = (pac[f >> 2] >> ((~f & 3) << 1)) & 3;
            if (p >= l_pac) c = 3 - c;
            v |= (uint64_t)c << (62 - 2 * b);
        }
        text[w] = v;
    }
}

}  // namespace arachne
