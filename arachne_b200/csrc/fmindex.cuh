// FM-index primitives, B200 layout.
//
// The reference keeps the BWT as 64-byte blocks of 128 symbols: 4 x u64 cumulative A/C/G/T counts + 8 x u32 words of
// 16 two-bit symbols (bwa/bwtindex.c:151-173, bwa/bwt.h:72-78), so one rank query reads one 64-byte block = two 32-byte
// DRAM sectors.  In HBM we keep the same information re-cut to the sector size: one 32-byte block per 64 symbols,
// 4 x u32 cumulative counts + 4 x u32 symbol words (same bit order: symbol i of a word at bits (15-i)*2).  A rank query
// is then exactly ONE 32-byte sector (one 256-bit access), half the DRAM traffic per query at the same index size
// (seq_len/2 bytes).  u32 counts hold for every base count < 2^32 (seq_len up to ~1.6e10 on uniform text; checked at load).
// The '$' is not stored: k -= (k >= primary) (bwa/bwt.c:114,177).
//
// Functions mirror bwa/bwt.c: occ4/2occ4 (:169-220), occ (:107-129), invPsi (:53-59), sa (:86-96), extend (:262-275),
// set_intv (bwt.h:80).  Counting uses popcount on 64-bit lanes instead of the reference's 256-entry byte LUT; the
// integers are identical.
#pragma once
#include "common.cuh"

namespace arachne {

struct Blk32 {  // 32 bytes = one DRAM sector: 64 BWT symbols
    uint32_t cnt[4];
    uint32_t w[4];
};
static_assert(sizeof(Blk32) == 32, "occ block must be one 32-byte sector");

struct FMIndex {
    const Blk32* blocks;
    const uint64_t* sa;  // sa[0] == (uint64_t)-1, sampled every sa_intv rows (bwa/bwt.c:62-84)
    uint64_t primary, seq_len, n_sa;
    uint64_t L2[5];
    int sa_intv;
    // bi-interval of every kmer_k-mer (16 bytes each, packed like the seeding kernel's list entries), or null: lets the
    // LAST-like pass (bwt_seed_strategy1) start kmer_k bases in instead of kmer_k-1 dependent bwt_extend calls (index.cuh)
    const void* kmer;      // level kmer_k
    const void* kmer_all;  // levels 1..kmer_k back to back: level L (4^L entries) starts at entry (4^L - 4) / 3
    int kmer_k;
    // direct forward extension of unique matches (seed_kernel.cuh): inverse SA (row of every text position) and the text
    // T.revcomp(T) itself, 32 bases per u64 (base p at bits 62 - 2*(p & 31) of word p >> 5); both null when HBM is short
    const uint64_t* isa;
    const uint64_t* text;
};

struct OccStat {  // roofline accounting in the reference's own units: bwt_2occ4 calls and the 64-byte blocks they touch
    unsigned calls = 0, blocks = 0;
};

AR_HD int ar_popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}

// Re-cut one reference block (4 x u64 counts + 8 words, given as 16 little-endian u32) into two sector blocks.
AR_HD void fm_recut_block(const uint32_t* raw16, Blk32* out2) {
    uint64_t c[4];
    for (int i = 0; i < 4; ++i) c[i] = (uint64_t)raw16[2 * i + 1] << 32 | raw16[2 * i];
    for (int i = 0; i < 4; ++i) { out2[0].cnt[i] = (uint32_t)c[i]; out2[0].w[i] = raw16[8 + i]; out2[1].w[i] = raw16[12 + i]; }
    const uint64_t M = 0x5555555555555555ull;
    uint32_t cc = 0, cg = 0, ct = 0;
    for (int i = 0; i < 2; ++i) {
        uint64_t x = (uint64_t)raw16[8 + 2 * i] << 32 | raw16[9 + 2 * i];
        uint64_t lo = x & M, hi = (x >> 1) & M;
        ct += ar_popc64(hi & lo); cg += ar_popc64(hi & ~lo); cc += ar_popc64(~hi & lo);
    }
    out2[1].cnt[0] = (uint32_t)c[0] + (64 - cc - cg - ct);
    out2[1].cnt[1] = (uint32_t)c[1] + cc; out2[1].cnt[2] = (uint32_t)c[2] + cg; out2[1].cnt[3] = (uint32_t)c[3] + ct;
}

AR_HD Blk32 fm_load_block(const FMIndex& fm, uint64_t b) {
#if defined(__CUDA_ARCH__)
    // one 256-bit load (LDG.E.256 on sm_100): the whole sector in a single L1/L2 request instead of two 128-bit ones
    Blk32 o;
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(o.cnt[0]), "=r"(o.cnt[1]), "=r"(o.cnt[2]), "=r"(o.cnt[3]), "=r"(o.w[0]), "=r"(o.w[1]), "=r"(o.w[2]), "=r"(o.w[3])
                 : "l"(fm.blocks + b));
    return o;
#else
    return fm.blocks[b];
#endif
}

// symbols [0, nvalid) of the block, nvalid in 1..64 -> per-base counts
AR_HD void fm_block_count(const Blk32& o, int nvalid, uint32_t c[4]) {
    const uint64_t M = 0x5555555555555555ull;
    uint32_t cc = 0, cg = 0, ct = 0;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        int v = nvalid - 32 * i;
        v = v < 0 ? 0 : (v > 32 ? 32 : v);
        uint64_t m = v == 0 ? 0ull : (M << ((32 - v) << 1)) & M;  // the v leading symbols of this 64-bit lane
        uint64_t x = (uint64_t)o.w[2 * i] << 32 | o.w[2 * i + 1];
        uint64_t lo = x & m, hi = (x >> 1) & m;
        ct += ar_popc64(hi & lo);
        cg += ar_popc64(hi & ~lo);
        cc += ar_popc64(~hi & lo);
    }
    c[1] = cc; c[2] = cg; c[3] = ct;
    c[0] = (uint32_t)nvalid - cc - cg - ct;
}

AR_HD void fm_occ4_blk(const Blk32& o, uint64_t kk, uint64_t cnt[4]) {
    uint32_t c[4];
    fm_block_count(o, (int)(kk & 63) + 1, c);
    cnt[0] = (uint64_t)o.cnt[0] + c[0]; cnt[1] = (uint64_t)o.cnt[1] + c[1]; cnt[2] = (uint64_t)o.cnt[2] + c[2]; cnt[3] = (uint64_t)o.cnt[3] + c[3];
}

// bwt_occ4, bwa/bwt.c:169-186
AR_HD void fm_occ4(const FMIndex& fm, uint64_t k, uint64_t cnt[4]) {
    if (k == (uint64_t)-1) { cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0; return; }
    k -= (k >= fm.primary);
    Blk32 o = fm_load_block(fm, k >> 6);
    fm_occ4_blk(o, k, cnt);
}

// blocks of the reference layout (64 bytes / 128 symbols) that bwt_2occ4(k, l) touches, bwa/bwt.c:189-220
AR_HD unsigned fm_ref_blocks(const FMIndex& fm, uint64_t k, uint64_t l) {
    if (k == (uint64_t)-1 || l == (uint64_t)-1) return (k != (uint64_t)-1) + (l != (uint64_t)-1);
    uint64_t k_ = k - (k >= fm.primary), l_ = l - (l >= fm.primary);
    return (l_ >> 7) != (k_ >> 7) ? 2 : 1;
}

// bwt_2occ4, bwa/bwt.c:189-220
AR_HD void fm_2occ4(const FMIndex& fm, uint64_t k, uint64_t l, uint64_t ck[4], uint64_t cl[4], OccStat* st = nullptr) {
    uint64_t k_ = k - (k >= fm.primary), l_ = l - (l >= fm.primary);
    if (st) { ++st->calls; st->blocks += fm_ref_blocks(fm, k, l); }
    if ((l_ >> 6) != (k_ >> 6) || k == (uint64_t)-1 || l == (uint64_t)-1) {
        fm_occ4(fm, k, ck);
        fm_occ4(fm, l, cl);
    } else {
        Blk32 o = fm_load_block(fm, k_ >> 6);
        fm_occ4_blk(o, k_, ck);
        fm_occ4_blk(o, l_, cl);
    }
}

// register-friendly selects (dynamic indexing of kernel-parameter or register arrays would spill to local memory)
AR_HD uint64_t fm_L2(const FMIndex& fm, int c) {
    return c == 0 ? fm.L2[0] : c == 1 ? fm.L2[1] : c == 2 ? fm.L2[2] : c == 3 ? fm.L2[3] : fm.L2[4];
}
template <class T> AR_HD T sel4(const T a[4], int c) { return c == 0 ? a[0] : c == 1 ? a[1] : c == 2 ? a[2] : a[3]; }

// bwt_set_intv, bwa/bwt.h:80
AR_HD void fm_set_intv(const FMIndex& fm, int c, Intv& ik) {
    ik.x0 = fm_L2(fm, c) + 1;
    ik.x2 = fm_L2(fm, c + 1) - fm_L2(fm, c);
    ik.x1 = fm_L2(fm, 3 - c) + 1;
    ik.info = 0;
}

// bwt_extend, bwa/bwt.c:262-275.  is_back=1 extends to the left of the match (prepends a base), 0 appends the
// complement on the reverse strand.  x(!is_back) is the coordinate that is looked up.
AR_HD void fm_extend(const FMIndex& fm, const Intv& ik, Intv ok[4], int is_back, OccStat* st = nullptr) {
    uint64_t tk[4], tl[4];
    uint64_t xa = is_back ? ik.x0 : ik.x1;  // x[!is_back]
    uint64_t xb = is_back ? ik.x1 : ik.x0;  // x[is_back]
    fm_2occ4(fm, xa - 1, xa - 1 + ik.x2, tk, tl, st);
    uint64_t na[4], nb[4], ns[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        na[i] = fm.L2[i] + 1 + tk[i];
        ns[i] = tl[i] - tk[i];
    }
    nb[3] = xb + (xa <= fm.primary && xa + ik.x2 - 1 >= fm.primary);
    nb[2] = nb[3] + ns[3];
    nb[1] = nb[2] + ns[2];
    nb[0] = nb[1] + ns[1];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (is_back) { ok[i].x0 = na[i]; ok[i].x1 = nb[i]; }
        else { ok[i].x1 = na[i]; ok[i].x0 = nb[i]; }
        ok[i].x2 = ns[i];
    }
}

// one LF step: bwt_invPsi, bwa/bwt.c:53-59 (B0 and occ come from the same sector)
AR_HD uint64_t fm_inv_psi(const FMIndex& fm, uint64_t k) {
    if (k == fm.primary) return 0;
    uint64_t x = k - (k > fm.primary);
    Blk32 o = fm_load_block(fm, x >> 6);
    int c = (sel4(o.w, (int)(x & 63) >> 4) >> ((~x & 15) << 1)) & 3;
    uint32_t cnt[4];
    fm_block_count(o, (int)(x & 63) + 1, cnt);
    return fm_L2(fm, c) + sel4(o.cnt, c) + sel4(cnt, c);
}

// bwt_sa, bwa/bwt.c:86-96.  Row-sampled SA: geometric walk length (mean ~ sa_intv-1).
AR_HD uint64_t fm_sa(const FMIndex& fm, uint64_t k) {
    uint64_t sa = 0, mask = (uint64_t)fm.sa_intv - 1;
    while (k & mask) {
        ++sa;
        k = fm_inv_psi(fm, k);
    }
    return sa + fm.sa[k / fm.sa_intv];
}

}  // namespace arachne
