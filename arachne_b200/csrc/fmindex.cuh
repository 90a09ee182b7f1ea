// FM-index primitives over the reference's on-disk layout, resident in HBM.
//
// Layout (bwa/bwtindex.c:151-173, bwa/bwt.h:72-78): the .bwt payload is a sequence of 64-byte blocks, one per
// 128 BWT symbols: 4 x u64 cumulative A/C/G/T counts followed by 8 x u32 words of 16 two-bit symbols
// (symbol i of a word at bits (15-i)*2).  We keep exactly that image, 64-byte aligned, so one occ query is one
// aligned 64-byte line: four 128-bit loads (ld.global.nc.v4) per touched block.  The '$' is not stored:
// k -= (k >= primary) (bwa/bwt.c:114,177).
//
// Functions mirror bwa/bwt.c: occ4/2occ4 (:169-220), occ (:107-129), invPsi (:53-59), sa (:86-96),
// extend (:262-275), set_intv (bwt.h:80).  Counting uses popcount on 64-bit lanes instead of the reference's
// 256-entry byte LUT; results are identical integers.
#pragma once
#include "common.cuh"

namespace arachne {

struct OccBlock {  // 64 bytes
    uint64_t cnt[4];
    uint32_t w[8];
};
static_assert(sizeof(OccBlock) == 64, "occ block must be one 64-byte line");

struct FMIndex {
    const OccBlock* blocks;
    const uint64_t* sa;  // sa[0] == (uint64_t)-1, sampled every sa_intv rows (bwa/bwt.c:62-84)
    uint64_t primary, seq_len, n_sa;
    uint64_t L2[5];
    int sa_intv;
};

AR_HD OccBlock fm_load_block(const FMIndex& fm, uint64_t b) {
#if defined(__CUDA_ARCH__)
    const uint4* p = reinterpret_cast<const uint4*>(fm.blocks + b);
    uint4 v0 = __ldg(p), v1 = __ldg(p + 1), v2 = __ldg(p + 2), v3 = __ldg(p + 3);
    OccBlock o;
    o.cnt[0] = (uint64_t)v0.y << 32 | v0.x; o.cnt[1] = (uint64_t)v0.w << 32 | v0.z;
    o.cnt[2] = (uint64_t)v1.y << 32 | v1.x; o.cnt[3] = (uint64_t)v1.w << 32 | v1.z;
    o.w[0] = v2.x; o.w[1] = v2.y; o.w[2] = v2.z; o.w[3] = v2.w;
    o.w[4] = v3.x; o.w[5] = v3.y; o.w[6] = v3.z; o.w[7] = v3.w;
    return o;
#else
    return fm.blocks[b];
#endif
}

AR_HD int ar_popc64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}

// symbols [0, nvalid) of the block, nvalid in 1..128 -> per-base counts
AR_HD void fm_block_count(const OccBlock& o, int nvalid, uint32_t c[4]) {
    const uint64_t M = 0x5555555555555555ull;
    uint32_t cc = 0, cg = 0, ct = 0;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
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

AR_HD void fm_occ4_blk(const OccBlock& o, uint64_t kk, uint64_t cnt[4]) {
    uint32_t c[4];
    fm_block_count(o, (int)(kk & 127) + 1, c);
    cnt[0] = o.cnt[0] + c[0]; cnt[1] = o.cnt[1] + c[1]; cnt[2] = o.cnt[2] + c[2]; cnt[3] = o.cnt[3] + c[3];
}

// bwt_occ4, bwa/bwt.c:169-186
AR_HD void fm_occ4(const FMIndex& fm, uint64_t k, uint64_t cnt[4]) {
    if (k == (uint64_t)-1) { cnt[0] = cnt[1] = cnt[2] = cnt[3] = 0; return; }
    k -= (k >= fm.primary);
    OccBlock o = fm_load_block(fm, k >> 7);
    fm_occ4_blk(o, k, cnt);
}

// bwt_2occ4, bwa/bwt.c:189-220.  *n_blocks receives the number of 64-byte lines touched (roofline accounting).
struct OccStat {  // roofline accounting: bwt_2occ4 calls and the 64-byte blocks they touch
    unsigned calls = 0, blocks = 0;
};

AR_HD void fm_2occ4(const FMIndex& fm, uint64_t k, uint64_t l, uint64_t ck[4], uint64_t cl[4], OccStat* st = nullptr) {
    uint64_t k_ = k - (k >= fm.primary), l_ = l - (l >= fm.primary);
    if (st) { ++st->calls; st->blocks += (k == (uint64_t)-1 || l == (uint64_t)-1) ? (k != l ? 1 : 0) : ((l_ >> 7) != (k_ >> 7) ? 2 : 1); }
    if ((l_ >> 7) != (k_ >> 7) || k == (uint64_t)-1 || l == (uint64_t)-1) {
        fm_occ4(fm, k, ck);
        fm_occ4(fm, l, cl);
    } else {
        OccBlock o = fm_load_block(fm, k_ >> 7);
        fm_occ4_blk(o, k_, ck);
        fm_occ4_blk(o, l_, cl);
    }
}

// bwt_set_intv, bwa/bwt.h:80
AR_HD void fm_set_intv(const FMIndex& fm, int c, Intv& ik) {
    ik.x0 = fm.L2[c] + 1;
    ik.x2 = fm.L2[c + 1] - fm.L2[c];
    ik.x1 = fm.L2[3 - c] + 1;
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

// one LF step: bwt_invPsi, bwa/bwt.c:53-59 (B0 and occ come from the same 64-byte line)
AR_HD uint64_t fm_inv_psi(const FMIndex& fm, uint64_t k) {
    if (k == fm.primary) return 0;
    uint64_t x = k - (k > fm.primary);
    OccBlock o = fm_load_block(fm, x >> 7);
    int c = (o.w[(x & 127) >> 4] >> ((~x & 15) << 1)) & 3;
    uint32_t cnt[4];
    fm_block_count(o, (int)(x & 127) + 1, cnt);
    // bwt_occ(k, c): k == seq_len cannot reach here with k != primary unless k is the last row; handled identically
    // because x = k - 1 then addresses the last stored symbol.
    return fm.L2[c] + o.cnt[c] + cnt[c];
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
