// FM-index construction: <prefix>.pac -> <prefix>.bwt + <prefix>.sa, byte-identical to what `bwa index` writes
// (bwa/bwtindex.c:61-123 bwt_pac2bwt, :151-173 bwt_bwtupdate_core, bwa/bwt.c:62-84 bwt_cal_sa, dumps bwt.c:385-409).
//
// The reference builds the BWT of T = forward strand + reverse complement (bntseq.c:301-307) with a single-threaded
// induced-sorting / rope builder and then recovers the sampled SA by walking the whole BWT once (seq_len dependent
// LF steps): hours at 3.1 Gb.  Here the suffix array is produced directly, batch by batch, and the BWT and the SA
// samples are read off it:
//
//   1. T is packed 32 bases per u64 (MSB first).  The key of suffix i is its first 29 bases (58 bits) followed by a 5-bit
//      tag min(n - i, 29): padding past the end is 'A' (the smallest base) and a shorter suffix gets the smaller tag, which
//      is exactly the order "'$' sorts before everything" that is_sa / the rope builder implement.
//   2. Suffixes are cut into batches by their first 6 bases (4096 bins, grouped up to `batch_cap` suffixes); a batch is
//      one LSD radix sort of (key, position).  Batches come out in suffix order, so SA rows are assigned by a running
//      offset and no batch is ever revisited: the full SA (50 GB at 3.1 Gb) never exists.
//   3. Suffixes that still tie after 29 bases are refined 29 bases at a time: re-key the tied ones at the next depth and
//      sort (group, key); groups shrink until none is left.  On iid text a handful of suffixes reach round 1.
//   4. Row r of the finished batch gives BWT symbol T[SA[r]-1] (SA[r]==0 marks `primary`) and, when r % 32 == 0, the SA
//      sample.  Row 0 is the empty suffix.
//   5. The occ-interleaved .bwt payload (4 x u64 running counts + 8 words per 128 symbols, trailing count record) is
//      assembled on the device from the BWT words with popcounts and one exclusive scan per base.
//
// The per-element logic below is AR_HD and the driver is a template over a backend (for_each / sort / select / scan), so
// the same code runs on the device (index_build.cu: kernels + CUB) and, in the CPU test tier only, on the host
// (tests/hostsim) where it is pinned against the files the reference's own `bwa index` writes.
#pragma once
#include "common.cuh"

namespace arachne {
namespace ib {

constexpr int KEY_BASES = 29;    // bases per refinement round
constexpr int FINE_K = 6;        // batch bins = first 6 bases
constexpr int N_BINS = 1 << (2 * FINE_K);
constexpr int SA_INTV = 32;      // bwtindex.c:309

struct TextView {
    const uint64_t* w;  // ceil(n/32)+2 words, zero padded
    uint64_t n;         // 2 * l_pac
};

AR_HD int pac_base(const uint8_t* pac, uint64_t i) { return pac[i >> 2] >> ((~i & 3) << 1) & 3; }  // bntseq.h _get_pac
AR_HD int strand_base(const uint8_t* pac, uint64_t l_pac, uint64_t p) {                             // bntseq.c:301-307
    return p < l_pac ? pac_base(pac, p) : 3 - pac_base(pac, 2 * l_pac - 1 - p);
}
AR_HD int text_base(const TextView& t, uint64_t p) { return (int)(t.w[p >> 5] >> ((~p & 31) << 1) & 3); }

// the 32 bases starting at pos (zero = 'A' past the end)
AR_HD uint64_t text_window(const TextView& t, uint64_t pos) {
    const uint64_t i = pos >> 5;
    const int o = (int)(pos & 31) * 2;
    const uint64_t hi = t.w[i];
    return o ? (hi << o | t.w[i + 1] >> (64 - o)) : hi;
}
AR_HD uint32_t sfx_bin(const TextView& t, uint64_t pos) { return (uint32_t)(text_window(t, pos) >> (64 - 2 * FINE_K)); }
AR_HD uint64_t sfx_key(const TextView& t, uint64_t pos) {
    if (pos >= t.n) return 0;
    const uint64_t rem = t.n - pos;
    return (text_window(t, pos) >> 6) << 5 | (rem < (uint64_t)KEY_BASES ? rem : (uint64_t)KEY_BASES);
}

// ---- element functors (one call per index; the backend decides where they run) ----
struct PackText {  // i = output word
    const uint8_t* pac; uint64_t l_pac, n; uint64_t* w; uint64_t n_words;
    AR_HD void operator()(uint64_t i) const {
        uint64_t v = 0;
        const uint64_t p0 = i << 5;
        if (p0 + 32 <= l_pac && (p0 & 3) == 0) {  // forward strand, byte aligned: 8 pac bytes, first base in the top bits
            const uint8_t* s = pac + (p0 >> 2);
            for (int k = 0; k < 8; ++k) v = v << 8 | s[k];
        } else {
            for (int k = 0; k < 32; ++k) { const uint64_t p = p0 + k; v = v << 2 | (uint64_t)(p < n ? strand_base(pac, l_pac, p) : 0); }
        }
        w[i] = v;
    }
};
struct MakeKeys {
    TextView t; const uint64_t* pos; uint64_t depth; uint64_t* key;
    AR_HD void operator()(uint64_t j) const { key[j] = sfx_key(t, pos[j] + depth); }
};
struct FlagTies {  // flag[j] = element j continues the run of j-1 (same group, same key)
    const uint64_t* key; const uint32_t* grp; uint8_t* flag; uint32_t* not_flag;
    AR_HD void operator()(uint64_t j) const {
        const bool f = j > 0 && key[j] == key[j - 1] && (!grp || grp[j] == grp[j - 1]);
        flag[j] = f; not_flag[j] = !f;
    }
};
struct MarkTied {  // tied[j] = j belongs to a run of length > 1
    const uint8_t* flag; uint64_t n; uint8_t* tied;
    AR_HD void operator()(uint64_t j) const { tied[j] = flag[j] | (j + 1 < n ? flag[j + 1] : 0); }
};
struct Iota32 { uint32_t* a; AR_HD void operator()(uint64_t j) const { a[j] = (uint32_t)j; } };
template <class T, class I> struct Gather { const T* src; const I* idx; T* dst; AR_HD void operator()(uint64_t j) const { dst[j] = src[idx[j]]; } };
template <class T, class I> struct Scatter { const T* src; const I* idx; T* dst; AR_HD void operator()(uint64_t j) const { dst[idx[j]] = src[j]; } };
struct GatherRun {  // run id of a slot = inclusive count of run heads up to it
    const uint32_t* scan; const uint32_t* slot; uint32_t* dst;
    AR_HD void operator()(uint64_t j) const { dst[j] = scan[slot[j]]; }
};
struct EmitRows {  // finished batch: row = row0 + j
    TextView t; const uint64_t* pos; uint64_t row0; uint8_t* rows; uint64_t* sa; uint64_t* primary;
    AR_HD void operator()(uint64_t j) const {
        const uint64_t r = row0 + j, p = pos[j];
        if (p == 0) { *primary = r; rows[r] = 0; } else rows[r] = (uint8_t)text_base(t, p - 1);
        if ((r & (SA_INTV - 1)) == 0) sa[r / SA_INTV] = p;
    }
};
// payload word offsets (u32 units): block b = symbols [128b, 128b+128): counts at 16b, words at 16b+8..; trailing counts
AR_HD uint64_t bwt_payload_words(uint64_t n) { return (n + 15) / 16 + ((n + 127) / 128 + 1) * 8; }
struct WriteWords {  // q = BWT word (16 symbols, first symbol in the top bits: bwtindex.c:117-118)
    const uint8_t* rows; uint64_t n, primary; uint32_t* out;
    AR_HD void operator()(uint64_t q) const {
        uint32_t v = 0;
        for (int k = 0; k < 16; ++k) {
            const uint64_t x = q * 16 + k;
            v = v << 2 | (x < n ? rows[x + (x >= primary)] : 0u);  // the row of '$' is not stored (bwtindex.c is_bwt / rope walk)
        }
        out[(q >> 3) * 16 + 8 + (q & 7)] = v;
    }
};
struct CountBlocks {  // b in [0, n_blk]: symbols of block b per base (b == n_blk -> 0)
    const uint32_t* out; uint64_t n, n_blk; uint64_t* cnt;  // cnt[c * (n_blk+1) + b]
    AR_HD void operator()(uint64_t b) const {
        uint64_t c1 = 0, c2 = 0, c3 = 0, valid = 0;
        if (b < n_blk) {
            valid = n - b * 128 < 128 ? n - b * 128 : 128;
            const uint64_t nw = (valid + 15) / 16;
            for (uint64_t k = 0; k < nw; ++k) {
                const uint32_t x = out[b * 16 + 8 + k], lo = x & 0x55555555u, hi = (x >> 1) & 0x55555555u;
#if defined(__CUDA_ARCH__)
                c3 += __popc(hi & lo); c2 += __popc(hi & ~lo); c1 += __popc(~hi & lo);
#else
                c3 += __builtin_popcount(hi & lo); c2 += __builtin_popcount(hi & ~lo); c1 += __builtin_popcount(~hi & lo);
#endif
            }
        }
        const uint64_t s = n_blk + 1;
        cnt[b] = valid - c1 - c2 - c3; cnt[s + b] = c1; cnt[2 * s + b] = c2; cnt[3 * s + b] = c3;
    }
};
struct WriteCounts {  // b in [0, n_blk]: running counts before block b; b == n_blk is the trailing record (bwtindex.c:168-169)
    const uint64_t* pre; uint64_t n, n_blk; uint32_t* out;
    AR_HD void operator()(uint64_t b) const {
        const uint64_t s = n_blk + 1, at = b < n_blk ? b * 16 : bwt_payload_words(n) - 8;
        for (int c = 0; c < 4; ++c) { const uint64_t v = pre[c * s + b]; out[at + 2 * c] = (uint32_t)v; out[at + 2 * c + 1] = (uint32_t)(v >> 32); }
    }
};

struct BuildStats {
    uint64_t n = 0, primary = 0, L2[5] = {0, 0, 0, 0, 0};
    uint64_t n_batches = 0, max_batch = 0, tied_round1 = 0, max_rounds = 0;
};

// Growable typed scratch owned by the backend.
template <class B, class T> struct Buf {
    B& be; T* p = nullptr; uint64_t cap = 0;
    explicit Buf(B& b) : be(b) {}
    ~Buf() { if (p) be.release(p); }
    bool need(uint64_t n) {
        if (n <= cap) return true;
        if (p) be.release(p);
        cap = n + n / 4 + 16;
        p = (T*)be.acquire(cap * sizeof(T));
        if (!p) cap = 0;
        return p != nullptr;
    }
};

// Refine the ties left in (key, pos)[0..m) after the depth-0 sort.  pos is permuted in place.
template <class B>
bool refine_ties(B& be, const TextView& t, const uint64_t* key, uint64_t* pos, uint64_t m, BuildStats& st) {
    Buf<B, uint8_t> flag(be), tied(be);
    Buf<B, uint32_t> nf(be), scan(be), c_slot(be), c_grp(be), p1(be), p2(be), g1(be), g2(be), sel(be), n_slot(be), n_grp(be), iota(be);
    Buf<B, uint64_t> c_pos(be), c_key(be), k1(be), s_pos(be), s_key(be);
    if (!flag.need(m) || !tied.need(m) || !nf.need(m) || !scan.need(m) || !sel.need(m)) return false;
    be.for_each(m, FlagTies{key, nullptr, flag.p, nf.p});
    be.for_each(m, MarkTied{flag.p, m, tied.p});
    uint64_t T = be.select_flagged(tied.p, sel.p, m);
    if (!be.ok()) return false;
    if (T == 0) return true;
    st.tied_round1 += T;
    be.inclusive_sum(nf.p, scan.p, m);
    if (!c_slot.need(T) || !c_grp.need(T) || !c_pos.need(T)) return false;
    be.copy32(c_slot.p, sel.p, T);
    be.for_each(T, GatherRun{scan.p, c_slot.p, c_grp.p});
    be.for_each(T, Gather<uint64_t, uint32_t>{pos, c_slot.p, c_pos.p});
    uint64_t depth = KEY_BASES, rounds = 0;
    while (T) {
        ++rounds;
        if (!c_key.need(T) || !k1.need(T) || !iota.need(T) || !p1.need(T) || !p2.need(T) || !g1.need(T) || !g2.need(T) || !s_pos.need(T) || !s_key.need(T) ||
            !flag.need(T) || !tied.need(T) || !nf.need(T) || !scan.need(T) || !sel.need(T) || !n_slot.need(T) || !n_grp.need(T))
            return false;
        be.for_each(T, MakeKeys{t, c_pos.p, depth, c_key.p});
        be.for_each(T, Iota32{iota.p});
        be.sort_pairs(c_key.p, k1.p, iota.p, p1.p, T, 63);              // by key
        be.for_each(T, Gather<uint32_t, uint32_t>{c_grp.p, p1.p, g1.p});
        be.sort_pairs(g1.p, g2.p, p1.p, p2.p, T, 32);                    // then, stably, by group
        be.for_each(T, Gather<uint64_t, uint32_t>{c_pos.p, p2.p, s_pos.p});
        be.for_each(T, Gather<uint64_t, uint32_t>{c_key.p, p2.p, s_key.p});
        // groups are contiguous slot ranges in slot order, so the i-th element in (group, key) order lands in the i-th slot
        be.for_each(T, Scatter<uint64_t, uint32_t>{s_pos.p, c_slot.p, pos});
        be.for_each(T, FlagTies{s_key.p, g2.p, flag.p, nf.p});
        be.for_each(T, MarkTied{flag.p, T, tied.p});
        const uint64_t T2 = be.select_flagged(tied.p, sel.p, T);
        if (!be.ok()) return false;
        if (T2) {
            be.inclusive_sum(nf.p, scan.p, T);
            be.for_each(T2, Gather<uint32_t, uint32_t>{c_slot.p, sel.p, n_slot.p});
            be.for_each(T2, Gather<uint32_t, uint32_t>{scan.p, sel.p, n_grp.p});
            be.for_each(T2, Gather<uint64_t, uint32_t>{s_pos.p, sel.p, c_pos.p});  // c_pos is free: s_pos holds this round's order
            be.copy32(c_slot.p, n_slot.p, T2);
            be.copy32(c_grp.p, n_grp.p, T2);
        }
        T = T2;
        depth += KEY_BASES;
        if (depth > t.n + KEY_BASES) { be.fail("suffix refinement did not converge"); return false; }
    }
    if (rounds > st.max_rounds) st.max_rounds = rounds;
    return be.ok();
}

// d_pac: l_pac/4+1 bytes on the backend.  Outputs (backend memory, caller releases): payload (u32 x bwt_payload_words(n)),
// sa (u64 x (n+32)/32, sa[0] = -1).
template <class B>
bool build_fm_index(B& be, const uint8_t* d_pac, uint64_t l_pac, uint64_t batch_cap, uint32_t** payload_out, uint64_t** sa_out, BuildStats& st) {
    const uint64_t n = 2 * l_pac;
    if (n == 0) { be.fail("empty reference"); return false; }
    if (batch_cap == 0) batch_cap = 1ull << 28;
    st.n = n;
    const uint64_t n_words = (n + 31) / 32 + 2;
    uint64_t* w = (uint64_t*)be.acquire(n_words * 8);
    if (!w) return false;
    be.for_each(n_words, PackText{d_pac, l_pac, n, w, n_words});
    TextView t{w, n};
    // batches: consecutive bins up to batch_cap suffixes
    uint64_t hist[N_BINS];
    be.histogram(t, hist);
    if (!be.ok()) return false;
    struct Batch { uint32_t b0, b1; uint64_t size; };
    Batch batches[N_BINS];
    uint64_t nb = 0, max_batch = 0, total = 0;
    for (uint32_t b = 0; b < (uint32_t)N_BINS;) {
        uint64_t sz = hist[b]; uint32_t e = b + 1;
        while (e < (uint32_t)N_BINS && sz + hist[e] <= batch_cap) sz += hist[e++];
        if (sz) { batches[nb++] = Batch{b, e, sz}; if (sz > max_batch) max_batch = sz; total += sz; }
        b = e;
    }
    if (total != n) { be.fail("bin histogram does not cover the text"); return false; }
    if (max_batch >= (1ull << 32)) { be.fail("a suffix bin exceeds 2^32 entries; not supported"); return false; }
    st.n_batches = nb; st.max_batch = max_batch;
    const uint64_t n_sa = (n + SA_INTV) / SA_INTV;
    uint8_t* rows = (uint8_t*)be.acquire(n + 1);
    uint64_t* sa = (uint64_t*)be.acquire(n_sa * 8);
    uint64_t* d_primary = (uint64_t*)be.acquire(8);
    uint64_t *keyA = (uint64_t*)be.acquire(max_batch * 8), *keyB = (uint64_t*)be.acquire(max_batch * 8);
    uint64_t *posA = (uint64_t*)be.acquire(max_batch * 8), *posB = (uint64_t*)be.acquire(max_batch * 8);
    if (!rows || !sa || !d_primary || !keyA || !keyB || !posA || !posB) return false;
    uint64_t row0 = 1;  // row 0 is the empty suffix
    for (uint64_t k = 0; k < nb; ++k) {
        const uint64_t m = be.select_bins(t, batches[k].b0, batches[k].b1, posA);
        if (!be.ok()) return false;
        if (m != batches[k].size) { be.fail("batch size mismatch"); return false; }
        be.for_each(m, MakeKeys{t, posA, 0, keyA});
        be.sort_pairs(keyA, keyB, posA, posB, m, 63);
        if (!refine_ties(be, t, keyB, posB, m, st)) return false;
        be.for_each(m, EmitRows{t, posB, row0, rows, sa, d_primary});
        row0 += m;
    }
    be.release(keyA); be.release(keyB); be.release(posA); be.release(posB);
    // row 0: '$' suffix, preceded by the last base; sa[0] = -1 (bwt.c:83)
    {
        uint8_t last = 0;  // read through the backend: the text lives there
        uint64_t wlast = 0;
        be.to_host(&wlast, w + ((n - 1) >> 5), 8);
        last = (uint8_t)(wlast >> ((~(n - 1) & 31) << 1) & 3);
        be.to_device(rows, &last, 1);
        const uint64_t minus1 = ~0ull;
        be.to_device(sa, &minus1, 8);
        be.to_host(&st.primary, d_primary, 8);
    }
    if (!be.ok()) return false;
    be.release(w); be.release(d_primary);
    const uint64_t n_blk = (n + 127) / 128, pw = bwt_payload_words(n);
    uint32_t* out = (uint32_t*)be.acquire(pw * 4);
    uint64_t* cnt = (uint64_t*)be.acquire((n_blk + 1) * 4 * 8);
    uint64_t* pre = (uint64_t*)be.acquire((n_blk + 1) * 4 * 8);
    if (!out || !cnt || !pre) return false;
    be.zero(out, pw * 4);
    be.for_each((n + 15) / 16, WriteWords{rows, n, st.primary, out});
    be.release(rows);
    be.for_each(n_blk + 1, CountBlocks{out, n, n_blk, cnt});
    for (int c = 0; c < 4; ++c) be.exclusive_sum(cnt + c * (n_blk + 1), pre + c * (n_blk + 1), n_blk + 1);
    be.for_each(n_blk + 1, WriteCounts{pre, n, n_blk, out});
    uint64_t tot[4];
    for (int c = 0; c < 4; ++c) be.to_host(&tot[c], pre + c * (n_blk + 1) + n_blk, 8);
    if (!be.ok()) return false;
    st.L2[0] = 0;
    for (int c = 0; c < 4; ++c) st.L2[c + 1] = st.L2[c] + tot[c];  // bwtindex.c:83-86
    if (st.L2[4] != n) { be.fail("BWT symbol counts do not add up"); return false; }
    be.release(cnt); be.release(pre);
    *payload_out = out; *sa_out = sa;
    return true;
}

}  // namespace ib
}  // namespace arachne
