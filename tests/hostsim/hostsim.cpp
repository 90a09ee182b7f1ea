// TEST-ONLY host simulation: compiles the product's thread-sequential device logic (csrc/*.cuh, AR_HD functions)
// with g++ so the CPU test tier can pin seeding / chaining / the extension driver / region de-duplication against
// the oracle without a GPU.  The warp-cooperative DP primitives cannot run on the host; here they are replaced by
// the oracle's scalar kernels (oracle/restate_ksw.c).  Nothing in the shipped library links this file.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#define ARACHNE_HOSTSIM 1
#include "../../arachne_b200/csrc/bwt_image.h"
#include "../../arachne_b200/csrc/gorand.cuh"
#include "smem.cuh"
#include "../../arachne_b200/csrc/ext_driver.cuh"
#include "../../arachne_b200/csrc/rescue.cuh"
#include "../../arachne_b200/csrc/cigar.cuh"
#include "../../include/arachne_cuda.h"

extern "C" {
typedef struct { int a, b, o_del, e_del, o_ins, e_ins; } orc_sc_t;
int orc_extend(const orc_sc_t*, int, const uint8_t*, int, const uint8_t*, int, int, int, int, int*, uint64_t*);
int orc_global(const orc_sc_t*, int, const uint8_t*, int, const uint8_t*, int, uint32_t*, int, int*);
void orc_local(const orc_sc_t*, int, const uint8_t*, int, const uint8_t*, int, int*);
int arachne_opts_default_host(arachne_mem_opt_t* o);
}

using namespace arachne;

struct HostIndex {
    std::vector<uint8_t> bwt, sa_raw, pac, image;
    std::vector<uint64_t> sa;
    std::vector<Contig> contigs;
    FMIndex fm;
    RefView ref;
};

static bool slurp(const std::string& p, std::vector<uint8_t>& v) {
    FILE* f = fopen(p.c_str(), "rb"); if (!f) return false;
    fseek(f, 0, SEEK_END); long n = ftell(f); fseek(f, 0, SEEK_SET);
    v.resize(n); size_t g = fread(v.data(), 1, n, f); fclose(f); return g == (size_t)n;
}

static long g_score_only_dp = 0;
struct HostCtx {
    uint64_t cells = 0, calls = 0;
    bool leader() const { return true; }
    void sync() const {}
    int bcast(int v) const { return v; }
    void fetch_ref(const RefView& ref, int64_t beg, int64_t end, uint8_t* dst) const { for (int64_t k = 0; k < end - beg; ++k) dst[k] = ref_base(ref, beg + k); }
    int ksw_extend(const MemOpt& o, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts, int w, int eb, int zd, int h0, int* qle,
                   int* tle, int* gtle, int* gscore, int* max_off) {
        std::vector<uint8_t> qq(qlen + 1), tt(tlen + 1);
        for (int i = 0; i < qlen; ++i) qq[i] = q[(long)i * qs];
        for (int i = 0; i < tlen; ++i) tt[i] = t[(long)i * ts];
        orc_sc_t s = {o.a, o.b, o.o_del, o.e_del, o.o_ins, o.e_ins};
        int out5[5]; uint64_t c = 0;
        int r = orc_extend(&s, qlen, qq.data(), tlen, tt.data(), w, eb, zd, h0, out5, &c);
        cells += c; ++calls;
        *qle = out5[0]; *tle = out5[1]; *gtle = out5[2]; *gscore = out5[3]; *max_off = out5[4];
        return r;
    }
    int ksw_global(const MemOpt& o, int qlen, const uint8_t* q, int qs, int tlen, const uint8_t* t, int ts, int w, uint32_t* cigar, int cap, int* nc) {
        if (!cigar) ++g_score_only_dp;  // the score-only use is mem_patch_reg's re-alignment of two colinear regions (csrc/extend.cuh::gen_score)
        std::vector<uint8_t> qq(qlen + 1), tt(tlen + 1);
        for (int i = 0; i < qlen; ++i) qq[i] = q[(long)i * qs];
        for (int i = 0; i < tlen; ++i) tt[i] = t[(long)i * ts];
        orc_sc_t s = {o.a, o.b, o.o_del, o.e_del, o.o_ins, o.e_ins};
        int n = 0;
        int r = orc_global(&s, qlen, qq.data(), tlen, tt.data(), w, cigar, cap, &n);
        if (nc) *nc = n;
        return r;
    }
    void ksw_local(const MemOpt& o, int qlen, const uint8_t* ms, int is_rev, int tlen, const uint8_t* t, int xtra, LocalAln* out) {
        std::vector<uint8_t> qq(qlen + 1);
        for (int i = 0; i < qlen; ++i) qq[i] = is_rev ? (ms[qlen - 1 - i] < 4 ? 3 - ms[qlen - 1 - i] : 4) : ms[i];
        orc_sc_t s = {o.a, o.b, o.o_del, o.e_del, o.o_ins, o.e_ins};
        int o7[7];
        orc_local(&s, qlen, qq.data(), tlen, t, xtra, o7);
        out->score = o7[0]; out->te = o7[1]; out->qe = o7[2]; out->score2 = o7[3]; out->te2 = o7[4]; out->tb = o7[5]; out->qb = o7[6];
    }
};

static void default_opts(MemOpt& o) {
    memset(&o, 0, sizeof(o));
    o.a = 1; o.b = 4; o.o_del = o.o_ins = 6; o.e_del = o.e_ins = 1; o.w = 100; o.T = 30; o.zdrop = 100; o.pen_unpaired = 17;
    o.pen_clip5 = o.pen_clip3 = 5; o.max_mem_intv = 20; o.min_seed_len = 19; o.split_width = 10; o.max_occ = 500; o.max_chain_gap = 10000;
    o.max_ins = 10000; o.mask_level = 0.5f; o.drop_ratio = 0.5f; o.XA_drop_ratio = 0.8f; o.split_factor = 1.5f; o.chunk_size = 10000000;
    o.n_threads = 1; o.max_XA_hits = 5; o.max_XA_hits_alt = 200; o.max_matesw = 50; o.mask_level_redun = 0.95f; o.max_chain_extend = 1 << 30;
    o.mapQ_coef_len = 50; o.mapQ_coef_fac = 3;
    for (int i = 0, k = 0; i < 4; ++i) { for (int j = 0; j < 4; ++j) o.mat[k++] = i == j ? o.a : -o.b; o.mat[k++] = -1; }
    for (int j = 0; j < 5; ++j) o.mat[20 + j] = -1;
}

extern "C" {

void* hs_load(const char* prefix_) {
    std::string prefix(prefix_);
    auto* h = new HostIndex();
    if (!slurp(prefix + ".bwt", h->bwt) || !slurp(prefix + ".sa", h->sa_raw) || !slurp(prefix + ".pac", h->pac)) { delete h; return nullptr; }
    h->pac.resize(h->pac.size() + 16, 0);  // the CIGAR tiers read the packed reference in 32-bit words (the device buffer is padded too)
    const uint64_t* bh = (const uint64_t*)h->bwt.data();
    FMIndex& fm = h->fm;
    fm.primary = bh[0]; fm.L2[0] = 0; for (int i = 1; i <= 4; ++i) fm.L2[i] = bh[i];
    fm.seq_len = fm.L2[4];
    if (!bwt_recut_image(h->bwt.data() + 40, h->bwt.size() - 40, fm.seq_len, h->image)) { delete h; return nullptr; }
    fm.blocks = (const Blk32*)h->image.data();
    const uint64_t* sh = (const uint64_t*)h->sa_raw.data();
    fm.sa_intv = (int)sh[5]; fm.n_sa = (fm.seq_len + fm.sa_intv) / fm.sa_intv;
    h->sa.resize(fm.n_sa); h->sa[0] = (uint64_t)-1; memcpy(h->sa.data() + 1, h->sa_raw.data() + 56, (fm.n_sa - 1) * 8);
    fm.sa = h->sa.data();
    FILE* fp = fopen((prefix + ".ann").c_str(), "r");
    long long lp; int ns; unsigned seed; char str[8192];
    if (!fp || fscanf(fp, "%lld%d%u", &lp, &ns, &seed) != 3) { delete h; return nullptr; }
    h->contigs.resize(ns);
    for (int i = 0; i < ns; ++i) {
        unsigned gi; long long off; int len, na, c;
        if (fscanf(fp, "%u%8191s", &gi, str) != 2) break;
        while ((c = fgetc(fp)) != '\n' && c != EOF) {}
        if (fscanf(fp, "%lld%d%d", &off, &len, &na) != 3) break;
        h->contigs[i] = {off, len, 0};
    }
    fclose(fp);
    h->ref.pac = h->pac.data(); h->ref.contigs = h->contigs.data(); h->ref.l_pac = lp; h->ref.n_seqs = ns;
    return h;
}
void hs_free(void* h) { delete (HostIndex*)h; }

void hs_occ(void* h_, uint64_t k, uint64_t l, uint64_t* ck, uint64_t* cl) { fm_2occ4(((HostIndex*)h_)->fm, k, l, ck, cl); }
uint64_t hs_sa(void* h_, uint64_t k) { return fm_sa(((HostIndex*)h_)->fm, k); }

int hs_seed(void* h_, int len, const uint8_t* seq, uint64_t* out, int cap) {
    HostIndex* h = (HostIndex*)h_; MemOpt opt; default_opts(opt);
    std::vector<Intv> scratch(3 * (len + 1)), o(cap);
    int n = collect_intv(opt, h->fm, len, seq, o.data(), cap, scratch.data());
    if (n > 0) memcpy(out, o.data(), (size_t)n * 32);
    return n;
}

// mem_chain (+flt): chains rows {pos,rid,n,w,kept,first_seed_row}; seeds rows {rbeg,qbeg,len,score}
int hs_chain(void* h_, int len, const uint8_t* seq, int do_flt, int64_t* chains_out, int cap_c, int64_t* seeds_out, int cap_s, int* n_seeds_out, float* frac_rep) {
    HostIndex* h = (HostIndex*)h_; MemOpt opt; default_opts(opt);
    int cap = 4 * (len + 1);
    std::vector<Intv> scratch(3 * (len + 1)), iv(cap);
    int n_intv = len >= opt.min_seed_len ? collect_intv(opt, h->fm, len, seq, iv.data(), cap, scratch.data()) : 0;
    std::vector<uint64_t> sav;
    for (int i = 0; i < n_intv; ++i) {
        const Intv& p = iv[i];
        int64_t step = p.x2 > (uint64_t)opt.max_occ ? (int64_t)(p.x2 / opt.max_occ) : 1; int count = 0;
        for (int64_t k = 0; k < (int64_t)p.x2 && count < opt.max_occ; k += step, ++count) sav.push_back(fm_sa(h->fm, p.x0 + k));
    }
    int nsa = (int)sav.size();
    std::vector<Seed> seeds(nsa + 1); std::vector<Chain> chains(nsa + 1); std::vector<BNode> nodes(nsa / 4 + 4); std::vector<int32_t> order(nsa + 1), kept(nsa + 1);
    float fr = 0;
    int n = chain_seeds(opt, h->ref, len, iv.data(), n_intv, sav.data(), seeds.data(), chains.data(), nodes.data(), nsa / 4 + 4, order.data(), &fr);
    if (n < 0) return n;
    if (do_flt) n = chain_filter(opt, n, order.data(), chains.data(), seeds.data(), kept.data());
    int ns = 0;
    for (int i = 0; i < n; ++i) {
        const Chain& c = chains[order[i]];
        if (i < cap_c) { int64_t* r = chains_out + 6 * i; r[0] = c.pos; r[1] = c.rid; r[2] = c.n; r[3] = c.w; r[4] = c.kept; r[5] = ns; }
        for (int j = 0, s = c.first_seed; j < c.n; ++j, s = seeds[s].next, ++ns)
            if (ns < cap_s) { int64_t* r = seeds_out + 4 * ns; r[0] = seeds[s].rbeg; r[1] = seeds[s].qbeg; r[2] = seeds[s].len; r[3] = seeds[s].score; }
    }
    *n_seeds_out = ns; if (frac_rep) *frac_rep = fr;
    return n;
}

// mem_align1_core
int hs_align1(void* h_, int len, const uint8_t* seq, arachne_alnreg_t* out, int cap_out, uint64_t* counters) {
    HostIndex* h = (HostIndex*)h_; MemOpt opt; default_opts(opt);
    int cap = 4 * (len + 1);
    std::vector<Intv> scratch(3 * (len + 1)), iv(cap);
    int n_intv = len >= opt.min_seed_len ? collect_intv(opt, h->fm, len, seq, iv.data(), cap, scratch.data()) : 0;
    if (n_intv < 0) return -1;
    std::vector<uint64_t> sav;
    for (int i = 0; i < n_intv; ++i) {
        const Intv& p = iv[i];
        int64_t step = p.x2 > (uint64_t)opt.max_occ ? (int64_t)(p.x2 / opt.max_occ) : 1; int count = 0;
        for (int64_t k = 0; k < (int64_t)p.x2 && count < opt.max_occ; k += step, ++count) sav.push_back(fm_sa(h->fm, p.x0 + k));
    }
    int nsa = (int)sav.size();
    std::vector<Seed> seeds(nsa + 1), ks(nsa + 1); std::vector<Chain> chains(nsa + 1); std::vector<BNode> nodes(nsa / 4 + 4);
    std::vector<int32_t> order(nsa + 1), kept(nsa + 1);
    float fr = 0;
    int n = chain_seeds(opt, h->ref, len, iv.data(), n_intv, sav.data(), seeds.data(), chains.data(), nodes.data(), nsa / 4 + 4, order.data(), &fr);
    if (n < 0) return -2;
    int nk = chain_filter(opt, n, order.data(), chains.data(), seeds.data(), kept.data());
    int nks = 0;
    for (int i = 0; i < nk; ++i) {
        Chain& c = chains[order[i]];
        int s = c.first_seed; c.first_seed = nks;
        for (int j = 0; j < c.n; ++j, s = seeds[s].next) ks[nks++] = seeds[s];
    }
    std::vector<AlnReg> av(nks + 1); std::vector<uint64_t> srt(nsa + 1);
    int win_cap = 2 * len + 1024 + 4 * opt.w;
    std::vector<uint8_t> win(win_cap);
    HostCtx ctx; int n_av = 0;
    {   // the shipped extension driver (csrc/ext_driver.cuh), its DP problems solved by the oracle's scalar ksw_extend2
        ExtState S; ext_state_init(S);
        ExtProb P; ExtRes R = {0, -1, -1, -1, -1, 0};
        std::vector<uint8_t> qq, tt;
        for (;;) {
            int rc = ext_advance(opt, h->ref, 0, len, nk, chains.data(), order.data(), ks.data(), av.data(), nks, srt.data(), S, &R, &P);
            if (rc < 0) return -10 + rc;
            if (rc == 0) break;
            qq.resize(P.qlen + 1); tt.resize(P.tlen + 1);
            for (int i = 0; i < P.qlen; ++i) qq[i] = seq[P.q0 + ((P.dirs & 1) ? -i : i)];
            for (int i = 0; i < P.tlen; ++i) tt[i] = ref_base(h->ref, P.tpos + ((P.dirs & 2) ? -(int64_t)i : (int64_t)i));
            int qle, tle, gtle, gscore, max_off;
            R.mx = ctx.ksw_extend(opt, P.qlen, qq.data(), 1, P.tlen, tt.data(), 1, P.w, P.end_bonus, opt.zdrop, P.h0, &qle, &tle, &gtle, &gscore, &max_off);
            R.max_j = qle - 1; R.max_i = tle - 1; R.max_ie = gtle - 1; R.gscore = gscore; R.max_off = max_off;
        }
        n_av = S.n_regions;
    }
    int m = sort_dedup_patch(ctx, opt, h->ref, seq, n_av, av.data(), win.data(), win_cap);
    if (m < 0) return -20 + m;
    for (int i = 0; i < m && i < cap_out; ++i) memcpy(&out[i], &av[i], sizeof(AlnReg));
    if (counters) { counters[0] = ctx.calls; counters[1] = ctx.cells; counters[2] = nsa; }
    return m;
}

// GoBwaMemMateSW: regions of both reads after rescue
int hs_pair(void* h_, int l1, const uint8_t* s1, int l2, const uint8_t* s2, int score_delta, arachne_alnreg_t* o1, int* n1_, arachne_alnreg_t* o2, int* n2_, int cap) {
    HostIndex* h = (HostIndex*)h_; MemOpt opt; default_opts(opt);
    std::vector<AlnReg> r1(cap + 64), r2(cap + 64);
    int n1 = l1 > 0 ? hs_align1(h_, l1, s1, (arachne_alnreg_t*)r1.data(), cap, nullptr) : 0;
    int n2 = l2 > 0 ? hs_align1(h_, l2, s2, (arachne_alnreg_t*)r2.data(), cap, nullptr) : 0;
    if (n1 < 0 || n2 < 0) return -1;
    HostCtx ctx;
    int win_cap = 2 * (l1 > l2 ? l1 : l2) + 1024 + 4 * opt.w;
    std::vector<uint8_t> win(win_cap);
    int rc = pair_rescue(ctx, opt, h->ref, score_delta, l1, s1, l2, s2, r1.data(), &n1, cap + 64, r2.data(), &n2, cap + 64, win.data(), win_cap);
    if (rc < 0) return rc;
    memcpy(o1, r1.data(), (size_t)n1 * sizeof(AlnReg)); memcpy(o2, r2.data(), (size_t)n2 * sizeof(AlnReg));
    *n1_ = n1; *n2_ = n2;
    return 0;
}

// mem_reg2aln + GetAlignments walk; mm gets n_mm ref loci then n_mm read offsets
int hs_reg2aln(void* h_, int len, const uint8_t* seq, const arachne_alnreg_t* reg, arachne_aln_t* out, uint32_t* cigar, int cap, int32_t* mm) {
    HostIndex* h = (HostIndex*)h_; MemOpt opt; default_opts(opt);
    HostCtx ctx;
    int win_cap = 2 * len + 1024 + 4 * opt.w;
    std::vector<uint8_t> win(win_cap);
    std::vector<int32_t> m(2 * len + 8);
    AlnRec a;
    int rc = reg2aln(ctx, opt, h->ref, len, seq, *(const AlnReg*)reg, &a, cigar, cap, m.data(), win.data(), win_cap);
    if (rc < 0) return rc;
    memcpy(out, &a, sizeof(a));
    for (int i = 0; i < a.n_mm; ++i) { mm[i] = m[i]; mm[a.n_mm + i] = m[len + i]; }
    return 0;
}

// ks_introsort (csrc/common.cuh) with the two region orders of mem_sort_dedup_patch: 0 = by re (alnreg_slt2, bwamem.c:390-392),
// 1 = by score desc, rb, qb (alnreg_slt, :393-395).  The permutation of equal keys is what the test compares with the reference's.
void hs_introsort_regs(int mode, int n, arachne_alnreg_t* a_) {
    AlnReg* a = (AlnReg*)a_;
    if (mode == 0) ks_introsort((size_t)n, a, [](const AlnReg& x, const AlnReg& y) { return x.re < y.re; });
    else ks_introsort((size_t)n, a, [](const AlnReg& x, const AlnReg& y) { return x.score > y.score || (x.score == y.score && (x.rb < y.rb || (x.rb == y.rb && x.qb < y.qb))); });
}

// The device's Go math/rand (csrc/gorand.cuh: stateless closed form for the first 273 outputs, materialised 607-word state after) as
// k_place_tag drives it: Float64 outputs 0..n-1 of NewSource(seed), switching to the stateful generator after `switch_at` outputs
// (0 = stateful from the start, as for pairs with many candidate combinations; >= 273 = only when the stateless window runs out).
int hs_gorand_float64(int64_t seed, int n, int switch_at, double* out) {
    std::vector<uint64_t> scratch(GORAND_LEN);
    GoRandDev r; r.init(seed);
    if (switch_at == 0) r.materialise(scratch.data());
    for (int i = 0; i < n; ++i) {
        if (!r.state && i == switch_at) r.materialise(scratch.data());
        bool need = false;
        double f = r.float64(&need);
        if (need) { r.materialise(scratch.data()); need = false; f = r.float64(&need); if (need) return -1; }
        out[i] = f;
    }
    return 0;
}

long hs_patch_dp_calls(void) { return g_score_only_dp; }  // how often region patching got as far as its global re-alignment

const void* hs_fm(void* h) { return &((HostIndex*)h)->fm; }  // for hostsim_seed.cpp
const void* hs_ref(void* h) { return &((HostIndex*)h)->ref; }  // for hostsim_cigar.cpp
}
