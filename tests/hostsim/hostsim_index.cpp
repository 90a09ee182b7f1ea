// TEST-ONLY host backend for the index-construction driver (csrc/index_build.cuh): the same per-element functors and the
// same batch / tie-refinement driver the device runs, with loops and std::stable_sort in place of kernels and CUB, so the
// CPU test tier can pin the algorithm against the files the reference's own `bwa index` writes.  Never shipped.
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <numeric>
#include <string>
#include <vector>
#include "../../arachne_b200/csrc/index_build.cuh"

using namespace arachne::ib;

struct HostBackend {
    bool good = true;
    std::string err;
    bool ok() const { return good; }
    void fail(const char* m) { if (good) err = m; good = false; }
    void* acquire(size_t bytes) { return calloc(bytes ? bytes : 1, 1); }
    void release(void* p) { free(p); }
    template <class F> void for_each(uint64_t n, F f) { for (uint64_t i = 0; i < n; ++i) f(i); }
    void histogram(const TextView& t, uint64_t* h) { memset(h, 0, N_BINS * 8); for (uint64_t p = 0; p < t.n; ++p) ++h[sfx_bin(t, p)]; }
    uint64_t select_bins(const TextView& t, uint32_t b0, uint32_t b1, uint64_t* out) {
        uint64_t m = 0;
        for (uint64_t p = 0; p < t.n; ++p) { uint32_t b = sfx_bin(t, p); if (b >= b0 && b < b1) out[m++] = p; }
        return m;
    }
    uint64_t select_flagged(const uint8_t* f, uint32_t* out, uint64_t n) { uint64_t m = 0; for (uint64_t i = 0; i < n; ++i) if (f[i]) out[m++] = (uint32_t)i; return m; }
    template <class K, class V> void sort_pairs(const K* kin, K* kout, const V* vin, V* vout, uint64_t n, int end_bit) {
        std::vector<uint64_t> idx(n);
        std::iota(idx.begin(), idx.end(), 0);
        const K mask = end_bit >= (int)sizeof(K) * 8 ? ~(K)0 : (((K)1 << end_bit) - 1);
        std::stable_sort(idx.begin(), idx.end(), [&](uint64_t a, uint64_t b) { return (kin[a] & mask) < (kin[b] & mask); });
        for (uint64_t i = 0; i < n; ++i) { kout[i] = kin[idx[i]]; vout[i] = vin[idx[i]]; }
    }
    void inclusive_sum(const uint32_t* in, uint32_t* out, uint64_t n) { uint32_t s = 0; for (uint64_t i = 0; i < n; ++i) { s += in[i]; out[i] = s; } }
    void exclusive_sum(const uint64_t* in, uint64_t* out, uint64_t n) { uint64_t s = 0; for (uint64_t i = 0; i < n; ++i) { uint64_t v = in[i]; out[i] = s; s += v; } }
    void copy32(uint32_t* d, const uint32_t* s, uint64_t n) { memmove(d, s, n * 4); }
    void to_host(void* d, const void* s, size_t b) { memcpy(d, s, b); }
    void to_device(void* d, const void* s, size_t b) { memcpy(d, s, b); }
    void zero(void* p, size_t b) { memset(p, 0, b); }
};

// pac: forward-strand packed bases (l_pac/4+1 bytes readable).  Writes the .bwt payload and the sampled SA into caller
// buffers sized bwt_payload_words(2*l_pac)*4 bytes and ((2*l_pac+32)/32)*8 bytes; meta = {primary, L2[1..4], n_batches,
// tied_suffixes, max_rounds}.
extern "C" int hostsim_index_build(const uint8_t* pac, uint64_t l_pac, uint64_t batch_cap, uint32_t* payload, uint64_t* sa, uint64_t* meta) {
    HostBackend be;
    std::vector<uint8_t> p(pac, pac + l_pac / 4 + 1);
    p.resize(p.size() + 8, 0);
    uint32_t* out = nullptr; uint64_t* s = nullptr;
    BuildStats st;
    if (!build_fm_index(be, p.data(), l_pac, batch_cap, &out, &s, st)) { fprintf(stderr, "hostsim_index_build: %s\n", be.err.c_str()); return -1; }
    memcpy(payload, out, bwt_payload_words(st.n) * 4);
    memcpy(sa, s, ((st.n + SA_INTV) / SA_INTV) * 8);
    meta[0] = st.primary; for (int i = 1; i <= 4; ++i) meta[i] = st.L2[i];
    meta[5] = st.n_batches; meta[6] = st.tied_round1; meta[7] = st.max_rounds;
    free(out); free(s);
    return 0;
}
