// Index handle: reads the files `bwa index` writes and keeps the FM-index image, sampled SA and packed reference
// resident in HBM.  Replaces bwa_idx_load / bwt_restore_bwt / bwt_restore_sa / bns_restore
// (bwa/bwa.c:262-294, bwa/bwt.c:421-462, bwa/bntseq.c:98-196).
#pragma once
#include <cuda_runtime.h>
#include <string>
#include <vector>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <condition_variable>
#include <thread>
#include <chrono>
#include "fmindex.cuh"
#include "bwt_image.h"
#include "../../include/arachne_cuda.h"

namespace arachne {
void set_error(const char* fmt, ...);
}

struct arachne_index {
    int device = 0;
    arachne::FMIndex fm{};   // device pointers
    arachne::RefView ref{};  // device pointers
    void *d_bwt = nullptr, *d_sa = nullptr, *d_pac = nullptr, *d_contigs = nullptr, *d_kmer = nullptr, *d_isa = nullptr, *d_text = nullptr;
    // runtime knobs (arachne_index_set_option): use the k-mer table in seeding; resident seeding blocks per SM
    int opt_seed_kmer = 1, opt_seed_blocks = 3, opt_seed_slist = -1, opt_seed_direct = 1;
    int opt_seed3_threads = 768;
    int opt_seed_impl = 2;  // 2: k_seed2 (one read per lane), 3: k_seed3 (states in shared memory, one queue per kind of step)
    size_t device_bytes = 0;
    std::vector<arachne::Contig> contigs;
    std::vector<std::string> names;
    int64_t l_pac = 0;
    // host copy of the packed reference (GetSeq-style host queries, compat shims)
    std::vector<uint8_t> h_pac;
    // per-index batch workspaces (device arenas, pinned scalars, stream, events): created on first use and reused by
    // every later batch.  The reference's workers share one read-only index and bring their own scratch (gobwa.go passes
    // buf=NULL); here the scratch is the handle's: N_WS slots, so that two host threads can be in flight on one handle and
    // one batch's host<->device copies overlap the other's kernels; further callers queue.  `mu` guards the small
    // shared state (options, stream choice, the pool of recycled pinned result sets).
    static constexpr int N_WS = 2;  // 3 in flight measured slower (persistent kernels of three batches oversubscribe the SMs)
    std::mutex mu;
    std::mutex ws_gate; std::condition_variable ws_cv; bool ws_busy[N_WS] = {};  // which batch workspace is leased (batch.inl::take_ws)
    int env_resc_mult = 4; bool env_cigar_tiers = true;  // ARACHNE_RESC_MULT / ARACHNE_CIGAR_TIERS, read once at load
    void* ws[N_WS] = {};
    std::vector<void*> priv_pool;
    void* user_stream = nullptr;
    bool has_user_stream = false;
};

namespace arachne {

#define AR_CUDA(x)                                                                         \
    do {                                                                                   \
        cudaError_t e_ = (x);                                                              \
        if (e_ != cudaSuccess) {                                                           \
            set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #x, cudaGetErrorString(e_)); \
            return ARACHNE_ERR_CUDA;                                                       \
        }                                                                                  \
    } while (0)

}  // namespace arachne
#include "index_kernels.cuh"
namespace arachne {

inline bool read_file(const std::string& path, std::vector<uint8_t>& out) {
    FILE* fp = fopen(path.c_str(), "rb");
    if (!fp) return false;
    fseek(fp, 0, SEEK_END);
    long n = ftell(fp);
    fseek(fp, 0, SEEK_SET);
    out.resize(n);
    size_t got = n ? fread(out.data(), 1, n, fp) : 0;
    fclose(fp);
    return got == (size_t)n;
}

inline int index_load(const char* prefix_, int device, arachne_index** out) {
    if (!prefix_ || !out) { set_error("arachne_index_load: null argument"); return ARACHNE_ERR_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        set_error("arachne_index_load: no CUDA device (this library has no CPU path)");
        return ARACHNE_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { set_error("arachne_index_load: device %d out of range (%d present)", device, ndev); return ARACHNE_ERR_ARG; }
    std::string prefix(prefix_);
    {   // bwa_idx_infer_prefix, bwa/bwa.c:218-243
        FILE* fp = fopen((prefix + ".64.bwt").c_str(), "rb");
        if (fp) { fclose(fp); prefix += ".64"; }
    }
    std::vector<uint8_t> bwt, sa, pac;
    if (!read_file(prefix + ".bwt", bwt) || bwt.size() < 40 + 64) { set_error("cannot read %s.bwt", prefix.c_str()); return ARACHNE_ERR_IO; }
    if (!read_file(prefix + ".sa", sa) || sa.size() < 56) { set_error("cannot read %s.sa", prefix.c_str()); return ARACHNE_ERR_IO; }
    if (!read_file(prefix + ".pac", pac) || pac.empty()) { set_error("cannot read %s.pac", prefix.c_str()); return ARACHNE_ERR_IO; }
    auto* idx = new arachne_index();
    idx->device = device;
    const uint64_t* bh = reinterpret_cast<const uint64_t*>(bwt.data());
    FMIndex& fm = idx->fm;
    fm.primary = bh[0];
    fm.L2[0] = 0;
    for (int i = 1; i <= 4; ++i) fm.L2[i] = bh[i];
    fm.seq_len = fm.L2[4];
    size_t bwt_bytes = bwt.size() - 40;
    // bwt_bwtupdate_core (bwtindex.c:151-173): 32 bytes of counts per started 128-symbol block, 4 bytes per started
    // 16-symbol word, one trailing 32-byte count record
    const size_t want_bytes = ((fm.seq_len + 127) / 128) * 32 + ((fm.seq_len + 15) / 16) * 4 + 32;
    if (bwt_bytes < want_bytes) { delete idx; set_error("%s.bwt is not an occ-interleaved BWT (run `bwa index`)", prefix.c_str()); return ARACHNE_ERR_IO; }
    const uint64_t* sh = reinterpret_cast<const uint64_t*>(sa.data());
    if (sh[0] != fm.primary || sh[6] != fm.seq_len) { delete idx; set_error("SA-BWT inconsistency in %s", prefix.c_str()); return ARACHNE_ERR_IO; }
    fm.sa_intv = (int)sh[5];
    fm.n_sa = (fm.seq_len + fm.sa_intv) / fm.sa_intv;
    if (sa.size() < 56 + (fm.n_sa - 1) * 8 || (fm.sa_intv & (fm.sa_intv - 1))) { delete idx; set_error("truncated %s.sa", prefix.c_str()); return ARACHNE_ERR_IO; }
    // .ann / .amb (text)
    {
        FILE* fp = fopen((prefix + ".ann").c_str(), "r");
        if (!fp) { delete idx; set_error("cannot read %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
        long long lp; int ns; unsigned seed;
        if (fscanf(fp, "%lld%d%u", &lp, &ns, &seed) != 3) { fclose(fp); delete idx; set_error("parse error in %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
        idx->l_pac = lp;
        idx->contigs.resize(ns); idx->names.resize(ns);
        char str[8192];
        for (int i = 0; i < ns; ++i) {
            unsigned gi; long long off; int len, nambs, c;
            if (fscanf(fp, "%u%8191s", &gi, str) != 2) { fclose(fp); delete idx; set_error("parse error in %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
            idx->names[i] = str;
            while ((c = fgetc(fp)) != '\n' && c != EOF) {}
            if (fscanf(fp, "%lld%d%d", &off, &len, &nambs) != 3) { fclose(fp); delete idx; set_error("parse error in %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
            idx->contigs[i].offset = off; idx->contigs[i].len = len; idx->contigs[i].is_alt = 0;
        }
        fclose(fp);
        fp = fopen((prefix + ".alt").c_str(), "r");  // bns_restore, bntseq.c:176-196: names listed in .alt are ALT contigs
        if (fp) {
            while (fscanf(fp, "%8191s", str) == 1) {
                if (str[0] != '@')
                    for (int i = 0; i < ns; ++i) if (idx->names[i] == str) idx->contigs[i].is_alt = 1;
                int c; while ((c = fgetc(fp)) != '\n' && c != EOF) {}
            }
            fclose(fp);
        }
    }
    if ((uint64_t)idx->l_pac * 2 != fm.seq_len || pac.size() < (size_t)(idx->l_pac / 4 + 1)) { delete idx; set_error("inconsistent .pac/.ann/.bwt for %s", prefix.c_str()); return ARACHNE_ERR_IO; }
    // re-cut the reference's 64-byte / 128-symbol blocks into 32-byte sector blocks (fmindex.cuh) and upload
    for (int i = 0; i < 4; ++i)
        if (fm.L2[i + 1] - fm.L2[i] >= 0xffffffffull) { delete idx; set_error("%s: a base count exceeds 2^32-1 (sector layout keeps u32 counts)", prefix.c_str()); return ARACHNE_ERR_IO; }
    cudaError_t e = cudaSetDevice(device);
    std::vector<uint8_t> image;
    if (!bwt_recut_image(bwt.data() + 40, bwt_bytes, fm.seq_len, image)) { delete idx; set_error("%s.bwt is truncated", prefix.c_str()); return ARACHNE_ERR_IO; }
    size_t img = image.size(), sab = fm.n_sa * 8, pacb = (size_t)(idx->l_pac / 4 + 1);
    std::vector<uint64_t> sav(fm.n_sa);
    sav[0] = (uint64_t)-1;
    memcpy(sav.data() + 1, sa.data() + 56, (fm.n_sa - 1) * 8);
    if (e == cudaSuccess) e = cudaMalloc(&idx->d_bwt, img);
    if (e == cudaSuccess) e = cudaMalloc(&idx->d_sa, sab);
    if (e == cudaSuccess) e = cudaMalloc(&idx->d_pac, pacb + 16);
    if (e == cudaSuccess) e = cudaMalloc(&idx->d_contigs, sizeof(Contig) * idx->contigs.size());
    if (e == cudaSuccess) e = cudaMemcpy(idx->d_bwt, image.data(), img, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(idx->d_sa, sav.data(), sab, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(idx->d_pac, 0, pacb + 16);
    if (e == cudaSuccess) e = cudaMemcpy(idx->d_pac, pac.data(), pacb, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(idx->d_contigs, idx->contigs.data(), sizeof(Contig) * idx->contigs.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
        set_error("index upload failed: %s", cudaGetErrorString(e));
        cudaFree(idx->d_bwt); cudaFree(idx->d_sa); cudaFree(idx->d_pac); cudaFree(idx->d_contigs);
        delete idx;
        return ARACHNE_ERR_CUDA;
    }
    fm.blocks = reinterpret_cast<const Blk32*>(idx->d_bwt);
    fm.sa = reinterpret_cast<const uint64_t*>(idx->d_sa);
    idx->ref.pac = reinterpret_cast<const uint8_t*>(idx->d_pac);
    idx->ref.contigs = reinterpret_cast<const Contig*>(idx->d_contigs);
    idx->ref.l_pac = idx->l_pac;
    idx->ref.n_seqs = (int)idx->contigs.size();
    idx->device_bytes = img + sab + pacb + sizeof(Contig) * idx->contigs.size();
    fm.kmer = nullptr; fm.kmer_all = nullptr; fm.kmer_k = 0;
    {   // k-mer table for the LAST-like seeding pass: K = min(14, log4(seq_len) - 1), off below 8 (ARACHNE_SEED_KMER=0 disables, =K forces)
        int K = 0;
        for (uint64_t v = fm.seq_len; v > 3; v >>= 2) ++K;
        K = K - 1 < 14 ? K - 1 : 14;
        if (const char* ev = getenv("ARACHNE_SEED_KMER")) K = atoi(ev);
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        const uint64_t n_ent = K >= 1 ? kmer_level_off(K + 1) : 0;
        if (K >= 8 && K <= 15 && n_ent * 16 < free_b / 4 && fm.seq_len < (1ull << 40) - 2) {
            void* dk = nullptr;
            e = cudaMalloc(&dk, n_ent * 16);
            if (e == cudaSuccess) {
                FMIndex f2 = fm;  // the SA is not used here
                for (int L = 0; L < K && e == cudaSuccess; ++L) {
                    k_kmer_level<<<148 * 8, 256>>>(f2, reinterpret_cast<uint4*>(dk), L);
                    e = cudaGetLastError();
                }
                if (e == cudaSuccess) e = cudaDeviceSynchronize();
            }
            if (e == cudaSuccess) {
                idx->d_kmer = dk; fm.kmer_all = dk; fm.kmer = reinterpret_cast<const uint4*>(dk) + kmer_level_off(K); fm.kmer_k = K;
                idx->device_bytes += n_ent * 16;
            } else { if (dk) cudaFree(dk); (void)cudaGetLastError(); }
        }
    }
    {   // densify the sampled SA when HBM has the room (ARACHNE_SA_INTV=1|2|4|8|16|32 overrides; 32 keeps the file's sampling)
        int d = 0;
        if (const char* ev = getenv("ARACHNE_SA_INTV")) d = atoi(ev);
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        if (d <= 0 || (d & (d - 1)) || d > fm.sa_intv)
            for (d = 1; d < fm.sa_intv && ((fm.seq_len + d) / d) * 8 > free_b * 2 / 5; d <<= 1) {}
        if (d < fm.sa_intv) {
            const uint64_t n_dense = (fm.seq_len + d) / d;
            void* dd = nullptr;
            e = cudaMalloc(&dd, n_dense * 8);
            if (e == cudaSuccess) {
                cudaDeviceProp prop;
                int sms = cudaGetDeviceProperties(&prop, device) == cudaSuccess ? prop.multiProcessorCount : 148;
                k_sa_densify<<<sms * 8, 256>>>(fm, reinterpret_cast<uint64_t*>(dd), d);
                e = cudaGetLastError();
                if (e == cudaSuccess) e = cudaDeviceSynchronize();
            }
            if (e == cudaSuccess) {
                cudaFree(idx->d_sa);
                idx->d_sa = dd;
                fm.sa = reinterpret_cast<const uint64_t*>(dd);
                fm.sa_intv = d; fm.n_sa = n_dense;
                idx->device_bytes += n_dense * 8 - sab;
            } else {  // keep the file's sampling: slower, same answers
                if (dd) cudaFree(dd);
                (void)cudaGetLastError();
            }
        }
    }
    fm.isa = nullptr; fm.text = nullptr;
    {   // inverse SA + both-strand text for the direct extension of unique matches (ARACHNE_SEED_DIRECT=0 disables): needs every
        // SA row in HBM and room for another 8 bytes per row next to ~48 GB kept free for the batch workspaces
        const char* ev = getenv("ARACHNE_SEED_DIRECT");
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        const uint64_t n_words = fm.seq_len / 32 + 8;
        const size_t reserve = total_b / 4 < ((size_t)48 << 30) ? total_b / 4 : ((size_t)48 << 30);
        if (!(ev && atoi(ev) == 0) && fm.sa_intv == 1 && (fm.seq_len + 1) * 8 + n_words * 8 + reserve < free_b) {
            void *di = nullptr, *dt = nullptr;
            e = cudaMalloc(&di, (fm.seq_len + 1) * 8);
            if (e == cudaSuccess) e = cudaMalloc(&dt, n_words * 8);
            if (e == cudaSuccess) {
                k_isa_fill<<<148 * 8, 256>>>(fm.sa, fm.seq_len + 1, reinterpret_cast<uint64_t*>(di));
                k_text_both<<<148 * 8, 256>>>(idx->ref.pac, (uint64_t)idx->l_pac, n_words, reinterpret_cast<uint64_t*>(dt));
                e = cudaGetLastError();
                if (e == cudaSuccess) e = cudaDeviceSynchronize();
            }
            if (e == cudaSuccess) {
                idx->d_isa = di; idx->d_text = dt;
                fm.isa = reinterpret_cast<const uint64_t*>(di); fm.text = reinterpret_cast<const uint64_t*>(dt);
                idx->device_bytes += (fm.seq_len + 1) * 8 + n_words * 8;
            } else { if (di) cudaFree(di); if (dt) cudaFree(dt); (void)cudaGetLastError(); }
        }
    }
    idx->h_pac.assign(pac.begin(), pac.begin() + pacb);
    *out = idx;
    return ARACHNE_OK;
}

}  // namespace arachne
