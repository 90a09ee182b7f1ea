// libarachne_cuda: kernels and batch orchestration of the candidate-alignment / read-placement hot path (sm_100a).
//
// Stages of one batch (reads = 2 x pairs), every stage one or a few kernels over the whole batch:
//   k_pack_reads, k_seed2, k_sort_intv   mem_collect_intv: SMEMs + re-seeding + LAST-like seeds -> SA intervals (seed_kernel.cuh)
//   k_expand, k_sa                       one SA row per seed occurrence (bwamem.c:278-283), bwt_sa (stage_kernels.cuh)
//   k_chain                              chaining, chain weights, chain filter (chain.cuh; kernel in stage_kernels.cuh)
//   k_ext_drive / k_ext_dp rounds        mem_chain2aln driver + ksw_extend2 as queues of uniform problems, region dedup (ext_stage.cuh)
//   k_rescue                             mem_matesw as GoBwaMemMateSW drives it (rescue.cuh, warp_ctx.cuh; kernel in pair_kernels.cuh)
//   k_cigar_fast / _band / k_cigar       mem_reg2aln + the GetAlignments walk (cigar_kernel.cuh, cigar.cuh; k_cigar in pair_kernels.cuh)
//   k_place_*                            tagBestAlignments .. optimizer .. MAPQ (place.cuh)
// The index is read-only and shared; all per-batch state lives in a device workspace that is reused (batch.inl).
#include <cuda_runtime.h>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_radix_sort.cuh>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <algorithm>
#include "index.cuh"
#include "warp_ctx.cuh"
#include "cigar.cuh"
#include "md5_seed.h"

namespace arachne {

static thread_local char g_err[1024] = "";
void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

static_assert(sizeof(MemOpt) == sizeof(arachne_mem_opt_t), "MemOpt layout");
static_assert(sizeof(AlnReg) == sizeof(arachne_alnreg_t), "AlnReg layout");

}  // namespace arachne
#include "counters.cuh"
namespace arachne {

// ------------------------------------------------------------------------------------------------ k_seed2
}  // namespace arachne
#include "place.cuh"
#include "records.cuh"
#include "cigar_kernel.cuh"
#include "seed_kernel.cuh"
#include "seed3_kernel.cuh"
namespace arachne {

}  // namespace arachne
#include "stage_kernels.cuh"
namespace arachne {

// ------------------------------------------------------------------------------------------------ extension rounds
}  // namespace arachne
#include "ext_stage.cuh"
namespace arachne {

}  // namespace arachne
#include "pair_kernels.cuh"
namespace arachne {

// ------------------------------------------------------------------------------------------------ stage kernels
__global__ void k_dpx_peak(int* out, int iters, int seed) {
    int a0 = threadIdx.x + seed, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
    const int d = -(seed | 1), z = blockIdx.x & 7;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
            a0 = __viaddmax_s32(a0, d, z); a1 = __viaddmax_s32(a1, d, z); a2 = __viaddmax_s32(a2, d, z); a3 = __viaddmax_s32(a3, d, z);
            a4 = __viaddmax_s32(a4, d, z); a5 = __viaddmax_s32(a5, d, z); a6 = __viaddmax_s32(a6, d, z); a7 = __viaddmax_s32(a7, d, z);
        }
    }
    if ((a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7) == 0x7fffffff) out[0] = a0;
}
// ------------------------------------------------------------------------------------------------ host side
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    DevBuf() = default;
    DevBuf(const DevBuf&) = delete;
    DevBuf& operator=(const DevBuf&) = delete;
    ~DevBuf() { release(); }  // stage entry points return early on errors: nothing may leak
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 256;
        if (cudaMalloc(&p, want) != cudaSuccess) { p = nullptr; return -1; }
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

static int validate_opts(const arachne_mem_opt_t* o) {
    if (!o) { set_error("null options"); return ARACHNE_ERR_ARG; }
    if (o->e_del <= 0 || o->e_ins <= 0 || o->a <= 0 || o->w <= 0 || o->max_occ <= 0 || o->min_seed_len <= 0) {
        set_error("options out of range"); return ARACHNE_ERR_OPTS;
    }
    // the extension kernel keeps H and E as bytes and the substitution scores as signed bytes (ext_stage.cuh); the callers
    // check a * read length <= 255 per batch / problem
    if (o->a > 127 || o->b < 0 || o->b > 128 || o->o_del < 0 || o->o_ins < 0) {
        set_error("scoring out of the kernels' range: need a <= 127, 0 <= b <= 128, gap opens >= 0"); return ARACHNE_ERR_OPTS;
    }
    for (int i = 0, k = 0; i < 5; ++i)
        for (int j = 0; j < 5; ++j, ++k) {
            int want = (i < 4 && j < 4) ? (i == j ? o->a : -o->b) : -1;
            if (o->mat[k] != want) { set_error("scoring matrix is not bwa_fill_scmat(a,b): not implemented by the kernels"); return ARACHNE_ERR_OPTS; }
        }
    return ARACHNE_OK;
}

static int sm_count(int device) {
    int n = 148;
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device);
    return n;
}

}  // namespace arachne

using namespace arachne;

// pinned host buffer (results land here so the device->host copies are true async DMA)
struct PinBuf {
    void* p = nullptr;
    size_t cap = 0;
    int ensure(size_t bytes) {
        if (bytes <= cap) return 0;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 4 + 4096;
        if (cudaMallocHost(&p, want) != cudaSuccess) { p = nullptr; return -1; }
        cap = want;
        return 0;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct BatchPriv {
    PinBuf reg_off, regs, alns, cigar, mm, cand_off, cands, lmp, rec_off, recs;
    arachne_index_t* owner = nullptr;
    void release() { reg_off.release(); regs.release(); alns.release(); cigar.release(); mm.release(); cand_off.release(); cands.release(); lmp.release(); rec_off.release(); recs.release(); }
};

// Per-index workspace: every device arena of the batch pipeline, grown on demand and kept between batches.
struct Workspace {
    DevBuf d_qpk, d_intv_cnt2, d_seq, d_off, d_scratch, d_arena, d_misc, d_intv_off, d_intv_n, d_nsa, d_sa_off, d_rows, d_cnt, d_tmp;
    DevBuf d_seeds, d_kseeds, d_chains, d_nodes, d_order, d_kept, d_nchain, d_nkseed, d_srt, d_regs, d_nregs, d_win, d_reg_off, d_out;
    DevBuf d_nregs_bak, d_reg_loc, d_blist, d_resc, d_reg_read, d_z, d_cig_scr, d_mm_scr, d_alns, d_cig, d_mm, d_hscr, d_rscr, d_xstate, d_xprob, d_xres, d_xrank, d_xperm, d_xctl;
    DevBuf d_bc_off, d_run_rfa, d_rand_seed, d_centro, d_pl_cnt, d_pl_scan, d_cand_off, d_lmp, d_cands, d_hot, d_aux, d_slow, d_slot_of_bc, d_bc_of_slot,
        d_first, d_keys_a, d_keys_b, d_vals_a, d_vals_b, d_seg_off, d_mol, d_list2, d_list3, d_zband, d_attach, d_rec_i, d_rec_d, d_rec_cnt, d_recs;
    cudaStream_t own_stream = nullptr;
    cudaEvent_t ev[12] = {};
    cudaEvent_t ev_block = nullptr;  // ARACHNE_BLOCKING_SYNC=1: host waits sleep on this event instead of spinning in cudaStreamSynchronize
    // (one spinning core per batch in flight is fine on an idle host; with 8 ranks x 2 batch threads on 16 cores it is not)
    cudaError_t sync(cudaStream_t st) {
        if (!ev_block) return cudaStreamSynchronize(st);
        cudaError_t e = cudaEventRecord(ev_block, st);
        return e != cudaSuccess ? e : cudaEventSynchronize(ev_block);
    }
    Counters* h_cnt = nullptr;              // pinned
    unsigned long long* h_scalars = nullptr;  // pinned, 8 slots
    std::vector<DevBuf*> all() {
        return {&d_qpk, &d_intv_cnt2, &d_seq, &d_off, &d_scratch, &d_arena, &d_misc, &d_intv_off, &d_intv_n, &d_nsa, &d_sa_off, &d_rows, &d_cnt, &d_tmp,
                &d_seeds, &d_kseeds, &d_chains, &d_nodes, &d_order, &d_kept, &d_nchain, &d_nkseed, &d_srt, &d_regs, &d_nregs,
                &d_win, &d_reg_off, &d_out, &d_nregs_bak, &d_reg_loc, &d_blist, &d_resc, &d_reg_read, &d_z, &d_cig_scr, &d_mm_scr, &d_alns, &d_cig, &d_mm, &d_hscr, &d_rscr, &d_xstate, &d_xprob, &d_xres, &d_xrank, &d_xperm, &d_xctl,
                &d_bc_off, &d_run_rfa, &d_rand_seed, &d_centro, &d_pl_cnt, &d_pl_scan, &d_cand_off, &d_lmp, &d_cands, &d_hot, &d_aux, &d_slow, &d_slot_of_bc,
                &d_bc_of_slot, &d_first, &d_keys_a, &d_keys_b, &d_vals_a, &d_vals_b, &d_seg_off, &d_mol, &d_list2, &d_list3, &d_zband, &d_attach, &d_rec_i, &d_rec_d,
                &d_rec_cnt, &d_recs};
    }
    int init() {
        if (cudaStreamCreateWithFlags(&own_stream, cudaStreamNonBlocking) != cudaSuccess) return -1;
        for (auto& e : ev) if (cudaEventCreate(&e) != cudaSuccess) return -1;
        if (cudaMallocHost((void**)&h_cnt, sizeof(Counters)) != cudaSuccess) return -1;
        if (cudaMallocHost((void**)&h_scalars, 64) != cudaSuccess) return -1;
        if (getenv("ARACHNE_BLOCKING_SYNC") && atoi(getenv("ARACHNE_BLOCKING_SYNC")) != 0 &&
            cudaEventCreateWithFlags(&ev_block, cudaEventBlockingSync | cudaEventDisableTiming) != cudaSuccess) return -1;
        return 0;
    }
    void destroy() {
        for (auto* b : all()) b->release();
        for (auto& e : ev) if (e) cudaEventDestroy(e);
        if (ev_block) cudaEventDestroy(ev_block);
        if (own_stream) cudaStreamDestroy(own_stream);
        if (h_cnt) cudaFreeHost(h_cnt);
        if (h_scalars) cudaFreeHost(h_scalars);
    }
};

extern "C" {

const char* arachne_last_error(void) { return g_err; }

int arachne_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
    return n;
}

int arachne_index_load(const char* prefix, int device, arachne_index_t** out) {
    const int rc = index_load(prefix, device, out);
    if (rc == ARACHNE_OK && getenv("ARACHNE_SEED_IMPL")) { const int v = atoi(getenv("ARACHNE_SEED_IMPL")); if (v == 2 || v == 3) (*out)->opt_seed_impl = v; }
    if (rc == ARACHNE_OK) {  // environment knobs are read here, once, not inside the batch call
        if (getenv("ARACHNE_RESC_MULT") && atoi(getenv("ARACHNE_RESC_MULT")) > 0) (*out)->env_resc_mult = atoi(getenv("ARACHNE_RESC_MULT"));
        if (getenv("ARACHNE_CIGAR_TIERS") && atoi(getenv("ARACHNE_CIGAR_TIERS")) == 0) (*out)->env_cigar_tiers = false;
    }
    if (rc == ARACHNE_OK && getenv("ARACHNE_SEED3_THREADS")) { const int v = atoi(getenv("ARACHNE_SEED3_THREADS")); if (v == 512 || v == 768 || v == 1024) (*out)->opt_seed3_threads = v; }
    return rc;
}

void arachne_index_free(arachne_index_t* idx) {
    if (!idx) return;
    cudaSetDevice(idx->device);
    for (int k = 0; k < arachne_index::N_WS; ++k)
        if (idx->ws[k]) { auto* w = reinterpret_cast<Workspace*>(idx->ws[k]); w->destroy(); delete w; idx->ws[k] = nullptr; }
    for (void* p : idx->priv_pool) { auto* bp = reinterpret_cast<BatchPriv*>(p); bp->release(); delete bp; }
    idx->priv_pool.clear();
    cudaFree(idx->d_bwt); cudaFree(idx->d_sa); cudaFree(idx->d_pac); cudaFree(idx->d_contigs); cudaFree(idx->d_kmer); cudaFree(idx->d_isa); cudaFree(idx->d_text);
    delete idx;
}
int64_t arachne_index_l_pac(const arachne_index_t* idx) { return idx ? idx->l_pac : -1; }
int32_t arachne_index_n_seqs(const arachne_index_t* idx) { return idx ? (int32_t)idx->contigs.size() : -1; }
size_t arachne_index_device_bytes(const arachne_index_t* idx) { return idx ? idx->device_bytes : 0; }
int64_t arachne_index_fetch_seq(const arachne_index_t* idx, int64_t* beg, int64_t mid, int64_t* end, int32_t* rid, uint8_t* out) {
    if (!idx || !beg || !end || !out) { set_error("arachne_index_fetch_seq: null argument"); return ARACHNE_ERR_ARG; }
    if (*end < *beg) { const int64_t t = *beg; *beg = *end; *end = t; }
    const int64_t l_pac = idx->l_pac;
    if (!(*beg <= mid && mid < *end) || mid < 0 || mid >= 2 * l_pac) {  // the reference asserts here (bntseq.c:427)
        set_error("arachne_index_fetch_seq: mid %lld outside [%lld, %lld) or the reference", (long long)mid, (long long)*beg, (long long)*end);
        return ARACHNE_ERR_ARG;
    }
    const bool is_rev = mid >= l_pac;                        // bns_depos, bntseq.h
    const int64_t pos_f = is_rev ? 2 * l_pac - 1 - mid : mid;
    int lo = 0, hi = (int)idx->contigs.size(), r = 0;        // bns_pos2rid, bntseq.c:346-363
    while (lo < hi) {
        const int m = (lo + hi) >> 1;
        if (pos_f >= idx->contigs[m].offset) { r = m; lo = m + 1; } else hi = m;
    }
    int64_t far_beg = idx->contigs[r].offset, far_end = far_beg + idx->contigs[r].len;
    if (is_rev) { const int64_t t = far_beg; far_beg = 2 * l_pac - far_end; far_end = 2 * l_pac - t; }
    if (*beg < far_beg) *beg = far_beg;
    if (*end > far_end) *end = far_end;
    if (rid) *rid = r;
    const uint8_t* pac = idx->h_pac.data();
    int64_t n = 0;
    for (int64_t p = *beg; p < *end; ++p) {                  // bns_get_seq, bntseq.c:398-419
        const int64_t f = p >= l_pac ? 2 * l_pac - 1 - p : p;
        const int b = pac[f >> 2] >> ((~f & 3) << 1) & 3;
        out[n++] = (uint8_t)(p >= l_pac ? 3 - b : b);
    }
    return n;
}
int arachne_index_contig(const arachne_index_t* idx, int32_t rid, const char** name, int64_t* offset, int32_t* len) {
    if (!idx || rid < 0 || rid >= (int32_t)idx->contigs.size()) { set_error("contig id out of range"); return ARACHNE_ERR_ARG; }
    if (name) *name = idx->names[rid].c_str();
    if (offset) *offset = idx->contigs[rid].offset;
    if (len) *len = idx->contigs[rid].len;
    return ARACHNE_OK;
}

int arachne_opts_default(arachne_mem_opt_t* o) {  // mem_opt_init, bwa/bwamem.c:48-84
    if (!o) { set_error("null options"); return ARACHNE_ERR_ARG; }
    memset(o, 0, sizeof(*o));
    o->a = 1; o->b = 4;
    o->o_del = o->o_ins = 6;
    o->e_del = o->e_ins = 1;
    o->w = 100; o->T = 30; o->zdrop = 100;
    o->pen_unpaired = 17;
    o->pen_clip5 = o->pen_clip3 = 5;
    o->max_mem_intv = 20;
    o->min_seed_len = 19; o->split_width = 10; o->max_occ = 500; o->max_chain_gap = 10000; o->max_ins = 10000;
    o->mask_level = 0.50f; o->drop_ratio = 0.50f; o->XA_drop_ratio = 0.80f; o->split_factor = 1.5f;
    o->chunk_size = 10000000; o->n_threads = 1;
    o->max_XA_hits = 5; o->max_XA_hits_alt = 200; o->max_matesw = 50;
    o->mask_level_redun = 0.95f; o->min_chain_weight = 0; o->max_chain_extend = 1 << 30;
    o->mapQ_coef_len = 50; o->mapQ_coef_fac = 3;  // (int)log(50)
    for (int i = 0, k = 0; i < 4; ++i) {
        for (int j = 0; j < 4; ++j) o->mat[k++] = i == j ? o->a : -o->b;
        o->mat[k++] = -1;
    }
    for (int j = 0; j < 5; ++j) o->mat[20 + j] = -1;
    return ARACHNE_OK;
}

// md5(name)[0:8] little-endian as int64: the seed tagBestAlignments hands to rand.NewSource (aligner.go:1413-1416).  csrc/md5_seed.h.
int64_t arachne_name_seed(const char* name, int32_t len) {
    if (!name || len < 0) return 1;
    return md5_name_seed(name, len);
}

// ---------------------------------------------------------------------------------------------- batch
#include "batch.inl"

#include "records_host.inl"

// ---------------------------------------------------------------------------------------------- stage entry points
// DPX issue peak: a register-resident chain-free stream of VIADDMNMX (the max-plus instruction the DP kernels are built on).
// 8 independent accumulators per thread, 1024 threads per SM resident.  Returns giga-ops/s (one op = one 32-bit viaddmax).
int arachne_dpx_peak(int device, double* gops, float* ms_out) {
    if (!gops) { set_error("arachne_dpx_peak: null argument"); return ARACHNE_ERR_ARG; }
    if (arachne_device_count() <= device) { set_error("no CUDA device %d (no CPU path)", device); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaSetDevice(device));
    int* d = nullptr; AR_CUDA(cudaMalloc(&d, 4096));
    const int iters = 4096, blocks = sm_count(device) * 4, threads = 256;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k_dpx_peak<<<blocks, threads>>>(d, iters, 1);
    cudaEventRecord(e0);
    for (int r = 0; r < 5; ++r) k_dpx_peak<<<blocks, threads>>>(d, iters, r + 2);
    cudaEventRecord(e1);
    AR_CUDA(cudaDeviceSynchronize());
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    *gops = 5.0 * (double)blocks * threads * iters * 16 / (ms * 1e-3) / 1e9;
    if (ms_out) *ms_out = ms / 5;
    cudaFree(d); cudaEventDestroy(e0); cudaEventDestroy(e1);
    return ARACHNE_OK;
}

int arachne_occ_batch(arachne_index_t* idx, int64_t n, const uint64_t* k, const uint64_t* l, uint64_t* cnt) {
    if (!idx || !k || !l || !cnt || n < 0) { set_error("arachne_occ_batch: bad argument"); return ARACHNE_ERR_ARG; }
    AR_CUDA(cudaSetDevice(idx->device));
    uint64_t *dk = nullptr, *dl = nullptr, *dc = nullptr;
    AR_CUDA(cudaMalloc(&dk, n * 8 + 8)); AR_CUDA(cudaMalloc(&dl, n * 8 + 8)); AR_CUDA(cudaMalloc(&dc, n * 64 + 8));
    AR_CUDA(cudaMemcpy(dk, k, n * 8, cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(dl, l, n * 8, cudaMemcpyHostToDevice));
    k_occ<<<256, 256>>>(idx->fm, (long)n, dk, dl, dc);
    AR_CUDA(cudaGetLastError());
    AR_CUDA(cudaMemcpy(cnt, dc, n * 64, cudaMemcpyDeviceToHost));
    cudaFree(dk); cudaFree(dl); cudaFree(dc);
    return ARACHNE_OK;
}

int arachne_sa_batch(arachne_index_t* idx, int64_t n, const uint64_t* k, uint64_t* sa) {
    if (!idx || !k || !sa || n < 0) { set_error("arachne_sa_batch: bad argument"); return ARACHNE_ERR_ARG; }
    AR_CUDA(cudaSetDevice(idx->device));
    uint64_t* dk = nullptr; Counters* dc = nullptr;
    AR_CUDA(cudaMalloc(&dk, n * 8 + 8)); AR_CUDA(cudaMalloc(&dc, sizeof(Counters) + 8));
    AR_CUDA(cudaMemset(dc, 0, sizeof(Counters) + 8));
    AR_CUDA(cudaMemcpy(dk, k, n * 8, cudaMemcpyHostToDevice));
    k_sa<<<256, 256>>>(idx->fm, (unsigned long long)n, dk, reinterpret_cast<unsigned long long*>(dc + 1), dc);
    AR_CUDA(cudaGetLastError());
    AR_CUDA(cudaMemcpy(sa, dk, n * 8, cudaMemcpyDeviceToHost));
    cudaFree(dk); cudaFree(dc);
    return ARACHNE_OK;
}

int arachne_seed_batch(arachne_index_t* idx, const arachne_mem_opt_t* opt_, int64_t n_reads, const uint8_t* seq, const int64_t* seq_off,
                       int64_t* intv_off, uint64_t* intv, int64_t intv_cap) {
    if (!idx || !seq || !seq_off || !intv_off || !intv) { set_error("arachne_seed_batch: null argument"); return ARACHNE_ERR_ARG; }
    int rc = validate_opts(opt_);
    if (rc) return rc;
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    AR_CUDA(cudaSetDevice(idx->device));
    int maxlen = 0;
    for (int64_t i = 0; i < n_reads; ++i) maxlen = std::max<int>(maxlen, (int)(seq_off[i + 1] - seq_off[i]));
    if (maxlen > ARACHNE_MAX_READ_LEN) { set_error("read longer than %d", ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG; }
    const S3Plan s3 = seed3_plan(maxlen);
    const int seed_impl = idx->opt_seed_impl;
    const int L1 = maxlen + 1, blocks = sm_count(idx->device) * 3, rows = 2 * L1;
    const size_t T = seed_impl == 3 ? (size_t)sm_count(idx->device) * s3.ns : (size_t)blocks * 256;
    int cap = 32;
    DevBuf d_seq, d_off, d_scr, d_arena, d_next, d_in, d_nsa, d_cnt;
    int64_t nb = seq_off[n_reads];
    DevBuf d_q, d_c2;
    auto rel = [&]() { for (DevBuf* b : {&d_seq, &d_off, &d_scr, &d_arena, &d_next, &d_in, &d_nsa, &d_cnt, &d_q, &d_c2}) b->release(); };
    if (d_seq.ensure(nb + 16) || d_off.ensure((n_reads + 1) * 8) || d_scr.ensure((size_t)T * rows * sizeof(uint4)) || d_next.ensure(64) ||
        d_in.ensure(n_reads * 4 + 4) || d_nsa.ensure(n_reads * 4 + 4) || d_cnt.ensure(sizeof(Counters))) { rel(); set_error("out of device memory"); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaMemcpy(d_seq.p, seq, nb, cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(d_off.p, seq_off, (n_reads + 1) * 8, cudaMemcpyHostToDevice));
    Counters hc;
    const ulonglong2* d_qpk = nullptr;
    const int qstride = seed_qstride(maxlen);
    const int seed_direct = idx->opt_seed_direct && idx->fm.isa && idx->fm.text;
    int slist_n = 0, nwq = 0;
    const size_t seed_smem = seed_smem_plan(maxlen, 3, -1, &slist_n, &nwq);
    {
        if (d_q.ensure((size_t)n_reads * qstride * sizeof(ulonglong2))) { rel(); set_error("out of device memory"); return ARACHNE_ERR_CUDA; }
        k_pack_reads<<<148 * 8, 256>>>(d_seq.as<uint8_t>(), d_off.as<int64_t>(), (int)n_reads, qstride, d_q.as<ulonglong2>());
        AR_CUDA(cudaGetLastError());
        d_qpk = d_q.as<ulonglong2>();
    }
    for (int attempt = 0;; ++attempt) {
        if (d_arena.ensure((size_t)n_reads * cap * sizeof(Intv) + 32)) { rel(); set_error("out of device memory"); return ARACHNE_ERR_CUDA; }
        AR_CUDA(cudaMemset(d_next.p, 0, 64)); AR_CUDA(cudaMemset(d_cnt.p, 0, sizeof(Counters)));
        AR_CUDA(cudaMemset(d_in.p, 0, n_reads * 4 + 4)); AR_CUDA(cudaMemset(d_nsa.p, 0, n_reads * 4 + 4));
        if (d_c2.ensure(n_reads * 8 + 8)) { rel(); set_error("out of device memory"); return ARACHNE_ERR_CUDA; }
        AR_CUDA(cudaMemset(d_c2.p, 0, n_reads * 8 + 8));
        SeedOut so; so.intv = d_arena.as<Intv>(); so.cnt2 = d_c2.as<int>(); so.cap = cap;
        if (seed_impl == 3) {
            const int nt = idx->opt_seed3_threads;
            auto kfn = nt == 512 ? k_seed3<512> : nt == 1024 ? k_seed3<1024> : k_seed3<768>;
            AR_CUDA(cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)s3.smem));
            kfn<<<sm_count(idx->device), nt, s3.smem>>>(idx->fm, opt, d_off.as<int64_t>(), (int)n_reads, maxlen, d_scr.as<uint4>(), rows, so,
                                    d_next.as<unsigned int>(), d_cnt.as<Counters>(), idx->opt_seed_kmer, s3.ns, d_qpk, qstride, s3.nwq, seed_direct);
        } else {
        AR_CUDA(cudaFuncSetAttribute(k_seed2<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)seed_smem));
        k_seed2<3><<<blocks, 256, seed_smem>>>(idx->fm, opt, d_off.as<int64_t>(), (int)n_reads, maxlen, d_scr.as<uint4>(), rows, so,
                                    d_next.as<unsigned int>(), d_cnt.as<Counters>(), idx->opt_seed_kmer, slist_n, d_qpk, qstride, nwq, seed_direct);
        }
        AR_CUDA(cudaGetLastError());
        AR_CUDA(cudaMemcpy(&hc, d_cnt.p, sizeof(hc), cudaMemcpyDeviceToHost));
        if (!hc.err) break;
        if (attempt >= 3) { rel(); set_error("more than %d intervals at read %d", cap, hc.err_read); return ARACHNE_ERR_OVERFLOW; }
        cap *= 4;
    }
    k_sort_intv<<<256, 128>>>((int)n_reads, d_arena.as<Intv>(), d_c2.as<int>(), cap, opt.max_occ, d_in.as<int>(), d_nsa.as<unsigned int>());
    AR_CUDA(cudaGetLastError());
    std::vector<int> in_(n_reads);
    AR_CUDA(cudaMemcpy(in_.data(), d_in.p, n_reads * 4, cudaMemcpyDeviceToHost));
    std::vector<Intv> a((size_t)n_reads * cap + 1);
    AR_CUDA(cudaMemcpy(a.data(), d_arena.p, (size_t)n_reads * cap * sizeof(Intv), cudaMemcpyDeviceToHost));
    int64_t o = 0;
    for (int64_t r = 0; r < n_reads; ++r) {
        intv_off[r] = o;
        if (o + in_[r] > intv_cap) { rel(); set_error("interval capacity too small"); return ARACHNE_ERR_OVERFLOW; }
        memcpy(intv + 4 * o, a.data() + (size_t)r * cap, (size_t)in_[r] * sizeof(Intv));
        o += in_[r];
    }
    intv_off[n_reads] = o;
    rel();
    return ARACHNE_OK;
}

int arachne_ksw_extend_batch(int device, const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* q, const int64_t* q_off, const uint8_t* t,
                             const int64_t* t_off, const int32_t* h0, const int32_t* w, int end_bonus, int zdrop, int32_t* out, uint64_t* cells,
                             float* kernel_ms, int repeat) {
    if (!q || !q_off || !t || !t_off || !h0 || !w || !out || n < 0 || n > 0x7fffffff) { set_error("arachne_ksw_extend_batch: bad argument"); return ARACHNE_ERR_ARG; }
    int rc = validate_opts(opt_);
    if (rc) return rc;
    int maxq = 1;
    bool wide = false;  // some problem's scores do not fit a byte: the 13-bit column codec
    std::vector<ExtProb> probs((size_t)n);
    std::vector<int> order((size_t)n);
    for (int64_t i = 0; i < n; ++i) {
        const int64_t ql = q_off[i + 1] - q_off[i], tl = t_off[i + 1] - t_off[i];
        if (ql > ARACHNE_MAX_READ_LEN || ql < 1 || h0[i] <= 0 || tl < 0 || tl > 0x7fffffff) {
            set_error("problem %lld: qlen must be 1..%d and h0 > 0 (ksw.c:385)", (long long)i, ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG;
        }
        if ((long)h0[i] + ql * opt_->a > 8191) { set_error("problem %lld: h0 + qlen * a = %ld exceeds the kernel's 13-bit score fields", (long long)i, (long)h0[i] + ql * opt_->a); return ARACHNE_ERR_ARG; }
        if ((long)h0[i] + ql * opt_->a > 255) wide = true;
        maxq = std::max<int>(maxq, (int)ql);
        ExtProb& P = probs[(size_t)i];
        P.read = (int32_t)i; P.q0 = 0; P.qlen = (int32_t)ql; P.tlen = (int32_t)tl; P.w = w[i]; P.h0 = h0[i]; P.end_bonus = end_bonus; P.dirs = 0; P.tpos = t_off[i]; P.qoff = q_off[i];
        order[(size_t)i] = (int)i;
    }
    // the queue k_ext_dp consumes is sorted by query length, longest first (k_ext_binbase / k_ext_scatter do this in the pipeline)
    std::stable_sort(order.begin(), order.end(), [&](int x, int y) { return probs[(size_t)x].qlen > probs[(size_t)y].qlen; });
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    if (arachne_device_count() <= device) { set_error("no CUDA device %d (no CPU path)", device); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaSetDevice(device));
    std::vector<ExtProb> queue((size_t)n);
    for (int64_t i = 0; i < n; ++i) queue[(size_t)i] = probs[(size_t)order[(size_t)i]];
    DevBuf dq, dt, dp, dres, dc, dctl;
    if (dq.ensure(q_off[n] + 16) || dt.ensure(t_off[n] + 16) || dp.ensure(n * sizeof(ExtProb) + 8) ||
        dres.ensure(n * sizeof(ExtRes) + 8) || dc.ensure(sizeof(Counters)) || dctl.ensure(16)) { set_error("out of device memory"); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaMemcpy(dq.p, q, q_off[n], cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(dt.p, t, t_off[n], cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(dp.p, queue.data(), n * sizeof(ExtProb), cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemset(dc.p, 0, sizeof(Counters)));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    if (repeat < 1) repeat = 1;
    ExtDpArgs DA;
    DA.seq = dq.as<uint8_t>(); DA.tbytes = dt.as<uint8_t>(); DA.queue = dp.as<ExtProb>(); DA.res = dres.as<ExtRes>();
    DA.n_items = dctl.as<unsigned>(); DA.head = dctl.as<unsigned>() + 1; DA.ncols = maxq + 2; DA.zdrop = zdrop; DA.one = 1; DA.cnt = dc.as<Counters>();
    const size_t smem = (size_t)DA.ncols * XDP_BLOCK * sizeof(uint32_t);
    int per_sm = 1;
    auto dp_kernel = wide ? k_ext_dp<false, true> : k_ext_dp<false, false>;
    AR_CUDA(cudaFuncSetAttribute(dp_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    AR_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, dp_kernel, XDP_BLOCK, smem));
    if (per_sm < 1) per_sm = 1;
    const int blocks = sm_count(device) * per_sm;
    RefView no_ref; memset(&no_ref, 0, sizeof(no_ref));
    const unsigned ctl0[2] = {(unsigned)n, 0u};
    for (int it = 0; it <= repeat; ++it) {  // one untimed launch, then `repeat` timed ones
        if (it == 1) { AR_CUDA(cudaMemset(dc.p, 0, sizeof(Counters))); cudaEventRecord(e0); }
        AR_CUDA(cudaMemcpyAsync(dctl.p, ctl0, 8, cudaMemcpyHostToDevice));
        dp_kernel<<<blocks, XDP_BLOCK, smem>>>(opt, no_ref, DA);
        AR_CUDA(cudaGetLastError());
    }
    cudaEventRecord(e1);
    AR_CUDA(cudaDeviceSynchronize());
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    if (kernel_ms) *kernel_ms = ms / repeat;
    std::vector<ExtRes> res((size_t)n);
    AR_CUDA(cudaMemcpy(res.data(), dres.p, n * sizeof(ExtRes), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < n; ++i) {
        const ExtRes& R = res[(size_t)i];
        int32_t* o = out + 6 * i;
        o[0] = R.mx; o[1] = R.max_j + 1; o[2] = R.max_i + 1; o[3] = R.max_ie + 1; o[4] = R.gscore; o[5] = R.max_off;
    }
    if (cells) {
        Counters hc;
        AR_CUDA(cudaMemcpy(&hc, dc.p, sizeof(hc), cudaMemcpyDeviceToHost));
        *cells = hc.ext_cells / (unsigned)repeat;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return ARACHNE_OK;
}

int arachne_ksw_global_batch(int device, const arachne_mem_opt_t* opt_, int64_t n, const uint8_t* q, const int64_t* q_off, const uint8_t* t,
                             const int64_t* t_off, const int32_t* w, int32_t* scores, int64_t* cigar_off, uint32_t* cigar, int64_t cigar_cap) {
    if (!q || !q_off || !t || !t_off || !w || !scores || n < 0) { set_error("arachne_ksw_global_batch: null argument"); return ARACHNE_ERR_ARG; }
    int rc = validate_opts(opt_);
    if (rc) return rc;
    MemOpt opt; memcpy(&opt, opt_, sizeof(opt));
    if (arachne_device_count() <= device) { set_error("no CUDA device %d (no CPU path)", device); return ARACHNE_ERR_CUDA; }
    AR_CUDA(cudaSetDevice(device));
    long zmax = 1; int cmax = 8;
    for (int64_t i = 0; i < n; ++i) {
        long ql = q_off[i + 1] - q_off[i], tl = t_off[i + 1] - t_off[i];
        if (ql > ARACHNE_MAX_READ_LEN || ql < 1) { set_error("problem %lld: qlen must be 1..%d", (long long)i, ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG; }
        long ncol = std::min<long>(ql, 2L * w[i] + 1);
        zmax = std::max(zmax, ncol * tl);
        cmax = std::max<int>(cmax, (int)(ql + tl + 2));
    }
    const int blocks = sm_count(device) * 2;
    const long nwarps = (long)blocks * EXT_WARPS;
    DevBuf dq, dqo, dt, dto, dw, ds, dz, dcg, dn;
    bool want = cigar_off != nullptr;
    if (dq.ensure(q_off[n] + 16) || dqo.ensure((n + 1) * 8) || dt.ensure(t_off[n] + 16) || dto.ensure((n + 1) * 8) || dw.ensure(n * 4 + 4) ||
        ds.ensure(n * 4 + 4) || (want && (dz.ensure((size_t)nwarps * zmax) || dcg.ensure((size_t)n * cmax * 4 + 4) || dn.ensure(n * 4 + 4)))) {
        set_error("out of device memory"); return ARACHNE_ERR_CUDA;
    }
    AR_CUDA(cudaMemcpy(dq.p, q, q_off[n], cudaMemcpyHostToDevice)); AR_CUDA(cudaMemcpy(dqo.p, q_off, (n + 1) * 8, cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(dt.p, t, t_off[n], cudaMemcpyHostToDevice)); AR_CUDA(cudaMemcpy(dto.p, t_off, (n + 1) * 8, cudaMemcpyHostToDevice));
    AR_CUDA(cudaMemcpy(dw.p, w, n * 4, cudaMemcpyHostToDevice));
    k_ksw_global<<<blocks, EXT_WARPS * 32>>>(opt, (long)n, dq.as<uint8_t>(), dqo.as<int64_t>(), dt.as<uint8_t>(), dto.as<int64_t>(), dw.as<int>(),
                                             ds.as<int>(), want ? dz.as<uint8_t>() : nullptr, zmax, want ? dcg.as<uint32_t>() : nullptr, cmax,
                                             want ? dn.as<int>() : nullptr);
    AR_CUDA(cudaGetLastError());
    AR_CUDA(cudaMemcpy(scores, ds.p, n * 4, cudaMemcpyDeviceToHost));
    if (want) {
        std::vector<int> nc(n); std::vector<uint32_t> cg((size_t)n * cmax);
        AR_CUDA(cudaMemcpy(nc.data(), dn.p, n * 4, cudaMemcpyDeviceToHost));
        AR_CUDA(cudaMemcpy(cg.data(), dcg.p, (size_t)n * cmax * 4, cudaMemcpyDeviceToHost));
        int64_t o = 0;
        for (int64_t i = 0; i < n; ++i) {
            cigar_off[i] = o;
            if (o + nc[i] > cigar_cap) { set_error("cigar pool too small"); return ARACHNE_ERR_OVERFLOW; }
            memcpy(cigar + o, cg.data() + (size_t)i * cmax, (size_t)nc[i] * 4);
            o += nc[i];
        }
        cigar_off[n] = o;
    }
    for (DevBuf* b : {&dq, &dqo, &dt, &dto, &dw, &ds, &dz, &dcg, &dn}) b->release();
    return ARACHNE_OK;
}

}  // extern "C"

#include "noalloc.inl"
