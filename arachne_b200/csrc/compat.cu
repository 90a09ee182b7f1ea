// The gobwa-shaped layer (include/arachne_compat.h): per-pair / per-candidate calls of the reference's Go host served from GPU
// batches that are formed from the calls in flight at the same moment (leader / follower dynamic batching).  Host code only:
// everything device-side goes through arachne_align_batch().
#include <cuda_runtime.h>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>
#include "../../include/arachne_compat.h"

namespace arachne { void set_error(const char* fmt, ...); }
using arachne::set_error;

namespace {

// One pair's results: header, then regions of read 1, regions of read 2, their alignment records and CIGAR words, in one malloc.
struct PairBlock {
    uint64_t magic;
    int32_t n1, n2;
    arachne_alnreg_t* regs;   // n1 + n2
    arachne_aln_t* alns;      // n1 + n2; cigar_off re-based into this block's cigar array
    uint32_t* cigar;
};
constexpr uint64_t PAIR_MAGIC = 0x41524143484e4550ull;  // "ARACHNEP"

struct Request {
    const uint8_t *r1, *r2;
    int l1, l2, delta;
    PairBlock* block = nullptr;
    int rc = 0;
    bool taken = false;  // some leader has put this request into its batch
    bool done = false;
};

}  // namespace

struct arachne_compat {
    arachne_index_t* idx;
    arachne_mem_opt_t opt;
    int max_batch, max_wait_us;
    std::mutex mu;
    std::condition_variable cv_done, cv_more;
    std::vector<Request*> pending;
    bool leader_waiting = false;
    std::map<uintptr_t, PairBlock*> live;  // first region address -> block (blocks without regions are not listed)
    arachne_compat_stats_t st{};
};

// Runs one batch for the given requests and fills their blocks.  Called without the lock held.
static void run_batch(arachne_compat* c, std::vector<Request*>& reqs) {
    const int64_t n = (int64_t)reqs.size();
    std::vector<int64_t> off((size_t)(2 * n + 1));
    int64_t nb = 0;
    for (int64_t p = 0; p < n; ++p) { off[2 * p] = nb; nb += reqs[p]->l1; off[2 * p + 1] = nb; nb += reqs[p]->l2; }
    off[2 * n] = nb;
    std::vector<uint8_t> seq((size_t)nb + 16);
    for (int64_t p = 0; p < n; ++p) {
        if (reqs[p]->l1) memcpy(seq.data() + off[2 * p], reqs[p]->r1, (size_t)reqs[p]->l1);
        if (reqs[p]->l2) memcpy(seq.data() + off[2 * p + 1], reqs[p]->r2, (size_t)reqs[p]->l2);
    }
    arachne_batch_in_t in;
    memset(&in, 0, sizeof(in));
    in.n_pairs = n; in.seq = seq.data(); in.seq_off = off.data(); in.score_delta = reqs[0]->delta;
    in.stages = ARACHNE_STAGE_REGIONS | ARACHNE_STAGE_RESCUE | ARACHNE_STAGE_CIGAR;
    arachne_batch_out_t* out = nullptr;
    int rc = arachne_align_batch(c->idx, &c->opt, &in, &out);
    for (int64_t p = 0; p < n && rc == ARACHNE_OK; ++p) {
        const int64_t g0 = out->reg_off[2 * p], g1 = out->reg_off[2 * p + 1], g2 = out->reg_off[2 * p + 2];
        const int64_t nr = g2 - g0;
        int64_t ncig = 0;
        for (int64_t g = g0; g < g2; ++g) ncig += out->alns[g].n_cigar;
        const size_t bytes = sizeof(PairBlock) + (size_t)nr * (sizeof(arachne_alnreg_t) + sizeof(arachne_aln_t)) + (size_t)ncig * 4 + 16;
        auto* b = static_cast<PairBlock*>(malloc(bytes));
        if (!b) { rc = ARACHNE_ERR_CUDA; set_error("arachne_compat: out of host memory"); break; }
        b->magic = PAIR_MAGIC; b->n1 = (int32_t)(g1 - g0); b->n2 = (int32_t)(g2 - g1);
        b->regs = reinterpret_cast<arachne_alnreg_t*>(b + 1);
        b->alns = reinterpret_cast<arachne_aln_t*>(b->regs + nr);
        b->cigar = reinterpret_cast<uint32_t*>(b->alns + nr);
        if (nr) { memcpy(b->regs, out->regs + g0, (size_t)nr * sizeof(arachne_alnreg_t)); memcpy(b->alns, out->alns + g0, (size_t)nr * sizeof(arachne_aln_t)); }
        int64_t co = 0;
        for (int64_t k = 0; k < nr; ++k) {
            const int nc = b->alns[k].n_cigar;
            if (nc) memcpy(b->cigar + co, out->cigar + b->alns[k].cigar_off, (size_t)nc * 4);
            b->alns[k].cigar_off = co;
            co += nc;
        }
        reqs[p]->block = b;
    }
    if (out) arachne_batch_free(out);
    if (rc != ARACHNE_OK)  // all or nothing: a caller that gets an error code never sees (and could never release) a block
        for (Request* r : reqs) { free(r->block); r->block = nullptr; }
    for (Request* r : reqs) r->rc = rc;
}

extern "C" {

int arachne_compat_open(arachne_index_t* idx, const arachne_mem_opt_t* opt, int max_batch_pairs, int max_wait_us, arachne_compat_t** out) {
    if (!idx || !out) { set_error("arachne_compat_open: null argument"); return ARACHNE_ERR_ARG; }
    auto* c = new arachne_compat();
    c->idx = idx;
    if (opt) c->opt = *opt; else arachne_opts_default(&c->opt);
    c->max_batch = max_batch_pairs > 0 ? max_batch_pairs : 4096;
    c->max_wait_us = max_wait_us > 0 ? max_wait_us : 200;
    *out = c;
    return ARACHNE_OK;
}

void arachne_compat_close(arachne_compat_t* c) {
    if (!c) return;
    for (auto& kv : c->live) free(kv.second);
    delete c;
}

int arachne_compat_mem_mate_sw(arachne_compat_t* c, const uint8_t* r1, int l1, const uint8_t* r2, int l2, int score_delta,
                               arachne_alnreg_v* out1, arachne_alnreg_v* out2, void** owner) {
    if (!c || !out1 || !out2 || !owner || l1 < 0 || l2 < 0 || (l1 && !r1) || (l2 && !r2)) { set_error("arachne_compat_mem_mate_sw: bad argument"); return ARACHNE_ERR_ARG; }
    if (l1 > ARACHNE_MAX_READ_LEN || l2 > ARACHNE_MAX_READ_LEN) { set_error("read longer than %d", ARACHNE_MAX_READ_LEN); return ARACHNE_ERR_ARG; }
    Request req;
    req.r1 = r1; req.r2 = r2; req.l1 = l1; req.l2 = l2; req.delta = score_delta;
    std::unique_lock<std::mutex> lk(c->mu);
    c->pending.push_back(&req);
    ++c->st.n_calls;
    while (!req.done) {
        if (!c->leader_waiting && !req.taken) {
            // Leader: give the other workers max_wait_us to join (or until the batch is full), then take every waiting request
            // with this call's score_delta -- in the reference it is one constant, aligner.go:454 -- and run them as one batch.
            c->leader_waiting = true;
            const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(c->max_wait_us);
            while ((int)c->pending.size() < c->max_batch && c->cv_more.wait_until(lk, deadline) != std::cv_status::timeout) {}
            std::vector<Request*> mine, rest;
            for (Request* r : c->pending) ((r->delta == req.delta && (int)mine.size() < c->max_batch) ? mine : rest).push_back(r);
            for (Request* r : mine) r->taken = true;
            c->pending.swap(rest);
            c->leader_waiting = false;
            ++c->st.n_batches;
            if (mine.size() > c->st.max_batch) c->st.max_batch = mine.size();
            const bool left_behind = !c->pending.empty();
            lk.unlock();
            if (left_behind) c->cv_done.notify_all();  // somebody left behind must become the next leader
            run_batch(c, mine);                                // the handle has two workspaces: the next leader's batch overlaps this one
            lk.lock();
            for (Request* r : mine) {
                if (r->block && r->block->n1 + r->block->n2 > 0) { c->live[reinterpret_cast<uintptr_t>(r->block->regs)] = r->block; }
                if (r->block) ++c->st.n_live_blocks;
                r->done = true;
            }
            c->cv_done.notify_all();
        } else {
            if ((int)c->pending.size() >= c->max_batch) c->cv_more.notify_one();
            c->cv_done.wait(lk);
        }
    }
    lk.unlock();
    if (req.rc != ARACHNE_OK) return req.rc;
    PairBlock* b = req.block;
    out1->n = out1->m = (size_t)b->n1; out1->a = b->n1 ? b->regs : nullptr;
    out2->n = out2->m = (size_t)b->n2; out2->a = b->n2 ? b->regs + b->n1 : nullptr;
    *owner = b;
    return ARACHNE_OK;
}

int arachne_compat_reg2aln(arachne_compat_t* c, const arachne_alnreg_t* reg, arachne_go_aln_t* out) {
    if (!c || !reg || !out) { set_error("arachne_compat_reg2aln: null argument"); return ARACHNE_ERR_ARG; }
    PairBlock* b = nullptr;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        auto it = c->live.upper_bound(reinterpret_cast<uintptr_t>(reg));
        if (it != c->live.begin()) { --it; b = it->second; }
    }
    const uintptr_t addr = reinterpret_cast<uintptr_t>(reg), base = b ? reinterpret_cast<uintptr_t>(b->regs) : 0;
    const long k = b && addr >= base && (addr - base) % sizeof(arachne_alnreg_t) == 0 ? (long)((addr - base) / sizeof(arachne_alnreg_t)) : -1;
    if (!b || k < 0 || k >= b->n1 + b->n2) { set_error("arachne_compat_reg2aln: not a live region of this batcher"); return ARACHNE_ERR_ARG; }
    const arachne_aln_t& a = b->alns[k];
    const unsigned is_alt = ((unsigned)reg->n_comp_is_alt >> 30) & 1u;
    out->pos = a.pos; out->rid = a.rid; out->flag = 0;
    out->flag2 = (uint32_t)(a.is_rev != 0) | is_alt << 1 | (uint32_t)a.nm << 10;  // is_rev:1 is_alt:1 mapq:8 NM:22
    out->n_cigar = a.n_cigar; out->cigar = b->cigar + a.cigar_off; out->XA = nullptr;
    out->score = a.score; out->sub = a.sub; out->alt_sc = reg->alt_sc;
    return ARACHNE_OK;
}

void arachne_compat_release(arachne_compat_t* c, void* owner) {
    auto* b = static_cast<PairBlock*>(owner);
    if (!c || !b || b->magic != PAIR_MAGIC) return;
    {
        std::lock_guard<std::mutex> lk(c->mu);
        if (b->n1 + b->n2 > 0) c->live.erase(reinterpret_cast<uintptr_t>(b->regs));
        --c->st.n_live_blocks;
    }
    b->magic = 0;
    free(b);
}

uint8_t* arachne_compat_fetch_seq(const arachne_index_t* idx, int64_t* beg, int64_t mid, int64_t* end, int* rid) {
    if (!idx || !beg || !end) { set_error("arachne_compat_fetch_seq: null argument"); return nullptr; }
    const int64_t span = *end > *beg ? *end - *beg : *beg - *end;
    auto* buf = static_cast<uint8_t*>(malloc((size_t)span + 1));
    if (!buf) { set_error("arachne_compat_fetch_seq: out of host memory"); return nullptr; }
    int32_t r = -1;
    if (arachne_index_fetch_seq(idx, beg, mid, end, &r, buf) < 0) { free(buf); return nullptr; }
    if (rid) *rid = r;
    return buf;
}

void arachne_compat_stats(arachne_compat_t* c, arachne_compat_stats_t* out) {
    if (!c || !out) return;
    std::lock_guard<std::mutex> lk(c->mu);
    *out = c->st;
}

}  // extern "C"
