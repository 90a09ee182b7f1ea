// TEST-ONLY: the gobwa-shaped layer (csrc/compat.cu, pure host code) compiled for a machine without a GPU, with the batch entry points it
// calls -- arachne_align_batch, arachne_batch_free, arachne_opts_default, arachne_index_fetch_seq -- served by the reference's own C code
// (oracle/_ref: ref_batch = gobwa.go's call sequence per pair).  What this exercises is the layer's OWN logic: leader / follower batching
// of concurrent per-pair calls, result blocks and their lifetime, the region-address lookup of GoBwaSmithWaterman -- functionally against
// the reference and, built with -fsanitize=thread, for data races (scripts/host_sanitize.sh).  Nothing in the shipped library links this file.
#include <atomic>
#include <chrono>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include "../../arachne_b200/csrc/compat.cu"

extern "C" {
typedef struct { int64_t n_reads, n_regs, n_cigar; int64_t* reg_off; void* regs; int64_t* aln; uint32_t* cigar; } ref_batch_t;  // oracle/ref_shim.c
ref_batch_t* ref_batch(void* idx, int64_t n_pairs, const uint8_t* seq, const int64_t* off, int score_delta, int do_rescue, int do_cigar, int n_threads);
void ref_batch_free(ref_batch_t* b);
void* ref_idx_load(const char* prefix);
void ref_idx_free(void* idx);
int ref_fetch_seq(void* idx, int64_t* beg, int64_t mid, int64_t* end, uint8_t* out, int cap);
void ref_opt_default(void* out);
int ref_sizeof_opt(void);
}

namespace arachne {
static thread_local char g_err[512];
void set_error(const char* fmt, ...) { va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof g_err, fmt, ap); va_end(ap); }
}  // namespace arachne

static std::atomic<int> g_in_flight{0}, g_max_in_flight{0}, g_fail_next{0};

extern "C" {

const char* arachne_last_error(void) { return arachne::g_err; }

// the "index handle" of the mock is the reference's bwaidx_t
arachne_index_t* mock_index_open(const char* prefix) { return reinterpret_cast<arachne_index_t*>(ref_idx_load(prefix)); }
void mock_index_close(arachne_index_t* idx) { ref_idx_free(idx); }
int mock_max_batches_in_flight(void) { return g_max_in_flight.load(); }
void mock_fail_next_batches(int n) { g_fail_next = n; }  // the next n batches return ARACHNE_ERR_CUDA

int arachne_opts_default(arachne_mem_opt_t* o) {
    if ((size_t)ref_sizeof_opt() != sizeof(arachne_mem_opt_t)) return ARACHNE_ERR_OPTS;
    ref_opt_default(o);
    return ARACHNE_OK;
}

int arachne_align_batch(arachne_index_t* idx, const arachne_mem_opt_t*, const arachne_batch_in_t* in, arachne_batch_out_t** out_) {
    const int now = ++g_in_flight;
    for (int m = g_max_in_flight.load(); now > m && !g_max_in_flight.compare_exchange_weak(m, now);) {}
    std::this_thread::sleep_for(std::chrono::milliseconds(3));  // a batch takes a while on the device too: callers pile up behind it
    int rc = ARACHNE_OK;
    if (g_fail_next.load() > 0 && g_fail_next.fetch_sub(1) > 0) { arachne::set_error("mock: batch failed on request"); rc = ARACHNE_ERR_CUDA; }
    if (rc == ARACHNE_OK) {
        ref_batch_t* b = ref_batch(idx, in->n_pairs, in->seq, in->seq_off, in->score_delta, (in->stages & ARACHNE_STAGE_RESCUE) != 0, 1, 1);
        auto* o = static_cast<arachne_batch_out_t*>(calloc(1, sizeof(arachne_batch_out_t)));
        o->n_reads = b->n_reads; o->n_regs = b->n_regs; o->n_cigar = b->n_cigar;
        o->reg_off = static_cast<int64_t*>(malloc((size_t)(b->n_reads + 1) * 8)); memcpy(o->reg_off, b->reg_off, (size_t)(b->n_reads + 1) * 8);
        o->regs = static_cast<arachne_alnreg_t*>(malloc((size_t)b->n_regs * sizeof(arachne_alnreg_t) + 8)); memcpy(o->regs, b->regs, (size_t)b->n_regs * sizeof(arachne_alnreg_t));
        o->alns = static_cast<arachne_aln_t*>(calloc((size_t)b->n_regs + 1, sizeof(arachne_aln_t)));
        for (int64_t g = 0; g < b->n_regs; ++g) {
            const int64_t* a = b->aln + 8 * g;  // pos rid is_rev NM n_cigar cigar_off score sub
            arachne_aln_t& x = o->alns[g];
            x.pos = a[0]; x.rid = (int32_t)a[1]; x.is_rev = (int32_t)a[2]; x.nm = (int32_t)a[3]; x.n_cigar = (int32_t)a[4]; x.cigar_off = a[5]; x.score = (int32_t)a[6]; x.sub = (int32_t)a[7];
        }
        o->cigar = static_cast<uint32_t*>(malloc((size_t)b->n_cigar * 4 + 8)); memcpy(o->cigar, b->cigar, (size_t)b->n_cigar * 4);
        ref_batch_free(b);
        *out_ = o;
    }
    --g_in_flight;
    return rc;
}

void arachne_batch_free(arachne_batch_out_t* o) { if (!o) return; free(o->reg_off); free(o->regs); free(o->alns); free(o->cigar); free(o); }

int64_t arachne_index_fetch_seq(const arachne_index_t* idx, int64_t* beg, int64_t mid, int64_t* end, int32_t* rid, uint8_t* out) {
    const int64_t span = *end > *beg ? *end - *beg : *beg - *end;
    const int r = ref_fetch_seq(const_cast<arachne_index_t*>(idx), beg, mid, end, out, (int)span);
    if (rid) *rid = r;
    return *end - *beg;
}

}  // extern "C"
