// Barcode-grouped FASTQ ingest: what src/fastqreader does in front of the aligner, as a multi-threaded native
// reader that hands out batches in exactly the layout arachne_align_batch() takes.
//   ReadOneLine      src/fastqreader/reader.go:129-198   one R1 + one R2 record per call, header -> ParseHeader
//   ParseHeader      :95-123   name = first field minus its 2-character mate suffix; barcode = BX:Z:(\S+)\s; valid = VX:i:([01])\s
//   ReadBarcodeSet   :208-291  consecutive records of one barcode, capped at 30000, 201-record continuation units
//   worthRunningRFA  src/aligner/aligner.go:1018-1030
//   FastZipReader    src/fastqreader/zipread.go:69-76 (`gunzip -c` child process; here zlib in a thread per file)
// Pipeline: one decompress/split thread per file cuts its stream into chunks of a fixed number of records (so chunk i of R1
// and chunk i of R2 hold the same pairs); `n_threads` parser threads turn chunk pairs into codes 0..4, qualities, names,
// barcodes and md5 name seeds; the caller's thread (arachne_ingest_next) walks the parsed chunks in order and applies the
// work-unit rule.  Host code only.
#include <zlib.h>
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>
#include "../../include/arachne_ingest.h"

namespace arachne { void set_error(const char* fmt, ...); }
using arachne::set_error;

namespace {

constexpr int CHUNK_RECORDS = 4096;
constexpr int UNIT_CAP = 30000, UNIT_CONT = 200;  // reader.go:236,266

struct RawChunk { std::string text; int n_records = 0; bool last = false; long first_line = 0; };

struct Parsed {  // one chunk pair
    int n = 0;
    std::vector<uint8_t> seq; std::vector<char> qual; std::vector<int32_t> len;  // 2 per pair
    std::vector<int64_t> soff;  // first base of each pair in seq / qual
    std::string names, barcodes, rgs; std::vector<int32_t> name_end, bc_end, rg_end;
    std::vector<uint8_t> valid, same_prev; std::vector<int64_t> seed;  // same_prev[r]: record r has the barcode of record r - 1 of this chunk
    bool last = false; int rc = 0; std::string err;
};

// Bounded in-order hand-off between pipeline steps.
template <class T> struct Channel {
    std::mutex mu; std::condition_variable cv; std::deque<T> q; size_t cap; bool closed = false;
    explicit Channel(size_t c) : cap(c) {}
    bool push(T v) { std::unique_lock<std::mutex> l(mu); cv.wait(l, [&] { return q.size() < cap || closed; }); if (closed) return false; q.push_back(std::move(v)); cv.notify_all(); return true; }
    bool pop(T& v) { std::unique_lock<std::mutex> l(mu); cv.wait(l, [&] { return !q.empty() || closed; }); if (q.empty()) return false; v = std::move(q.front()); q.pop_front(); cv.notify_all(); return true; }
    void close() { std::lock_guard<std::mutex> l(mu); closed = true; cv.notify_all(); }
};

const uint8_t* nt4_table() {  // nst_nt4_table, bwa/bntseq.c:47-64 (SequenceConvert, gobwa.go:159-167)
    static uint8_t t[256]; static bool init = false;
    if (!init) { memset(t, 4, 256); t[(int)'-'] = 5; const char* s = "ACGT"; for (int i = 0; i < 4; ++i) { t[(int)s[i]] = (uint8_t)i; t[(int)s[i] + 32] = (uint8_t)i; } init = true; }
    return t;
}

inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }

// ParseHeader on the header line without its '@' (the line includes its newline, as the reference's ReadString leaves it).
// Returns false when there is no BX:Z: tag followed by whitespace: the reference then leaves name and barcode empty.
bool parse_header(const char* s, size_t n, std::string* name, std::string* bc, bool* valid, std::string* rg) {
    // read group: the last whitespace-separated field when there are at least two (reader.go:143-152)
    rg->clear();
    {
        size_t e = n; while (e && is_space(s[e - 1])) --e;
        size_t b = e; while (b && !is_space(s[b - 1])) --b;
        size_t f = 0; while (f < n && is_space(s[f])) ++f;
        if (b > f) rg->assign(s + b, e - b);  // a field before it exists
    }
    name->clear(); bc->clear(); *valid = false;
    size_t f0 = 0; while (f0 < n && is_space(s[f0])) ++f0;
    size_t f1 = f0; while (f1 < n && !is_space(s[f1])) ++f1;
    const char* p = nullptr;
    for (size_t i = 0; i + 5 <= n; ++i) if (!memcmp(s + i, "BX:Z:", 5)) {  // leftmost match of BX:Z:(\S+)\s
        size_t e = i + 5; while (e < n && !is_space(s[e])) ++e;
        if (e > i + 5 && e < n) { p = s + i + 5; bc->assign(p, e - (i + 5)); break; }
    }
    if (!p) return false;
    if (f1 - f0 >= 2) name->assign(s + f0, f1 - f0 - 2);
    for (size_t i = 0; i + 7 <= n; ++i) if (!memcmp(s + i, "VX:i:", 5) && (s[i + 5] == '0' || s[i + 5] == '1') && is_space(s[i + 6])) { *valid = s[i + 5] == '1'; break; }
    return true;
}

}  // namespace

extern "C" int64_t arachne_name_seed(const char* name, int32_t len);

static void delete_batch(void* p);
struct arachne_ingest {
    std::string path[2];
    int n_threads = 4;
    std::unique_ptr<Channel<RawChunk>> raw[2];
    std::vector<std::thread> threads;
    // parsed chunks, re-ordered by index
    std::mutex take_mu;  // chunk i of both files is popped by one thread at a time
    std::mutex pmu; std::condition_variable pcv; std::map<long, Parsed> parsed; long next_parse = 0, next_take = 0; bool parse_done = false;
    int split_rc[2] = {0, 0}; std::string split_err[2];
    // assembler state (ReadBarcodeSet's)
    Parsed cur; int cur_pos = 0; bool have_cur = false, eof = false;
    std::string last_barcode; bool have_last = false;
    long n_pairs_total = 0;
    std::mutex pool_mu; std::vector<void*> pool;  // released batch buffers: their (already touched) memory is reused
    ~arachne_ingest() { for (void* p : pool) delete_batch(p); for (auto& c : raw) if (c) c->close(); { std::lock_guard<std::mutex> l(pmu); parse_done = true; next_take = 1L << 60; pcv.notify_all(); } for (auto& t : threads) if (t.joinable()) t.join(); }
};

namespace {

void split_thread(arachne_ingest* I, int which) {
    gzFile f = gzopen(I->path[which].c_str(), "rb");  // transparently reads plain files too
    if (!f) { I->split_rc[which] = ARACHNE_ERR_IO; I->split_err[which] = "cannot open " + I->path[which]; I->raw[which]->close(); return; }
    gzbuffer(f, 1 << 20);
    std::string carry; std::vector<char> buf(8 << 20);
    RawChunk ch; long line = 0; ch.first_line = 0;
    int lines_in_chunk = 0;
    for (;;) {
        const int got = gzread(f, buf.data(), (unsigned)buf.size());
        if (got < 0) { I->split_rc[which] = ARACHNE_ERR_IO; I->split_err[which] = "read error in " + I->path[which]; break; }
        if (got == 0) break;
        const char* p = buf.data(); const char* end = p + got;
        const char* seg = p;  // start of the bytes not yet appended to the chunk: one append per chunk boundary / buffer, not per line
        while (p < end) {
            const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)(end - p)));
            if (!nl) break;
            p = nl + 1; ++line;
            if (++lines_in_chunk == 4 * CHUNK_RECORDS) {
                ch.text.append(seg, (size_t)(p - seg)); seg = p;
                ch.n_records = CHUNK_RECORDS;
                if (!I->raw[which]->push(std::move(ch))) { gzclose(f); return; }
                ch = RawChunk(); ch.first_line = line; lines_in_chunk = 0;
            }
        }
        ch.text.append(seg, (size_t)(end - seg));
    }
    gzclose(f);
    if (!ch.text.empty() && ch.text.back() != '\n') { ch.text.push_back('\n'); ++lines_in_chunk; }
    ch.n_records = lines_in_chunk / 4;  // an incomplete trailing record is dropped, as the reference's EOF mid-record is (reader.go:170-181)
    ch.last = true;
    I->raw[which]->push(std::move(ch));
    I->raw[which]->close();
}

void parse_pair(const RawChunk& a, const RawChunk& b, Parsed* P) {
    const uint8_t* nt4 = nt4_table();
    const int n = a.n_records < b.n_records ? a.n_records : b.n_records;
    P->n = n; P->last = a.last || b.last;
    P->len.reserve(2 * n); P->name_end.reserve(n); P->bc_end.reserve(n); P->rg_end.reserve(n); P->valid.reserve(n); P->seed.reserve(n);
    const char* pa = a.text.data(); const char* pb = b.text.data();
    std::string name, bc, rg; bool valid;
    auto next_line = [](const char*& p, size_t* len) { const char* s = p; const char* nl = static_cast<const char*>(memchr(p, '\n', (size_t)1 << 40)); *len = (size_t)(nl - s); p = nl + 1; return s; };
    for (int r = 0; r < n; ++r) {
        size_t l;
        const char* h1 = next_line(pa, &l); const size_t hl = l;
        const char* s1 = next_line(pa, &l); const size_t sl1 = l; next_line(pa, &l); const char* q1 = next_line(pa, &l); const size_t ql1 = l;
        next_line(pb, &l);
        const char* s2 = next_line(pb, &l); const size_t sl2 = l; next_line(pb, &l); const char* q2 = next_line(pb, &l); const size_t ql2 = l;
        if (hl == 0 || h1[0] != '@') { P->rc = ARACHNE_ERR_IO; char m[128]; snprintf(m, sizeof(m), "malformed FASTQ: record header expected at line %ld of R1", a.first_line + 4L * r + 1); P->err = m; return; }
        parse_header(h1 + 1, hl, &name, &bc, &valid, &rg);  // hl - 1 header bytes + the newline
        P->names += name; P->name_end.push_back((int32_t)P->names.size());
        {
            const int pb0 = r >= 2 ? P->bc_end[r - 2] : 0, pb1 = r >= 1 ? P->bc_end[r - 1] : 0;
            P->same_prev.push_back(r >= 1 && (size_t)(pb1 - pb0) == bc.size() && !memcmp(P->barcodes.data() + pb0, bc.data(), bc.size()));
        }
        P->barcodes += bc; P->bc_end.push_back((int32_t)P->barcodes.size());
        P->rgs += rg; P->rg_end.push_back((int32_t)P->rgs.size());
        P->valid.push_back(valid);
        P->seed.push_back(arachne_name_seed(name.data(), (int32_t)name.size()));
        const size_t o = P->seq.size();
        P->soff.push_back((int64_t)o);
        P->seq.resize(o + sl1 + sl2); P->qual.resize(o + sl1 + sl2);
        for (size_t k = 0; k < sl1; ++k) P->seq[o + k] = nt4[(uint8_t)s1[k]];
        for (size_t k = 0; k < sl2; ++k) P->seq[o + sl1 + k] = nt4[(uint8_t)s2[k]];
        // qualities are carried as read; a quality line shorter than its sequence is padded with '!' (the reference would index past it)
        for (size_t k = 0; k < sl1; ++k) P->qual[o + k] = k < ql1 ? q1[k] : '!';
        for (size_t k = 0; k < sl2; ++k) P->qual[o + sl1 + k] = k < ql2 ? q2[k] : '!';
        P->len.push_back((int32_t)sl1); P->len.push_back((int32_t)sl2);
    }
}

void parse_thread(arachne_ingest* I) {
    for (;;) {
        RawChunk a, b; long idx;
        {   // chunk i of both files must be taken together and in order
            std::lock_guard<std::mutex> t(I->take_mu);
            if (I->parse_done) return;
            const bool ga = I->raw[0]->pop(a), gb = I->raw[1]->pop(b);
            if (!ga || !gb) {
                I->parse_done = true;
                Parsed e; e.last = true; e.rc = I->split_rc[0] ? I->split_rc[0] : I->split_rc[1]; e.err = I->split_rc[0] ? I->split_err[0] : I->split_err[1];
                std::lock_guard<std::mutex> l(I->pmu);
                I->parsed[I->next_parse++] = std::move(e); I->pcv.notify_all(); return;
            }
            idx = I->next_parse++;
            if (a.last || b.last) { I->parse_done = true; I->raw[0]->close(); I->raw[1]->close(); }
        }
        Parsed P; parse_pair(a, b, &P);
        std::unique_lock<std::mutex> l(I->pmu);
        I->pcv.wait(l, [&] { return idx - I->next_take < 4L * I->n_threads + 8; });  // bound the re-order window
        I->parsed[idx] = std::move(P);
        I->pcv.notify_all();
    }
}

// the next record of the stream, or false at the end
bool next_record(arachne_ingest* I, int* rc) {
    while (!I->have_cur || I->cur_pos >= I->cur.n) {
        if (I->have_cur && I->cur.last) { I->eof = true; return false; }
        std::unique_lock<std::mutex> l(I->pmu);
        I->pcv.wait(l, [&] { return I->parsed.count(I->next_take) > 0; });
        I->cur = std::move(I->parsed[I->next_take]); I->parsed.erase(I->next_take); ++I->next_take;
        I->pcv.notify_all();
        l.unlock();
        I->cur_pos = 0; I->have_cur = true;
        if (I->cur.rc) { *rc = I->cur.rc; set_error("%s", I->cur.err.c_str()); I->eof = true; return false; }
    }
    return true;
}

struct BatchBuf {
    std::vector<uint8_t> seq; std::vector<char> qual; std::vector<int64_t> off{0}; std::vector<int64_t> unit_off{0};
    std::vector<uint8_t> unique, run_rfa, valid; std::vector<int64_t> seed;
    std::string names; std::vector<int64_t> name_off{0}; std::string barcodes; std::vector<int64_t> bc_off{0}; std::string rgs; std::vector<int64_t> rg_off{0};
    arachne_ingest_batch_t pub;
    arachne_ingest* owner = nullptr;
    void reset() {
        seq.clear(); qual.clear(); off.assign(1, 0); unit_off.assign(1, 0); unique.clear(); run_rfa.clear(); valid.clear(); seed.clear();
        names.clear(); name_off.assign(1, 0); barcodes.clear(); bc_off.assign(1, 0); rgs.clear(); rg_off.assign(1, 0);
    }
};

}  // namespace

static void delete_batch(void* p) { delete static_cast<BatchBuf*>(p); }

extern "C" {

int arachne_ingest_open(const char* r1, const char* r2, int n_threads, arachne_ingest_t** out) {
    if (!r1 || !r2 || !out) { set_error("arachne_ingest_open: null argument"); return ARACHNE_ERR_ARG; }
    for (const char* p : {r1, r2}) { FILE* f = fopen(p, "rb"); if (!f) { set_error("cannot open %s", p); return ARACHNE_ERR_IO; } fclose(f); }
    auto* I = new arachne_ingest();
    I->path[0] = r1; I->path[1] = r2; I->n_threads = n_threads > 0 ? n_threads : 4;
    for (int k = 0; k < 2; ++k) I->raw[k].reset(new Channel<RawChunk>((size_t)I->n_threads * 2 + 4));
    for (int k = 0; k < 2; ++k) I->threads.emplace_back(split_thread, I, k);
    for (int k = 0; k < I->n_threads; ++k) I->threads.emplace_back(parse_thread, I);
    *out = I;
    return ARACHNE_OK;
}

void arachne_ingest_close(arachne_ingest_t* I) { delete I; }

int arachne_ingest_next(arachne_ingest_t* I, int64_t target_pairs, arachne_ingest_batch_t** out) {
    if (!I || !out) { set_error("arachne_ingest_next: null argument"); return ARACHNE_ERR_ARG; }
    *out = nullptr;
    if (I->eof && !(I->have_cur && I->cur_pos < I->cur.n)) return ARACHNE_INGEST_EOF;
    BatchBuf* B = nullptr;
    {
        std::lock_guard<std::mutex> l(I->pool_mu);
        if (!I->pool.empty()) { B = static_cast<BatchBuf*>(I->pool.back()); I->pool.pop_back(); }
    }
    if (!B) B = new BatchBuf(); else B->reset();
    B->owner = I;
    {   // growing a 100 MB vector by doubling copies it twice over and page-faults all of it: size from the request instead
        const size_t want = (size_t)std::min<int64_t>(target_pairs, 4000000) + 65536;
        B->seq.reserve(want * 320); B->qual.reserve(want * 320); B->off.reserve(2 * want + 2); B->seed.reserve(want); B->name_off.reserve(want + 1); B->names.reserve(want * 24);
    }
    int rc = 0;
    int64_t n_pairs = 0;
    // ReadBarcodeSet, one call per loop trip (reader.go:208-291); a record that opens the next unit is simply not consumed yet
    for (;;) {
        int index = 0; bool ended_by_change = false, at_eof = false;
        std::string first_bc;
        const int64_t unit_start = n_pairs;
        while (index < UNIT_CAP) {
            if (!next_record(I, &rc)) { at_eof = true; break; }
            const Parsed& P = I->cur; const int r = I->cur_pos;
            const int b0 = r ? P.bc_end[r - 1] : 0;
            const size_t bl = (size_t)(P.bc_end[r] - b0);
            if (index == 0) first_bc.assign(P.barcodes.data() + b0, bl);
            else if (bl != first_bc.size() || memcmp(P.barcodes.data() + b0, first_bc.data(), bl)) { ended_by_change = true; break; }
            // how many records of this chunk can be taken in one go: the run of this barcode (same_prev marks a record whose barcode
            // equals its predecessor's), cut by the cap and by the continuation rule (a record at index >= 200 ends the unit)
            const bool cont = I->have_last && first_bc == I->last_barcode;
            int lim = UNIT_CAP - index;
            if (cont) { const int to_break = index >= UNIT_CONT ? 1 : UNIT_CONT - index + 1; lim = lim < to_break ? lim : to_break; }
            int k = 1;
            while (k < lim && r + k < P.n && P.same_prev[r + k]) ++k;
            // append records [r, r + k)
            const size_t so = (size_t)P.soff[r], se = r + k < P.n ? (size_t)P.soff[r + k] : P.seq.size();
            B->seq.insert(B->seq.end(), P.seq.begin() + so, P.seq.begin() + se);
            B->qual.insert(B->qual.end(), P.qual.begin() + so, P.qual.begin() + se);
            for (int t = 2 * r; t < 2 * (r + k); ++t) B->off.push_back(B->off.back() + P.len[t]);
            const int n0 = r ? P.name_end[r - 1] : 0; const int64_t nb0 = (int64_t)B->names.size() - n0;
            B->names.append(P.names.data() + n0, (size_t)(P.name_end[r + k - 1] - n0));
            for (int t = r; t < r + k; ++t) B->name_off.push_back(nb0 + P.name_end[t]);
            B->seed.insert(B->seed.end(), P.seed.begin() + r, P.seed.begin() + r + k);
            {
                const int g0 = r ? P.rg_end[r - 1] : 0; const int64_t gb0 = (int64_t)B->rgs.size() - g0;
                B->rgs.append(P.rgs.data() + g0, (size_t)(P.rg_end[r + k - 1] - g0));
                for (int t = r; t < r + k; ++t) B->rg_off.push_back(gb0 + P.rg_end[t]);
            }
            if (index == 0) {
                B->barcodes += first_bc; B->bc_off.push_back((int64_t)B->barcodes.size());
                B->valid.push_back(P.valid[r]);
            }
            I->cur_pos += k; n_pairs += k; index += k;
            if (cont && index - 1 >= UNIT_CONT && index - 1 > 0) break;  // "abnormal break" (reader.go:266-270): the record at index >= 200 stays in the unit
        }
        if (rc) { delete B; return rc; }
        const int64_t n_unit = n_pairs - unit_start;
        if (n_unit == 0) break;  // end of input
        I->last_barcode = first_bc; I->have_last = true;
        const bool full = ended_by_change || at_eof;  // reader.go:283-289
        B->unit_off.push_back(n_pairs);
        B->unique.push_back(full);
        B->run_rfa.push_back(full && first_bc.find('-') != std::string::npos && n_unit >= 5);  // worthRunningRFA
        if (at_eof || n_pairs >= target_pairs) break;
    }
    if (n_pairs == 0) { delete B; return ARACHNE_INGEST_EOF; }
    I->n_pairs_total += n_pairs;
    arachne_ingest_batch_t& o = B->pub;
    o.n_pairs = n_pairs; o.n_units = (int64_t)B->unique.size();
    o.seq = B->seq.data(); o.seq_off = B->off.data(); o.qual = B->qual.data(); o.unit_off = B->unit_off.data();
    o.unique = B->unique.data(); o.run_rfa = B->run_rfa.data(); o.valid = B->valid.data(); o.rand_seed = B->seed.data();
    o.names = B->names.data(); o.name_off = B->name_off.data(); o.barcodes = B->barcodes.data(); o.barcode_char_off = B->bc_off.data();
    o.read_groups = B->rgs.data(); o.read_group_off = B->rg_off.data(); o.priv = B;
    *out = &o;
    return ARACHNE_OK;
}

void arachne_ingest_batch_free(arachne_ingest_batch_t* b) {
    if (!b) return;
    auto* B = static_cast<BatchBuf*>(b->priv);
    arachne_ingest* I = B->owner;
    if (I) {
        std::lock_guard<std::mutex> l(I->pool_mu);
        if (I->pool.size() < 3) { I->pool.push_back(B); return; }
    }
    delete B;
}

}  // extern "C"
