// Barcode-grouped FASTQ ingest: what src/fastqreader does in front of the aligner, as a multi-threaded native
// reader that hands out batches in exactly the layout arachne_align_batch() takes.
//   ReadOneLine      src/fastqreader/reader.go:129-198   one R1 + one R2 record per call, header -> ParseHeader
//   ParseHeader      :95-123   name = first field minus its 2-character mate suffix; barcode = BX:Z:(\S+)\s; valid = VX:i:([01])\s
//   ReadBarcodeSet   :208-291  consecutive records of one barcode, capped at 30000, 201-record continuation units
//   worthRunningRFA  src/aligner/aligner.go:1018-1030
//   FastZipReader    src/fastqreader/zipread.go:69-76 (`gunzip -c` child process; here zlib in a thread per file)
// Pipeline: one source/split thread per file cuts its stream into chunks of a fixed number of records (so chunk i of R1
// and chunk i of R2 hold the same pairs); `n_threads` parser threads turn chunk pairs into codes 0..4, qualities, names,
// barcodes and md5 name seeds; the caller's thread (arachne_ingest_next) walks the parsed chunks in order, applies the
// work-unit rule record run by record run and copies each chunk's taken records in ONE piece per array.
// Sources: plain text, gzip (one zlib stream per file: ~250 MB/s, the format's limit) and BGZF / bgzip (gzip members of <= 64 KB
// that carry their own size: a file's blocks are inflated by a team of threads, so compressed input scales with the cores).
// Host code only.
#include <zlib.h>
#include <time.h>
#if defined(__x86_64__)
#include <emmintrin.h>
#include <tmmintrin.h>
#endif
#include <algorithm>
#include <atomic>
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
#include "md5_seed.h"

namespace arachne { void set_error(const char* fmt, ...); }
using arachne::set_error;

namespace {

constexpr int CHUNK_RECORDS = 4096;
constexpr int UNIT_CAP = 30000, UNIT_CONT = 200;  // reader.go:236,266

double thread_cpu() { timespec t; clock_gettime(CLOCK_THREAD_CPUTIME_ID, &t); return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec; }

// Growable array without value-initialisation: the big per-batch / per-chunk buffers are written exactly once.
template <class T> struct Raw {
    T* p = nullptr; size_t n = 0, cap = 0;
    Raw() = default;
    Raw(const Raw&) = delete; Raw& operator=(const Raw&) = delete;
    Raw(Raw&& o) noexcept : p(o.p), n(o.n), cap(o.cap) { o.p = nullptr; o.n = o.cap = 0; }
    Raw& operator=(Raw&& o) noexcept { if (this != &o) { free(p); p = o.p; n = o.n; cap = o.cap; o.p = nullptr; o.n = o.cap = 0; } return *this; }
    ~Raw() { free(p); }
    void reserve(size_t want) {
        if (want <= cap) return;
        size_t c = cap + cap / 2; if (c < want) c = want;
        T* q = static_cast<T*>(realloc(p, c * sizeof(T))); if (!q) throw std::bad_alloc();
        p = q; cap = c;
    }
    T* grow(size_t k) { reserve(n + k); T* at = p + n; n += k; return at; }  // room for k more, returns where they start
    void push(const T& v) { if (n == cap) reserve(n + 1); p[n++] = v; }
    void clear() { n = 0; }
    T& back() { return p[n - 1]; }
};

// A chunk of one file: its text and where every line ends (offset of the '\n'), found once by the splitter.
struct RawChunk { Raw<char> text; Raw<uint32_t> nl; int n_records = 0; bool last = false; long first_line = 0; };

struct Parsed {  // one chunk pair
    int n = 0;
    Raw<uint8_t> seq; Raw<char> qual;          // R1, R2 of record 0, R1, R2 of record 1, ...
    std::vector<int32_t> len;                   // 2 per record
    std::vector<int64_t> soff;                  // n + 1: first base of each record in seq / qual
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
    static const struct T { uint8_t t[256]; T() { memset(t, 4, 256); t[(int)'-'] = 5; const char* s = "ACGT"; for (int i = 0; i < 4; ++i) { t[(int)s[i]] = (uint8_t)i; t[(int)s[i] + 32] = (uint8_t)i; } } } tab;
    return tab.t;
}

// text -> codes for n bases.  With SSSE3 (checked at run time) 16 bases per step: the low nibble of A/C/G/T/a/c/g/t picks the code and the
// letter it must be from two 16-entry tables, any other byte is 4, '-' is 5; the scalar loop is the definition.
void bases_to_codes_scalar(const char* s, size_t n, uint8_t* out) { const uint8_t* t = nt4_table(); for (size_t k = 0; k < n; ++k) out[k] = t[(uint8_t)s[k]]; }
#if defined(__x86_64__)
__attribute__((target("ssse3"))) void bases_to_codes_ssse3(const char* s, size_t n, uint8_t* out) {
    const __m128i code_of = _mm_setr_epi8(4, 0, 4, 1, 3, 4, 4, 2, 4, 4, 4, 4, 4, 4, 4, 4);                      // nibble 1 A, 3 C, 4 T, 7 G
    const __m128i letter_of = _mm_setr_epi8(-1, 'A', -1, 'C', 'T', -1, -1, 'G', -1, -1, -1, -1, -1, -1, -1, -1);
    const __m128i low4 = _mm_set1_epi8(0x0f), upper = _mm_set1_epi8((char)0xdf), four = _mm_set1_epi8(4), five = _mm_set1_epi8(5), dash = _mm_set1_epi8('-');
    size_t k = 0;
    for (; k + 16 <= n; k += 16) {
        const __m128i c = _mm_loadu_si128(reinterpret_cast<const __m128i*>(s + k));
        const __m128i nib = _mm_and_si128(c, low4);
        const __m128i is_base = _mm_cmpeq_epi8(_mm_and_si128(c, upper), _mm_shuffle_epi8(letter_of, nib));
        __m128i code = _mm_or_si128(_mm_and_si128(is_base, _mm_shuffle_epi8(code_of, nib)), _mm_andnot_si128(is_base, four));
        const __m128i is_dash = _mm_cmpeq_epi8(c, dash);
        code = _mm_or_si128(_mm_andnot_si128(is_dash, code), _mm_and_si128(is_dash, five));
        _mm_storeu_si128(reinterpret_cast<__m128i*>(out + k), code);
    }
    bases_to_codes_scalar(s + k, n - k, out + k);
}
using bases_fn = void (*)(const char*, size_t, uint8_t*);
bases_fn pick_bases_to_codes() { return __builtin_cpu_supports("ssse3") ? bases_to_codes_ssse3 : bases_to_codes_scalar; }
#else
using bases_fn = void (*)(const char*, size_t, uint8_t*);
bases_fn pick_bases_to_codes() { return bases_to_codes_scalar; }
#endif

// Two notions of white space meet in ParseHeader: strings.Fields splits at unicode.IsSpace (of the ASCII range: space, \t \n \v \f \r), the
// regular expressions' \s and \S are RE2's (space, \t \n \f \r: a vertical tab is NOT white space there).
inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }   // strings.Fields
inline bool re_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f'; }                // regexp \s

// ParseHeader on the header line without its '@' (the n bytes include the newline, as the reference's ReadString leaves it), results
// appended to the chunk's string pools.  Without a `BX:Z:(\S+)\s` match the reference leaves name and barcode empty and valid false.
struct HeaderOut { std::string *names, *barcodes, *rgs; };
inline bool parse_header(const char* s, size_t n, const HeaderOut& o, bool* valid) {
    // read group: the last whitespace-separated field when there are at least two (reader.go:143-152)
    size_t f0 = 0; while (f0 < n && is_space(s[f0])) ++f0;
    {
        size_t e = n; while (e && is_space(s[e - 1])) --e;
        size_t b = e; while (b && !is_space(s[b - 1])) --b;
        if (b > f0) o.rgs->append(s + b, e - b);  // a field before it exists
    }
    *valid = false;
    const char* bx = nullptr; size_t bx_len = 0;
    for (const char* at = s; at + 5 <= s + n;) {  // leftmost match of BX:Z:(\S+)\s
        const char* hit = static_cast<const char*>(memmem(at, (size_t)(s + n - at), "BX:Z:", 5));
        if (!hit) break;
        const char* v = hit + 5; const char* e = v; while (e < s + n && !re_space(*e)) ++e;
        if (e > v && e < s + n) { bx = v; bx_len = (size_t)(e - v); break; }
        at = hit + 1;
    }
    if (!bx) return false;
    o.barcodes->append(bx, bx_len);
    size_t f1 = f0; while (f1 < n && !is_space(s[f1])) ++f1;
    if (f1 - f0 >= 2) o.names->append(s + f0, f1 - f0 - 2);
    for (const char* at = s; at + 7 <= s + n;) {  // leftmost match of VX:i:([01])\s
        const char* hit = static_cast<const char*>(memmem(at, (size_t)(s + n - at), "VX:i:", 5));
        if (!hit || hit + 7 > s + n) break;
        if ((hit[5] == '0' || hit[5] == '1') && re_space(hit[6])) { *valid = hit[5] == '1'; break; }
        at = hit + 1;
    }
    return true;
}

// ---- byte sources ----------------------------------------------------------------------------------------------------------
struct Source {
    virtual ~Source() {}
    virtual long read(char* dst, size_t cap, std::string* err) = 0;  // bytes produced (0 = end), < 0 = error
};

struct PlainSource : Source {  // uncompressed text straight from the file
    FILE* f = nullptr; std::string path;
    ~PlainSource() override { if (f) fclose(f); }
    long read(char* dst, size_t cap, std::string* err) override {
        const size_t got = fread(dst, 1, cap, f);
        if (got == 0 && ferror(f)) { *err = "read error in " + path; return -1; }
        return (long)got;
    }
};

struct GzSource : Source {  // gzip (one or several members) through zlib's gz layer
    gzFile f = nullptr; std::string path;
    ~GzSource() override { if (f) gzclose(f); }
    long read(char* dst, size_t cap, std::string* err) override {
        const int got = gzread(f, dst, (unsigned)std::min<size_t>(cap, 1u << 30));
        if (got < 0) { *err = "read error in " + path; return -1; }
        return got;
    }
};

// BGZF (SAM spec 4.1; what bgzip writes): every gzip member has an extra field `BC` with its own compressed size and ends in
// CRC32 + ISIZE, so a group of members is read in one piece, their output offsets follow from the ISIZEs, and a team of threads
// inflates them side by side straight into the caller's buffer.
struct BgzfSource : Source {
    FILE* f = nullptr; std::string path; int team = 4;
    std::vector<uint8_t> comp; size_t comp_lo = 0, comp_hi = 0;  // unconsumed compressed bytes
    std::vector<char> spill; size_t spill_lo = 0;                 // output of a block that did not fit the caller's buffer
    bool at_eof = false;
    ~BgzfSource() override { if (f) fclose(f); }
    static bool is_bgzf(const uint8_t* h, size_t n) {
        return n >= 18 && h[0] == 0x1f && h[1] == 0x8b && h[2] == 8 && (h[3] & 4) && h[10] == 6 && h[11] == 0 && h[12] == 'B' && h[13] == 'C' && h[14] == 2 && h[15] == 0;
    }
    struct Block { const uint8_t* src; uint32_t csize, isize; char* dst; };
    static int inflate_block(const Block& b) {
        z_stream z; memset(&z, 0, sizeof(z));
        if (inflateInit2(&z, -15) != Z_OK) return -1;
        z.next_in = const_cast<Bytef*>(b.src + 18); z.avail_in = b.csize - 18 - 8;
        z.next_out = reinterpret_cast<Bytef*>(b.dst); z.avail_out = b.isize;
        const int rc = inflate(&z, Z_FINISH);
        const bool ok = rc == Z_STREAM_END && z.total_out == b.isize;
        inflateEnd(&z);
        if (!ok) return -1;
        uint32_t crc; memcpy(&crc, b.src + b.csize - 8, 4);
        return (uint32_t)crc32(crc32(0L, Z_NULL, 0), reinterpret_cast<const Bytef*>(b.dst), b.isize) == crc ? 0 : -1;
    }
    long read(char* dst, size_t cap, std::string* err) override {
        for (;;) {
            if (spill_lo < spill.size()) {
                const size_t k = std::min(cap, spill.size() - spill_lo);
                memcpy(dst, spill.data() + spill_lo, k); spill_lo += k;
                return (long)k;
            }
            // top up the compressed window
            if (!at_eof && comp_hi - comp_lo < (4u << 20)) {
                if (comp_lo) { memmove(comp.data(), comp.data() + comp_lo, comp_hi - comp_lo); comp_hi -= comp_lo; comp_lo = 0; }
                if (comp.size() < (16u << 20)) comp.resize(16u << 20);
                const size_t got = fread(comp.data() + comp_hi, 1, comp.size() - comp_hi, f);
                comp_hi += got;
                if (got == 0) { if (ferror(f)) { *err = "read error in " + path; return -1; } at_eof = true; }
            }
            if (comp_hi == comp_lo) return 0;
            // the run of complete blocks whose output fits
            std::vector<Block> blocks; size_t at = comp_lo, out = 0;
            bool room = true;
            while (comp_hi - at >= 18) {
                const uint8_t* h = comp.data() + at;
                if (!is_bgzf(h, comp_hi - at)) { *err = "not a BGZF block in " + path; return -1; }
                const uint32_t csize = (uint32_t)(h[16] | h[17] << 8) + 1;
                if (csize < 26) { *err = "corrupt BGZF block in " + path; return -1; }
                if (comp_hi - at < csize) break;
                uint32_t isize; memcpy(&isize, h + csize - 4, 4);
                if (isize > 65536) { *err = "corrupt BGZF block in " + path; return -1; }
                if (out + isize > cap) { room = false; break; }
                blocks.push_back(Block{h, csize, isize, dst + out});
                out += isize; at += csize;
            }
            if (blocks.empty()) {
                if (!room) {  // one block larger than the caller's room: through the spill buffer
                    const uint8_t* h = comp.data() + at; const uint32_t csize = (uint32_t)(h[16] | h[17] << 8) + 1;
                    uint32_t isize; memcpy(&isize, h + csize - 4, 4);
                    spill.resize(isize); spill_lo = 0;
                    if (inflate_block(Block{h, csize, isize, spill.data()})) { *err = "corrupt BGZF block in " + path; return -1; }
                    comp_lo = at + csize;
                    continue;
                }
                if (at_eof) { *err = "truncated BGZF file " + path; return -1; }
                continue;  // need more compressed bytes
            }
            std::atomic<size_t> next{0}; std::atomic<int> bad{0};
            auto work = [&] { for (size_t i; (i = next.fetch_add(1)) < blocks.size();) if (inflate_block(blocks[i])) bad = 1; };
            const int n_thr = (int)std::min<size_t>((size_t)team, (blocks.size() + 3) / 4);
            std::vector<std::thread> helpers;
            for (int t = 1; t < n_thr; ++t) helpers.emplace_back(work);
            work();
            for (auto& t : helpers) t.join();
            if (bad) { *err = "corrupt BGZF block in " + path; return -1; }
            comp_lo = at;
            if (out == 0) continue;  // only empty blocks (the end-of-file marker)
            return (long)out;
        }
    }
};

std::unique_ptr<Source> open_source(const std::string& path, int team, std::string* err) {
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) { *err = "cannot open " + path; return nullptr; }
    uint8_t head[18]; const size_t got = fread(head, 1, sizeof(head), f);
    if (BgzfSource::is_bgzf(head, got)) {
        rewind(f);
        auto s = std::make_unique<BgzfSource>(); s->f = f; s->path = path; s->team = team > 1 ? team : 1;
        return s;
    }
    if (!(got >= 2 && head[0] == 0x1f && head[1] == 0x8b)) {
        rewind(f); setvbuf(f, nullptr, _IONBF, 0);  // large reads go straight to the caller's buffer
        auto s = std::make_unique<PlainSource>(); s->f = f; s->path = path;
        return s;
    }
    fclose(f);
    auto s = std::make_unique<GzSource>(); s->path = path;
    s->f = gzopen(path.c_str(), "rb");
    if (!s->f) { *err = "cannot open " + path; return nullptr; }
    gzbuffer(s->f, 1 << 20);
    return s;
}

}  // namespace

static void delete_batch(void* p);
struct arachne_ingest {
    std::string path[2];
    int n_threads = 4;
    std::unique_ptr<Channel<RawChunk>> raw[2];
    std::vector<std::thread> threads;
    // parsed chunks, re-ordered by index
    std::mutex take_mu;  // chunk i of both files is popped by one thread at a time
    std::mutex pmu; std::condition_variable pcv; std::map<long, Parsed> parsed; long next_parse = 0, next_take = 0; std::atomic<bool> parse_done{false};  // set under take_mu by the parsers, under pmu at close
    int split_rc[2] = {0, 0}; std::string split_err[2];
    // assembler state (ReadBarcodeSet's)
    Parsed cur; int cur_pos = 0; bool have_cur = false, eof = false;
    std::string last_barcode; bool have_last = false;
    long n_pairs_total = 0;
    // thread CPU seconds per role: ARACHNE_INGEST_STATS=1 prints them at close (which step of the pipeline is the serial one)
    std::mutex stat_mu; double cpu_split[2] = {0, 0}, cpu_parse = 0, cpu_asm = 0;
    std::mutex pool_mu; std::vector<void*> pool;  // released batch buffers: their (already touched) memory is reused
    // chunk and parse buffers go round in circles too: a fresh 2 MB allocation is an mmap and ~500 page faults every time
    std::mutex recycle_mu; std::vector<RawChunk> free_chunks; std::vector<Parsed> free_parsed;
    RawChunk take_chunk() { std::lock_guard<std::mutex> l(recycle_mu); RawChunk c; if (!free_chunks.empty()) { c = std::move(free_chunks.back()); free_chunks.pop_back(); } c.text.clear(); c.nl.clear(); c.n_records = 0; c.last = false; c.first_line = 0; return c; }
    void give_chunk(RawChunk&& c) { std::lock_guard<std::mutex> l(recycle_mu); if (free_chunks.size() < 64) free_chunks.push_back(std::move(c)); }
    Parsed take_parsed() { std::lock_guard<std::mutex> l(recycle_mu); Parsed p; if (!free_parsed.empty()) { p = std::move(free_parsed.back()); free_parsed.pop_back(); } return p; }
    void give_parsed(Parsed&& p) { std::lock_guard<std::mutex> l(recycle_mu); if (free_parsed.size() < 64) free_parsed.push_back(std::move(p)); }
    ~arachne_ingest() {
        for (void* p : pool) delete_batch(p);
        for (auto& c : raw) if (c) c->close();
        { std::lock_guard<std::mutex> l(pmu); parse_done = true; next_take = 1L << 60; pcv.notify_all(); }
        for (auto& t : threads) if (t.joinable()) t.join();
        if (const char* e = getenv("ARACHNE_INGEST_STATS")) if (*e && *e != '0')
            fprintf(stderr, "[arachne ingest] %ld pairs; thread CPU s: source+split R1 %.3f R2 %.3f, parse (all %d threads) %.3f, assemble %.3f\n",
                    n_pairs_total, cpu_split[0], cpu_split[1], n_threads, cpu_parse, cpu_asm);
    }
};

namespace {

// Offsets of the next newlines of text[from, to) appended to nl until it holds `want` entries; returns how far it scanned.
// 16 bytes per step (SSE2 is baseline on x86-64): lines of 1..150 bytes make a memchr call per line the splitter's main cost.
size_t scan_newlines(const char* text, size_t from, size_t to, Raw<uint32_t>* nl, size_t want) {
    size_t i = from;
#if defined(__SSE2__)
    const __m128i key = _mm_set1_epi8('\n');
    for (; i + 16 <= to && nl->n < want; i += 16) {
        unsigned m = (unsigned)_mm_movemask_epi8(_mm_cmpeq_epi8(_mm_loadu_si128(reinterpret_cast<const __m128i*>(text + i)), key));
        while (m) {
            const unsigned b = (unsigned)__builtin_ctz(m); m &= m - 1;
            nl->push((uint32_t)(i + b));
            if (nl->n == want) return i + b + 1;
        }
    }
#endif
    for (; i < to && nl->n < want; ++i) if (text[i] == '\n') { nl->push((uint32_t)i); if (nl->n == want) return i + 1; }
    return i;
}

void split_thread(arachne_ingest* I, int which) {
    std::string err;
    std::unique_ptr<Source> src = open_source(I->path[which], I->n_threads, &err);
    if (!src) { I->split_rc[which] = ARACHNE_ERR_IO; I->split_err[which] = err; I->raw[which]->close(); return; }
    // The source writes straight into the chunk being filled; what follows the chunk's last line moves to the next chunk.
    constexpr size_t READ = 1u << 20, LINES = 4 * (size_t)CHUNK_RECORDS;
    RawChunk ch = I->take_chunk(); long line = 0; size_t scanned = 0;
    for (;;) {
        ch.text.reserve(ch.text.n + READ); ch.nl.reserve(LINES);
        const long got = src->read(ch.text.p + ch.text.n, READ, &err);
        if (got < 0) {  // no "last" chunk: the parser that finds this channel closed early reports the error to the caller
            I->split_rc[which] = ARACHNE_ERR_IO; I->split_err[which] = err; I->raw[which]->close(); I->cpu_split[which] = thread_cpu(); return;
        }
        if (got == 0) break;
        ch.text.n += (size_t)got;
        for (;;) {
            scanned = scan_newlines(ch.text.p, scanned, ch.text.n, &ch.nl, LINES);
            if (ch.nl.n < LINES) break;
            RawChunk next = I->take_chunk();
            const size_t rest = ch.text.n - scanned;
            next.text.reserve(rest + READ); memcpy(next.text.p, ch.text.p + scanned, rest); next.text.n = rest;
            line += (long)LINES; next.first_line = line;
            ch.text.n = scanned; ch.n_records = CHUNK_RECORDS;
            if (!I->raw[which]->push(std::move(ch))) { I->cpu_split[which] = thread_cpu(); return; }
            ch = std::move(next); scanned = 0;
        }
    }
    src.reset();
    if (ch.text.n && ch.text.p[ch.text.n - 1] != '\n') { ch.text.push('\n'); ch.nl.push((uint32_t)(ch.text.n - 1)); }
    ch.n_records = (int)(ch.nl.n / 4);  // an incomplete trailing record is dropped, as the reference's EOF mid-record is (reader.go:170-181)
    ch.last = true;
    I->raw[which]->push(std::move(ch));
    I->raw[which]->close();
    I->cpu_split[which] = thread_cpu();
}

void parse_pair(const RawChunk& a, const RawChunk& b, Parsed* P) {
    static const bases_fn to_codes = pick_bases_to_codes();
    const int n = a.n_records < b.n_records ? a.n_records : b.n_records;
    P->n = n; P->last = a.last || b.last; P->rc = 0; P->err.clear();
    P->len.resize(2 * (size_t)n); P->soff.resize((size_t)n + 1);
    P->name_end.resize(n); P->bc_end.resize(n); P->rg_end.resize(n); P->valid.resize(n); P->same_prev.resize(n); P->seed.resize(n);
    P->seq.clear(); P->qual.clear(); P->names.clear(); P->barcodes.clear(); P->rgs.clear();
    P->seq.reserve(a.text.n + b.text.n); P->qual.reserve(a.text.n + b.text.n);  // upper bounds: no growth below
    P->names.reserve((size_t)n * 32); P->barcodes.reserve((size_t)n * 20); P->rgs.reserve((size_t)n * 12);
    const HeaderOut ho{&P->names, &P->barcodes, &P->rgs};
    const char* ta = a.text.p; const char* tb = b.text.p; const uint32_t* la = a.nl.p; const uint32_t* lb = b.nl.p;
    size_t o = 0;
    for (int r = 0; r < n; ++r) {
        // line k of the chunk is text[nl[k - 1] + 1, nl[k])
        const size_t ha = r ? (size_t)la[4 * r - 1] + 1 : 0, hb = r ? (size_t)lb[4 * r - 1] + 1 : 0;
        const char* h1 = ta + ha; const size_t hl = la[4 * r] - ha;
        const char* s1 = ta + la[4 * r] + 1; const size_t sl1 = la[4 * r + 1] - la[4 * r] - 1;
        const char* q1 = ta + la[4 * r + 2] + 1; const size_t ql1 = la[4 * r + 3] - la[4 * r + 2] - 1;
        (void)hb;
        const char* s2 = tb + lb[4 * r] + 1; const size_t sl2 = lb[4 * r + 1] - lb[4 * r] - 1;
        const char* q2 = tb + lb[4 * r + 2] + 1; const size_t ql2 = lb[4 * r + 3] - lb[4 * r + 2] - 1;
        if (hl == 0 || h1[0] != '@') { P->rc = ARACHNE_ERR_IO; char m[128]; snprintf(m, sizeof(m), "malformed FASTQ: record header expected at line %ld of R1", a.first_line + 4L * r + 1); P->err = m; return; }
        const size_t n0 = P->names.size(), b0 = P->barcodes.size();
        bool valid;
        parse_header(h1 + 1, hl, ho, &valid);  // hl - 1 header bytes + the newline
        P->name_end[r] = (int32_t)P->names.size(); P->bc_end[r] = (int32_t)P->barcodes.size(); P->rg_end[r] = (int32_t)P->rgs.size();
        {
            const size_t bl = P->barcodes.size() - b0, pb0 = r >= 2 ? (size_t)P->bc_end[r - 2] : 0;
            P->same_prev[r] = r >= 1 && b0 - pb0 == bl && !memcmp(P->barcodes.data() + pb0, P->barcodes.data() + b0, bl);
        }
        P->valid[r] = valid;
        P->seed[r] = arachne::md5_name_seed(P->names.data() + n0, (int64_t)(P->names.size() - n0));
        P->soff[r] = (int64_t)o;
        uint8_t* sq = P->seq.p + o; char* ql = P->qual.p + o;
        to_codes(s1, sl1, sq); to_codes(s2, sl2, sq + sl1);
        // qualities are carried as read; a quality line shorter than its sequence is padded with '!' (the reference would index past it)
        if (ql1 >= sl1) memcpy(ql, q1, sl1); else { memcpy(ql, q1, ql1); memset(ql + ql1, '!', sl1 - ql1); }
        if (ql2 >= sl2) memcpy(ql + sl1, q2, sl2); else { memcpy(ql + sl1, q2, ql2); memset(ql + sl1 + ql2, '!', sl2 - ql2); }
        P->len[2 * r] = (int32_t)sl1; P->len[2 * r + 1] = (int32_t)sl2;
        o += sl1 + sl2;
    }
    P->soff[n] = (int64_t)o; P->seq.n = P->qual.n = o;
}

void parse_thread(arachne_ingest* I) {
    struct Tally { arachne_ingest* I; ~Tally() { std::lock_guard<std::mutex> l(I->stat_mu); I->cpu_parse += thread_cpu(); } } tally{I};
    for (;;) {
        RawChunk a, b; long idx;
        {   // chunk i of both files must be taken together and in order
            std::lock_guard<std::mutex> t(I->take_mu);
            if (I->parse_done) return;
            const bool ga = I->raw[0]->pop(a), gb = I->raw[1]->pop(b);
            if (!ga || !gb) {
                I->parse_done = true; I->raw[0]->close(); I->raw[1]->close();  // the other file's splitter may be waiting for room
                Parsed e; e.last = true; e.rc = I->split_rc[0] ? I->split_rc[0] : I->split_rc[1]; e.err = I->split_rc[0] ? I->split_err[0] : I->split_err[1];
                std::lock_guard<std::mutex> l(I->pmu);
                I->parsed[I->next_parse++] = std::move(e); I->pcv.notify_all(); return;
            }
            idx = I->next_parse++;
            if (a.last || b.last) { I->parse_done = true; I->raw[0]->close(); I->raw[1]->close(); }
        }
        Parsed P = I->take_parsed(); parse_pair(a, b, &P);
        I->give_chunk(std::move(a)); I->give_chunk(std::move(b));
        std::unique_lock<std::mutex> l(I->pmu);
        I->pcv.wait(l, [&] { return idx - I->next_take < 4L * I->n_threads + 8; });  // bound the re-order window
        I->parsed[idx] = std::move(P);
        I->pcv.notify_all();
    }
}

struct BatchBuf {
    Raw<uint8_t> seq; Raw<char> qual; Raw<int64_t> off, unit_off, seed, name_off, bc_off, rg_off;
    Raw<uint8_t> unique, run_rfa, valid;
    Raw<char> names, barcodes, rgs;
    arachne_ingest_batch_t pub;
    arachne_ingest* owner = nullptr;
    void reset() {
        seq.clear(); qual.clear(); unique.clear(); run_rfa.clear(); valid.clear(); seed.clear(); names.clear(); barcodes.clear(); rgs.clear();
        off.clear(); off.push(0); unit_off.clear(); unit_off.push(0); name_off.clear(); name_off.push(0); bc_off.clear(); bc_off.push(0); rg_off.clear(); rg_off.push(0);
    }
};

// Records [lo, hi) of the current chunk go to the batch: one copy per array.
void append_records(BatchBuf* B, const Parsed& P, int lo, int hi) {
    if (hi <= lo) return;
    const size_t so = (size_t)P.soff[lo], nb = (size_t)P.soff[hi] - so;
    memcpy(B->seq.grow(nb), P.seq.p + so, nb);
    memcpy(B->qual.grow(nb), P.qual.p + so, nb);
    {
        int64_t at = B->off.back(); int64_t* w = B->off.grow(2 * (size_t)(hi - lo)); const int32_t* len = P.len.data() + 2 * (size_t)lo;
        for (int t = 0; t < 2 * (hi - lo); ++t) { at += len[t]; w[t] = at; }
    }
    auto strings = [&](Raw<char>& pool, Raw<int64_t>& offs, const std::string& src, const std::vector<int32_t>& end) {
        const int32_t s0 = lo ? end[lo - 1] : 0; const int64_t rebase = (int64_t)pool.n - s0;
        const size_t bytes = (size_t)(end[hi - 1] - s0);
        memcpy(pool.grow(bytes), src.data() + s0, bytes);
        int64_t* w = offs.grow((size_t)(hi - lo));
        for (int t = lo; t < hi; ++t) w[t - lo] = rebase + end[t];
    };
    strings(B->names, B->name_off, P.names, P.name_end);
    strings(B->rgs, B->rg_off, P.rgs, P.rg_end);
    memcpy(B->seed.grow((size_t)(hi - lo)), P.seed.data() + lo, (size_t)(hi - lo) * sizeof(int64_t));
}

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
    const double cpu0 = thread_cpu();
    BatchBuf* B = nullptr;
    {
        std::lock_guard<std::mutex> l(I->pool_mu);
        if (!I->pool.empty()) { B = static_cast<BatchBuf*>(I->pool.back()); I->pool.pop_back(); }
    }
    if (!B) B = new BatchBuf();
    B->reset();
    B->owner = I;
    {   // growing a 100 MB array by steps copies it several times over: size it from the request instead
        const size_t want = (size_t)std::min<int64_t>(target_pairs, 4000000) + 65536;
        B->seq.reserve(want * 320); B->qual.reserve(want * 320); B->off.reserve(2 * want + 2); B->seed.reserve(want); B->name_off.reserve(want + 1);
        B->names.reserve(want * 24); B->rg_off.reserve(want + 1);
    }
    int rc = 0;
    int64_t n_pairs = 0;
    int pend_lo = I->cur_pos;  // records [pend_lo, cur_pos) of the current chunk are taken but not copied yet
    auto flush = [&] { if (I->have_cur) append_records(B, I->cur, pend_lo, I->cur_pos); pend_lo = I->cur_pos; };
    // the next record of the stream is I->cur[I->cur_pos]; false at the end of the input
    auto next_record = [&]() -> bool {
        while (!I->have_cur || I->cur_pos >= I->cur.n) {
            flush();
            if (I->have_cur && I->cur.last) { I->eof = true; return false; }
            std::unique_lock<std::mutex> l(I->pmu);
            I->pcv.wait(l, [&] { return I->parsed.count(I->next_take) > 0; });
            Parsed done = std::move(I->cur);
            I->cur = std::move(I->parsed[I->next_take]); I->parsed.erase(I->next_take); ++I->next_take;
            I->pcv.notify_all();
            l.unlock();
            if (done.seq.cap) I->give_parsed(std::move(done));
            I->cur_pos = 0; pend_lo = 0; I->have_cur = true;
            if (I->cur.rc) { rc = I->cur.rc; set_error("%s", I->cur.err.c_str()); I->eof = true; I->cur.n = 0; return false; }
        }
        return true;
    };
    // ReadBarcodeSet, one call per loop trip (reader.go:208-291); a record that opens the next unit is simply not consumed yet
    for (;;) {
        int index = 0; bool ended_by_change = false, at_eof = false;
        std::string first_bc;
        const int64_t unit_start = n_pairs;
        bool cont = false;
        while (index < UNIT_CAP) {
            if (!next_record()) { at_eof = true; break; }
            const Parsed& P = I->cur; const int r = I->cur_pos;
            const int b0 = r ? P.bc_end[r - 1] : 0;
            const size_t bl = (size_t)(P.bc_end[r] - b0);
            if (index == 0) {
                first_bc.assign(P.barcodes.data() + b0, bl);
                cont = I->have_last && first_bc == I->last_barcode;
                memcpy(B->barcodes.grow(bl), first_bc.data(), bl); B->bc_off.push((int64_t)B->barcodes.n);
                B->valid.push(P.valid[r]);
            } else if (bl != first_bc.size() || memcmp(P.barcodes.data() + b0, first_bc.data(), bl)) { ended_by_change = true; break; }
            // how many records of this chunk can be taken in one go: the run of this barcode (same_prev marks a record whose barcode
            // equals its predecessor's), cut by the cap and by the continuation rule (a record at index >= 200 ends the unit)
            int lim = UNIT_CAP - index;
            if (cont) { const int to_break = index >= UNIT_CONT ? 1 : UNIT_CONT - index + 1; lim = lim < to_break ? lim : to_break; }
            int k = 1;
            while (k < lim && r + k < P.n && P.same_prev[r + k]) ++k;
            I->cur_pos += k; n_pairs += k; index += k;
            if (cont && index - 1 >= UNIT_CONT) break;  // "abnormal break" (reader.go:266-270): the record at index >= 200 stays in the unit
        }
        if (rc) { delete B; return rc; }
        const int64_t n_unit = n_pairs - unit_start;
        if (n_unit == 0) break;  // end of input
        I->last_barcode = first_bc; I->have_last = true;
        const bool full = ended_by_change || at_eof;  // reader.go:283-289
        B->unit_off.push(n_pairs);
        B->unique.push(full);
        B->run_rfa.push(full && first_bc.find('-') != std::string::npos && n_unit >= 5);  // worthRunningRFA
        if (at_eof || n_pairs >= target_pairs) break;
    }
    flush();
    I->cpu_asm += thread_cpu() - cpu0;
    if (n_pairs == 0) { delete B; return ARACHNE_INGEST_EOF; }
    I->n_pairs_total += n_pairs;
    arachne_ingest_batch_t& o = B->pub;
    o.n_pairs = n_pairs; o.n_units = (int64_t)B->unique.n;
    o.seq = B->seq.p; o.seq_off = B->off.p; o.qual = B->qual.p; o.unit_off = B->unit_off.p;
    o.unique = B->unique.p; o.run_rfa = B->run_rfa.p; o.valid = B->valid.p; o.rand_seed = B->seed.p;
    o.names = B->names.p; o.name_off = B->name_off.p; o.barcodes = B->barcodes.p; o.barcode_char_off = B->bc_off.p;
    o.read_groups = B->rgs.p; o.read_group_off = B->rg_off.p; o.priv = B;
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
