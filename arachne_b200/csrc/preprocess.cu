// FASTQ preprocessing (include/arachne_preprocess.h): format sniffing, header standardisation and the barcode sort
// that src/preprocess delegates to three samtools processes.  Host code reads / rewrites / writes; the ordering is a stable LSD radix
// sort on the device over the barcode packed big-endian into 64-bit words (byte order == strcmp order, shorter strings first).
//   findFastqFormat            src/preprocess/standardize.go:98-125
//   the patterns               :16-25
//   standardizeFrom*           :128-186
//   writeR1/R2FastqRecord      :73-95
//   runSamtoolsPipelineWithPipes  src/preprocess/preprocess.go:43-110
#include <cuda_runtime.h>
#include <cub/cub.cuh>
#include <zlib.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <unordered_set>
#include <vector>
#include "../../include/arachne_preprocess.h"

namespace arachne { void set_error(const char* fmt, ...); }
using arachne::set_error;

namespace {

inline bool ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\v' || c == '\f'; }
inline bool re_ws(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f'; }  // Go's \s
inline bool digit(char c) { return c >= '0' && c <= '9'; }

// ---- the patterns of standardize.go:16-25, written out (RE2 leftmost semantics: the first position where the pattern fits) ----
// `TAG(\S+)\s`: the value of the first occurrence of `lit` that is followed by at least one non-space and then a space.
bool find_tag_value(const std::string& s, const char* lit, std::string* val) {
    const size_t ll = strlen(lit);
    for (size_t p = s.find(lit); p != std::string::npos; p = s.find(lit, p + 1)) {
        size_t e = p + ll;
        while (e < s.size() && !re_ws(s[e])) ++e;
        if (e > p + ll && e < s.size()) { if (val) val->assign(s, p + ll, e - p - ll); return true; }
    }
    return false;
}
// `VX:i:([01])\s`
bool find_vx(const std::string& s, char* v) {
    for (size_t p = s.find("VX:i:"); p != std::string::npos; p = s.find("VX:i:", p + 1))
        if (p + 6 < s.size() && (s[p + 5] == '0' || s[p + 5] == '1') && re_ws(s[p + 6])) { if (v) *v = s[p + 5]; return true; }
    return false;
}
// `BX:Z:(A\d{2}C\d{2}B\d{2}D\d{2})\s`
bool find_haplotag(const std::string& s, std::string* val) {
    for (size_t p = s.find("BX:Z:"); p != std::string::npos; p = s.find("BX:Z:", p + 1)) {
        const size_t b = p + 5;
        if (b + 12 >= s.size()) return false;
        bool ok = re_ws(s[b + 12]);
        for (int k = 0; k < 4 && ok; ++k) ok = s[b + 3 * k] == "ACBD"[k] && digit(s[b + 3 * k + 1]) && digit(s[b + 3 * k + 2]);
        if (ok) { if (val) val->assign(s, b, 12); return true; }
    }
    return false;
}
// `#([0-9]+_[0-9]+_[0-9]+)\s`
bool find_stlfr(const std::string& s, std::string* val) {
    for (size_t p = s.find('#'); p != std::string::npos; p = s.find('#', p + 1)) {
        size_t e = p + 1;
        int part = 0;
        for (; part < 3; ++part) {
            const size_t d0 = e;
            while (e < s.size() && digit(s[e])) ++e;
            if (e == d0) break;
            if (part < 2) { if (e < s.size() && s[e] == '_') ++e; else break; }
        }
        // `[0-9]+` is greedy but the pattern can also stop early; with `\s` required right after, only the full run can match
        if (part == 3 && e < s.size() && re_ws(s[e])) { if (val) val->assign(s, p + 1, e - p - 1); return true; }
    }
    return false;
}
// `:([ATCGN]+)\s`
bool find_tellseq(const std::string& s, std::string* val) {
    for (size_t p = s.find(':'); p != std::string::npos; p = s.find(':', p + 1)) {
        size_t e = p + 1;
        while (e < s.size() && (s[e] == 'A' || s[e] == 'T' || s[e] == 'C' || s[e] == 'G' || s[e] == 'N')) ++e;
        if (e > p + 1 && e < s.size() && re_ws(s[e])) { if (val) val->assign(s, p + 1, e - p - 1); return true; }
    }
    return false;
}

int header_format(const std::string& h) {  // the if-chain of findFastqFormat, standardize.go:113-121
    if (find_tag_value(h, "BX:Z:", nullptr) && find_vx(h, nullptr)) return ARACHNE_FQ_STANDARD;
    if (find_haplotag(h, nullptr)) return ARACHNE_FQ_HAPLOTAGGING;
    if (find_stlfr(h, nullptr)) return ARACHNE_FQ_STLFR;
    if (find_tellseq(h, nullptr)) return ARACHNE_FQ_TELLSEQ;
    return ARACHNE_FQ_UNKNOWN;
}

// A header line (no '@', no newline) -> the string the patterns see: the first field without its mate suffix, the rest of the line,
// a newline.  (ParseHeader, reader.go:101-102, cuts the suffix the same way; the patterns need the barcode to END the name.)
struct Header { std::string name; std::string rest; int mate = 0; };
Header split_header(const char* s, size_t n) {
    Header h;
    size_t e = 0;
    while (e < n && !ws(s[e])) ++e;
    size_t ne = e;
    if (ne >= 2 && s[ne - 2] == '/' && (s[ne - 1] == '1' || s[ne - 1] == '2')) { h.mate = s[ne - 1] - '0'; ne -= 2; }
    h.name.assign(s, ne);
    h.rest.assign(s + e, n - e);
    return h;
}
std::string pattern_text(const Header& h) { return h.name + h.rest + "\n"; }

// fields of the comment that are SAM tags (XX:T:value), in order: what `samtools import -T '*'` keeps
bool is_sam_tag(const char* s, size_t n) {
    auto al = [](char c) { return (c >= 'A' && c <= 'Z') || (c >= 'a' && c <= 'z'); };
    return n >= 5 && al(s[0]) && (al(s[1]) || digit(s[1])) && s[2] == ':' && strchr("AifZHB", s[3]) != nullptr && s[4] == ':';
}
void sam_tags(const std::string& rest, std::vector<std::pair<size_t, size_t>>* out) {
    out->clear();
    size_t p = 0;
    while (p < rest.size()) {
        while (p < rest.size() && ws(rest[p])) ++p;
        size_t e = p;
        while (e < rest.size() && !ws(rest[e])) ++e;
        if (e > p && is_sam_tag(rest.data() + p, e - p)) out->push_back({p, e - p});
        p = e;
    }
}

// ---- whole-file reader: 4-line records, plain or gzip ----
struct Rec { uint64_t hdr, seq, qual; uint32_t hdr_n, seq_n, qual_n; };
struct FastqFile {
    std::string text; std::vector<Rec> recs; std::string err;
    void load(const char* path, int64_t max_records) {
        gzFile f = gzopen(path, "rb");
        if (!f) { err = std::string("cannot open ") + path; return; }
        gzbuffer(f, 1 << 20);
        std::vector<char> buf(1 << 22);
        int n;
        // (sniffing reads a prefix only: stop once enough lines are in)
        size_t lines = 0;
        while ((n = gzread(f, buf.data(), (unsigned)buf.size())) > 0) {
            text.append(buf.data(), (size_t)n);
            if (max_records > 0) { for (int i = 0; i < n; ++i) lines += buf[i] == '\n'; if ((int64_t)lines >= 4 * max_records) break; }
        }
        if (n < 0) { int e; err = std::string("read error in ") + path + ": " + gzerror(f, &e); }
        gzclose(f);
        if (!err.empty()) return;
        size_t p = 0;
        const size_t N = text.size();
        auto line = [&](uint64_t* beg, uint32_t* len) -> bool {
            if (p >= N) return false;
            const char* nl = static_cast<const char*>(memchr(text.data() + p, '\n', N - p));
            size_t e = nl ? (size_t)(nl - text.data()) : N;
            *beg = p; size_t l = e - p;
            if (l && text[e - 1] == '\r') --l;
            *len = (uint32_t)l; p = nl ? e + 1 : N;
            return true;
        };
        while (p < N && (max_records <= 0 || (int64_t)recs.size() < max_records)) {
            Rec r; uint64_t b; uint32_t l;
            if (!line(&r.hdr, &r.hdr_n)) break;
            if (r.hdr_n == 0 && p >= N) break;  // trailing blank line
            if (r.hdr_n == 0 || text[r.hdr] != '@') { err = std::string(path) + ": record " + std::to_string(recs.size() + 1) + " does not start with '@'"; return; }
            if (!line(&r.seq, &r.seq_n) || !line(&b, &l) || !line(&r.qual, &r.qual_n)) {
                if (max_records > 0) break;  // a prefix may end inside a record
                err = std::string(path) + ": truncated record " + std::to_string(recs.size() + 1); return;
            }
            if (l == 0 || text[b] != '+') { err = std::string(path) + ": record " + std::to_string(recs.size() + 1) + " has no '+' line"; return; }
            r.hdr += 1; r.hdr_n -= 1;  // without '@'
            recs.push_back(r);
        }
    }
};

struct Writer {
    gzFile gz = nullptr; FILE* fp = nullptr; std::string buf;
    bool open(const char* path, int level) {
        const size_t n = strlen(path);
        if (n > 3 && strcmp(path + n - 3, ".gz") == 0) { char mode[8]; snprintf(mode, sizeof mode, "wb%d", level); gz = gzopen(path, mode); if (gz) gzbuffer(gz, 1 << 20); return gz != nullptr; }
        fp = fopen(path, "wb");
        return fp != nullptr;
    }
    bool flush() {
        bool ok = true;
        if (!buf.empty()) ok = gz ? gzwrite(gz, buf.data(), (unsigned)buf.size()) == (int)buf.size() : fwrite(buf.data(), 1, buf.size(), fp) == buf.size();
        buf.clear();
        return ok;
    }
    bool close() { bool ok = flush(); if (gz) ok = gzclose(gz) == Z_OK && ok; if (fp) ok = fclose(fp) == 0 && ok; gz = nullptr; fp = nullptr; return ok; }
};

__global__ void __launch_bounds__(256) k_pp_gather(const uint64_t* __restrict__ words, const uint32_t* __restrict__ perm, uint64_t* out, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = words[perm[i]];
}
__global__ void __launch_bounds__(256) k_pp_iota(uint32_t* perm, uint32_t n) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) perm[i] = i;
}

#define PP_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { set_error("preprocess: %s: %s", #x, cudaGetErrorString(e_)); rc = ARACHNE_ERR_CUDA; goto done; } } while (0)

// Stable sort of n keys of n_words big-endian 64-bit words each (word 0 most significant; keys[w * n + i]).  perm_out[rank] = input index.
int device_sort(int device, const std::vector<uint64_t>& keys, int n_words, uint32_t n, std::vector<uint32_t>* perm_out) {
    int rc = ARACHNE_OK;
    uint64_t *d_keys = nullptr, *d_in = nullptr, *d_out = nullptr; uint32_t *d_p0 = nullptr, *d_p1 = nullptr; void* d_tmp = nullptr;
    size_t tmp = 0;
    const unsigned grid = (n + 255) / 256;
    perm_out->resize(n);
    if (n == 0) return rc;
    PP_CUDA(cudaSetDevice(device));
    PP_CUDA(cudaMalloc(&d_keys, (size_t)n_words * n * 8)); PP_CUDA(cudaMalloc(&d_in, (size_t)n * 8)); PP_CUDA(cudaMalloc(&d_out, (size_t)n * 8));
    PP_CUDA(cudaMalloc(&d_p0, (size_t)n * 4)); PP_CUDA(cudaMalloc(&d_p1, (size_t)n * 4));
    PP_CUDA(cudaMemcpy(d_keys, keys.data(), (size_t)n_words * n * 8, cudaMemcpyHostToDevice));
    PP_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp, d_in, d_out, d_p0, d_p1, (int)n));
    PP_CUDA(cudaMalloc(&d_tmp, tmp + 16));
    k_pp_iota<<<grid, 256>>>(d_p0, n);
    for (int w = n_words - 1; w >= 0; --w) {  // least significant word first; every pass is stable
        k_pp_gather<<<grid, 256>>>(d_keys + (size_t)w * n, d_p0, d_in, n);
        PP_CUDA(cub::DeviceRadixSort::SortPairs(d_tmp, tmp, d_in, d_out, d_p0, d_p1, (int)n));
        std::swap(d_p0, d_p1);
    }
    PP_CUDA(cudaGetLastError());
    PP_CUDA(cudaMemcpy(perm_out->data(), d_p0, (size_t)n * 4, cudaMemcpyDeviceToHost));
done:
    cudaFree(d_keys); cudaFree(d_in); cudaFree(d_out); cudaFree(d_p0); cudaFree(d_p1); cudaFree(d_tmp);
    return rc;
}

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

extern "C" {

int arachne_fastq_header_format(const char* header) {
    if (!header) return ARACHNE_FQ_UNKNOWN;
    size_t n = strlen(header);
    while (n && (header[n - 1] == '\n' || header[n - 1] == '\r')) --n;
    return header_format(pattern_text(split_header(header, n)));
}

int arachne_fastq_format(const char* r1, const char* r2, int* format) {
    if (!r1 || !r2 || !format) { set_error("arachne_fastq_format: null argument"); return ARACHNE_ERR_ARG; }
    FastqFile a, b;
    a.load(r1, 200); b.load(r2, 1);
    if (!a.err.empty() || !b.err.empty()) { set_error("%s", (a.err.empty() ? b.err : a.err).c_str()); return ARACHNE_ERR_IO; }
    *format = ARACHNE_FQ_UNKNOWN;
    for (const Rec& r : a.recs) {
        const int f = header_format(pattern_text(split_header(a.text.data() + r.hdr, r.hdr_n)));
        if (f != ARACHNE_FQ_UNKNOWN) { *format = f; break; }
    }
    return ARACHNE_OK;
}

int arachne_preprocess(const char* r1, const char* r2, const char* out_r1, const char* out_r2, const arachne_preprocess_opt_t* opt_,
                       arachne_preprocess_stats_t* stats) {
    if (!r1 || !r2 || !out_r1 || !out_r2) { set_error("arachne_preprocess: null argument"); return ARACHNE_ERR_ARG; }
    arachne_preprocess_opt_t opt; memset(&opt, 0, sizeof opt); opt.standardize = 1; opt.sort = 1;
    if (opt_) opt = *opt_;
    const std::string tag = std::string(opt.tag && *opt.tag ? opt.tag : "BX");
    if (tag.size() != 2) { set_error("arachne_preprocess: a SAM tag has two characters, got '%s'", tag.c_str()); return ARACHNE_ERR_ARG; }
    const int level = opt.compress_level > 0 ? (opt.compress_level > 9 ? 9 : opt.compress_level) : 4;
    arachne_preprocess_stats_t st; memset(&st, 0, sizeof st);
    double t0 = now_ms();
    // ---- read both files (a thread each) ----
    FastqFile f[2];
    { std::thread t([&] { f[1].load(r2, 0); }); f[0].load(r1, 0); t.join(); }
    for (int k = 0; k < 2; ++k) if (!f[k].err.empty()) { set_error("%s", f[k].err.c_str()); return ARACHNE_ERR_IO; }
    if (f[0].recs.size() != f[1].recs.size()) { set_error("arachne_preprocess: %zu records in %s but %zu in %s", f[0].recs.size(), r1, f[1].recs.size(), r2); return ARACHNE_ERR_IO; }
    const size_t n = f[0].recs.size();
    if (n >= 0xffffffffull) { set_error("arachne_preprocess: too many pairs for one pass"); return ARACHNE_ERR_ARG; }
    // ---- format (findFastqFormat) ----
    int format = ARACHNE_FQ_UNKNOWN;
    for (size_t i = 0; i < n && i < 200 && format == ARACHNE_FQ_UNKNOWN; ++i)
        format = header_format(pattern_text(split_header(f[0].text.data() + f[0].recs[i].hdr, f[0].recs[i].hdr_n)));
    st.format = format;
    const bool convert = opt.standardize && format != ARACHNE_FQ_STANDARD;
    if (convert && format == ARACHNE_FQ_UNKNOWN) { set_error("arachne_preprocess: %s is in none of the known formats (standard, haplotagging, stLFR, TELLseq)", r1); return ARACHNE_ERR_ARG; }
    // ---- new headers (without '@', mate suffix and newline: "<name>" and "\tTAG...") + the sort key of every pair ----
    std::vector<std::string> name[2], tags[2];
    std::vector<std::string> key(n);
    for (int k = 0; k < 2; ++k) { name[k].resize(n); tags[k].resize(n); }
    const std::string tag_lit = tag + ":";
    auto do_range = [&](size_t lo, size_t hi, int64_t* n_invalid, int64_t* n_no_tag) {
        std::vector<std::pair<size_t, size_t>> tg;
        for (size_t i = lo; i < hi; ++i) {
            for (int k = 0; k < 2; ++k) {
                const Rec& r = f[k].recs[i];
                Header h = split_header(f[k].text.data() + r.hdr, r.hdr_n);
                std::string out_tags, bc;
                if (convert) {
                    // the barcode and its validity (standardize.go:128-186).  The broken source tests Contains("00") / Contains("N") with the
                    // sense inverted relative to stLFR's `!invalid`; here VX:i:1 always means "no invalid segment".
                    const std::string txt = pattern_text(h);
                    bool valid = false;
                    if (format == ARACHNE_FQ_HAPLOTAGGING) { if (find_haplotag(txt, &bc) || find_tag_value(txt, "BX:Z:", &bc)) valid = bc.find("00") == std::string::npos; }
                    else if (format == ARACHNE_FQ_STLFR) {
                        if (find_stlfr(txt, &bc)) valid = !(bc.compare(0, 2, "0_") == 0 || bc.find("_0_") != std::string::npos || (bc.size() >= 2 && bc.compare(bc.size() - 2, 2, "_0") == 0));
                    } else if (find_tellseq(txt, &bc)) valid = bc.find('N') == std::string::npos;
                    if (bc.empty()) valid = false;
                    out_tags = "\tBX:Z:" + bc + "\tVX:i:" + (valid ? "1" : "0");
                    if (k == 0) { *n_invalid += !valid; if (tag == "BX") key[i] = bc; else if (tag == "VX") key[i] = valid ? "1" : "0"; }
                } else {
                    sam_tags(h.rest, &tg);
                    for (auto& t : tg) {
                        out_tags += '\t'; out_tags.append(h.rest, t.first, t.second);
                        if (k == 0 && key[i].empty() && h.rest.compare(t.first, 3, tag_lit) == 0) key[i].assign(h.rest, t.first + 5, t.second - 5);
                        if (k == 0 && t.second == 6 && h.rest.compare(t.first, 6, "VX:i:0") == 0) ++*n_invalid;
                    }
                }
                if (k == 0 && key[i].empty()) ++*n_no_tag;
                name[k][i] = std::move(h.name); tags[k][i] = std::move(out_tags);
            }
        }
    };
    {
        const int nt = 8;
        std::vector<std::thread> th; std::vector<int64_t> inv(nt, 0), nok(nt, 0);
        for (int t = 0; t < nt; ++t) th.emplace_back(do_range, n * t / nt, n * (t + 1) / nt, &inv[t], &nok[t]);
        for (auto& t : th) t.join();
        for (int t = 0; t < nt; ++t) { st.n_invalid += inv[t]; st.n_no_tag += nok[t]; }
    }
    st.n_pairs = (int64_t)n;
    st.ms_read = (float)(now_ms() - t0); t0 = now_ms();
    // ---- order ----
    std::vector<uint32_t> perm(n);
    for (size_t i = 0; i < n; ++i) perm[i] = (uint32_t)i;
    if (opt.sort && n) {
        size_t maxlen = 0;
        for (const std::string& s : key) maxlen = s.size() > maxlen ? s.size() : maxlen;
        if (maxlen > 32) { set_error("arachne_preprocess: a %s value of %zu bytes; the sort takes at most 32", tag.c_str(), maxlen); return ARACHNE_ERR_ARG; }
        const int nw = maxlen ? (int)((maxlen + 7) / 8) : 1;
        std::vector<uint64_t> words((size_t)nw * n, 0);
        for (size_t i = 0; i < n; ++i)
            for (size_t c = 0; c < key[i].size(); ++c) words[(c >> 3) * n + i] |= (uint64_t)(uint8_t)key[i][c] << (56 - 8 * (c & 7));
        const int rc = device_sort(opt.device, words, nw, (uint32_t)n, &perm);
        if (rc != ARACHNE_OK) return rc;
    }
    if (opt.sort) { for (size_t i = 0; i < n; ++i) if (!key[perm[i]].empty() && (i == 0 || key[perm[i]] != key[perm[i - 1]])) ++st.n_barcodes; }
    else { std::unordered_set<std::string> seen; for (const std::string& k : key) if (!k.empty()) seen.insert(k); st.n_barcodes = (int64_t)seen.size(); }
    st.ms_sort = (float)(now_ms() - t0); t0 = now_ms();
    // ---- write (a thread per file) ----
    bool ok[2] = {true, true};
    auto write_file = [&](int k, const char* path) {
        Writer w;
        if (!w.open(path, level)) { ok[k] = false; return; }
        for (size_t j = 0; j < n && ok[k]; ++j) {
            const size_t i = perm[j];
            const Rec& r = f[k].recs[i];
            w.buf += '@'; w.buf += name[k][i]; w.buf += k == 0 ? "/1" : "/2"; w.buf += tags[k][i]; w.buf += '\n';
            w.buf.append(f[k].text, r.seq, r.seq_n); w.buf += "\n+\n"; w.buf.append(f[k].text, r.qual, r.qual_n); w.buf += '\n';
            if (w.buf.size() > (1u << 22)) ok[k] = w.flush();
        }
        ok[k] = w.close() && ok[k];
    };
    { std::thread t(write_file, 1, out_r2); write_file(0, out_r1); t.join(); }
    if (!ok[0] || !ok[1]) { set_error("arachne_preprocess: cannot write %s", ok[0] ? out_r2 : out_r1); return ARACHNE_ERR_IO; }
    st.ms_write = (float)(now_ms() - t0);
    if (stats) *stats = st;
    return ARACHNE_OK;
}

}  // extern "C"
