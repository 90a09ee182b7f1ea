// arachne_index_build: the device backend of index_build.cuh (grid-stride kernels + CUB sort/select/scan) and the file
// I/O around it.  Replaces the BWT/SA half of `bwa index` (bwa/bwtindex.c:244-300: "Construct BWT", "Update BWT",
// "Construct SA from BWT and Occ"); the FASTA packing half (bntseq.c:275-328) stays the reference's.
#include <cuda_runtime.h>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <thrust/iterator/counting_iterator.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include "index_build.cuh"
#include "../../include/arachne_cuda.h"

namespace arachne {
void set_error(const char* fmt, ...);

namespace ib {

template <class F>
__global__ void __launch_bounds__(256) k_ib_for_each(uint64_t n, F f) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < n; i += (uint64_t)gridDim.x * blockDim.x) f(i);
}

// 4096-bin histogram of the first 6 bases of every suffix: per-block shared-memory counters, one flush per block
__global__ void __launch_bounds__(512) k_ib_hist(TextView t, unsigned long long* hist) {
    __shared__ unsigned int sh[N_BINS];
    for (int i = threadIdx.x; i < N_BINS; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    // each block takes contiguous spans of 2^20 positions so a u32 counter cannot overflow between flushes
    const uint64_t span = 1ull << 20, n_spans = (t.n + span - 1) / span;
    for (uint64_t s = blockIdx.x; s < n_spans; s += gridDim.x) {
        const uint64_t lo = s * span, hi = lo + span < t.n ? lo + span : t.n;
        for (uint64_t p = lo + threadIdx.x; p < hi; p += blockDim.x) atomicAdd(&sh[sfx_bin(t, p)], 1u);
        __syncthreads();
        for (int i = threadIdx.x; i < N_BINS; i += blockDim.x) {
            const unsigned int v = sh[i];
            if (v) { atomicAdd(&hist[i], (unsigned long long)v); sh[i] = 0; }
        }
        __syncthreads();
    }
}

struct InBins {
    TextView t; uint32_t b0, b1;
    __host__ __device__ bool operator()(uint64_t pos) const { const uint32_t b = sfx_bin(t, pos); return b >= b0 && b < b1; }
};

struct DeviceBackend {
    cudaStream_t st = nullptr;
    int n_sm = 148;
    bool good = true;
    void* tmp = nullptr; size_t tmp_bytes = 0;
    unsigned long long* d_count = nullptr;
    size_t live_bytes = 0, peak_bytes = 0;
    std::vector<std::pair<void*, size_t>> live;

    bool ok() const { return good; }
    void fail(const char* msg) { if (good) set_error("arachne_index_build: %s", msg); good = false; }
    bool ck(cudaError_t e, const char* what) {
        if (e != cudaSuccess && good) { set_error("arachne_index_build: %s -> %s", what, cudaGetErrorString(e)); good = false; }
        return e == cudaSuccess;
    }
    bool init(int device) {
        if (!ck(cudaSetDevice(device), "cudaSetDevice")) return false;
        cudaDeviceProp p;
        if (ck(cudaGetDeviceProperties(&p, device), "cudaGetDeviceProperties")) n_sm = p.multiProcessorCount;
        ck(cudaStreamCreate(&st), "cudaStreamCreate");
        d_count = (unsigned long long*)acquire(8);
        return good;
    }
    void* acquire(size_t bytes) {
        if (!good) return nullptr;
        void* p = nullptr;
        if (!ck(cudaMalloc(&p, bytes ? bytes : 1), "cudaMalloc")) return nullptr;
        live.push_back({p, bytes}); live_bytes += bytes; if (live_bytes > peak_bytes) peak_bytes = live_bytes;
        return p;
    }
    void release(void* p) {
        for (size_t i = 0; i < live.size(); ++i)
            if (live[i].first == p) { cudaStreamSynchronize(st); cudaFree(p); live_bytes -= live[i].second; live.erase(live.begin() + i); return; }
    }
    void forget(void* p) {  // ownership passes to the caller
        for (size_t i = 0; i < live.size(); ++i) if (live[i].first == p) { live.erase(live.begin() + i); return; }
    }
    void shutdown() {
        if (st) cudaStreamSynchronize(st);
        for (auto& l : live) cudaFree(l.first);
        live.clear();
        if (tmp) cudaFree(tmp);
        tmp = nullptr;
        if (st) cudaStreamDestroy(st);
        st = nullptr;
    }
    bool need_tmp(size_t bytes) {
        if (bytes <= tmp_bytes) return true;
        if (tmp) { cudaStreamSynchronize(st); cudaFree(tmp); tmp = nullptr; tmp_bytes = 0; }
        if (!ck(cudaMalloc(&tmp, bytes + bytes / 8 + 256), "cudaMalloc(temp)")) return false;
        tmp_bytes = bytes + bytes / 8 + 256;
        return true;
    }
    template <class F> void for_each(uint64_t n, F f) {
        if (!good || n == 0) return;
        const uint64_t want = (n + 255) / 256, cap = (uint64_t)n_sm * 32;
        k_ib_for_each<<<(unsigned)(want < cap ? want : cap), 256, 0, st>>>(n, f);
        ck(cudaGetLastError(), "kernel launch");
    }
    void histogram(const TextView& t, uint64_t* hist_host) {
        unsigned long long* d = (unsigned long long*)acquire(N_BINS * 8);
        if (!d) return;
        ck(cudaMemsetAsync(d, 0, N_BINS * 8, st), "memset");
        k_ib_hist<<<n_sm * 2, 512, 0, st>>>(t, d);
        ck(cudaGetLastError(), "k_ib_hist");
        to_host(hist_host, d, N_BINS * 8);
        release(d);
    }
    uint64_t read_count() {
        unsigned long long c = 0;
        to_host(&c, d_count, 8);
        return c;
    }
    uint64_t select_bins(const TextView& t, uint32_t b0, uint32_t b1, uint64_t* out) {
        if (!good) return 0;
        thrust::counting_iterator<uint64_t> it(0);
        InBins pred{t, b0, b1};
        size_t bytes = 0;
        ck(cub::DeviceSelect::If(nullptr, bytes, it, out, d_count, (::cuda::std::int64_t)t.n, pred, st), "DeviceSelect::If(size)");
        if (!need_tmp(bytes)) return 0;
        ck(cub::DeviceSelect::If(tmp, bytes, it, out, d_count, (::cuda::std::int64_t)t.n, pred, st), "DeviceSelect::If");
        return read_count();
    }
    uint64_t select_flagged(const uint8_t* flags, uint32_t* out, uint64_t n) {
        if (!good || n == 0) return 0;
        thrust::counting_iterator<uint32_t> it(0);
        size_t bytes = 0;
        ck(cub::DeviceSelect::Flagged(nullptr, bytes, it, flags, out, d_count, (::cuda::std::int64_t)n, st), "DeviceSelect::Flagged(size)");
        if (!need_tmp(bytes)) return 0;
        ck(cub::DeviceSelect::Flagged(tmp, bytes, it, flags, out, d_count, (::cuda::std::int64_t)n, st), "DeviceSelect::Flagged");
        return read_count();
    }
    template <class K, class V> void sort_pairs(const K* kin, K* kout, const V* vin, V* vout, uint64_t n, int end_bit) {
        if (!good || n == 0) return;
        size_t bytes = 0;
        ck(cub::DeviceRadixSort::SortPairs(nullptr, bytes, kin, kout, vin, vout, (::cuda::std::int64_t)n, 0, end_bit, st), "SortPairs(size)");
        if (!need_tmp(bytes)) return;
        ck(cub::DeviceRadixSort::SortPairs(tmp, bytes, kin, kout, vin, vout, (::cuda::std::int64_t)n, 0, end_bit, st), "SortPairs");
    }
    void inclusive_sum(const uint32_t* in, uint32_t* out, uint64_t n) {
        if (!good || n == 0) return;
        size_t bytes = 0;
        ck(cub::DeviceScan::InclusiveSum(nullptr, bytes, in, out, (::cuda::std::int64_t)n, st), "InclusiveSum(size)");
        if (!need_tmp(bytes)) return;
        ck(cub::DeviceScan::InclusiveSum(tmp, bytes, in, out, (::cuda::std::int64_t)n, st), "InclusiveSum");
    }
    void exclusive_sum(const uint64_t* in, uint64_t* out, uint64_t n) {
        if (!good || n == 0) return;
        size_t bytes = 0;
        ck(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (::cuda::std::int64_t)n, st), "ExclusiveSum(size)");
        if (!need_tmp(bytes)) return;
        ck(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (::cuda::std::int64_t)n, st), "ExclusiveSum");
    }
    void copy32(uint32_t* dst, const uint32_t* src, uint64_t n) {
        if (good && n) ck(cudaMemcpyAsync(dst, src, n * 4, cudaMemcpyDeviceToDevice, st), "copy32");
    }
    void to_host(void* dst, const void* src, size_t bytes) {
        if (!good) return;
        ck(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, st), "D2H");
        ck(cudaStreamSynchronize(st), "sync");
    }
    void to_device(void* dst, const void* src, size_t bytes) {
        if (!good) return;
        ck(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, st), "H2D");
        ck(cudaStreamSynchronize(st), "sync");
    }
    void zero(void* p, size_t bytes) { if (good) ck(cudaMemsetAsync(p, 0, bytes, st), "memset"); }
};

static double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

// stream a device buffer to an open file through a bounded pinned staging buffer
static bool dump_device(DeviceBackend& be, FILE* fp, const void* d, size_t bytes) {
    const size_t chunk = 64u << 20;
    void* h = nullptr;
    if (!be.ck(cudaMallocHost(&h, chunk), "cudaMallocHost")) return false;
    bool okay = true;
    for (size_t off = 0; off < bytes && okay; off += chunk) {
        const size_t nb = bytes - off < chunk ? bytes - off : chunk;
        be.to_host(h, (const char*)d + off, nb);
        okay = be.ok() && fwrite(h, 1, nb, fp) == nb;
    }
    cudaFreeHost(h);
    return okay;
}

}  // namespace ib
}  // namespace arachne

extern "C" int arachne_index_build(const char* prefix_, int device, int64_t batch_cap, arachne_index_build_info_t* info) {
    using namespace arachne;
    using namespace arachne::ib;
    if (!prefix_) { set_error("arachne_index_build: null prefix"); return ARACHNE_ERR_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { set_error("arachne_index_build: no CUDA device (this library has no CPU path)"); return ARACHNE_ERR_CUDA; }
    if (device < 0 || device >= ndev) { set_error("arachne_index_build: device %d out of range", device); return ARACHNE_ERR_ARG; }
    const std::string prefix(prefix_);
    const double t0 = now_s();
    // l_pac from .ann (bns_dump, bntseq.c:75), bases from .pac (forward strand; last byte = l_pac % 4, bntseq.c:311-318)
    long long l_pac = 0;
    {
        FILE* fp = fopen((prefix + ".ann").c_str(), "r");
        if (!fp || fscanf(fp, "%lld", &l_pac) != 1 || l_pac <= 0) { if (fp) fclose(fp); set_error("cannot read l_pac from %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
        fclose(fp);
    }
    std::vector<uint8_t> pac;
    {
        FILE* fp = fopen((prefix + ".pac").c_str(), "rb");
        if (!fp) { set_error("cannot read %s.pac", prefix.c_str()); return ARACHNE_ERR_IO; }
        fseek(fp, 0, SEEK_END);
        const long sz = ftell(fp);
        fseek(fp, 0, SEEK_SET);
        pac.resize(sz > 0 ? sz : 0);
        const size_t got = sz > 0 ? fread(pac.data(), 1, sz, fp) : 0;
        fclose(fp);
        const size_t data = (size_t)((l_pac >> 2) + ((l_pac & 3) ? 1 : 0));
        if (got != (size_t)sz || (size_t)sz < data + 1 || pac[sz - 1] != (uint8_t)(l_pac % 4)) {
            set_error("%s.pac does not hold %lld bases (a forward-strand .pac as `bwa index` leaves it is required)", prefix.c_str(), l_pac);
            return ARACHNE_ERR_IO;
        }
    }
    DeviceBackend be;
    int rc = ARACHNE_OK;
    uint32_t* payload = nullptr; uint64_t* sa = nullptr;
    BuildStats st;
    double t1 = t0, t2 = t0;
    if (!be.init(device)) rc = ARACHNE_ERR_CUDA;
    if (rc == ARACHNE_OK) {
        const size_t pacb = (size_t)(l_pac / 4 + 1);
        uint8_t* d_pac = (uint8_t*)be.acquire(pacb + 8);
        if (d_pac) { be.zero(d_pac, pacb + 8); be.to_device(d_pac, pac.data(), pacb < pac.size() ? pacb : pac.size()); }
        t1 = now_s();
        if (!d_pac || !build_fm_index(be, d_pac, (uint64_t)l_pac, (uint64_t)(batch_cap > 0 ? batch_cap : 0), &payload, &sa, st)) rc = ARACHNE_ERR_CUDA;
        if (rc == ARACHNE_OK) cudaStreamSynchronize(be.st);
        t2 = now_s();
    }
    if (rc == ARACHNE_OK) {
        const uint64_t n = st.n, n_sa = (n + SA_INTV) / SA_INTV, sa_intv = SA_INTV;
        FILE* fb = fopen((prefix + ".bwt").c_str(), "wb");
        FILE* fs = fopen((prefix + ".sa").c_str(), "wb");
        bool okay = fb && fs;
        if (okay) {  // bwt_dump_bwt / bwt_dump_sa, bwt.c:385-409
            okay = fwrite(&st.primary, 8, 1, fb) == 1 && fwrite(st.L2 + 1, 8, 4, fb) == 4 && dump_device(be, fb, payload, bwt_payload_words(n) * 4);
            okay = okay && fwrite(&st.primary, 8, 1, fs) == 1 && fwrite(st.L2 + 1, 8, 4, fs) == 4 && fwrite(&sa_intv, 8, 1, fs) == 1 && fwrite(&n, 8, 1, fs) == 1 &&
                   dump_device(be, fs, sa + 1, (n_sa - 1) * 8);
        }
        if (fb) okay = (fclose(fb) == 0) && okay;
        if (fs) okay = (fclose(fs) == 0) && okay;
        if (!okay) { if (be.ok()) set_error("cannot write %s.bwt / .sa", prefix.c_str()); rc = be.ok() ? ARACHNE_ERR_IO : ARACHNE_ERR_CUDA; }
    }
    if (info) {
        memset(info, 0, sizeof(*info));
        info->seq_len = (int64_t)st.n; info->primary = (int64_t)st.primary;
        info->n_batches = (int64_t)st.n_batches; info->max_batch = (int64_t)st.max_batch;
        info->tied_suffixes = (int64_t)st.tied_round1; info->max_rounds = (int64_t)st.max_rounds;
        info->device_peak_bytes = (int64_t)be.peak_bytes;
        info->seconds_read = t1 - t0; info->seconds_build = t2 - t1; info->seconds_write = now_s() - t2;
    }
    be.shutdown();
    return rc;
}

// ---------------------------------------------------------------------------------------------- pack step of `bwa index`
// bns_fasta2bntseq(for_only = 1) + bns_dump (bwa/bntseq.c:229-328, 66-96): FASTA (plain or gzip) -> <prefix>.pac / .ann / .amb, byte for byte
// what the reference leaves after `bwa index` -- including its replacement of every ambiguous base by lrand48() & 3 under srand48(11)
// (bntseq.c:261,290-291; the 48-bit LCG is restated here, not taken from libc, so the files do not depend on the platform).  Records are read
// with kseq.h's rules (bwa/kseq.h:176-215): name = text up to the first whitespace, comment = rest of the header line, sequence = every
// following line without its line end until a line starts with '>' or '@'.  Host only.
#include <zlib.h>
extern "C" int arachne_index_pack(const char* fasta, const char* prefix_, int64_t* l_pac_out) {
    using arachne::set_error;
    if (!fasta || !prefix_) { set_error("arachne_index_pack: null argument"); return ARACHNE_ERR_ARG; }
    gzFile f = gzopen(fasta, "rb");
    if (!f) { set_error("cannot open %s", fasta); return ARACHNE_ERR_IO; }
    gzbuffer(f, 1 << 20);
    const std::string prefix(prefix_);
    struct Ann { std::string name, anno; int64_t offset; int32_t len, n_ambs; };
    struct Hole { int64_t offset; int32_t len; char amb; };
    std::vector<Ann> anns; std::vector<Hole> holes; std::vector<uint8_t> pac;
    int64_t l_pac = 0;
    uint64_t rng = ((uint64_t)11 << 16) | 0x330e;  // srand48(11)
    uint8_t nt4[256]; memset(nt4, 4, 256); nt4[(int)'-'] = 5;
    for (int i = 0; i < 4; ++i) { nt4[(int)"ACGT"[i]] = (uint8_t)i; nt4[(int)"acgt"[i]] = (uint8_t)i; }
    // kseq.h's reader, character for character (kseq.h:95-141 ks_getuntil2, :176-215 kseq_read): which bytes end up in a record's sequence
    // depends on details such as "a trailing CR is dropped when the ACCUMULATED string is longer than one byte" and "only the first byte
    // of a line is tested for '>', '+', '@'", and every byte that gets in is a base of the packed reference.
    struct KStream {
        gzFile f; std::vector<char> buf; int beg = 0, end = 0; bool at_eof = false, io_error = false;
        bool refill() {
            if (at_eof) return false;
            const int got = gzread(f, buf.data(), (unsigned)buf.size());
            if (got < 0) io_error = true;
            beg = 0; end = got > 0 ? got : 0;
            if (end == 0) { at_eof = true; return false; }
            return true;
        }
        int getc() { if (beg >= end && !refill()) return -1; return (unsigned char)buf[beg++]; }
        // Appends to str up to (not including) the next delimiter -- '\n' when `line`, any isspace byte otherwise -- and consumes the
        // delimiter (*dret receives it).  False when nothing was left to read.
        bool until(bool line, std::string& str, int* dret) {
            bool gotany = false;
            for (;;) {
                if (beg >= end && !refill()) break;
                int i = beg;
                if (line) { const void* nl = memchr(buf.data() + beg, '\n', (size_t)(end - beg)); i = nl ? (int)(static_cast<const char*>(nl) - buf.data()) : end; }
                else while (i < end && !isspace((unsigned char)buf[i])) ++i;
                gotany = true;
                str.append(buf.data() + beg, (size_t)(i - beg));
                beg = i + 1;
                if (i < end) { if (dret) *dret = (unsigned char)buf[i]; break; }
            }
            if (!gotany && at_eof) return false;
            if (line && str.size() > 1 && str.back() == '\r') str.pop_back();
            return true;
        }
    } ks;
    ks.f = f; ks.buf.resize(4 << 20);
    auto add_record = [&](const std::string& name, const std::string& comment, const std::string& seq) {  // add1, bntseq.c:227-273
        Ann a; a.name = name; a.anno = comment.empty() ? std::string("(null)") : comment;
        a.offset = anns.empty() ? 0 : anns.back().offset + anns.back().len; a.len = 0; a.n_ambs = 0;
        anns.push_back(a);
        Ann& p = anns.back();
        int lasts = 0;
        for (size_t k = 0; k < seq.size(); ++k) {
            const int ch = (unsigned char)seq[k];
            int c = nt4[ch];
            if (c >= 4) {
                if (lasts == ch && !holes.empty()) ++holes.back().len;
                else { holes.push_back({p.offset + p.len, 1, (char)ch}); ++p.n_ambs; }
                rng = (0x5DEECE66Dull * rng + 0xB) & ((1ull << 48) - 1);  // lrand48()
                c = (int)((rng >> 17) & 3);
            }
            lasts = ch;
            if ((l_pac & 3) == 0) pac.push_back(0);
            pac[(size_t)(l_pac >> 2)] |= (uint8_t)(c << ((~l_pac & 3) << 1));
            ++l_pac; ++p.len;
        }
    };
    int last_char = 0, c = 0;
    std::string name, comment, seq, qual;
    for (;;) {  // one kseq_read per trip
        if (last_char == 0) {
            while ((c = ks.getc()) != -1 && c != '>' && c != '@') {}
            if (c == -1) break;
            last_char = c;
        }
        name.clear(); comment.clear(); seq.clear(); qual.clear();
        if (!ks.until(false, name, &c)) break;
        if (c != '\n') ks.until(true, comment, nullptr);
        while ((c = ks.getc()) != -1 && c != '>' && c != '+' && c != '@') {
            if (c == '\n') continue;
            seq.push_back((char)c);
            ks.until(true, seq, nullptr);
        }
        if (c == '>' || c == '@') last_char = c;
        if (c == '+') {  // a FASTQ record: its quality string must have the sequence's length, else the reference stops reading here
            while ((c = ks.getc()) != -1 && c != '\n') {}
            if (c == -1) break;
            while (ks.until(true, qual, nullptr) && qual.size() < seq.size()) {}
            last_char = 0;
            if (qual.size() != seq.size()) break;
        }
        add_record(name, comment, seq);
        if (c == -1) break;
    }
    if (ks.io_error) { gzclose(f); set_error("read error in %s", fasta); return ARACHNE_ERR_IO; }
    gzclose(f);
    if (anns.empty()) { set_error("%s holds no FASTA record", fasta); return ARACHNE_ERR_IO; }
    FILE* fp = fopen((prefix + ".pac").c_str(), "wb");
    if (!fp) { set_error("cannot write %s.pac", prefix.c_str()); return ARACHNE_ERR_IO; }
    fwrite(pac.data(), 1, (size_t)((l_pac >> 2) + ((l_pac & 3) == 0 ? 0 : 1)), fp);
    if (l_pac % 4 == 0) { const uint8_t z = 0; fwrite(&z, 1, 1, fp); }
    const uint8_t ct = (uint8_t)(l_pac % 4); fwrite(&ct, 1, 1, fp);
    fclose(fp);
    fp = fopen((prefix + ".ann").c_str(), "w");
    if (!fp) { set_error("cannot write %s.ann", prefix.c_str()); return ARACHNE_ERR_IO; }
    fprintf(fp, "%lld %d %u\n", (long long)l_pac, (int)anns.size(), 11u);
    for (const Ann& a : anns) { fprintf(fp, "%d %s", 0, a.name.c_str()); if (!a.anno.empty()) fprintf(fp, " %s\n", a.anno.c_str()); else fprintf(fp, "\n"); fprintf(fp, "%lld %d %d\n", (long long)a.offset, a.len, a.n_ambs); }
    fclose(fp);
    fp = fopen((prefix + ".amb").c_str(), "w");
    if (!fp) { set_error("cannot write %s.amb", prefix.c_str()); return ARACHNE_ERR_IO; }
    fprintf(fp, "%lld %d %u\n", (long long)l_pac, (int)anns.size(), (unsigned)holes.size());
    for (const Hole& h : holes) fprintf(fp, "%lld %d %c\n", (long long)h.offset, h.len, h.amb);
    fclose(fp);
    if (l_pac_out) *l_pac_out = l_pac;
    return ARACHNE_OK;
}
