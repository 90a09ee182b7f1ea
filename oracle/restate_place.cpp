// TEST INFRASTRUCTURE ONLY (see restate_place.h).  CPU restatement of the reference's Go placement code:
//   GetChains / GetAlignments          src/aligner/aligner.go:1633-1715, 1484-1631
//   scoreAlignment, isPair             :556-581, 1032-1063
//   tagBestAlignments                  :1397-1481
//   inferMolecules .. scrapMolecules   :1300-1393, 991-1016, 527-545
//   optimizer.Optimize / GenerateMove  src/optimizer/optimizer.go:15-27, aligner.go:1065-1097
//   fastScore / acceptMove             :1109-1237, 1261-1298
//   estimateMapQualities (+helpers)    :643-755, 797-922, 978-989
//   OrderedMap semantics               src/aligner/ordered_map.go, ordered_alignment_map.go
// Written to mirror the Go data structures one to one (ordered maps, pointer graph) so that it can be read against the Go
// text; the CUDA stage (csrc/place.cuh) is organised completely differently and is checked against this file.
#include "restate_place.h"
#include "gorand_table.h"
#include <cfloat>
#include <cmath>
#include <cstring>
#include <algorithm>
#include <string>
#include <unordered_map>
#include <vector>
#include <set>
#include <tuple>

namespace {

// ------------------------------------------------------------------------------------------------ MD5 (RFC 1321)
struct Md5 {
    uint32_t a0 = 0x67452301, b0 = 0xefcdab89, c0 = 0x98badcfe, d0 = 0x10325476;
    static uint32_t rol(uint32_t x, int c) { return (x << c) | (x >> (32 - c)); }
    void block(const uint8_t* p) {
        static const int s[64] = {7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 7, 12, 17, 22, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20, 5, 9, 14, 20,
                                  4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 4, 11, 16, 23, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21, 6, 10, 15, 21};
        static uint32_t K[64];
        static bool init = false;
        if (!init) { for (int i = 0; i < 64; ++i) K[i] = (uint32_t)(long long)std::floor(std::fabs(std::sin((double)(i + 1))) * 4294967296.0); init = true; }
        uint32_t M[16];
        for (int i = 0; i < 16; ++i) M[i] = (uint32_t)p[4 * i] | (uint32_t)p[4 * i + 1] << 8 | (uint32_t)p[4 * i + 2] << 16 | (uint32_t)p[4 * i + 3] << 24;
        uint32_t A = a0, B = b0, C = c0, D = d0;
        for (int i = 0; i < 64; ++i) {
            uint32_t F; int g;
            if (i < 16) { F = (B & C) | (~B & D); g = i; }
            else if (i < 32) { F = (D & B) | (~D & C); g = (5 * i + 1) & 15; }
            else if (i < 48) { F = B ^ C ^ D; g = (3 * i + 5) & 15; }
            else { F = C ^ (B | ~D); g = (7 * i) & 15; }
            F = F + A + K[i] + M[g];
            A = D; D = C; C = B; B = B + rol(F, s[i]);
        }
        a0 += A; b0 += B; c0 += C; d0 += D;
    }
    void digest(const uint8_t* msg, int64_t len, uint8_t out[16]) {
        int64_t i = 0;
        for (; i + 64 <= len; i += 64) block(msg + i);
        uint8_t tail[128]; int64_t r = len - i;
        memset(tail, 0, sizeof(tail)); memcpy(tail, msg + i, (size_t)r); tail[r] = 0x80;
        int tl = r < 56 ? 64 : 128;
        uint64_t bits = (uint64_t)len * 8;
        for (int k = 0; k < 8; ++k) tail[tl - 8 + k] = (uint8_t)(bits >> (8 * k));
        block(tail); if (tl == 128) block(tail + 64);
        uint32_t v[4] = {a0, b0, c0, d0};
        for (int k = 0; k < 16; ++k) out[k] = (uint8_t)(v[k >> 2] >> (8 * (k & 3)));
    }
};

// ------------------------------------------------------------------------------------------------ Go math/rand (rng.go)
const uint64_t kCooked[607] = GORAND_TABLE_INIT;

struct GoRand {
    uint64_t vec[607];
    int tap, feed;
    static int32_t seedrand(int32_t x) {  // x[n+1] = 48271 * x[n] mod (2**31 - 1)
        int32_t hi = x / 44488, lo = x % 44488;
        x = 48271 * lo - 3399 * hi;
        if (x < 0) x += 2147483647;
        return x;
    }
    explicit GoRand(int64_t seed) {
        tap = 0; feed = 607 - 273;
        seed = seed % 2147483647;
        if (seed < 0) seed += 2147483647;
        if (seed == 0) seed = 89482311;
        int32_t x = (int32_t)seed;
        for (int i = -20; i < 607; ++i) {
            x = seedrand(x);
            if (i >= 0) {
                uint64_t u = (uint64_t)x << 40;
                x = seedrand(x); u ^= (uint64_t)x << 20;
                x = seedrand(x); u ^= (uint64_t)x;
                vec[i] = u ^ kCooked[i];
            }
        }
    }
    uint64_t Uint64() {
        if (--tap < 0) tap += 607;
        if (--feed < 0) feed += 607;
        uint64_t x = vec[feed] + vec[tap];
        vec[feed] = x;
        return x;
    }
    double Float64() {
        for (;;) {
            double f = (double)(int64_t)(Uint64() & 0x7fffffffffffffffull) / 9223372036854775808.0;
            if (f != 1.0) return f;
        }
    }
};

// ------------------------------------------------------------------------------------------------ ordered maps (ordered_map.go)
template <class V>
struct OrderedMap {
    std::unordered_map<int, int> index;
    std::vector<int> reverse_index;
    std::vector<V> store;
    bool Has(int key) const { return index.count(key) != 0; }
    V Get(int key) const { auto it = index.find(key); return it == index.end() ? V() : store[it->second]; }
    void Set(int key, V val) {
        auto it = index.find(key);
        if (it != index.end()) store[it->second] = val;
        else { index[key] = (int)store.size(); reverse_index.push_back(key); store.push_back(val); }
    }
    void Delete(int key) {  // swap the last element into the hole (ordered_map.go:39-51)
        auto it = index.find(key);
        if (it == index.end()) return;
        int i = it->second;
        if (store.size() > 1) {
            store[i] = store.back();
            index[reverse_index.back()] = i;
            reverse_index[i] = reverse_index.back();
        }
        store.pop_back(); reverse_index.pop_back();
        index.erase(key);
    }
    int Len() const { return (int)reverse_index.size(); }
};

struct Alignment {  // aligner.go:65-114 (fields that influence results)
    int cand = -1;  // batch-global output slot
    int id = 0;
    bool read1 = false, is_proper = false;
    int soft_clipped = 0, soft_clipped_length = 0;
    int read_len = 0;
    int mapq = 0;
    double molecule_difference = 0;
    int contig = -1;  // rid; -1 == "" (placeholder)
    int64_t pos = -1, aend = 0;
    int score = 0, mismatches = 0, matches = 0, indels = 0;
    int read_id = 0, mate_id = 0;
    Alignment* mate_alignment = nullptr;
    bool reversed = false;
    int molecule_id = -1;
    bool active = false, bwa_pick = false;
    double log_alignment_probability = 0;
    double sum_move_probability_change = 1.0;
    double molecule_confidence = 0.00075 * 0.025;
    bool active_molecule = false;
    // MapQData
    Alignment* second_best = nullptr;
    double second_best_score = 0, as_score = 0;
    double raw_mapq = 0;
    int reg = -1;
    // record assembly (markDuplicates, CheckSplitReads, AppendBam)
    int readmap_s = 0, readmap_e = 0;
    bool duplicate = false;
    Alignment* secondary = nullptr;  // the split alignment chosen for this (active) alignment
    Alignment* primary = nullptr;    // set on a split alignment: the active alignment it belongs to
    bool mapq_data_replaced = false; // split.mapq_data = &MapQData{second_best_score, score} (split.go:157)
    std::vector<int> mmLocs, mmReadLocs;
    int n_cigar = 0; const uint32_t* cigar = nullptr;  // mem_reg2aln's CIGAR, len<<4|op
};

struct CandidateMolecule {  // aligner.go:178-193
    int id = 0;
    int chrom = -1;
    int64_t start = 0, stop = 0;
    OrderedMap<OrderedMap<Alignment*>*> alignments;
    OrderedMap<Alignment*> best_alignment_for_read, active_alignments;
    bool active_molecule = false;
    double molecule_confidence = 1.0, differences = 0;
    int soft_clipped = 0;
    std::vector<OrderedMap<Alignment*>*> owned;
    ~CandidateMolecule() { for (auto* p : owned) delete p; }
};

double g_improper = -4.0;  // *improper_pair_penalty (set per orc_place call; the oracle is single threaded)

bool isPair(const Alignment* read1, const Alignment* read2) {  // aligner.go:1032-1063
    if (read1->reversed == read2->reversed || read1->contig != read2->contig) return false;
    const Alignment *forward, *reverse;
    if (read1->reversed) { forward = read2; reverse = read1; } else { forward = read1; reverse = read2; }
    int64_t dist = reverse->pos - forward->pos;
    return dist >= -35 && dist < 750;
}

double scoreAlignment(const Alignment* aln, const Alignment* mate, double log_molecule_penalty) {  // aligner.go:556-581
    double score = 0.0;
    if (aln) {
        score += (double)(aln->mismatches * -2 + aln->indels * -3);
        if (aln->soft_clipped > 0) { score -= 5.0 * (double)aln->soft_clipped; score -= (double)aln->soft_clipped_length * 0.5; }
    }
    if (mate) {
        score += (double)(mate->mismatches * -2 + mate->indels * -3);
        if (mate->soft_clipped > 0) { score -= 5.0 * (double)mate->soft_clipped; score -= (double)mate->soft_clipped_length * 0.5; }
    }
    if (!mate || !aln || !isPair(aln, mate)) score += g_improper;
    if (aln && !aln->active_molecule) score += log_molecule_penalty;
    return score;
}

double psuedoCountAlignmentScore(const Alignment* aln, double log_molecule_penalty) {  // aligner.go:547-554
    double score = 0.0;
    score -= 10.0;
    score -= ((double)aln->read_len - 25.0) * 0.5;
    score += log_molecule_penalty;
    return score;
}

// ------------------------------------------------------------------------------------------------ reference access (GetSeq)
inline int pac_base(const uint8_t* pac, int64_t l) { return pac[l >> 2] >> ((~l & 3) << 1) & 3; }  // bntseq.c:225

}  // namespace

namespace {

struct Reg {  // mem_alnreg_t, bwa/bwamem.h:86-104
    int64_t rb, re; int qb, qe, rid, score, truesc, sub, alt_sc, csub, sub_n, w, seedcov, secondary, secondary_all, seedlen0, n_comp_is_alt;
    float frac_rep; uint64_t hash;
};
static_assert(sizeof(Reg) == 88, "mem_alnreg_t");

// InterpretAlign (gobwa.go:339-371): Offset / Alignment_end of a region
void interpret_align(const orc_ref_t* ref, const Reg& r, int64_t* offset, int64_t* aend) {
    const int64_t coff = ref->ctg_off[r.rid], l_pac = ref->l_pac;
    *offset = r.rb < l_pac ? r.rb - coff : l_pac * 2 - 1 - r.rb - coff;
    *aend = r.re < l_pac ? r.re - coff : l_pac * 2 - 1 - r.re - coff;
}

// GoBwaReference.GetSeq (gobwa.go:50-80) for an interval inside one contig: bases as codes 0..3 in read orientation
void get_seq(const orc_ref_t* ref, int rid, int64_t start, int64_t end, bool reversed, std::vector<uint8_t>& out) {
    const int64_t off = ref->ctg_off[rid];
    int64_t n = end - start;
    out.assign((size_t)std::max<int64_t>(n, 0), 0);
    // bns_fetch_seq clamps [beg,end) to the contig holding the midpoint (bntseq.c:421-446)
    int64_t b = start + off, e = end + off;
    const int64_t far_beg = off, far_end = off + ref->ctg_len[rid];
    if (b < far_beg) b = far_beg;
    if (e > far_end) e = far_end;
    const int64_t m = e - b;
    for (int64_t i = 0; i < m; ++i) {
        int c = pac_base(ref->pac, b + i);
        if (reversed) { if (m - i - 1 < n) out[(size_t)(m - i - 1)] = (uint8_t)(3 - c); }
        else if (i < n) out[(size_t)i] = (uint8_t)c;
    }
}

struct Walk { int matches, mismatches, indels, soft_clipped, soft_clipped_length; std::vector<int> mmLocs, mmReadLocs; };

// the CIGAR / reference walk of GetAlignments (aligner.go:1505-1569)
void walk_alignment(const orc_ref_t* ref, int l_query, const uint8_t* readSeq, int64_t chain_pos, int64_t chain_aend, int rid, bool reversed,
                    int edit_distance, int n_cigar, const uint32_t* raw_cigar, Walk* w) {
    int matches = 0, indels = 0, indel_length = 0, soft_clipping = 0, soft_clipping_length = 0;
    int64_t refStart = chain_pos, refEnd = chain_aend;
    if (reversed) { refStart = chain_aend + 1; refEnd = chain_pos + 1; }
    std::vector<uint8_t> refSeq;
    get_seq(ref, rid, refStart, refEnd, reversed, refSeq);
    int refSeqOffset = 0, readOffset = 0;
    w->mmLocs.clear(); w->mmReadLocs.clear();
    int k = reversed ? n_cigar - 1 : 0, inc = reversed ? -1 : 1;
    for (; k < n_cigar && k >= 0; k += inc) {
        int op = raw_cigar[k] & 0xf, len = (int)(raw_cigar[k] >> 4);
        if (op == 0) {
            matches += len;
            for (int m = 0; m < len; ++m) {
                if (refSeqOffset + m >= (int)refSeq.size()) continue;
                if (readOffset + m < l_query && refSeq[refSeqOffset + m] != readSeq[readOffset + m]) {
                    if (reversed) w->mmLocs.push_back((int)refEnd - (refSeqOffset + m));
                    else w->mmLocs.push_back(refSeqOffset + (int)refStart + m);
                    w->mmReadLocs.push_back(readOffset + m);
                }
            }
            refSeqOffset += len; readOffset += len;
        } else if (op == 1) { indels += 1; indel_length += len; readOffset += len; }
        else if (op == 2) { indels += 1; indel_length += len; refSeqOffset += len; }
        else if (op == 3) { soft_clipping += 1; soft_clipping_length += len; readOffset += len; }
    }
    int mismatches = edit_distance - indel_length;
    matches -= mismatches;
    if (mismatches < 0) mismatches = 0;
    w->matches = matches; w->mismatches = mismatches; w->indels = indels; w->soft_clipped = soft_clipping; w->soft_clipped_length = soft_clipping_length;
}

// ------------------------------------------------------------------------------------------------ sort.Sort as Go 1.19+ implements it
// Pattern-defeating quicksort of package sort (zsortinterface.go: pdqsort, insertionSort, heapSort, partition, partitionEqual,
// partialInsertionSort, breakPatterns, choosePivot, order2, median, medianAdjacent, reverseRange), restated from the published algorithm.
// Less(i, j) = data[i].pos < data[j].pos (ByPosition, aligner.go:203-206).  Not executable-pinned: there is no Go toolchain here.
int g_tie_mode = 0;
namespace gosort {
typedef std::vector<Alignment*> V;
inline bool less(const V& d, int i, int j) { return d[i]->pos < d[j]->pos; }
inline void swp(V& d, int i, int j) { std::swap(d[i], d[j]); }
void insertionSort(V& d, int a, int b) { for (int i = a + 1; i < b; ++i) for (int j = i; j > a && less(d, j, j - 1); --j) swp(d, j, j - 1); }
void siftDown(V& d, int lo, int hi, int first) {
    int root = lo;
    for (;;) {
        int child = 2 * root + 1;
        if (child >= hi) return;
        if (child + 1 < hi && less(d, first + child, first + child + 1)) ++child;
        if (!less(d, first + root, first + child)) return;
        swp(d, first + root, first + child);
        root = child;
    }
}
void heapSort(V& d, int a, int b) {
    int first = a, lo = 0, hi = b - a;
    for (int i = (hi - 1) / 2; i >= 0; --i) siftDown(d, i, hi, first);
    for (int i = hi - 1; i >= 0; --i) { swp(d, first, first + i); siftDown(d, lo, i, first); }
}
int partition(V& d, int a, int b, int pivot, bool* already) {
    swp(d, a, pivot);
    int i = a + 1, j = b - 1;
    while (i <= j && less(d, i, a)) ++i;
    while (i <= j && !less(d, j, a)) --j;
    if (i > j) { swp(d, j, a); *already = true; return j; }
    swp(d, i, j); ++i; --j;
    for (;;) {
        while (i <= j && less(d, i, a)) ++i;
        while (i <= j && !less(d, j, a)) --j;
        if (i > j) break;
        swp(d, i, j); ++i; --j;
    }
    swp(d, j, a);
    *already = false;
    return j;
}
int partitionEqual(V& d, int a, int b, int pivot) {
    swp(d, a, pivot);
    int i = a + 1, j = b - 1;
    for (;;) {
        while (i <= j && !less(d, a, i)) ++i;
        while (i <= j && less(d, a, j)) --j;
        if (i > j) break;
        swp(d, i, j); ++i; --j;
    }
    return i;
}
bool partialInsertionSort(V& d, int a, int b) {
    const int maxSteps = 5, shortestShifting = 50;
    int i = a + 1;
    for (int j = 0; j < maxSteps; ++j) {
        while (i < b && !less(d, i, i - 1)) ++i;
        if (i == b) return true;
        if (b - a < shortestShifting) return false;
        swp(d, i, i - 1);
        if (i - a >= 2) for (int k = i - 1; k >= 1; --k) { if (!less(d, k, k - 1)) break; swp(d, k, k - 1); }
        if (b - i >= 2) for (int k = i + 1; k < b; ++k) { if (!less(d, k, k - 1)) break; swp(d, k, k - 1); }
    }
    return false;
}
inline int bitsLen(unsigned x) { int n = 0; while (x) { ++n; x >>= 1; } return n; }
void breakPatterns(V& d, int a, int b) {
    int length = b - a;
    if (length >= 8) {
        uint64_t random = (uint64_t)length;
        unsigned modulus = 1u << bitsLen((unsigned)length);
        int idx = a + (length / 4) * 2 - 1;
        for (int i = 0; i < 3; ++i) {
            random ^= random << 13; random ^= random >> 7; random ^= random << 17;
            int other = (int)((unsigned)random & (modulus - 1));
            if (other >= length) other -= length;
            swp(d, idx - 1 + i, a + other);
        }
    }
}
inline void order2(V& d, int& a, int& b, int* swaps) { if (less(d, b, a)) { ++*swaps; std::swap(a, b); } }
inline int median(V& d, int a, int b, int c, int* swaps) { order2(d, a, b, swaps); order2(d, b, c, swaps); order2(d, a, b, swaps); return b; }
inline int medianAdjacent(V& d, int a, int* swaps) { return median(d, a - 1, a, a + 1, swaps); }
enum { unknownHint, increasingHint, decreasingHint };
int choosePivot(V& d, int a, int b, int* hint) {
    const int shortestNinther = 50, maxSwaps = 4 * 3;
    int l = b - a, swaps = 0;
    int i = a + l / 4 * 1, j = a + l / 4 * 2, k = a + l / 4 * 3;
    if (l >= 8) {
        if (l >= shortestNinther) { i = medianAdjacent(d, i, &swaps); j = medianAdjacent(d, j, &swaps); k = medianAdjacent(d, k, &swaps); }
        j = median(d, i, j, k, &swaps);
    }
    if (swaps == 0) *hint = increasingHint; else if (swaps == maxSwaps) *hint = decreasingHint; else *hint = unknownHint;
    return j;
}
void reverseRange(V& d, int a, int b) { int i = a, j = b - 1; while (i < j) { swp(d, i, j); ++i; --j; } }
void pdqsort(V& d, int a, int b, int limit) {
    const int maxInsertion = 12;
    bool wasBalanced = true, wasPartitioned = true;
    for (;;) {
        int length = b - a;
        if (length <= maxInsertion) { insertionSort(d, a, b); return; }
        if (limit == 0) { heapSort(d, a, b); return; }
        if (!wasBalanced) { breakPatterns(d, a, b); --limit; }
        int hint, pivot = choosePivot(d, a, b, &hint);
        if (hint == decreasingHint) { reverseRange(d, a, b); pivot = (b - 1) - (pivot - a); hint = increasingHint; }
        if (wasBalanced && wasPartitioned && hint == increasingHint) { if (partialInsertionSort(d, a, b)) return; }
        if (a > 0 && !less(d, a - 1, pivot)) { a = partitionEqual(d, a, b, pivot); continue; }
        bool already;
        int mid = partition(d, a, b, pivot, &already);
        wasPartitioned = already;
        int leftLen = mid - a, rightLen = b - mid, balanceThreshold = length / 8;
        if (leftLen < rightLen) { wasBalanced = leftLen >= balanceThreshold; pdqsort(d, a, mid, limit); a = mid + 1; }
        else { wasBalanced = rightLen >= balanceThreshold; pdqsort(d, mid + 1, b, limit); b = mid; }
    }
}
void sort(V& d) { int n = (int)d.size(); pdqsort(d, 0, n, bitsLen((unsigned)n)); }
}  // namespace gosort

// ------------------------------------------------------------------------------------------------ tagBestAlignments (aligner.go:1397-1481)
std::vector<std::vector<Alignment*>> tagBestAlignments(std::vector<std::vector<Alignment*>>& alignments, const int64_t* rand_seed) {
    std::vector<std::vector<Alignment*>> positions;
    std::unordered_map<int, int> contigs;
    std::vector<bool> touched(alignments.size(), false);
    for (size_t read_id = 0; read_id < alignments.size(); ++read_id) {
        auto& alignmentArray = alignments[read_id];
        bool was_touched = touched[read_id];
        double bestScore = -DBL_MAX;
        Alignment *bestAlignment = nullptr, *bestMate = nullptr;
        int64_t seed = 1;
        if (!alignmentArray.empty()) seed = rand_seed[read_id >> 1];  // md5(read_name)[0:8]; R1 and R2 share ReadInfo
        GoRand random(seed);
        for (Alignment* alignment : alignmentArray) {
            auto& mateAlignments = alignments[alignment->mate_id];
            for (Alignment* mateAlignment : mateAlignments) {
                double totalScore = scoreAlignment(alignment, mateAlignment, 0.0) + (random.Float64() / 2.0);
                if (totalScore > bestScore) { bestScore = totalScore; bestAlignment = alignment; bestMate = mateAlignment; }
            }
            if (mateAlignments.empty()) {
                double score = (double)alignment->score + random.Float64() / 2.0;
                if (score > bestScore) { bestScore = score; bestAlignment = alignment; }
            }
            auto it = contigs.find(alignment->contig);
            if (it != contigs.end()) positions[it->second].push_back(alignment);
            else { contigs[alignment->contig] = (int)positions.size(); positions.push_back({alignment}); }
        }
        if (!was_touched && bestAlignment) {
            bestAlignment->active = true; bestAlignment->bwa_pick = true;
            if (bestMate) {
                if (isPair(bestAlignment, bestMate)) { bestAlignment->is_proper = true; bestMate->is_proper = true; }
                bestMate->active = true; bestMate->bwa_pick = true;
                touched[bestMate->read_id] = true;
            }
        }
    }
    // sort.Sort(ByPosition) (aligner.go:1477-1479): Go's pdqsort, unstable on equal pos -- UNPINNED against Go execution.  g_tie_mode picks
    // what equal positions do here: 0 = keep insertion order (what the device does), 1 = reverse it, 2 = gosort::pdqsort below (a restatement of
    // Go 1.24's sort.Sort from its published algorithm, deterministic: its only randomness is seeded by the slice length).
    // tests/test_cpu_tier.py::test_place_tie_order_is_not_observable compares the three.
    for (auto& p : positions) {
        if (g_tie_mode == 2) { gosort::sort(p); continue; }
        std::stable_sort(p.begin(), p.end(), [](const Alignment* a, const Alignment* b) { return a->pos < b->pos; });
        if (g_tie_mode == 1)
            for (size_t i = 0; i < p.size();) { size_t j = i; while (j < p.size() && p[j]->pos == p[i]->pos) ++j; std::reverse(p.begin() + i, p.begin() + j); i = j; }
    }
    return positions;
}

// ------------------------------------------------------------------------------------------------ molecules
std::vector<CandidateMolecule*> inferMolecules(std::vector<std::vector<Alignment*>>& positions) {  // aligner.go:1300-1338
    std::vector<CandidateMolecule*> toReturn;
    int molecule_num = 0;
    CandidateMolecule* cur = nullptr;
    auto add = [](CandidateMolecule* m, Alignment* a) {
        OrderedMap<Alignment*>* am = m->alignments.Get(a->read_id);
        if (am) am->Set(a->id, a);
        else { am = new OrderedMap<Alignment*>(); m->owned.push_back(am); am->Set(a->id, a); m->alignments.Set(a->read_id, am); }
    };
    for (auto& list : positions) {
        for (size_t i = 0; i < list.size(); ++i) {
            if (i == 0 || list[i]->pos - list[i - 1]->pos > 50000) {
                if (i > 0) cur->stop = list[i - 1]->pos;
                cur = new CandidateMolecule();
                cur->chrom = list[i]->contig; cur->start = list[i]->pos; cur->id = molecule_num;
                add(cur, list[i]);
                toReturn.push_back(cur);
                ++molecule_num;
            }
            add(cur, list[i]);
        }
        if (!list.empty()) cur->stop = list.back()->pos;
    }
    return toReturn;
}

void markBestAlignmentForReadInMolecule(std::vector<CandidateMolecule*>& molecules) {  // aligner.go:1340-1393
    for (CandidateMolecule* molecule : molecules) {
        OrderedMap<Alignment*> active_alignments, best_alignment_for_read;
        for (int read_id : molecule->alignments.reverse_index) {
            OrderedMap<Alignment*>* alns = molecule->alignments.Get(read_id);
            double best_score = -DBL_MAX;
            Alignment* best_alignment = nullptr;
            for (int alignment_id : alns->reverse_index) {
                Alignment* alignment = alns->Get(alignment_id);
                OrderedMap<Alignment*>* mate_alignments = molecule->alignments.Get(alignment->mate_id);
                if (mate_alignments && mate_alignments->Len() > 0) {
                    for (int mid : mate_alignments->reverse_index) {
                        Alignment* mate_alignment = mate_alignments->Get(mid);
                        double score = scoreAlignment(alignment, mate_alignment, 0.0);
                        if (score > best_score) { best_score = score; best_alignment = alignment; }
                    }
                } else if (alignment->log_alignment_probability > best_score) {
                    best_score = alignment->log_alignment_probability; best_alignment = alignment;
                }
                if (alignment->active) active_alignments.Set(read_id, alignment);
            }
            if (best_alignment->active) active_alignments.Set(read_id, best_alignment);
            best_alignment_for_read.Set(read_id, best_alignment);
        }
        molecule->active_alignments = active_alignments;
        molecule->best_alignment_for_read = best_alignment_for_read;
    }
}

std::vector<CandidateMolecule*> scrapMolecules(std::vector<CandidateMolecule*>& mols, std::vector<CandidateMolecule*>& dropped) {  // aligner.go:991-1016
    std::vector<CandidateMolecule*> toReturn;
    int count = 0;
    for (CandidateMolecule* m : mols) {
        const bool keep = m->active_alignments.Len() > 0;
        for (int read_id : m->alignments.reverse_index) {
            OrderedMap<Alignment*>* am = m->alignments.Get(read_id);
            for (int aid : am->reverse_index) am->Get(aid)->molecule_id = keep ? count : -1;
        }
        if (keep) { toReturn.push_back(m); ++count; } else dropped.push_back(m);
    }
    return toReturn;
}

void setMoleculeDifferences(std::vector<CandidateMolecule*>& mols) {  // aligner.go:527-545 (setBad == false at every call site)
    for (CandidateMolecule* m : mols) {
        int differences = 0;
        for (Alignment* a : m->active_alignments.store) differences += a->mismatches;
        m->differences = (double)differences / (double)m->active_alignments.Len();
        for (Alignment* a : m->active_alignments.store) a->molecule_difference = m->differences;
    }
}

}  // namespace

namespace {

struct Move {  // aligner.go:1099-1107
    double score_change = 0, alignment_change = 0;
    CandidateMolecule *source = nullptr, *sink = nullptr;
    std::vector<int> toDelete;
    std::vector<Alignment*> toSet;
    int num_moved = 0;
};

bool isActiveMolecule(const CandidateMolecule* mol, int read_change) {  // aligner.go:1239-1249
    double active = (double)(mol->active_alignments.Len() + read_change);
    double potential = (double)mol->best_alignment_for_read.Len();
    if (active <= 4) return false;
    if (active / potential < 0.1) return false;
    return true;
}

double fastScore(CandidateMolecule* sourceMolecule, CandidateMolecule* sinkMolecule, double log_unpaired_probability, Move* mv) {  // aligner.go:1109-1237
    double change = 0, alignment_change = 0;
    int num = 0;
    mv->toDelete.clear(); mv->toSet.clear();
    for (Alignment* sourceAlignment : sourceMolecule->active_alignments.store) {
        int read_id = sourceAlignment->read_id;
        Alignment* sinkAlignment = sinkMolecule->best_alignment_for_read.Get(read_id);
        if (sinkAlignment) {
            int mate_id = sourceAlignment->mate_id;
            Alignment* sourceMate = sourceMolecule->active_alignments.Get(mate_id);
            bool source_has_mate = sourceMate != nullptr;
            bool source_has_mate_pair = source_has_mate && isPair(sourceAlignment, sourceMate);
            Alignment* mate = sinkMolecule->best_alignment_for_read.Get(mate_id);
            bool sink_has_mate_pair = mate != nullptr && isPair(sinkAlignment, mate) && source_has_mate;
            if (!source_has_mate_pair || (source_has_mate && sink_has_mate_pair)) { mv->toDelete.push_back(read_id); mv->toSet.push_back(sinkAlignment); }
            alignment_change += sinkAlignment->log_alignment_probability - sourceAlignment->log_alignment_probability;
            // the mismatch-locus bookkeeping of the reference (:1157-1181) only feeds debug prints and panics
            if (source_has_mate_pair && !sink_has_mate_pair && sourceMolecule->id != sinkMolecule->id) alignment_change += log_unpaired_probability / 2.0;
            else if (!source_has_mate_pair && sink_has_mate_pair && sourceMolecule->id != sinkMolecule->id) alignment_change -= log_unpaired_probability / 2.0;
            ++num;
        }
    }
    bool source_active_before = isActiveMolecule(sourceMolecule, 0), source_active_after = isActiveMolecule(sourceMolecule, -num);
    if (!source_active_after && source_active_before && sourceMolecule->id != sinkMolecule->id) change -= (double)sourceMolecule->best_alignment_for_read.Len() * -0.5;
    bool sink_active_before = isActiveMolecule(sinkMolecule, 0), sink_active_after = isActiveMolecule(sinkMolecule, num);
    if (sink_active_after && !sink_active_before && sourceMolecule->id != sinkMolecule->id) change += (double)sinkMolecule->best_alignment_for_read.Len() * -0.5;
    if (sourceMolecule->active_alignments.Len() - num == 0 && num > 0 && sourceMolecule->id != sinkMolecule->id) change -= -3.0;
    if (sinkMolecule->active_alignments.Len() == 0 && num > 0 && sourceMolecule->id != sinkMolecule->id) change += -3.0;
    change += alignment_change;
    mv->source = sourceMolecule; mv->sink = sinkMolecule; mv->num_moved = num; mv->score_change = change; mv->alignment_change = alignment_change;
    return change;
}

void acceptMove(Move& move) {  // aligner.go:1261-1298
    for (size_t i = 0; i < move.toDelete.size(); ++i) {
        int read_id = move.toDelete[i];
        Alignment* sinkAlignment = move.toSet[i];
        Alignment* sourceAlignment = move.source->active_alignments.Get(read_id);
        move.source->active_alignments.Delete(read_id);
        move.sink->active_alignments.Set(read_id, sinkAlignment);
        sourceAlignment->active = false;
        sinkAlignment->active = true;
    }
}

struct Optimizer { std::vector<CandidateMolecule*>* candidate_molecules; int currentMoleculeMoveSource = 0; double log_unpaired_probability; };

void GenerateMove(Optimizer& o) {  // aligner.go:1065-1097 (the accept_move closure is never called)
    auto& mols = *o.candidate_molecules;
    CandidateMolecule* sourceMolecule = mols[o.currentMoleculeMoveSource];
    if (sourceMolecule->active_alignments.Len() == 0) { o.currentMoleculeMoveSource = (o.currentMoleculeMoveSource + 1) % (int)mols.size(); return; }
    Move best_move; best_move.score_change = -DBL_MAX;
    for (int i = 0; i < (int)mols.size(); ++i) {
        if (i == o.currentMoleculeMoveSource) continue;
        Move move;
        double score = fastScore(sourceMolecule, mols[i], o.log_unpaired_probability, &move);
        if ((score > best_move.score_change ||
             (score == best_move.score_change && move.sink->active_alignments.Len() > best_move.sink->active_alignments.Len())) && move.num_moved > 0)
            best_move = move;
    }
    double best_score = best_move.score_change;
    if (best_score > 0 || (best_score == 0 && best_move.sink->active_alignments.Len() > sourceMolecule->active_alignments.Len())) acceptMove(best_move);
    o.currentMoleculeMoveSource = (o.currentMoleculeMoveSource + 1) % (int)mols.size();
}

void Optimize(Optimizer& o, int temperature_steps, int steps_per_temp) {  // optimizer.go:15-27
    for (int t = 0; t < temperature_steps; ++t)
        for (int s = 0; s < steps_per_temp; ++s) GenerateMove(o);
}

double moleculeConfidence(const CandidateMolecule* mol, int num_active) { return (double)num_active / (double)mol->best_alignment_for_read.Len(); }

void moleculeMapqProbabilitySums(std::vector<CandidateMolecule*>& mols, double log_unpaired_probability) {  // aligner.go:697-720
    for (size_t a = 0; a < mols.size(); ++a)
        for (size_t b = 0; b < mols.size(); ++b) {
            if (a == b) continue;
            std::vector<Alignment*> sourceAlignments;
            for (Alignment* aln : mols[a]->active_alignments.store)
                if (mols[b]->best_alignment_for_read.Get(aln->read_id)) sourceAlignments.push_back(aln);
            Move mv;
            double sourceSinkChange = fastScore(mols[a], mols[b], log_unpaired_probability, &mv);
            double p = std::pow(10.0, sourceSinkChange);
            for (Alignment* aln : sourceAlignments) aln->sum_move_probability_change += p;
        }
}

void updateAlignmentsMoleculeStatus(std::vector<std::vector<Alignment*>>& alignments, std::vector<CandidateMolecule*>* mols) {  // aligner.go:643-675
    if (!mols) return;
    for (CandidateMolecule* m : *mols) {  // setMoleculeConfidences :978-989
        m->molecule_confidence = moleculeConfidence(m, m->active_alignments.Len());
        for (Alignment* a : m->active_alignments.store) { if (a->soft_clipped > 0) m->soft_clipped++; a->molecule_confidence = m->molecule_confidence; }
    }
    setMoleculeDifferences(*mols);
    for (auto& arr : alignments)
        for (Alignment* a : arr) {
            bool is_molecule_active = false;
            if (a->molecule_id != -1) {
                CandidateMolecule* m = (*mols)[a->molecule_id];
                is_molecule_active = m->active_alignments.Len() - m->soft_clipped > 4 && m->molecule_confidence > 0.1;
                a->active_molecule = is_molecule_active;
            }
            if (is_molecule_active) (*mols)[a->molecule_id]->active_molecule = true;
        }
}

double calculateLogMoleculePenalty(std::vector<CandidateMolecule*>* mols, double referenceLength) {  // aligner.go:722-755
    double dnaLength = 1000.0;
    if (!mols || mols->empty()) return 0.0;
    for (CandidateMolecule* mol : *mols) {
        if (mol->active_molecule) {
            int64_t smallest = INT64_MAX, biggest = -1;
            for (Alignment* a : mol->active_alignments.store) { if (a->pos > biggest) biggest = a->pos; if (a->pos < smallest) smallest = a->pos; }
            if (biggest >= smallest) dnaLength += (double)(biggest - smallest) + 1000.0;
        } else {
            for (Alignment* a : mol->active_alignments.store) dnaLength += (double)(a->aend - a->pos) * 2.0;
        }
    }
    return std::log10(dnaLength / referenceLength * 0.05);
}

double estimateMapQualities(std::vector<std::vector<Alignment*>>& alignments, std::vector<CandidateMolecule*>* mols, double log_unpaired_probability,
                            const orc_ref_t* ref) {  // aligner.go:797-922
    if (mols) moleculeMapqProbabilitySums(*mols, log_unpaired_probability);
    updateAlignmentsMoleculeStatus(alignments, mols);
    const double lmp = calculateLogMoleculePenalty(mols, 3200000000.0);
    for (size_t read_id = 0; read_id < alignments.size(); ++read_id) {
        auto& arr = alignments[read_id];
        std::vector<double> scores;
        if (!arr.empty()) {  // appendPsuedocountAlignmentScore :677-695
            auto& mateArray = alignments[arr[0]->mate_id];
            double bestSingle = -DBL_MAX;
            for (Alignment* m : mateArray) { double s = scoreAlignment(nullptr, m, lmp); if (s > bestSingle) bestSingle = s; }
            if (!mateArray.empty()) scores.push_back(bestSingle + psuedoCountAlignmentScore(arr[0], lmp));
            else scores.push_back(psuedoCountAlignmentScore(arr[0], lmp));
        }
        double total_probability = 0;
        for (Alignment* a : arr)
            for (Alignment* m : alignments[a->mate_id])
                if (a->active && m->active) { a->mate_alignment = m; m->mate_alignment = a; }
        for (Alignment* a : arr) {
            auto& mateArray = alignments[a->mate_id];
            double best_score = -DBL_MAX;
            for (Alignment* m : mateArray) { double s = scoreAlignment(a, m, lmp); if (s > best_score) best_score = s; }
            if (mateArray.empty()) best_score = scoreAlignment(a, nullptr, lmp);
            scores.push_back(best_score);
        }
        double second_best_raw_score = scores.empty() ? 0.0 : scores[0];
        double second_best_log_probability = -1000.0;
        Alignment* second_best_alignment = nullptr;
        for (Alignment* a : arr)
            for (Alignment* m : alignments[a->mate_id]) {
                double s = scoreAlignment(a, m, lmp);
                if (!a->active && s > second_best_log_probability) {
                    second_best_log_probability = s;
                    second_best_raw_score = scoreAlignment(a, m, 0.0);
                    second_best_alignment = a;
                    a->mate_alignment = m;
                }
            }
        for (Alignment* a : arr)
            if (a->active) {
                a->second_best = second_best_alignment;
                a->second_best_score = second_best_raw_score;
                a->as_score = scoreAlignment(a, a->mate_alignment, 0.0);
            }
        std::sort(scores.begin(), scores.end());
        for (int i = (int)scores.size() - 1; i >= 0 && (int)scores.size() - i <= 15; --i) total_probability += std::pow(10.0, scores[i]);
        for (Alignment* a : arr) {
            double score = scoreAlignment(a, a->mate_alignment, lmp);
            double mapq = -10.0 * std::log10(1.0 - std::pow(10.0, score) / total_probability);
            double moleculeMapq = -10.0 * std::log10(1.0 - (1.0 / a->sum_move_probability_change));
            // math.Min propagates NaN; std::fmin does not
            auto gomin = [](double x, double y) { if (std::isnan(x) || std::isnan(y)) return (double)NAN; return x < y ? x : y; };
            mapq = gomin(mapq, moleculeMapq);
            mapq = gomin(60.0, mapq);
            int64_t start = -1, end = -1;
            if (a->contig >= 0 && ref->centro_start) { start = ref->centro_start[a->contig]; end = ref->centro_end[a->contig]; }
            if (a->pos > start && a->pos <= end) mapq = 0.0;
            a->raw_mapq = mapq;
            a->mapq = std::isnan(mapq) ? INT32_MIN : (int)mapq;
        }
    }
    return lmp;
}

// ------------------------------------------------------------------------------------------------ markDuplicates (aligner.go:611-641)
void markDuplicates(std::vector<std::vector<Alignment*>>& alignments) {
    struct Tuple { bool read1, reversed; int contig; int64_t pos; int mateContig; int64_t matePos;
                   bool operator<(const Tuple& o) const { return std::tie(read1, reversed, contig, pos, mateContig, matePos) < std::tie(o.read1, o.reversed, o.contig, o.pos, o.mateContig, o.matePos); } };
    std::set<Tuple> seen;
    for (auto& arr : alignments)
        for (Alignment* a : arr)
            if (a->active) {
                const Alignment* m = a->mate_alignment;
                Tuple t{a->read1, a->reversed, a->contig, a->pos, m->contig, m->pos};
                if (!seen.insert(t).second) a->duplicate = true;
            }
}

// ------------------------------------------------------------------------------------------------ split reads (split.go:31-160)
Alignment* GetSplitAlignment(Alignment* primary, std::vector<Alignment*>& alignments, const orc_ref_t* ref, double* second_best_out) {
    *second_best_out = 0.0;
    if (primary->pos == -1) return nullptr;
    int Ps = primary->readmap_s, Pe = primary->readmap_e;
    if (Ps > Pe) std::swap(Ps, Pe);
    if (Pe - Ps > primary->read_len - 15) return nullptr;
    struct Cand { Alignment* a; double score; };
    std::vector<Cand> candidates;
    for (Alignment* c : alignments) {
        if (c->active || c->pos == -1) continue;
        int Ss = c->readmap_s, Se = c->readmap_e, overlap;
        if (Ss > Se) std::swap(Ss, Se);
        if ((Ps < Ss && Pe > Se) || (Ss < Ps && Se > Pe)) continue;
        else if (Ps < Ss) overlap = Pe - Ss;
        else overlap = Se - Ps;
        if (overlap < (Se - Ss) / 2) {
            c->is_proper = isPair(c, primary->mate_alignment);
            if (c->score >= 36 || c->is_proper) candidates.push_back({c, (double)c->score});
        }
    }
    if (candidates.empty()) return nullptr;
    // sort.Sort(SortSplitScoring) by score descending: insertion sort (stable) up to 12 elements, pdqsort above -- PARITY UNPINNED for
    // ties among more than 12 candidates; a stable sort is used here
    std::stable_sort(candidates.begin(), candidates.end(), [](const Cand& x, const Cand& y) { return x.score > y.score; });
    Alignment* c = candidates[0].a;
    double mapq;
    double second_best = scoreAlignment(primary, nullptr, 0.0) + psuedoCountAlignmentScore(candidates[0].a, 0.0);
    if (candidates.size() > 1) { mapq = candidates[0].score - candidates[1].score; second_best = scoreAlignment(primary, candidates[1].a, 0.0); }
    else mapq = candidates[0].score;
    int64_t start = -1, end = -1;
    if (c->contig >= 0 && ref->centro_start) { start = ref->centro_start[c->contig]; end = ref->centro_end[c->contig]; }
    if (c->pos > start && c->pos <= end) mapq = 0.0;
    if (mapq > 60) mapq = 60;
    c->mapq = (int)mapq;
    *second_best_out = second_best;
    return c;
}

void CheckSplitReads(std::vector<std::vector<Alignment*>>& reads, const orc_ref_t* ref) {
    for (auto& arr : reads) {
        Alignment* active = nullptr;
        for (Alignment* a : arr) if (a->active) { active = a; break; }
        double second_best;
        Alignment* split = GetSplitAlignment(active, arr, ref, &second_best);
        active->secondary = split;
        if (split) {
            split->second_best = nullptr; split->second_best_score = second_best;
            split->as_score = scoreAlignment(split, active->mate_alignment, 0.0);
            split->mapq_data_replaced = true;
            split->primary = active;
        }
    }
}

// ------------------------------------------------------------------------------------------------ AppendBam (bamwriter.go:283-568), fields only
struct RecOut { orc_record_t* recs; int64_t cap; int64_t n; const uint8_t* attach_bx; };

int append_record(Alignment* aln, Alignment* primary, int64_t read, bool attach_bx, RecOut* X) {
    if (X->n >= X->cap) return -1;
    orc_record_t& R = X->recs[X->n++];
    memset(&R, 0, sizeof(R));
    int flags = 0;
    if (!aln->is_proper && aln->score - 17 < 19) { aln->pos = -1; aln->mapq = 0; }
    R.mate_rid = -1; R.mate_pos = -1; R.tlen = 0;
    const Alignment* pm = primary->mate_alignment;
    flags |= 1;  // mate_id >= 0 always
    if (aln->is_proper) {
        if (aln == primary) flags |= 0x2;
        else if (isPair(aln, pm)) flags |= 0x2;
    }
    if (pm->pos == -1 || (!primary->is_proper && pm->score - 17 < 19)) { flags |= 0x8; }
    else { if (pm->reversed) flags |= 0x20; R.mate_rid = pm->contig; R.mate_pos = pm->pos; }
    flags |= aln->read1 ? 0x40 : 0x80;
    if (aln->duplicate) flags |= 0x400;
    if (pm->pos == -1) { R.mate_rid = -1; R.tlen = 0; }
    else if (aln == primary) {
        if (aln->contig == aln->mate_alignment->contig && (primary->is_proper || pm->score - 17 >= 19)) {
            if (aln->reversed) R.tlen = -(aln->aend - aln->mate_alignment->pos);
            else R.tlen = aln->mate_alignment->aend - aln->pos;
        }
    }
    if (aln != primary) flags |= 256;
    R.rid = aln->contig; R.mapq = aln->mapq;
    if (aln->pos == -1) { flags |= 0x4; R.mapq = 0; R.rid = -1; }
    if (aln->reversed) flags |= 0x10;
    R.flag = flags; R.read = (int32_t)read; R.reg = aln->reg; R.primary = aln == primary; R.duplicate = aln->duplicate; R.reversed = aln->reversed;
    R.pos = aln->pos;
    // SEQ / QUAL range and hard clipping of split records (HardClip, bamwriter.go:664-689, on the forward-strand CIGAR)
    R.clip_start = 0; R.clip_end = aln->read_len; R.hard = 0;
    if (aln != primary) {
        R.hard = 1;
        if (aln->n_cigar >= 1 && (aln->cigar[0] & 0xf) == 3) R.clip_start = (int)(aln->cigar[0] >> 4);
        if (aln->n_cigar >= 2 && (aln->cigar[aln->n_cigar - 1] & 0xf) == 3) R.clip_end -= (int)(aln->cigar[aln->n_cigar - 1] >> 4);
    }
    // tags
    R.as = (int32_t)aln->as_score; R.xs = (int32_t)aln->second_best_score;
    const Alignment* sb = aln->mapq_data_replaced ? nullptr : aln->second_best;
    R.xc_reg = sb ? sb->reg : -1; R.has_xc = sb != nullptr;
    R.xm = sb && sb->active_molecule; R.am = aln->active_molecule; R.xt = sb && aln->molecule_id == sb->molecule_id;
    const Alignment* sa = aln->secondary ? aln->secondary : aln->primary;
    R.sa_reg = -1;
    if (sa && sa->pos > -1) {
        R.sa_reg = sa->reg; R.sa_rid = sa->contig; R.sa_pos = sa->pos; R.sa_rev = sa->reversed; R.sa_hard = aln->secondary != nullptr; R.sa_mapq = sa->mapq;
        int indel = 0;
        for (int k = 0; k < sa->n_cigar; ++k) { const int op = sa->cigar[k] & 0xf; if (op == 1 || op == 2) indel += (int)(sa->cigar[k] >> 4); }
        R.sa_nm = (int32_t)sa->mmLocs.size() + indel;
    }
    R.attach_bx = attach_bx; R.has_dm = attach_bx && aln->active_molecule; R.dm = aln->molecule_difference;
    return 0;
}

int emit_records(std::vector<std::vector<Alignment*>>& alignments, int64_t read0, bool attach_bx, RecOut* X) {  // DoDumpToBam, bamwriter.go:635-658
    for (size_t rl = 0; rl < alignments.size(); ++rl)
        for (Alignment* a : alignments[rl])
            if (a->active) {
                if (append_record(a, a, read0 + (int64_t)rl, attach_bx, X)) return -1;
                if (a->secondary && append_record(a->secondary, a, read0 + (int64_t)rl, attach_bx, X)) return -1;
            }
    return 0;
}

}  // namespace

extern "C" {

void orc_md5(const uint8_t* msg, int64_t len, uint8_t out[16]) { Md5 m; m.digest(msg, len, out); }

int64_t orc_name_seed(const char* name, int len) {  // aligner.go:1414-1416
    uint8_t d[16]; orc_md5((const uint8_t*)name, len, d);
    uint64_t v = 0;
    for (int k = 0; k < 8; ++k) v |= (uint64_t)d[k] << (8 * k);
    return (int64_t)v;
}

void orc_gorand_float64(int64_t seed, int n, double* out) { GoRand r(seed); for (int i = 0; i < n; ++i) out[i] = r.Float64(); }

int orc_walk(const orc_ref_t* ref, int l_query, const uint8_t* query, const void* reg_, const int64_t* a8, const uint32_t* cigar, int* out5, int32_t* mm) {
    const Reg& r = *(const Reg*)reg_;
    int64_t cpos, caend; interpret_align(ref, r, &cpos, &caend);
    Walk w;
    walk_alignment(ref, l_query, query, cpos, caend, (int)a8[1], a8[2] != 0, (int)((uint32_t)a8[3] & 0x3fffffu), (int)a8[4], cigar + a8[5], &w);
    out5[0] = w.matches; out5[1] = w.mismatches; out5[2] = w.indels; out5[3] = w.soft_clipped; out5[4] = w.soft_clipped_length;
    int n = (int)w.mmLocs.size();
    for (int i = 0; i < n; ++i) { mm[i] = w.mmLocs[i]; mm[n + i] = w.mmReadLocs[i]; }
    return n;
}

static int place_impl(const orc_ref_t* ref, int64_t n_pairs, const uint8_t* seq, const int64_t* off, int64_t n_bc, const int64_t* bc_off, const uint8_t* run_rfa,
              const int64_t* rand_seed, const int64_t* reg_off, const void* regs_, const int64_t* aln, const uint32_t* cigar, double improper_penalty,
              int delta, int64_t* cand_off, orc_cand_t* cands, int64_t cap, double* lmp_out, RecOut* X) {
    const Reg* regs = (const Reg*)regs_;
    g_improper = improper_penalty;
    int64_t n_cand = 0;
    cand_off[0] = 0;
    const int64_t one_bc[2] = {0, n_pairs};
    if (n_bc <= 0) { n_bc = 1; bc_off = one_bc; }
    for (int64_t b = 0; b < n_bc; ++b) {
        const int64_t p0 = bc_off[b], p1 = bc_off[b + 1];
        const int n_reads = (int)(2 * (p1 - p0));
        std::vector<std::vector<Alignment*>> alignments(n_reads), full(n_reads);
        std::vector<Alignment*> pool;
        // ---- GetChains + GetAlignments (aligner.go:1633-1715, 1484-1631) ----
        int hit_num = 0;
        for (int rl = 0; rl < n_reads; ++rl) {
            const int64_t r = 2 * p0 + rl;
            const int l_query = (int)(off[r + 1] - off[r]);
            const uint8_t* readSeq = seq + off[r];
            const int64_t g0 = reg_off[r], g1 = reg_off[r + 1];
            int bestScore = 0;
            for (int64_t g = g0; g < g1; ++g) if (regs[g].score > bestScore) bestScore = regs[g].score;
            auto make = [&](int64_t g) {
                Alignment* a = new Alignment(); pool.push_back(a);
                a->read_id = rl; a->mate_id = rl ^ 1; a->read1 = (rl & 1) == 0; a->read_len = l_query;
                if (g >= 0) {
                    const Reg& rg = regs[g]; const int64_t* a8 = aln + 8 * g;
                    int64_t cpos, caend; interpret_align(ref, rg, &cpos, &caend);
                    const bool reversed = a8[2] != 0;
                    Walk w;
                    walk_alignment(ref, l_query, readSeq, cpos, caend, (int)a8[1], reversed, (int)((uint32_t)a8[3] & 0x3fffffu), (int)a8[4], cigar + a8[5], &w);
                    a->id = hit_num; a->reg = (int)g; a->contig = (int)a8[1]; a->reversed = reversed; a->score = rg.score;
                    a->matches = w.matches; a->mismatches = w.mismatches; a->indels = w.indels; a->soft_clipped = w.soft_clipped;
                    a->soft_clipped_length = w.soft_clipped_length;
                    a->pos = cpos; a->aend = caend;
                    if (cpos != -1 && reversed) { a->pos = caend + 1; a->aend = cpos + 1; }
                    a->mmLocs = w.mmLocs; a->mmReadLocs = w.mmReadLocs; a->n_cigar = (int)a8[4]; a->cigar = cigar + a8[5];
                    a->readmap_s = rg.qb; a->readmap_e = rg.qe;
                } else {
                    a->id = (rl & 1) ? hit_num : 0;  // the read-1 placeholder omits hit_id (aligner.go:1665-1675)
                    a->pos = -1; a->aend = 0; a->contig = -1; a->score = 0;
                }
                a->log_alignment_probability = scoreAlignment(a, nullptr, 0.0) - g_improper;
                ++hit_num;
                return a;
            };
            if (g1 == g0) { Alignment* a = make(-1); alignments[rl].push_back(a); full[rl].push_back(a); }  // 0 >= 0 - delta: the placeholder always survives the filter
            for (int64_t g = g0; g < g1; ++g) {
                Alignment* a = make(g);
                full[rl].push_back(a);
                if (a->score >= bestScore - delta) alignments[rl].push_back(a);
            }
        }
        // batch-global candidate slots
        for (int rl = 0; rl < n_reads; ++rl) {
            for (Alignment* a : alignments[rl]) { if (n_cand >= cap) { for (auto* p : pool) delete p; return -1; } a->cand = (int)n_cand++; }
            cand_off[2 * p0 + rl + 1] = n_cand;
        }
        // ---- DoRFAForOneBarcode (aligner.go:440-501) ----
        auto positions = tagBestAlignments(alignments, rand_seed + p0);
        double lmp = 0;
        std::vector<CandidateMolecule*> all_mols, dropped;
        if (!(run_rfa && run_rfa[b])) {
            lmp = estimateMapQualities(alignments, nullptr, improper_penalty, ref);
        } else {
            all_mols = inferMolecules(positions);
            markBestAlignmentForReadInMolecule(all_mols);
            std::vector<CandidateMolecule*> mols = scrapMolecules(all_mols, dropped);
            setMoleculeDifferences(mols);
            Optimizer o; o.candidate_molecules = &mols; o.log_unpaired_probability = improper_penalty;
            Optimize(o, 2, 4 * (int)mols.size());
            lmp = estimateMapQualities(alignments, &mols, improper_penalty, ref);
        }
        if (lmp_out) lmp_out[b] = lmp;
        markDuplicates(alignments);
        for (int rl = 0; rl < n_reads; ++rl)
            for (Alignment* a : alignments[rl]) {
                orc_cand_t& c = cands[a->cand];
                c.reg = a->reg; c.read = (int32_t)(2 * p0 + rl); c.id = a->id; c.active = a->active; c.is_proper = a->is_proper; c.bwa_pick = a->bwa_pick;
                c.mapq = a->mapq; c.molecule_id = a->molecule_id; c.active_molecule = a->active_molecule;
                c.second_best = (a->active && a->second_best) ? a->second_best->cand : -1;
                c.mate = a->mate_alignment ? a->mate_alignment->cand : -1;
                c.duplicate = a->duplicate; c.pos = a->pos; c.aend = a->aend; c.log_alignment_probability = a->log_alignment_probability;
                c.as_score = a->as_score; c.xs_score = a->second_best_score; c.molecule_difference = a->molecule_difference;
                c.sum_move_probability_change = a->sum_move_probability_change; c.molecule_confidence = a->molecule_confidence; c.raw_mapq = a->raw_mapq;
            }
        // the candidate table above is the state after estimateMapQualities + markDuplicates; split selection and the writer edit
        // is_proper / mapq / pos of some alignments afterwards (split.go:96,146; bamwriter.go:287-290) -- visible in the records only
        if (X) {
            CheckSplitReads(full, ref);
            const bool attach_bx = X->attach_bx ? X->attach_bx[b] != 0 : (run_rfa && run_rfa[b]);
            if (emit_records(alignments, 2 * p0, attach_bx, X) < 0) { for (auto* m : all_mols) delete m; for (auto* p : pool) delete p; return -2; }
        }
        for (auto* m : all_mols) delete m;
        for (auto* p : pool) delete p;
    }
    return 0;
}

void orc_set_tie_mode(int mode) { g_tie_mode = mode; }

int orc_place(const orc_ref_t* ref, int64_t n_pairs, const uint8_t* seq, const int64_t* off, int64_t n_bc, const int64_t* bc_off, const uint8_t* run_rfa,
              const int64_t* rand_seed, const int64_t* reg_off, const void* regs_, const int64_t* aln, const uint32_t* cigar, double improper_penalty,
              int delta, int64_t* cand_off, orc_cand_t* cands, int64_t cap, double* lmp_out) {
    return place_impl(ref, n_pairs, seq, off, n_bc, bc_off, run_rfa, rand_seed, reg_off, regs_, aln, cigar, improper_penalty, delta, cand_off, cands, cap, lmp_out, nullptr);
}

int orc_place_records(const orc_ref_t* ref, int64_t n_pairs, const uint8_t* seq, const int64_t* off, int64_t n_bc, const int64_t* bc_off, const uint8_t* run_rfa,
                      const uint8_t* attach_bx, const int64_t* rand_seed, const int64_t* reg_off, const void* regs_, const int64_t* aln, const uint32_t* cigar,
                      double improper_penalty, int delta, int64_t* cand_off, orc_cand_t* cands, int64_t cap, double* lmp_out, orc_record_t* recs, int64_t rec_cap,
                      int64_t* n_recs) {
    RecOut X; X.recs = recs; X.cap = rec_cap; X.n = 0; X.attach_bx = attach_bx;
    const int rc = place_impl(ref, n_pairs, seq, off, n_bc, bc_off, run_rfa, rand_seed, reg_off, regs_, aln, cigar, improper_penalty, delta, cand_off, cands, cap, lmp_out, &X);
    *n_recs = X.n;
    return rc;
}

}  // extern "C"
