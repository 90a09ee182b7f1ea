// Common PODs and helpers shared by every kernel of libarachne_cuda.
//
// Thread-sequential logic is written as AR_HD (host+device) so the *same source* can also be compiled by
// g++ into the test-only host simulation (tests/hostsim), which lets the CPU-only test tier pin the control
// logic against the oracle before a GPU is involved.  The shipped library never calls the host build of
// these functions: every entry point launches kernels and fails loudly without a device.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define AR_HD __host__ __device__ __forceinline__
#define AR_HDN __host__ __device__ __noinline__
#define AR_D __device__ __forceinline__
#else
#define AR_HD inline
#define AR_HDN inline
#endif

namespace arachne {

// ---- options: field-for-field the reference's mem_opt_t (bwa/bwamem.h:27-66) so the compat shim can hand it to Go ----
struct MemOpt {
    int a, b;
    int o_del, e_del;
    int o_ins, e_ins;
    int pen_unpaired;
    int pen_clip5, pen_clip3;
    int w;
    int zdrop;
    uint64_t max_mem_intv;
    int T;
    int flag;
    int min_seed_len;
    int min_chain_weight;
    int max_chain_extend;
    float split_factor;
    int split_width;
    int max_occ;
    int max_chain_gap;
    int n_threads;
    int chunk_size;
    float mask_level;
    float drop_ratio;
    float XA_drop_ratio;
    float mask_level_redun;
    float mapQ_coef_len;
    int mapQ_coef_fac;
    int max_ins;
    int max_matesw;
    int max_XA_hits, max_XA_hits_alt;
    int8_t mat[25];
};

// ---- alignment region: byte-for-byte mem_alnreg_t (bwa/bwamem.h:86-104), 88 bytes ----
struct AlnReg {
    int64_t rb, re;
    int qb, qe;
    int rid;
    int score;
    int truesc;
    int sub;
    int alt_sc;
    int csub;
    int sub_n;
    int w;
    int seedcov;
    int secondary;
    int secondary_all;
    int seedlen0;
    int n_comp_is_alt;  // n_comp:30, is_alt:2
    float frac_rep;
    uint64_t hash;
};
static_assert(sizeof(AlnReg) == 88, "AlnReg must match mem_alnreg_t");

// bi-directional SA interval, bwa/bwt.h:60-62
struct Intv {
    uint64_t x0, x1, x2, info;
};

// seed inside a chain (bwa/bwamem.c:168-172) + link to the next seed of the same chain
struct Seed {
    int64_t rbeg;
    int32_t qbeg, len;
    int32_t score;
    int32_t next;
};

// chain header (bwa/bwamem.c:174-180); seeds live in a per-read Seed array, linked through Seed::next while
// chaining and stored contiguously ([first, first+n)) after filtering
struct Chain {
    int64_t pos;
    int32_t rid;
    int32_t n;
    int32_t first_seed, last_seed;
    int32_t w;       // weight (29 bits in the reference)
    int32_t first;   // index of the first shadowed chain
    int8_t kept, is_alt;
    int16_t pad;
    float frac_rep;
    int32_t qbeg0, qend_last;  // chn_beg / chn_end cache
};

// contig annotation needed on the device (bwa/bntseq.h:41-48)
struct Contig {
    int64_t offset;
    int32_t len;
    int32_t is_alt;
};

struct RefView {
    const uint8_t* pac;  // 2-bit packed forward reference, base l at bits ((~l&3)<<1) of byte l>>2 (bntseq.c:225)
    const Contig* contigs;
    int64_t l_pac;
    int32_t n_seqs;
};

// ---- klib introsort, exact (bwa/ksort.h:150-226): the permutation it produces on ties is observable ----
template <class T, class Less>
AR_HD void ks_insertsort(T* s, T* t, Less lt) {
    for (T* i = s + 1; i < t; ++i)
        for (T* j = i; j > s && lt(*j, *(j - 1)); --j) {
            T tmp = *j; *j = *(j - 1); *(j - 1) = tmp;
        }
}

template <class T, class Less>
AR_HD void ks_combsort(size_t n, T* a, Less lt) {
    const double shrink = 1.2473309501039786540366528676643;
    int do_swap;
    size_t gap = n;
    do {
        if (gap > 2) {
            gap = (size_t)(gap / shrink);
            if (gap == 9 || gap == 10) gap = 11;
        }
        do_swap = 0;
        for (T* i = a; i < a + n - gap; ++i) {
            T* j = i + gap;
            if (lt(*j, *i)) { T tmp = *i; *i = *j; *j = tmp; do_swap = 1; }
        }
    } while (do_swap || gap > 2);
    if (gap != 1) ks_insertsort(a, a + n, lt);
}

template <class T, class Less>
AR_HDN void ks_introsort(size_t n, T* a, Less lt) {
    struct Frame { T *left, *right; int depth; };
    Frame stack[72];  // the reference sizes it 8*d+2 but never holds more than ~d frames (the larger side is pushed)
    Frame* top = stack;
    int d;
    T rp, tmp;
    T *s, *t, *i, *j, *k;
    if (n < 1) return;
    if (n == 2) {
        if (lt(a[1], a[0])) { tmp = a[0]; a[0] = a[1]; a[1] = tmp; }
        return;
    }
    for (d = 2; (1ul << d) < n; ++d) {}
    s = a; t = a + (n - 1); d <<= 1;
    for (;;) {
        if (s < t) {
            if (--d == 0) {
                ks_combsort((size_t)(t - s + 1), s, lt);
                t = s;
                continue;
            }
            i = s; j = t; k = i + ((j - i) >> 1) + 1;
            if (lt(*k, *i)) {
                if (lt(*k, *j)) k = j;
            } else k = lt(*j, *i) ? i : j;
            rp = *k;
            if (k != t) { tmp = *k; *k = *t; *t = tmp; }
            for (;;) {
                do ++i; while (lt(*i, rp));
                do --j; while (i <= j && lt(rp, *j));
                if (j <= i) break;
                tmp = *i; *i = *j; *j = tmp;
            }
            tmp = *i; *i = *t; *t = tmp;
            if (i - s > t - i) {
                if (i - s > 16) { top->left = s; top->right = i - 1; top->depth = d; ++top; }
                s = t - i > 16 ? i + 1 : t;
            } else {
                if (t - i > 16) { top->left = i + 1; top->right = t; top->depth = d; ++top; }
                t = i - s > 16 ? i - 1 : s;
            }
        } else {
            if (top == stack) {
                ks_insertsort(a, a + n, lt);
                return;
            }
            --top; s = top->left; t = top->right; d = top->depth;
        }
    }
}

template <class T> AR_HD T ar_min(T a, T b) { return a < b ? a : b; }
template <class T> AR_HD T ar_max(T a, T b) { return a > b ? a : b; }

}  // namespace arachne
