// Go's math/rand source (math/rand/rng.go: additive lagged Fibonacci, length 607, tap 273) for tagBestAlignments'
// tie-breaking noise: rand.New(rand.NewSource(md5(read_name)[0:8])).Float64()/2 (src/aligner/aligner.go:1413-1418,1431,1442).
//
// Go seeds the 607-word state from a Lehmer LCG (x <- 48271 x mod 2^31-1, three draws per word) XOR a fixed table, then
// output n is vec[feed_n] + vec[tap_n] written back at feed_n.  For the first 273 outputs neither operand has been
// overwritten yet, so output n = word(333-n) + word(606-n), and word(j) has a closed form: the three LCG draws behind it
// are seed' * 48271^(21+3j+k) mod (2^31-1).  A read pair needs |cands(R1)| x |cands(R2)| outputs -- almost always a
// handful -- so the common path is stateless: two table look-ups and six modular multiplications per output.  Pairs
// needing more than 273 outputs materialise the 607-word state in a global scratch slot and run the recurrence.
#pragma once
#include "common.cuh"
#include "gorand_table.h"

namespace arachne {

// AR_GR_FN / AR_GR_M: under nvcc exactly the qualifiers this code has always had; the TEST-ONLY host simulation (tests/hostsim, which
// defines ARACHNE_HOSTSIM) compiles the same generator with plain inline functions and host copies of the two tables.
#if defined(__CUDACC__)
__device__ const uint64_t d_gorand_cooked[607] = GORAND_TABLE_INIT;
__device__ const uint32_t d_gorand_pw[3 * 607] = GORAND_PW_INIT;
#define AR_GR_FN __device__ __forceinline__
#define AR_GR_M __device__
#elif defined(ARACHNE_HOSTSIM)
static const uint64_t d_gorand_cooked[607] = GORAND_TABLE_INIT;
static const uint32_t d_gorand_pw[3 * 607] = GORAND_PW_INIT;
#define AR_GR_FN inline
#define AR_GR_M
#endif
#define GORAND_COOKED d_gorand_cooked
#define GORAND_PW d_gorand_pw

constexpr int GORAND_LEN = 607, GORAND_TAP = 273;
constexpr uint32_t GORAND_M31 = 2147483647u;

AR_HD uint32_t gorand_norm_seed(int64_t seed) {  // rngSource.Seed: seed % (2^31-1), made positive, 0 -> 89482311
    seed = seed % (int64_t)GORAND_M31;
    if (seed < 0) seed += GORAND_M31;
    if (seed == 0) seed = 89482311;
    return (uint32_t)seed;
}

AR_HD uint32_t gorand_mulmod(uint32_t a, uint32_t b) {  // a*b mod 2^31-1 (== Schrage's seedrand for a step of 48271)
    uint64_t p = (uint64_t)a * b;
    p = (p & GORAND_M31) + (p >> 31);
    p = (p & GORAND_M31) + (p >> 31);
    return p == GORAND_M31 ? 0u : (uint32_t)p;
}

#if defined(__CUDACC__) || defined(ARACHNE_HOSTSIM)
AR_GR_FN uint64_t gorand_word(uint32_t s, int j) {  // seeded state word j
    const uint64_t x0 = gorand_mulmod(s, GORAND_PW[3 * j]), x1 = gorand_mulmod(s, GORAND_PW[3 * j + 1]), x2 = gorand_mulmod(s, GORAND_PW[3 * j + 2]);
    return ((x0 << 40) ^ (x1 << 20) ^ x2) ^ GORAND_COOKED[j];
}

AR_GR_FN double gorand_to_float(uint64_t u) { return (double)(int64_t)(u & 0x7fffffffffffffffull) / 9223372036854775808.0; }

struct GoRandDev {
    uint32_t s;
    int n;            // outputs consumed
    uint64_t* state;  // 607 words of global scratch, or nullptr while stateless
    int tap, feed;
    AR_GR_M void init(int64_t seed) { s = gorand_norm_seed(seed); n = 0; state = nullptr; tap = 0; feed = GORAND_LEN - GORAND_TAP; }
    // switch to the stateful generator positioned after the n outputs already taken
    AR_GR_M void materialise(uint64_t* scratch) {
        state = scratch;
        for (int j = 0; j < GORAND_LEN; ++j) state[j] = gorand_word(s, j);
        tap = 0; feed = GORAND_LEN - GORAND_TAP;
        const int done = n; n = 0;
        for (int i = 0; i < done; ++i) next_u64();
    }
    AR_GR_M uint64_t next_u64() {
        if (state) {
            if (--tap < 0) tap += GORAND_LEN;
            if (--feed < 0) feed += GORAND_LEN;
            uint64_t x = state[feed] + state[tap];
            state[feed] = x; ++n;
            return x;
        }
        const uint64_t x = gorand_word(s, 333 - n) + gorand_word(s, 606 - n);
        ++n;
        return x;
    }
    AR_GR_M bool stateless_ok(int more) const { return state != nullptr || n + more <= GORAND_TAP; }
    // rand.Float64: retry when the quotient rounds to 1.0 (a retry consumes an output, so it is only taken statefully or
    // while the stateless window lasts)
    AR_GR_M double float64(bool* need_state) {
        for (;;) {
            if (!state && n >= GORAND_TAP) { *need_state = true; return 0.0; }
            double f = gorand_to_float(next_u64());
            if (f != 1.0) return f;
        }
    }
};
#endif

}  // namespace arachne
