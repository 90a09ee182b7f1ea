// md5(name)[0:8], little-endian, as int64: the seed tagBestAlignments hands to rand.NewSource for a read pair
// (src/aligner/aligner.go:1413-1416).  RFC 1321, host code; rounds unrolled by group so the per-step function and message index
// are compile-time constants -- the FASTQ ingest computes one digest per read pair (csrc/ingest.cu).
#pragma once
#include <cstdint>
#include <cstring>

namespace arachne {

namespace md5_detail {

inline uint32_t rotl(uint32_t x, int r) { return (x << r) | (x >> (32 - r)); }

constexpr uint32_t K[64] = {
    0xd76aa478, 0xe8c7b756, 0x242070db, 0xc1bdceee, 0xf57c0faf, 0x4787c62a, 0xa8304613, 0xfd469501, 0x698098d8, 0x8b44f7af, 0xffff5bb1, 0x895cd7be,
    0x6b901122, 0xfd987193, 0xa679438e, 0x49b40821, 0xf61e2562, 0xc040b340, 0x265e5a51, 0xe9b6c7aa, 0xd62f105d, 0x02441453, 0xd8a1e681, 0xe7d3fbc8,
    0x21e1cde6, 0xc33707d6, 0xf4d50d87, 0x455a14ed, 0xa9e3e905, 0xfcefa3f8, 0x676f02d9, 0x8d2a4c8a, 0xfffa3942, 0x8771f681, 0x6d9d6122, 0xfde5380c,
    0xa4beea44, 0x4bdecfa9, 0xf6bb4b60, 0xbebfbc70, 0x289b7ec6, 0xeaa127fa, 0xd4ef3085, 0x04881d05, 0xd9d4d039, 0xe6db99e5, 0x1fa27cf8, 0xc4ac5665,
    0xf4292244, 0x432aff97, 0xab9423a7, 0xfc93a039, 0x655b59c3, 0x8f0ccc92, 0xffeff47d, 0x85845dd1, 0x6fa87e4f, 0xfe2ce6e0, 0xa3014314, 0x4e0811a1,
    0xf7537e82, 0xbd3af235, 0x2ad7d2bb, 0xeb86d391};

// one 64-byte block; w = the block as 16 little-endian words (x86-64 and the CUDA hosts we build for are little-endian)
inline void block(uint32_t h[4], const uint32_t w[16]) {
    uint32_t a = h[0], b = h[1], c = h[2], d = h[3];
#define AR_MD5_STEP(F, G, I, R)                                          \
    {                                                                    \
        const uint32_t t_ = a + (F) + K[I] + w[G];                       \
        a = d; d = c; c = b; b += rotl(t_, R);                           \
    }
#define AR_MD5_QUAD(F, G0, G1, G2, G3, I, R0, R1, R2, R3) \
    AR_MD5_STEP(F, G0, I, R0) AR_MD5_STEP(F, G1, I + 1, R1) AR_MD5_STEP(F, G2, I + 2, R2) AR_MD5_STEP(F, G3, I + 3, R3)
#define AR_F1 ((b & c) | (~b & d))
#define AR_F2 ((d & b) | (~d & c))
#define AR_F3 (b ^ c ^ d)
#define AR_F4 (c ^ (b | ~d))
    AR_MD5_QUAD(AR_F1, 0, 1, 2, 3, 0, 7, 12, 17, 22)    AR_MD5_QUAD(AR_F1, 4, 5, 6, 7, 4, 7, 12, 17, 22)
    AR_MD5_QUAD(AR_F1, 8, 9, 10, 11, 8, 7, 12, 17, 22)  AR_MD5_QUAD(AR_F1, 12, 13, 14, 15, 12, 7, 12, 17, 22)
    AR_MD5_QUAD(AR_F2, 1, 6, 11, 0, 16, 5, 9, 14, 20)   AR_MD5_QUAD(AR_F2, 5, 10, 15, 4, 20, 5, 9, 14, 20)
    AR_MD5_QUAD(AR_F2, 9, 14, 3, 8, 24, 5, 9, 14, 20)   AR_MD5_QUAD(AR_F2, 13, 2, 7, 12, 28, 5, 9, 14, 20)
    AR_MD5_QUAD(AR_F3, 5, 8, 11, 14, 32, 4, 11, 16, 23) AR_MD5_QUAD(AR_F3, 1, 4, 7, 10, 36, 4, 11, 16, 23)
    AR_MD5_QUAD(AR_F3, 13, 0, 3, 6, 40, 4, 11, 16, 23)  AR_MD5_QUAD(AR_F3, 9, 12, 15, 2, 44, 4, 11, 16, 23)
    AR_MD5_QUAD(AR_F4, 0, 7, 14, 5, 48, 6, 10, 15, 21)  AR_MD5_QUAD(AR_F4, 12, 3, 10, 1, 52, 6, 10, 15, 21)
    AR_MD5_QUAD(AR_F4, 8, 15, 6, 13, 56, 6, 10, 15, 21) AR_MD5_QUAD(AR_F4, 4, 11, 2, 9, 60, 6, 10, 15, 21)
#undef AR_F1
#undef AR_F2
#undef AR_F3
#undef AR_F4
#undef AR_MD5_QUAD
#undef AR_MD5_STEP
    h[0] += a; h[1] += b; h[2] += c; h[3] += d;
}

}  // namespace md5_detail

inline int64_t md5_name_seed(const char* name, int64_t len) {
    using namespace md5_detail;
    uint32_t h[4] = {0x67452301u, 0xefcdab89u, 0x98badcfeu, 0x10325476u};
    const uint64_t n = (uint64_t)len;
    uint64_t done = 0;
    uint32_t w[16];
    for (; done + 64 <= n; done += 64) { memcpy(w, name + done, 64); block(h, w); }
    uint8_t tail[128];
    const uint64_t rest = n - done, tail_len = rest + 9 <= 64 ? 64 : 128;
    memset(tail, 0, sizeof(tail));
    if (rest) memcpy(tail, name + done, rest);
    tail[rest] = 0x80;
    const uint64_t bits = n * 8;
    memcpy(tail + tail_len - 8, &bits, 8);
    for (uint64_t o = 0; o < tail_len; o += 64) { memcpy(w, tail + o, 64); block(h, w); }
    return (int64_t)((uint64_t)h[1] << 32 | h[0]);
}

}  // namespace arachne
