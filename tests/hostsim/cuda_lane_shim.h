// TEST-ONLY: the handful of CUDA constructs that kernels made of INDEPENDENT lanes use (k_seed2, k_pack_reads, k_sort_intv, the index
// builders of index_kernels.cuh, the lane-per-region CIGAR tiers k_cigar_fast / k_cigar_band: no lane reads another lane's registers or shared memory), defined so that g++ compiles those kernels
// unchanged and the host runs them ONE LANE AT A TIME.  A kernel "launch" is a plain call with threadIdx / blockIdx / blockDim / gridDim set
// by the caller; warp barriers are no-ops and a warp vote is the lane's own predicate, which is exact for one lane.  Never shipped.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>

struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
struct ulonglong2 { unsigned long long x, y; };
struct int2 { int x, y; };
inline int2 make_int2(int x, int y) { return int2{x, y}; }
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { return ulonglong2{x, y}; }
struct LaneDim3 { unsigned x = 1, y = 1, z = 1; };
struct LaneIdx3 { unsigned x = 0, y = 0, z = 0; };
static LaneIdx3 threadIdx, blockIdx;
static LaneDim3 blockDim, gridDim;

#define __device__
#define __host__
#define __global__
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)
#define __shared__            /* per-invocation storage: one lane is its whole block */

inline void __syncwarp() {}
inline void __syncthreads() {}
inline bool __all_sync(unsigned, bool p) { return p; }
template <class T, class V> inline T atomicAdd(T* p, V v) { const T old = *p; *p = (T)(old + (T)v); return old; }
inline int atomicCAS(int* p, int cmp, int val) { const int old = *p; if (old == cmp) *p = val; return old; }
template <class T> inline T __ldg(const T* p) { return *p; }
inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
inline int __clzll(long long x) { return x == 0 ? 64 : __builtin_clzll((unsigned long long)x); }
inline unsigned __funnelshift_l(unsigned lo, unsigned hi, unsigned shift) { return (unsigned)((((unsigned long long)hi << 32 | lo) << (shift & 31)) >> 32); }
inline unsigned __byte_perm(unsigned lo, unsigned hi, unsigned sel) {  // PRMT, default mode
    const unsigned long long src = (unsigned long long)hi << 32 | lo;
    unsigned r = 0;
    for (int i = 0; i < 4; ++i) {
        const unsigned n = (sel >> (4 * i)) & 0xfu, b = (unsigned)(src >> (8 * (n & 7u))) & 0xffu;
        r |= ((n & 8u) ? ((b & 0x80u) ? 0xffu : 0u) : b) << (8 * i);
    }
    return r;
}
using std::min;
using std::max;
