// TEST-ONLY: what cuda_lane_shim.h and cuda_warp_shim.h share -- vector types, empty qualifiers, plain atomics and the integer intrinsics
// the emulated kernels use, each stated in C.  Never shipped.
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstring>

struct uint2 { unsigned x, y; };
struct uint4 { unsigned x, y, z, w; };
struct ulonglong2 { unsigned long long x, y; };
struct int2 { int x, y; };
inline uint2 make_uint2(unsigned x, unsigned y) { return uint2{x, y}; }
inline uint4 make_uint4(unsigned x, unsigned y, unsigned z, unsigned w) { return uint4{x, y, z, w}; }
inline ulonglong2 make_ulonglong2(unsigned long long x, unsigned long long y) { return ulonglong2{x, y}; }
inline int2 make_int2(int x, int y) { return int2{x, y}; }
struct LaneDim3 { unsigned x = 1, y = 1, z = 1; };
struct LaneIdx3 { unsigned x = 0, y = 0, z = 0; };

#define __device__
#define __host__
#define __global__ inline   /* a kernel header may be compiled into several test TUs */
#define __forceinline__ inline
#define __noinline__
#define __launch_bounds__(...)

template <class T, class V> inline T atomicAdd(T* p, V v) { const T old = *p; *p = (T)(old + (T)v); return old; }
template <class T, class V> inline T atomicSub(T* p, V v) { const T old = *p; *p = (T)(old - (T)v); return old; }
template <class T, class V> inline T atomicExch(T* p, V v) { const T old = *p; *p = (T)v; return old; }
template <class T, class V> inline T atomicMin(T* p, V v) { const T old = *p; if ((T)v < old) *p = (T)v; return old; }
template <class T, class V> inline T atomicMax(T* p, V v) { const T old = *p; if ((T)v > old) *p = (T)v; return old; }
inline int atomicCAS(int* p, int cmp, int val) { const int old = *p; if (old == cmp) *p = val; return old; }
template <class T> inline T __ldg(const T* p) { return *p; }
inline int __clz(int x) { return x == 0 ? 32 : __builtin_clz((unsigned)x); }
inline int __clzll(long long x) { return x == 0 ? 64 : __builtin_clzll((unsigned long long)x); }
inline int __popc(unsigned x) { return __builtin_popcount(x); }
inline int __ffs(unsigned x) { return __builtin_ffs((int)x); }
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
// DPX (CUDA math API): three-operand add / min / max with optional clamp at zero; 32-bit wrap-around addition like the hardware's
inline int dpx_add(int a, int b) { return (int)((unsigned)a + (unsigned)b); }
inline int __viaddmax_s32(int a, int b, int c) { const int s = dpx_add(a, b); return s > c ? s : c; }
inline int __viaddmax_s32_relu(int a, int b, int c) { const int m = __viaddmax_s32(a, b, c); return m > 0 ? m : 0; }
inline int __viaddmin_s32_relu(int a, int b, int c) { const int s = dpx_add(a, b), m = s < c ? s : c; return m > 0 ? m : 0; }
inline int __vimax3_s32(int a, int b, int c) { const int m = a > b ? a : b; return m > c ? m : c; }
using std::min;
using std::max;
