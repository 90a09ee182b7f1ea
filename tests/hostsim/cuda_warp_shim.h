// TEST-ONLY: one WARP -- or one whole BLOCK of up to 8 warps -- of a kernel on the host, for kernels whose lanes cooperate (shuffles, ballots,
// warp reductions, one lane writing another lane's shared-memory column between warp barriers: k_ext_dp).  Every lane is a fiber (ucontext) of
// one host thread with its own stack = its own registers; a collective (__shfl_sync, __ballot_sync, __reduce_*_sync, __syncwarp -- always with
// the full mask in the kernels run here -- and __syncthreads) is a rendezvous of its group, the lane's warp or the block: a lane publishes its
// operand, yields until the whole group has arrived, and reads the others'.  Between collectives a lane runs alone, which is one of the
// interleavings the hardware allows; shared memory is ordinary memory, ordered by the rendezvous exactly where the kernel has a barrier.  The
// order in which lanes run between two barriers is selectable (lane_order): running the tests under more than one order catches a barrier
// missing before or after a leader's write, and under run_block the warps of a block interleave at every collective (queue heads, pool
// cursors and claim counters are then really shared).  It is not a race detector.  g++ compiles the kernel source unchanged.  Never shipped.
#pragma once
#include <ucontext.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <functional>
#include <memory>
#include <vector>
#include "cuda_shim_common.h"

namespace warpsim {
constexpr int W = 32;
constexpr int MAXB = 256;  // run_block: a whole block of up to 256 lanes = 8 warps
struct Lane { ucontext_t ctx; std::unique_ptr<char[]> stack; LaneIdx3 tid; bool done = false; };  // stacks are not zero-filled: only touched pages cost
struct Group {                      // the lanes that meet in one kind of collective: a warp, or the whole block
    unsigned long long xchg[2][MAXB];  // operands of collective number g live in xchg[g & 1]
    long gen = 0;                   // collectives completed by all lanes of the group
    int arrived = 0, size = 0;
};
struct Warp {                       // (a block of one or more warps)
    Lane lane[MAXB];
    int n = W;                      // lanes running
    ucontext_t sched;
    int cur = 0;
    Group warp[MAXB / W], block;
    long lane_gen[2][MAXB];         // collectives each lane has entered: [0] warp level, [1] block level
    long progress = 0;              // rendezvous entered or completed, lanes finished: a pass without any is a deadlock
    std::function<void()> body;
};
inline Warp*& cur_warp() { static Warp* w = nullptr; return w; }
// Which lane runs first between two collectives: 0 = lane 0 upwards (a leader's writes land before the others read), 1 = lane 31 downwards (the
// others read before the leader writes), 2 = the starting lane rotates, >= 3 = a random permutation per pass (the number seeds it).  A kernel whose barriers are complete gives the same results under all
// three; one that leans on lock-step execution of a converged warp does not (hs_warp_order(), WARPSIM_ORDER).
inline int& lane_order() { static int o = getenv("WARPSIM_ORDER") ? atoi(getenv("WARPSIM_ORDER")) : 0; return o; }
inline void yield() { Warp* w = cur_warp(); swapcontext(&w->lane[w->cur].ctx, &w->sched); }
// enter collective: publish v, wait for the others; returns the buffer holding this collective's operands (indexed by lane within the group)
inline const unsigned long long* rendezvous(unsigned long long v, bool warp_level = true) {
    Warp* w = cur_warp();
    const int me = w->cur;
    Group& G = warp_level ? w->warp[me / W] : w->block;
    const int idx = warp_level ? me % W : me;
    const long g = w->lane_gen[warp_level ? 0 : 1][me]++;
    if (g != G.gen) { fprintf(stderr, "warpsim: lane %d entered %s collective %ld while its group is at %ld (divergent collective)\n", me, warp_level ? "warp" : "block", g, G.gen); abort(); }
    G.xchg[g & 1][idx] = v;
    ++w->progress;
    if (++G.arrived == G.size) { G.arrived = 0; ++G.gen; }
    while (G.gen <= g) yield();
    return G.xchg[g & 1];
}
inline int lane_in_warp() { return cur_warp()->cur % W; }
inline void lane_entry() { Warp* w = cur_warp(); w->body(); w->lane[w->cur].done = true; ++w->progress; yield(); }
// run `body` as n lanes with threadIdx.x = first_tid .. first_tid + n - 1
inline void run_lanes(unsigned first_tid, int n, size_t stack_bytes, const std::function<void()>& body) {
    Warp* w = new Warp();
    cur_warp() = w;
    w->body = body; w->n = n;
    w->block.size = n;
    for (int g = 0; g * W < n; ++g) w->warp[g].size = n - g * W < W ? n - g * W : W;
    for (int l = 0; l < n; ++l) {
        Lane& L = w->lane[l];
        L.stack.reset(new char[stack_bytes]); L.tid.x = first_tid + (unsigned)l; w->lane_gen[0][l] = w->lane_gen[1][l] = 0;
        getcontext(&L.ctx);
        L.ctx.uc_stack.ss_sp = L.stack.get(); L.ctx.uc_stack.ss_size = stack_bytes; L.ctx.uc_link = &w->sched;
        makecontext(&L.ctx, lane_entry, 0);
    }
    long pass = 0; int perm[MAXB];
    for (;;) {
        int live = 0;
        if (lane_order() >= 3) {  // a fresh random permutation per pass; the order number is the seed
            static unsigned long long rng = 0;
            if (!rng) rng = 0x9e3779b97f4a7c15ull * (unsigned long long)lane_order();
            for (int i = 0; i < n; ++i) perm[i] = i;
            for (int i = n - 1; i > 0; --i) { rng = rng * 6364136223846793005ull + 1442695040888963407ull; std::swap(perm[i], perm[(rng >> 33) % (unsigned)(i + 1)]); }
        }
        for (int i = 0; i < n; ++i) {
            const int l = lane_order() == 1 ? n - 1 - i : lane_order() == 2 ? (int)((i + pass * 7) % n) : lane_order() >= 3 ? perm[i] : i;
            if (w->lane[l].done) continue;
            ++live; w->cur = l;
            swapcontext(&w->sched, &w->lane[l].ctx);
        }
        ++pass;
        if (!live) break;
        if (w->progress == 0) {
            int finished = 0;
            for (int l = 0; l < n; ++l) finished += w->lane[l].done;
            fprintf(stderr, "warpsim: deadlock: %d of %d lanes wait in collectives that the others (%d finished) will never enter\n", live, n, finished); abort();
        }
        w->progress = 0;
    }
    cur_warp() = nullptr;
    delete w;
}
inline void run_warp(unsigned first_tid, const std::function<void()>& body) { run_lanes(first_tid, W, 1 << 20, body); }
// a whole block: warp collectives meet within each warp of 32, __syncthreads across all lanes; the warps interleave at every collective
inline void run_block(int n_threads, const std::function<void()>& body, size_t stack_bytes = 1 << 17) {
    if (n_threads > MAXB) { fprintf(stderr, "warpsim: run_block of %d lanes\n", n_threads); abort(); }
    run_lanes(0, n_threads, stack_bytes, body);
}
}  // namespace warpsim

static LaneIdx3 blockIdx;
static LaneDim3 blockDim, gridDim;
#define threadIdx (warpsim::cur_warp()->lane[warpsim::cur_warp()->cur].tid)
#ifndef __shared__
#define __shared__            /* `extern __shared__ T name[]` names a global array the test defines.  A kernel with block-scope __shared__
                                 variables needs one copy per BLOCK: a test that runs such a kernel (one block at a time) defines __shared__
                                 as `static` before including this header (hostsim_place.cpp) */
#endif

inline void __syncwarp(unsigned = 0xffffffffu) { warpsim::rendezvous(0); }
inline void __syncthreads() { warpsim::rendezvous(0, false); }   // all lanes that run: the warp under run_warp (a kernel of one warp per block, or whose warps never meet), the block under run_block
inline unsigned __ballot_sync(unsigned, bool p) {
    const unsigned long long* x = warpsim::rendezvous(p ? 1u : 0u);
    unsigned m = 0;
    for (int l = 0; l < warpsim::W; ++l) m |= (unsigned)(x[l] & 1u) << l;
    return m;
}
inline bool __all_sync(unsigned m, bool p) { return __ballot_sync(m, p) == 0xffffffffu; }
inline bool __any_sync(unsigned m, bool p) { return __ballot_sync(m, p) != 0u; }
template <class T> inline T __shfl_sync(unsigned, T v, int src) {
    static_assert(sizeof(T) <= 8, "shuffle operand");
    unsigned long long raw = 0; memcpy(&raw, &v, sizeof(T));
    const unsigned long long* x = warpsim::rendezvous(raw);
    T out; memcpy(&out, &x[src & 31], sizeof(T));
    return out;
}
inline int __reduce_max_sync(unsigned, int v) {
    const unsigned long long* x = warpsim::rendezvous((unsigned long long)(long long)v);
    int m = (int)(long long)x[0];
    for (int l = 1; l < warpsim::W; ++l) m = std::max(m, (int)(long long)x[l]);
    return m;
}
inline unsigned __reduce_add_sync(unsigned, unsigned v) {
    const unsigned long long* x = warpsim::rendezvous(v);
    unsigned m = 0;
    for (int l = 0; l < warpsim::W; ++l) m += (unsigned)x[l];
    return m;
}
template <class T> inline T __shfl_up_sync(unsigned, T v, unsigned delta) {  // lanes below delta keep their own value
    static_assert(sizeof(T) <= 8, "shuffle operand");
    unsigned long long raw = 0; memcpy(&raw, &v, sizeof(T));
    const int me = warpsim::lane_in_warp();
    const unsigned long long* x = warpsim::rendezvous(raw);
    T out; memcpy(&out, &x[me >= (int)delta ? me - (int)delta : me], sizeof(T));
    return out;
}
inline int __reduce_min_sync(unsigned, int v) {
    const unsigned long long* x = warpsim::rendezvous((unsigned long long)(long long)v);
    int m = (int)(long long)x[0];
    for (int l = 1; l < warpsim::W; ++l) m = std::min(m, (int)(long long)x[l]);
    return m;
}
template <class T> inline T __shfl_down_sync(unsigned, T v, unsigned delta) {  // lanes whose source is beyond the warp keep their own value
    static_assert(sizeof(T) <= 8, "shuffle operand");
    unsigned long long raw = 0; memcpy(&raw, &v, sizeof(T));
    const int me = warpsim::lane_in_warp();
    const unsigned long long* x = warpsim::rendezvous(raw);
    T out; memcpy(&out, &x[me + (int)delta < warpsim::W ? me + (int)delta : me], sizeof(T));
    return out;
}
