// Kernels around k_ext_dp for arachne_ksw_extend_static() (noalloc.inl): validate the caller's problems, counting-sort them by query length into
// the static scratch's queue (longest first), and write the results back in the caller's order.  In a header of their own so that tests/hostsim
// can compile the same text for the host.
#pragma once
#include "ext_stage.cuh"

namespace arachne {

struct XsCtl {                // control block, cleared by every call
    unsigned n_items, head;   // k_ext_dp's queue length / queue head
    int status, pad;          // 0, or 1 + index of the lowest-numbered problem the kernels refused
    unsigned bin[256];        // problems per query length
    unsigned base[256];       // first queue slot of a query length (longest first)
    unsigned cursor[256];     // slots of a query length handed out so far
    Counters cnt;
};

struct XsIn {
    long n; const int64_t* q_off; const int64_t* t_off; const int32_t* h0; const int32_t* w;
    int end_bonus, max_qlen, a;
};

__device__ __forceinline__ bool xs_valid(const XsIn& I, long i, int* ql, int* tl) {
    const int64_t q = I.q_off[i + 1] - I.q_off[i], t = I.t_off[i + 1] - I.t_off[i];
    *ql = (int)q; *tl = (int)t;
    // ksw.c:385 asserts h0 > 0; the 13-bit column codec holds scores up to 8191
    return q >= 1 && q <= I.max_qlen && t >= 0 && t <= 0x7fffffff && I.h0[i] > 0 && (long)I.h0[i] + q * I.a <= 8191;
}

__global__ void __launch_bounds__(256) k_xs_count(XsIn I, XsCtl* c, int32_t* out) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < I.n; i += (long)gridDim.x * blockDim.x) {
        int ql, tl;
        if (xs_valid(I, i, &ql, &tl)) atomicAdd(&c->bin[ql], 1u);
        else {
            atomicMax(&c->status, (int)(0x7fffffff - i));  // lowest index wins; k_xs_finish turns it into 1 + index
            for (int k = 0; k < 6; ++k) out[6 * i + k] = -1;
        }
    }
}

__global__ void __launch_bounds__(256) k_xs_base(XsCtl* c) {
    __shared__ unsigned sh[256];
    const int t = threadIdx.x;
    const unsigned v = c->bin[255 - t];  // thread t owns query length 255 - t: longest first
    sh[t] = v;
    __syncthreads();
    for (int d = 1; d < 256; d <<= 1) {
        const unsigned x = t >= d ? sh[t - d] : 0u;
        __syncthreads();
        sh[t] += x;
        __syncthreads();
    }
    c->base[255 - t] = sh[t] - v;
    if (t == 255) { c->n_items = sh[255]; c->head = 0; }
}

__global__ void __launch_bounds__(256) k_xs_queue(XsIn I, XsCtl* c, ExtProb* queue) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < I.n; i += (long)gridDim.x * blockDim.x) {
        int ql, tl;
        if (!xs_valid(I, i, &ql, &tl)) continue;
        ExtProb P;
        P.read = (int32_t)i; P.q0 = 0; P.qlen = ql; P.tlen = tl; P.w = I.w[i]; P.h0 = I.h0[i]; P.end_bonus = I.end_bonus; P.dirs = 0;
        P.tpos = I.t_off[i]; P.qoff = I.q_off[i];
        queue[c->base[ql] + atomicAdd(&c->cursor[ql], 1u)] = P;  // the order inside one length does not matter: results go to res[P.read]
    }
}

__global__ void __launch_bounds__(256) k_xs_finish(XsIn I, XsCtl* c, const ExtRes* res, int32_t* out, int32_t* status, unsigned long long* cells) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < I.n; i += (long)gridDim.x * blockDim.x) {
        int ql, tl;
        if (!xs_valid(I, i, &ql, &tl)) continue;
        const ExtRes R = res[i];
        int32_t* o = out + 6 * i;
        o[0] = R.mx; o[1] = R.max_j + 1; o[2] = R.max_i + 1; o[3] = R.max_ie + 1; o[4] = R.gscore; o[5] = R.max_off;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const int s = c->status;
        *status = s ? (int32_t)(0x7fffffff - s) + 1 : 0;
        if (cells) *cells = c->cnt.ext_cells;
    }
}

}  // namespace arachne
