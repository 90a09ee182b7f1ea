// Per-batch work counters and the first-error slot every kernel of the batch reports into (read back by batch.inl).
#pragma once
#include <cstdint>
namespace arachne {

struct Counters {
    unsigned long long occ, occ_blocks, sa_lookups, lf_steps, ext_calls, ext_cells, glb_cells;
    int err;       // first internal error code (negative) raised by a kernel
    int err_read;
};

__device__ __forceinline__ void raise(Counters* c, int code, int read) {
    if (atomicCAS(&c->err, 0, code) == 0) c->err_read = read;
}

}  // namespace arachne
