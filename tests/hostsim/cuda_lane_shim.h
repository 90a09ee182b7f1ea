// TEST-ONLY: the handful of CUDA constructs that kernels made of INDEPENDENT lanes use (k_seed2, k_pack_reads, k_sort_intv, the index
// builders of index_kernels.cuh, the lane-per-region CIGAR tiers k_cigar_fast / k_cigar_band: no lane reads another lane's registers or
// shared memory), defined so that g++ compiles those kernels unchanged and the host runs them ONE LANE AT A TIME.  A kernel "launch" is a
// plain call with threadIdx / blockIdx / blockDim / gridDim set by the caller; warp barriers are no-ops and a warp vote is the lane's own
// predicate, which is exact for one lane.  (Kernels whose lanes DO talk to each other run under cuda_warp_shim.h.)  Never shipped.
#pragma once
#include "cuda_shim_common.h"

static LaneIdx3 threadIdx, blockIdx;
static LaneDim3 blockDim, gridDim;
#define __shared__            /* per-invocation storage: one lane is its whole block */

inline void __syncwarp() {}
inline void __syncthreads() {}
inline bool __all_sync(unsigned, bool p) { return p; }
