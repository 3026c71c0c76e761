// Block-level scan helpers shared by the scan / sort / bucketize kernels.
#pragma once
#include "common.cuh"

namespace tdfo {

// Exclusive scan of one value per thread over a 256-thread block.
// `warp_totals` is shared memory for 8 values of T.  Returns the exclusive
// prefix; `total` receives the block total in every thread.  Ends with the
// shared buffer free for reuse (trailing __syncthreads).
template <typename T>
__device__ __forceinline__ T block_exclusive_scan_256(T v, T* warp_totals, T& total) {
    const int lane = lane_id();
    const int warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T up = __shfl_up_sync(kFullMask, incl, o);
        if (lane >= o) incl += up;
    }
    if (lane == 31) warp_totals[warp] = incl;
    __syncthreads();
    T before = 0, all = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const T t = warp_totals[w];
        if (w < warp) before += t;
        all += t;
    }
    total = all;
    __syncthreads();
    return before + incl - v;
}

}  // namespace tdfo
