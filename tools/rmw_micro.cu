// Micro-benchmark (not product code): ceiling for "read a random 512 B row, modify, write it back".
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

// LPR lanes per row, CPL 16-byte chunks per lane, U rows in flight per group
template <int LPR, int CPL, int U, int MINB, bool SORTED>
__global__ void __launch_bounds__(256, MINB) k_rmw(uint4* __restrict__ w, uint32_t rows, uint32_t n, uint32_t stride_mul) {
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31, sub = lane % LPR, grp = lane / LPR;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i0 = warp * (RPW * U); i0 < n; i0 += nwarps * (RPW * U)) {
        uint4 v[U][CPL];
        size_t base[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t i = i0 + u * RPW + grp;
            // SORTED: increasing rows with random gaps (what the update kernel sees); else a permutation
            const uint32_t row = SORTED ? (uint32_t)(((uint64_t)i * rows) / n) : (uint32_t)(((uint64_t)i * stride_mul) % rows);
            base[u] = (size_t)row * 32;
#pragma unroll
            for (int k = 0; k < CPL; ++k) v[u][k] = __ldcg(w + base[u] + sub + k * LPR);
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                uint4 x = v[u][k];
                x.x += 1; x.y ^= x.x; x.z += x.y; x.w ^= 3;
                w[base[u] + sub + k * LPR] = x;
            }
    }
}

int main() {
    const uint32_t rows = 26000000u, n = 7264000u;
    uint4* w; CK(cudaMalloc(&w, (size_t)rows * 512)); CK(cudaMemset(w, 1, (size_t)rows * 512));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double bytes = (double)n * 1024;
    auto timeit = [&](const char* name, auto launch) {
        for (int i = 0; i < 2; ++i) launch();
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 10; ++i) launch();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 10;
        printf("%-44s %8.3f ms  %7.1f GB/s (read+write)\n", name, ms, bytes / ms / 1e6);
    };
    const uint32_t mul = 2654435761u % rows | 1;
    timeit("perm  LPR32 CPL1 U4 3cta", [&] { k_rmw<32, 1, 4, 3, false><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("perm  LPR32 CPL1 U8 3cta", [&] { k_rmw<32, 1, 8, 3, false><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("perm  LPR32 CPL1 U8 6cta", [&] { k_rmw<32, 1, 8, 6, false><<<148 * 6, 256>>>(w, rows, n, mul); });
    timeit("perm  LPR8  CPL4 U2 4cta", [&] { k_rmw<8, 4, 2, 4, false><<<148 * 4, 256>>>(w, rows, n, mul); });
    timeit("perm  LPR8  CPL4 U4 3cta", [&] { k_rmw<8, 4, 4, 3, false><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR32 CPL1 U8 3cta", [&] { k_rmw<32, 1, 8, 3, true><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR32 CPL1 U8 6cta", [&] { k_rmw<32, 1, 8, 6, true><<<148 * 6, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR32 CPL1 U16 3cta", [&] { k_rmw<32, 1, 16, 3, true><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR8  CPL4 U2 4cta", [&] { k_rmw<8, 4, 2, 4, true><<<148 * 4, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR8  CPL4 U4 3cta", [&] { k_rmw<8, 4, 4, 3, true><<<148 * 3, 256>>>(w, rows, n, mul); });
    timeit("sorted LPR8  CPL4 U4 6cta", [&] { k_rmw<8, 4, 4, 6, true><<<148 * 6, 256>>>(w, rows, n, mul); });
    return 0;
}
