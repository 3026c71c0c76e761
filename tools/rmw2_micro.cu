// Micro-benchmark (not product code): RMW of random 512 B rows + one gradient row (L2-resident
// slice) per row + 4-byte state RMW: how close to the pure RMW ceiling can the update get?
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

template <int LPR, int CPL, int U, int MINB, int THREADS, bool GRAD, bool STATE, bool KEYS>
__global__ void __launch_bounds__(THREADS, MINB) k_upd(uint4* __restrict__ w, float* __restrict__ h, const uint4* __restrict__ grad,
                                                    const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                                    uint32_t rows, uint32_t n, uint32_t gmask_rows) {
    constexpr int RPW = 32 / LPR;
    const int lane = threadIdx.x & 31, sub = lane % LPR, grp = lane / LPR;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint32_t nwarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t i0 = warp * (RPW * U); i0 < n; i0 += nwarps * (RPW * U)) {
        uint4 v[U][CPL], g[U][CPL];
        float hs[U];
        size_t base[U]; uint32_t rowv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint32_t i = min(i0 + u * RPW + grp, n - 1);
            uint32_t row, gr;
            if (KEYS) { row = __ldg(keys + i); gr = __ldg(vals + i) & gmask_rows; }
            else { row = (uint32_t)(((uint64_t)i * rows) / n); gr = (i * 2654435761u) & gmask_rows; }
            rowv[u] = row;
            base[u] = (size_t)row * 32;
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                v[u][k] = __ldcg(w + base[u] + sub + k * LPR);
                if (GRAD) g[u][k] = __ldg(grad + (size_t)gr * 32 + sub + k * LPR);
            }
            if (STATE) hs[u] = __ldcg(h + row);
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            float ss = 0.f;
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                uint4 x = v[u][k];
                if (GRAD) { x.x += g[u][k].x; x.y ^= g[u][k].y; x.z += g[u][k].z; x.w ^= g[u][k].w; ss += __uint_as_float(g[u][k].x); }
                else { x.x += 1; x.y ^= x.x; x.z += x.y; x.w ^= 3; }
                v[u][k] = x;
            }
            if (STATE) {
#pragma unroll
                for (int o = 1; o < LPR; o <<= 1) ss += __shfl_xor_sync(0xffffffffu, ss, o);
                if (sub == 0) h[rowv[u]] = hs[u] + ss;
            }
#pragma unroll
            for (int k = 0; k < CPL; ++k) w[base[u] + sub + k * LPR] = v[u][k];
        }
    }
}

__global__ void fill_keys(uint32_t* keys, uint32_t* vals, uint32_t rows, uint32_t n) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        keys[i] = (uint32_t)(((uint64_t)i * rows) / n);
        vals[i] = i * 2654435761u;
    }
}

int main() {
    const uint32_t rows = 26000000u, n = 7264000u;
    uint4* w; CK(cudaMalloc(&w, (size_t)rows * 512)); CK(cudaMemset(w, 1, (size_t)rows * 512));
    float* h; CK(cudaMalloc(&h, (size_t)rows * 4)); CK(cudaMemset(h, 0, (size_t)rows * 4));
    const uint32_t grows = 16384;   // 8 MB gradient slice: stays in L2
    uint4* grad; CK(cudaMalloc(&grad, (size_t)grows * 512)); CK(cudaMemset(grad, 1, (size_t)grows * 512));
    uint32_t *keys, *vals; CK(cudaMalloc(&keys, n * 4)); CK(cudaMalloc(&vals, n * 4));
    fill_keys<<<592, 256>>>(keys, vals, rows, n);
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double bytes = (double)n * 1024;
    auto timeit = [&](const char* name, auto launch) {
        for (int i = 0; i < 2; ++i) launch();
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 10; ++i) launch();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1)); CK(cudaGetLastError());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 10;
        printf("%-50s %8.3f ms  %7.1f GB/s (row read+write only)\n", name, ms, bytes / ms / 1e6);
    };
#define RUN(name, LPR, CPL, U, MINB, TH, G, S, K) timeit(name, [&] { k_upd<LPR, CPL, U, MINB, TH, G, S, K><<<148 * MINB, TH>>>(w, h, grad, keys, vals, rows, n, grows - 1); })
    RUN("rmw only        L8 C4 U4 3x256", 8, 4, 4, 3, 256, false, false, false);
    RUN("rmw+state       L8 C4 U4 3x256", 8, 4, 4, 3, 256, false, true, false);
    RUN("rmw+grad        L8 C4 U2 3x256", 8, 4, 2, 3, 256, true, false, false);
    RUN("rmw+grad        L8 C4 U2 4x256", 8, 4, 2, 4, 256, true, false, false);
    RUN("rmw+grad        L8 C4 U4 2x256", 8, 4, 4, 2, 256, true, false, false);
    RUN("rmw+grad+state  L8 C4 U2 3x256", 8, 4, 2, 3, 256, true, true, false);
    RUN("rmw+grad+state  L8 C4 U2 4x256", 8, 4, 2, 4, 256, true, true, false);
    RUN("rmw+grad+state+keys L8 C4 U2 3x256", 8, 4, 2, 3, 256, true, true, true);
    RUN("rmw+grad+state+keys L8 C4 U2 4x256", 8, 4, 2, 4, 256, true, true, true);
    RUN("rmw+grad+state+keys L8 C4 U1 4x256", 8, 4, 1, 4, 256, true, true, true);
    RUN("rmw+grad+state+keys L8 C4 U1 6x256", 8, 4, 1, 6, 256, true, true, true);
    RUN("rmw+grad+state+keys L16 C2 U2 4x256", 16, 2, 2, 4, 256, true, true, true);
    RUN("rmw+grad+state+keys L16 C2 U4 3x256", 16, 2, 4, 3, 256, true, true, true);
    return 0;
}
