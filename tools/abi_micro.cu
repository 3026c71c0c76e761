// Calls the shipped library kernels through the C ABI on the micro-benchmark's inputs.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>
#include "../include/tdfo_b200.h"
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)
#define TK(x) do { int rc = (x); if (rc) { printf("tdfo %d %s at %d\n", rc, tdfo_last_error(), __LINE__); exit(1);} } while (0)

int main(int argc, char** argv) {
    const int F = 26, P = 20; const int64_t B = 16384; const int64_t rows_per = 1000000;
    const int64_t bags = (int64_t)F * B, L = bags * P; const int64_t rows = F * rows_per;
    void* w; CK(cudaMalloc(&w, (size_t)rows * 512)); CK(cudaMemset(w, 0, (size_t)rows * 512));
    std::vector<int64_t> hidx(L), hoff(bags + 1), hbase(F), hrows(F), hcol(F);
    uint64_t x = 88172645463325252ull;
    for (int64_t i = 0; i < L; ++i) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; hidx[i] = (int64_t)(x % rows_per); }
    for (int64_t i = 0; i <= bags; ++i) hoff[i] = i * P;
    for (int f = 0; f < F; ++f) { hbase[f] = f * rows_per; hrows[f] = rows_per; hcol[f] = f * 128; }
    int64_t *idx, *off, *base, *frows, *col; float* out; int32_t* flag;
    CK(cudaMalloc(&idx, L * 8)); CK(cudaMalloc(&off, (bags + 1) * 8)); CK(cudaMalloc(&out, bags * 512));
    CK(cudaMalloc(&base, F * 8)); CK(cudaMalloc(&frows, F * 8)); CK(cudaMalloc(&col, F * 8)); CK(cudaMalloc(&flag, 4)); CK(cudaMemset(flag, 0, 4));
    CK(cudaMemcpy(idx, hidx.data(), L * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(off, hoff.data(), (bags + 1) * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(base, hbase.data(), F * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(frows, hrows.data(), F * 8, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(col, hcol.data(), F * 8, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double bytes = (double)L * 520 + bags * 520.0;
    auto timeit = [&](const char* name, auto launch) {
        for (int i = 0; i < 3; ++i) launch();
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 10; ++i) launch();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 10;
        printf("%-44s %8.3f ms  %7.1f GB/s\n", name, ms, bytes / ms / 1e6);
    };
    timeit("lib forward, out [B, F*D]", [&] { TK(tdfo_embedding_forward(w, TDFO_F32, rows, 128, idx, off, nullptr, base, frows, col, F, B, L, TDFO_POOL_SUM, out, F * 128, nullptr, nullptr, flag, nullptr)); });
    // same work as ONE feature over F*B bags: contiguous output rows, no per-feature division effects
    { int64_t one_rows = rows; int64_t zero = 0; CK(cudaMemcpy(frows, &one_rows, 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(col, &zero, 8, cudaMemcpyHostToDevice)); }
    timeit("lib forward, 1 feature, out [F*B, D]", [&] { TK(tdfo_embedding_forward(w, TDFO_F32, rows, 128, idx, off, nullptr, base, frows, col, 1, bags, L, TDFO_POOL_SUM, out, 128, nullptr, nullptr, flag, nullptr)); });
    return 0;
}
