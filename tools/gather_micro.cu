// Micro-benchmark (not product code): what does a random 512-byte-row gather reach on B200,
// and which in-flight mechanism gets there?  nvcc -arch=sm_100a -O3 -o gather_micro gather_micro.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <vector>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1);} } while (0)

__device__ __forceinline__ uint4 ldg_stream(const uint4* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ float4 f4(uint4 r) { return make_float4(__uint_as_float(r.x), __uint_as_float(r.y), __uint_as_float(r.z), __uint_as_float(r.w)); }

// ---- K0: no dependent index load: rows derived from a hash; U loads in flight per warp
template <int U>
__global__ void k_ideal(const uint4* __restrict__ w, uint32_t rows, int64_t bags, int P, float4* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t g = warp; g < bags; g += nwarps) {
        float4 acc = make_float4(0, 0, 0, 0);
        for (int j0 = 0; j0 < P; j0 += U) {
            uint4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                uint32_t h = (uint32_t)(g * P + j0 + u) * 2654435761u;
                h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
                uint32_t row = h % rows;
                v[u] = (j0 + u < P) ? ldg_stream(w + (size_t)row * 32 + lane) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) { float4 x = f4(v[u]); acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
        }
        out[g * 32 + lane] = acc;
    }
}

// ---- K1: the shipped structure: warp per bag, offsets -> indices -> rows (U=8)
__global__ void __launch_bounds__(256) k_v1(const uint4* __restrict__ w, const int64_t* __restrict__ idx, const int64_t* __restrict__ off,
                                             int64_t bags, float4* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t g = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (g >= bags) return;
    const int64_t s = off[g], e = off[g + 1];
    float4 acc = make_float4(0, 0, 0, 0);
    for (int64_t base = s; base < e; base += 32) {
        const int n = (int)min((int64_t)32, e - base);
        uint32_t my = lane < n ? (uint32_t)idx[base + lane] : 0;
        for (int j0 = 0; j0 < n; j0 += 8) {
            uint4 v[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                uint32_t row = __shfl_sync(0xffffffffu, my, (j0 + u) & 31);
                v[u] = (j0 + u < n) ? ldg_stream(w + (size_t)row * 32 + lane) : make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) { float4 x = f4(v[u]); acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
        }
    }
    out[g * 32 + lane] = acc;
}

// ---- K2: persistent warps, next bag's offsets+indices prefetched into registers while the
// current bag's rows are in flight; U rows per batch
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) k_prefetch(const uint4* __restrict__ w, const int64_t* __restrict__ idx,
                                                         const int64_t* __restrict__ off, int64_t bags, float4* __restrict__ out) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    int64_t g = warp;
    if (g >= bags) return;
    int64_t s = off[g], e = off[g + 1];
    uint32_t my = (s + lane < e) ? (uint32_t)idx[s + lane] : 0;
    while (g < bags) {
        const int64_t gn = g + nwarps;
        int64_t sn = 0, en = 0;
        uint32_t myn = 0;
        if (gn < bags) { sn = off[gn]; en = off[gn + 1]; myn = (sn + lane < en) ? (uint32_t)idx[sn + lane] : 0; }
        float4 acc = make_float4(0, 0, 0, 0);
        for (int64_t base = s; base < e; base += 32) {
            const int n = (int)min((int64_t)32, e - base);
            if (base != s) my = lane < n ? (uint32_t)idx[base + lane] : 0;
            for (int j0 = 0; j0 < n; j0 += U) {
                uint4 v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    uint32_t row = __shfl_sync(0xffffffffu, my, (j0 + u) & 31);
                    v[u] = (j0 + u < n) ? ldg_stream(w + (size_t)row * 32 + lane) : make_uint4(0, 0, 0, 0);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) { float4 x = f4(v[u]); acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
            }
        }
        out[g * 32 + lane] = acc;
        g = gn; s = sn; e = en; my = myn;
    }
}

// ---- K3: rows land in shared memory through per-lane bulk async copies (UBLKCP) + mbarrier,
// two stages per warp: bag k+1 in flight while bag k is summed from smem.
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(b)), "r"(count)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nWAIT_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra DONE_%=;\nbra WAIT_%=;\nDONE_%=:\n}" :: "r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

template <int WARPS>
__global__ void __launch_bounds__(WARPS * 32, 1) k_bulk(const uint4* __restrict__ w, const int64_t* __restrict__ idx,
                                                         const int64_t* __restrict__ off, int64_t bags, float4* __restrict__ out) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    uint4* stage[2] = {reinterpret_cast<uint4*>(smem + (size_t)wid * 32768), reinterpret_cast<uint4*>(smem + (size_t)wid * 32768 + 16384)};
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + (size_t)WARPS * 32768) + wid * 2;
    if (lane == 0) { mbar_init(bars, 1); mbar_init(bars + 1, 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncwarp();
    const int64_t warp = (int64_t)blockIdx.x * WARPS + wid;
    const int64_t nwarps = (int64_t)gridDim.x * WARPS;
    int64_t g = warp;
    if (g >= bags) return;
    uint32_t phase[2] = {0, 0};
    auto issue = [&](int st, int64_t s, int64_t e) -> int {   // first chunk (<=32 rows) of a bag
        const int n = (int)min((int64_t)32, e - s);
        if (lane == 0) mbar_expect(bars + st, (uint32_t)n * 512u);
        __syncwarp();
        if (lane < n) {
            const uint32_t row = (uint32_t)idx[s + lane];
            bulk_g2s(stage[st] + lane * 32, w + (size_t)row * 32, 512, bars + st);
        }
        return n;
    };
    int64_t s = off[g], e = off[g + 1];
    int st = 0;
    int n = issue(st, s, e);
    int64_t gn = g + nwarps;
    int64_t sn = 0, en = 0;
    if (gn < bags) { sn = off[gn]; en = off[gn + 1]; }
    while (g < bags) {
        int nn = 0;
        if (gn < bags) nn = issue(st ^ 1, sn, en);
        const int64_t gnn = gn + nwarps;
        int64_t snn = 0, enn = 0;
        if (gnn < bags) { snn = off[gnn]; enn = off[gnn + 1]; }
        mbar_wait(bars + st, phase[st]); phase[st] ^= 1;
        float4 acc = make_float4(0, 0, 0, 0);
        for (int j = 0; j < n; ++j) { float4 x = f4(stage[st][j * 32 + lane]); acc.x += x.x; acc.y += x.y; acc.z += x.z; acc.w += x.w; }
        // (bags longer than 32 are not exercised by this micro-benchmark)
        out[g * 32 + lane] = acc;
        __syncwarp();
        g = gn; s = sn; e = en; n = nn; gn = gnn; sn = snn; en = enn; st ^= 1;
    }
}

__global__ void fill_random(float* p, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        uint32_t h = (uint32_t)i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        p[i] = (float)(h & 0xffffff) / 16777216.f - 0.5f;
    }
}

int main(int argc, char** argv) {
    const int F = 26, P = 20; const int64_t B = 16384; const uint32_t rows_per = 1000000;
    const int64_t bags = (int64_t)F * B, L = bags * P; const uint32_t rows = F * rows_per;
    uint4* w; CK(cudaMalloc(&w, (size_t)rows * 512));
    if (argc > 1) { printf("table filled with pseudo-random floats\n"); fill_random<<<148 * 16, 256>>>((float*)w, (size_t)rows * 128); CK(cudaDeviceSynchronize()); }
    else { printf("table zero-filled\n"); CK(cudaMemset(w, 0, (size_t)rows * 512)); }
    std::vector<int64_t> hidx(L), hoff(bags + 1);
    uint64_t x = 88172645463325252ull;
    for (int64_t i = 0; i < L; ++i) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; hidx[i] = (int64_t)((i / (B * P)) * rows_per + x % rows_per); }
    for (int64_t i = 0; i <= bags; ++i) hoff[i] = i * P;
    int64_t *idx, *off; float4* out;
    CK(cudaMalloc(&idx, L * 8)); CK(cudaMalloc(&off, (bags + 1) * 8)); CK(cudaMalloc(&out, bags * 512));
    CK(cudaMemcpy(idx, hidx.data(), L * 8, cudaMemcpyHostToDevice)); CK(cudaMemcpy(off, hoff.data(), (bags + 1) * 8, cudaMemcpyHostToDevice));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    const double bytes = (double)L * 520 + bags * 520.0;
    auto timeit = [&](const char* name, auto launch) {
        for (int i = 0; i < 3; ++i) launch();
        CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 10; ++i) launch();
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        CK(cudaGetLastError());
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 10;
        printf("%-34s %8.3f ms  %7.1f GB/s\n", name, ms, bytes / ms / 1e6);
    };
    timeit("ideal U=4  occ=full", [&] { k_ideal<4><<<148 * 8, 256>>>(w, rows, bags, P, out); });
    timeit("ideal U=8  grid=148*6", [&] { k_ideal<8><<<148 * 6, 256>>>(w, rows, bags, P, out); });
    timeit("ideal U=10 grid=148*4", [&] { k_ideal<10><<<148 * 4, 256>>>(w, rows, bags, P, out); });
    timeit("ideal U=20 grid=148*2", [&] { k_ideal<20><<<148 * 2, 256>>>(w, rows, bags, P, out); });
    timeit("ideal U=20 grid=148*3", [&] { k_ideal<20><<<148 * 3, 256>>>(w, rows, bags, P, out); });
    timeit("v1 (shipped)", [&] { k_v1<<<(unsigned)((bags + 7) / 8), 256>>>(w, idx, off, bags, out); });
    timeit("prefetch U=8 minb=3", [&] { k_prefetch<8, 3><<<148 * 3, 256>>>(w, idx, off, bags, out); });
    timeit("prefetch U=4 minb=6", [&] { k_prefetch<4, 6><<<148 * 6, 256>>>(w, idx, off, bags, out); });
    timeit("prefetch U=10 minb=3", [&] { k_prefetch<10, 3><<<148 * 3, 256>>>(w, idx, off, bags, out); });
    timeit("prefetch U=20 minb=2", [&] { k_prefetch<20, 2><<<148 * 2, 256>>>(w, idx, off, bags, out); });
    {
        constexpr int W = 6; const size_t sm = (size_t)W * 32768 + W * 16;
        CK(cudaFuncSetAttribute(k_bulk<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        timeit("bulk smem 6 warps x 2 stages", [&] { k_bulk<W><<<148, W * 32, sm>>>(w, idx, off, bags, out); });
    }
    {
        constexpr int W = 3; const size_t sm = (size_t)W * 32768 + W * 16;
        CK(cudaFuncSetAttribute(k_bulk<W>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
        timeit("bulk smem 3 warps x 2 st, 2 CTA/SM", [&] { k_bulk<W><<<148 * 2, W * 32, sm>>>(w, idx, off, bags, out); });
    }
    // plain streaming copy for reference
    {
        float4* a; float4* b; size_t n = (size_t)1 << 28;  // 4 GiB each
        CK(cudaMalloc(&a, n * 16)); CK(cudaMalloc(&b, n * 16));
        CK(cudaMemset(a, 0, n * 16));
        for (int i = 0; i < 2; ++i) CK(cudaMemcpy(b, a, n * 16, cudaMemcpyDeviceToDevice));
        CK(cudaEventRecord(e0));
        for (int i = 0; i < 5; ++i) CK(cudaMemcpyAsync(b, a, n * 16, cudaMemcpyDeviceToDevice));
        CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); ms /= 5;
        printf("%-34s %8.3f ms  %7.1f GB/s (read+write)\n", "cudaMemcpy D2D 4 GiB", ms, 2.0 * n * 16 / ms / 1e6);
    }
    return 0;
}
