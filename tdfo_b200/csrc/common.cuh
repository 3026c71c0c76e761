// Shared device/host helpers for the tdfo_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <atomic>
#include "../../include/tdfo_b200.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "tdfo_b200 kernels are written for sm_100a (B200) only"
#endif

namespace tdfo {

constexpr int kNumSMs = 148;          // B200: 2 dies x 74 SMs
constexpr unsigned kFullMask = 0xffffffffu;

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define TDFO_CHECK_ARG(cond, ...)                         \
    do {                                                  \
        if (!(cond)) {                                    \
            ::tdfo::set_error(__VA_ARGS__);               \
            return TDFO_ERR_INVALID_ARG;                  \
        }                                                 \
    } while (0)

#define TDFO_CUDA(call)                                                    \
    do {                                                                   \
        cudaError_t e__ = (call);                                          \
        if (e__ != cudaSuccess) return ::tdfo::cuda_fail(e__, #call);      \
    } while (0)

#define TDFO_LAUNCH_CHECK(name)                                            \
    do {                                                                   \
        cudaError_t e__ = cudaGetLastError();                              \
        if (e__ != cudaSuccess) return ::tdfo::cuda_fail(e__, name);       \
    } while (0)

inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
inline bool aligned(const void* p, size_t a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; }

// Launch-count bookkeeping: bench.py reports how many of OUR kernels ran.
// (atomic: the ABI allows several host threads to call into the library at once)
extern std::atomic<unsigned long long> g_launch_count;
inline void count_launch(int n = 1) { g_launch_count.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// ---------------------------------------------------------------- device side
__device__ __forceinline__ int lane_id() { return threadIdx.x & 31; }

// Memory access flavours.  These are compiler intrinsics on purpose: an `asm volatile` load
// inside a `cond ? load : 0` makes ptxas wait for each load before issuing the next one
// (seen in SASS as one destination register reused by every LDG), which serialises a gather.
// Table rows are read once per step: streaming (evict-first), no reuse expected.
__device__ __forceinline__ uint4 ldg_stream_128(const void* p) {
    return __ldcs(reinterpret_cast<const uint4*>(p));
}
// 128-bit load that may hit in L1/L2 (gradient rows re-read by several rows).
__device__ __forceinline__ uint4 ldg_128(const void* p) {
    return __ldg(reinterpret_cast<const uint4*>(p));
}
// Plain (coherent) 128-bit load.
__device__ __forceinline__ uint4 ld_128(const void* p) {
    return *reinterpret_cast<const uint4*>(p);
}
// Read-modify-write rows: coherent, cached in L2 only (each row is touched once per step).
__device__ __forceinline__ uint4 ld_row_128(const void* p) {
    return __ldcg(reinterpret_cast<const uint4*>(p));
}
// Streaming store (written once, not re-read by this kernel).
__device__ __forceinline__ void st_stream_128(void* p, uint4 v) {
    __stcs(reinterpret_cast<uint4*>(p), v);
}
__device__ __forceinline__ void st_128(void* p, uint4 v) {
    *reinterpret_cast<uint4*>(p) = v;
}

__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" :: "l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}

// A 16-byte chunk of a table row, widened to fp32 lanes.
template <typename T> struct Chunk;
template <> struct Chunk<float> {
    static constexpr int kElems = 4;
    __device__ __forceinline__ static void unpack(uint4 r, float (&f)[4]) {
        f[0] = __uint_as_float(r.x); f[1] = __uint_as_float(r.y);
        f[2] = __uint_as_float(r.z); f[3] = __uint_as_float(r.w);
    }
    __device__ __forceinline__ static uint4 pack(const float (&f)[4]) {
        return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]),
                          __float_as_uint(f[2]), __float_as_uint(f[3]));
    }
};
template <> struct Chunk<__nv_bfloat16> {
    static constexpr int kElems = 8;
    __device__ __forceinline__ static void unpack(uint4 r, float (&f)[8]) {
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // bf16 -> fp32 is a 16-bit shift
            f[2 * i]     = __uint_as_float(w[i] << 16);
            f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
        }
    }
    __device__ __forceinline__ static uint4 pack(const float (&f)[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {   // round-to-nearest-even
            __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
            w[i] = *reinterpret_cast<uint32_t*>(&h);
        }
        return make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// ------------------------------------------------------------------
// Single-pass prefix scan plumbing (decoupled look-back).
// One 64-bit status word per tile: bits 63..62 = flag, bits 61..0 = value.
// Tiles take their id from an atomic ticket, so every predecessor of a running
// tile is itself running or done: the spin below always makes progress.
constexpr uint64_t kScanFlagAgg  = 1ull << 62;   // value = this tile's own total
constexpr uint64_t kScanFlagIncl = 2ull << 62;   // value = total of tiles 0..t
constexpr uint64_t kScanValueMask = (1ull << 62) - 1;
constexpr int kScanWatchdog = 1 << 24;           // spins before giving up (bug guard, never hit in a correct run)

struct ScanScratch {           // lives at the start of a scan scratch buffer (zeroed before launch)
    unsigned int ticket;
    unsigned int pad[63];
};

// Called by ALL 32 lanes of ONE warp of the block. Returns the exclusive prefix
// (sum of aggregates of tiles 0..tile-1) in every lane.
__device__ __forceinline__ uint64_t scan_lookback(uint64_t* status, int tile, uint64_t aggregate,
                                                  int32_t* err_flag) {
    const int lane = lane_id();
    if (lane == 0)
        st_relaxed_u64(status + tile, (tile == 0 ? kScanFlagIncl : kScanFlagAgg) | aggregate);
    if (tile == 0) return 0;
    uint64_t exclusive = 0;
    int base = tile - 1;
    int spins = 0;
    while (true) {
        const int t = base - lane;
        uint64_t w = (t >= 0) ? ld_relaxed_u64(status + t) : kScanFlagIncl;
        while (__any_sync(kFullMask, (w >> 62) == 0)) {
            if ((w >> 62) == 0) w = ld_relaxed_u64(status + t);
            if (++spins > kScanWatchdog) {
                if (err_flag && lane == 0) atomicOr(err_flag, TDFO_FLAG_SCAN_TIMEOUT);
                break;
            }
        }
        const unsigned incl = __ballot_sync(kFullMask, (w >> 62) == 2);
        uint64_t v = w & kScanValueMask;
        if (incl) {
            const int first = __ffs(incl) - 1;    // nearest predecessor holding an inclusive prefix
            if (lane > first) v = 0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(kFullMask, v, o);
        exclusive += v;
        if (incl || spins > kScanWatchdog) break;
        base -= 32;
    }
    if (lane == 0) st_relaxed_u64(status + tile, kScanFlagIncl | ((exclusive + aggregate) & kScanValueMask));
    return exclusive;
}

}  // namespace tdfo
