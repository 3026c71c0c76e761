// Single-pass device scans: lengths -> offsets (complete cumsum) and run-length
// encoding of sorted row keys.  Both stream their input once (decoupled
// look-back, tile ids from an atomic ticket) and are bandwidth-trivial next to
// the gather; they exist so the step has no host round trip.
#include "scan.cuh"
#include <mutex>

namespace tdfo {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;
constexpr int kScanTile = kScanThreads * kScanItems;   // 2048

static inline int64_t scan_tiles(int64_t n) { return (n + kScanTile - 1) / kScanTile; }

// ---------------------------------------------------------------- cumsum
__global__ void __launch_bounds__(kScanThreads)
complete_cumsum_kernel(const int32_t* __restrict__ lengths, int64_t n, int64_t* __restrict__ offsets,
                       ScanScratch* hdr, uint64_t* status) {
    __shared__ uint64_t warp_totals[8];
    __shared__ uint64_t s_prefix;
    __shared__ int s_tile;
    if (threadIdx.x == 0) s_tile = (int)atomicAdd(&hdr->ticket, 1u);
    __syncthreads();
    const int tile = s_tile;
    const int64_t base = (int64_t)tile * kScanTile + (int64_t)threadIdx.x * kScanItems;
    int64_t v[kScanItems];
    uint64_t sum = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        v[i] = (base + i < n) ? (int64_t)lengths[base + i] : 0;
        sum += (uint64_t)v[i];
    }
    uint64_t total;
    uint64_t excl = block_exclusive_scan_256<uint64_t>(sum, warp_totals, total);
    if (threadIdx.x < 32) {
        const uint64_t p = scan_lookback(status, tile, total, nullptr);
        if (threadIdx.x == 0) s_prefix = p;
    }
    __syncthreads();
    uint64_t run = s_prefix + excl;
    if (tile == 0 && threadIdx.x == 0) offsets[0] = 0;
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) {
        run += (uint64_t)v[i];
        if (base + i < n) offsets[base + i + 1] = (int64_t)run;
    }
}

// ---------------------------------------------------------------- unique / RLE
// Tile = 8 warps x 16 steps x 32 keys.  A warp walks its 512 consecutive keys 32 at a time, so
// every load is one coalesced 128-byte request; run heads are found with a ballot, positions
// with popc, and the heads of a step are written to consecutive slots.
constexpr int kUniqSteps = 16;
constexpr int kUniqTile = kScanThreads * kUniqSteps;   // 4096

__global__ void __launch_bounds__(kScanThreads)
unique_segments_kernel(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, int64_t n,
                       uint32_t* __restrict__ unique_keys, uint32_t* __restrict__ seg_starts,
                       uint32_t* __restrict__ first_vals, uint32_t* __restrict__ num_unique, ScanScratch* hdr, uint64_t* status,
                       int32_t* err_flag, uint32_t* long_hint) {
    __shared__ uint32_t warp_totals[8];
    __shared__ uint64_t s_prefix;
    __shared__ int s_tile;
    if (threadIdx.x == 0) s_tile = (int)atomicAdd(&hdr->ticket, 1u);
    __syncthreads();
    const int tile = s_tile;
    const int lane = lane_id();
    const int warp = threadIdx.x >> 5;
    const int64_t wbase = (int64_t)tile * kUniqTile + (int64_t)warp * (kUniqSteps * 32);
    const uint32_t lt = (1u << lane) - 1u;
    uint32_t k[kUniqSteps], heads[kUniqSteps];
    uint32_t count = 0;
    uint32_t carry = (wbase > 0 && wbase <= n) ? keys[wbase - 1] : 0u;     // key just before this warp's span
#pragma unroll
    for (int i = 0; i < kUniqSteps; ++i) {
        const int64_t idx = wbase + i * 32 + lane;
        k[i] = (idx < n) ? keys[idx] : 0u;
        uint32_t prev = __shfl_up_sync(kFullMask, k[i], 1);
        if (lane == 0) prev = carry;
        const bool head = idx < n && (idx == 0 || k[i] != prev);
        heads[i] = __ballot_sync(kFullMask, head);
        // A run of more than 64 equal keys covers at least one whole 32-key step without a head in it: such steps
        // flag "there may be hot rows" (never missed, rarely false), which lets the update skip its hot-row scan.
        if (long_hint != nullptr && heads[i] == 0 && lane == 0 && wbase + i * 32 + 31 < n) *long_hint = 1u;
        count += __popc(heads[i]);
        carry = __shfl_sync(kFullMask, k[i], 31);
    }
    if (lane == 0) warp_totals[warp] = count;
    __syncthreads();
    uint32_t before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) {
        const uint32_t t = warp_totals[w];
        if (w < warp) before += t;
        total += t;
    }
    if (threadIdx.x < 32) {
        const uint64_t p = scan_lookback(status, tile, (uint64_t)total, err_flag);
        if (threadIdx.x == 0) s_prefix = p;
    }
    __syncthreads();
    uint32_t pos = (uint32_t)s_prefix + before;
#pragma unroll
    for (int i = 0; i < kUniqSteps; ++i) {
        if (heads[i] & (1u << lane)) {
            const int64_t idx = wbase + i * 32 + lane;
            const uint32_t p = pos + __popc(heads[i] & lt);
            unique_keys[p] = k[i];
            seg_starts[p] = (uint32_t)idx;
            if (first_vals) first_vals[p] = vals[idx];
        }
        pos += __popc(heads[i]);
    }
    const int64_t last_tile = (n - 1) / kUniqTile;
    if (tile == last_tile && threadIdx.x == 0) {
        const uint32_t u = (uint32_t)s_prefix + total;
        *num_unique = u;
        seg_starts[u] = (uint32_t)n;
    }
}

static int check_scan_scratch(const char* who, int64_t n, void* scratch, size_t scratch_bytes) {
    if (!scratch || !aligned(scratch, 256)) {
        set_error("%s: scratch is null or not 256-byte aligned", who);
        return TDFO_ERR_SCRATCH;
    }
    const size_t need = tdfo_scan_scratch_bytes(n);
    if (scratch_bytes < need) {
        set_error("%s: scratch has %zu bytes, needs %zu", who, scratch_bytes, need);
        return TDFO_ERR_SCRATCH;
    }
    return TDFO_OK;
}

int complete_cumsum_impl(const int32_t* lengths, int64_t n, int64_t* offsets, void* scratch,
                         cudaStream_t st) {
    if (n == 0) {
        TDFO_CUDA(cudaMemsetAsync(offsets, 0, sizeof(int64_t), st));
        return TDFO_OK;
    }
    const size_t bytes = tdfo_scan_scratch_bytes(n);
    TDFO_CUDA(cudaMemsetAsync(scratch, 0, bytes, st));
    auto* hdr = static_cast<ScanScratch*>(scratch);
    auto* status = reinterpret_cast<uint64_t*>(static_cast<char*>(scratch) + sizeof(ScanScratch));
    complete_cumsum_kernel<<<(unsigned)scan_tiles(n), kScanThreads, 0, st>>>(lengths, n, offsets, hdr, status);
    count_launch();
    TDFO_LAUNCH_CHECK("complete_cumsum");
    return TDFO_OK;
}

int unique_segments_impl(const uint32_t* sorted_keys, const uint32_t* sorted_vals, int64_t n,
                         uint32_t* unique_keys, uint32_t* seg_starts, uint32_t* first_vals,
                         uint32_t* num_unique, int32_t* err_flag, void* scratch, cudaStream_t st,
                         uint32_t* long_hint) {
    if (long_hint) TDFO_CUDA(cudaMemsetAsync(long_hint, 0, sizeof(uint32_t), st));
    if (n == 0) {
        TDFO_CUDA(cudaMemsetAsync(num_unique, 0, sizeof(uint32_t), st));
        TDFO_CUDA(cudaMemsetAsync(seg_starts, 0, sizeof(uint32_t), st));
        return TDFO_OK;
    }
    const size_t bytes = tdfo_scan_scratch_bytes(n);
    TDFO_CUDA(cudaMemsetAsync(scratch, 0, bytes, st));
    auto* hdr = static_cast<ScanScratch*>(scratch);
    auto* status = reinterpret_cast<uint64_t*>(static_cast<char*>(scratch) + sizeof(ScanScratch));
    unique_segments_kernel<<<(unsigned)((n + kUniqTile - 1) / kUniqTile), kScanThreads, 0, st>>>(
        sorted_keys, sorted_vals, n, unique_keys, seg_starts, first_vals, num_unique, hdr, status, err_flag, long_hint);
    count_launch();
    TDFO_LAUNCH_CHECK("unique_segments");
    return TDFO_OK;
}

}  // namespace tdfo

using namespace tdfo;

extern "C" size_t tdfo_scan_scratch_bytes(int64_t n) {
    if (n < 0) n = 0;
    return align_up(sizeof(ScanScratch) + (size_t)scan_tiles(n) * sizeof(uint64_t), 256);
}

extern "C" int tdfo_kjt_complete_cumsum(const int32_t* lengths, int64_t n, int64_t* offsets,
                                        void* scratch, size_t scratch_bytes, void* stream) {
    TDFO_CHECK_ARG(n >= 0, "complete_cumsum: n < 0");
    TDFO_CHECK_ARG(offsets && (lengths || n == 0), "complete_cumsum: null pointer");
    if (n > 0) {
        const int rc = check_scan_scratch("complete_cumsum", n, scratch, scratch_bytes);
        if (rc) return rc;
    }
    return complete_cumsum_impl(lengths, n, offsets, scratch, static_cast<cudaStream_t>(stream));
}

extern "C" int tdfo_unique_segments(const uint32_t* sorted_keys, const uint32_t* sorted_vals, int64_t n,
                                    uint32_t* unique_keys, uint32_t* seg_starts, uint32_t* first_vals,
                                    uint32_t* num_unique, int32_t* err_flag,
                                    void* scratch, size_t scratch_bytes, void* stream) {
    TDFO_CHECK_ARG(n >= 0 && n < (1ll << 32) - 1, "unique_segments: n out of range");
    TDFO_CHECK_ARG(seg_starts && num_unique && (n == 0 || (sorted_keys && unique_keys)),
                   "unique_segments: null pointer");
    TDFO_CHECK_ARG(first_vals == nullptr || sorted_vals != nullptr || n == 0,
                   "unique_segments: first_vals needs sorted_vals");
    if (n > 0) {
        const int rc = check_scan_scratch("unique_segments", n, scratch, scratch_bytes);
        if (rc) return rc;
    }
    return unique_segments_impl(sorted_keys, sorted_vals, n, unique_keys, seg_starts, first_vals, num_unique,
                                err_flag, scratch, static_cast<cudaStream_t>(stream), nullptr);
}


// ---------------------------------------------------------------- cumsum without caller scratch
//
// Workspace for tdfo_kjt_complete_cumsum_static: the ticket word and one look-back status word
// per 2048-length tile, for up to TDFO_CUMSUM_STATIC_MAX_N lengths.  It lives in the library
// image, so every GPU that loads the library carries these ~16 KB until the process exits.
namespace tdfo {
constexpr size_t kStaticScanBytes =
    sizeof(ScanScratch) + (size_t)(TDFO_CUMSUM_STATIC_MAX_N / kScanTile) * sizeof(uint64_t);
static_assert(TDFO_CUMSUM_STATIC_MAX_N % kScanTile == 0, "maximum must be a whole number of scan tiles");
static_assert(kStaticScanBytes % 256 == 0, "keep the workspace a multiple of its alignment");
}  // namespace tdfo

static __device__ __align__(256) unsigned char g_static_scan_scratch[tdfo::kStaticScanBytes];

namespace tdfo {
constexpr int kMaxDevices = 64;
static void* g_static_scan_ptr[kMaxDevices] = {};     // device address of the array, per device
static std::mutex g_static_scan_mutex;

// In host code `g_static_scan_scratch` is only a handle for the symbol, not something a kernel
// can dereference: the address on the current device has to be asked for, once per device.
static int static_scan_scratch(void** out) {
    int dev = -1;
    TDFO_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) {
        set_error("complete_cumsum_static: device ordinal %d is outside [0, %d)", dev, kMaxDevices);
        return TDFO_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_static_scan_mutex);
    if (g_static_scan_ptr[dev] == nullptr) {
        void* p = nullptr;
        TDFO_CUDA(cudaGetSymbolAddress(&p, g_static_scan_scratch));
        g_static_scan_ptr[dev] = p;
    }
    *out = g_static_scan_ptr[dev];
    return TDFO_OK;
}
}  // namespace tdfo

extern "C" int tdfo_kjt_complete_cumsum_static(const int32_t* lengths, int64_t n, int64_t* offsets,
                                               void* stream) {
    TDFO_CHECK_ARG(n >= 0, "complete_cumsum_static: n < 0");
    TDFO_CHECK_ARG(offsets && (lengths || n == 0), "complete_cumsum_static: null pointer");
    if (n > (int64_t)TDFO_CUMSUM_STATIC_MAX_N) {
        set_error("complete_cumsum_static: n = %lld exceeds the built-in workspace (at most %lld lengths); "
                  "use tdfo_kjt_complete_cumsum with caller scratch for larger inputs",
                  (long long)n, (long long)TDFO_CUMSUM_STATIC_MAX_N);
        return TDFO_ERR_SCRATCH;
    }
    void* scratch = nullptr;
    const int rc = static_scan_scratch(&scratch);
    if (rc) return rc;
    // Same kernel as tdfo_kjt_complete_cumsum.  complete_cumsum_impl first zeroes the part of the
    // workspace this n uses (ticket + status words) on `stream`: the previous call's look-back
    // words are still in there and would otherwise be read as finished prefixes.
    return complete_cumsum_impl(lengths, n, offsets, scratch, static_cast<cudaStream_t>(stream));
}
