// Stable LSD radix sort of (row key, bag id) uint32 pairs, one-sweep style:
// one histogram kernel for all digit positions, then ONE kernel per digit that
// ranks a 4096-key tile in shared memory and resolves its global offsets with a
// per-digit decoupled look-back over earlier tiles (tile ids come from an atomic
// ticket, so a tile only ever waits on tiles that are already running).
// Traffic per pass: read + write of the pairs (16 B per pair), no second read.
#include "scan.cuh"

namespace tdfo {

constexpr int kSortThreads = 256;
#ifndef TDFO_SORT_ITEMS
#define TDFO_SORT_ITEMS 16
#endif
constexpr int kSortItems = TDFO_SORT_ITEMS;
constexpr int kSortTile = kSortThreads * kSortItems;    // 4096 pairs per tile
constexpr int kRadixLog = 9;                            // up to 9 bits per pass
constexpr int kRadix = 1 << kRadixLog;                  // bins per pass (lookback / histogram row stride)
constexpr int kBinsPerThread = kRadix / kSortThreads;   // 2
constexpr int kMaxPasses = 4;
#ifndef TDFO_SORT_BALLOT_MATCH
#define TDFO_SORT_BALLOT_MATCH 1
#endif
#ifndef TDFO_SORT_MINB
#define TDFO_SORT_MINB 4
#endif

constexpr uint32_t kLbAgg  = 1u << 30;
constexpr uint32_t kLbIncl = 2u << 30;
constexpr uint32_t kLbMask = (1u << 30) - 1;
#ifndef TDFO_SORT_LB_BATCH
#define TDFO_SORT_LB_BATCH 8
#endif
constexpr int kLbBatch = TDFO_SORT_LB_BATCH;            // look-back words requested per round trip

struct SortPlan {
    int passes;
    int bits[kMaxPasses];
    int shift[kMaxPasses];
    int64_t tiles;
    size_t off_hist, off_ticket, off_lookback, total;
};

static SortPlan make_sort_plan(int64_t n, int key_bits) {
    SortPlan p{};
    p.passes = (key_bits + kRadixLog - 1) / kRadixLog;
    const int per = (key_bits + p.passes - 1) / p.passes;
    int s = 0;
    for (int i = 0; i < p.passes; ++i) {
        p.shift[i] = s;
        p.bits[i] = (key_bits - s < per) ? key_bits - s : per;
        s += p.bits[i];
    }
    p.tiles = (n + kSortTile - 1) / kSortTile;
    p.off_hist = 0;
    p.off_ticket = align_up(sizeof(uint32_t) * kRadix * kMaxPasses, 256);
    p.off_lookback = p.off_ticket + 256;
    p.total = align_up(p.off_lookback + (size_t)p.passes * p.tiles * kRadix * sizeof(uint32_t), 256);
    return p;
}

struct HistArgs {
    int passes;
    int bits[kMaxPasses];
    int shift[kMaxPasses];
};

// All digit histograms in one read of the keys.
__global__ void __launch_bounds__(kSortThreads)
radix_histogram_kernel(const uint32_t* __restrict__ keys, int64_t n, HistArgs h, uint32_t* ghist) {
    __shared__ uint32_t s_hist[kMaxPasses][kRadix];
    for (int i = threadIdx.x; i < kMaxPasses * kRadix; i += kSortThreads) (&s_hist[0][0])[i] = 0;
    __syncthreads();
    for (int64_t i = (int64_t)blockIdx.x * kSortThreads + threadIdx.x; i < n;
         i += (int64_t)gridDim.x * kSortThreads) {
        const uint32_t k = keys[i];
        for (int p = 0; p < h.passes; ++p)
            atomicAdd(&s_hist[p][(k >> h.shift[p]) & ((1u << h.bits[p]) - 1)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kMaxPasses * kRadix; i += kSortThreads) {
        const uint32_t c = (&s_hist[0][0])[i];
        if (c) atomicAdd(ghist + i, c);
    }
}

__global__ void __launch_bounds__(kSortThreads, TDFO_SORT_MINB)
radix_onesweep_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                      uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                      int64_t n, int shift, int bits,
                      const uint32_t* __restrict__ ghist,   // [kRadix] counts of this pass
                      uint32_t* lookback,                   // [tiles][kRadix], zeroed
                      unsigned int* ticket, int32_t* err_flag) {
    __shared__ uint16_t s_warp_hist[kSortThreads / 32][kRadix];   // counts <= 4096 fit 16 bits
    __shared__ uint32_t s_keys[kSortTile];
    __shared__ uint32_t s_vals[kSortTile];
    __shared__ uint32_t s_tile_start[kRadix];   // first slot of digit d in the tile-sorted order
    __shared__ uint32_t s_out_base[kRadix];     // global slot of tile-sorted slot s is s_out_base[d] + s
    __shared__ uint32_t s_scan[8];
    __shared__ int s_tile;

    const int tid = threadIdx.x;
    const int lane = tid & 31;
    const int warp = tid >> 5;
    const uint32_t dmask = (1u << bits) - 1u;
    const int nbins = 1 << bits;

    if (tid == 0) s_tile = (int)atomicAdd(ticket, 1u);
    for (int i = tid; i < (kSortThreads / 32) * kRadix / 2; i += kSortThreads)
        reinterpret_cast<uint32_t*>(&s_warp_hist[0][0])[i] = 0;
    __syncthreads();
    const int tile = s_tile;
    const int64_t tile_base = (int64_t)tile * kSortTile;
    const int count = (int)min((int64_t)kSortTile, n - tile_base);

    // ---- load (warp w owns tile slots [w*512, w*512+512), coalesced 128 B per step)
    uint32_t key[kSortItems], val[kSortItems];
    uint16_t rank[kSortItems];
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int slot = warp * (kSortItems * 32) + i * 32 + lane;
        if (slot < count) {
            key[i] = keys_in[tile_base + slot];
            val[i] = vals_in[tile_base + slot];
        } else {
            key[i] = 0xffffffffu;
            val[i] = 0;
        }
    }
    // ---- rank inside the warp, in slot order (stable)
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int slot = warp * (kSortItems * 32) + i * 32 + lane;
        const uint32_t d = (slot < count) ? ((key[i] >> shift) & dmask) : (uint32_t)kRadix;
#if defined(TDFO_SORT_TIMING_NO_RANK)           // timing experiment only: results are wrong
        const uint32_t peers = 1u << lane;
#elif TDFO_SORT_BALLOT_MATCH
        // lanes holding the same digit, from one ballot per digit bit (10 bits: 9 + the "no key" bit)
        uint32_t peers = kFullMask;
#pragma unroll
        for (int bit = 0; bit <= kRadixLog; ++bit) {
            const uint32_t vote = __ballot_sync(kFullMask, (d >> bit) & 1u);
            peers &= ((d >> bit) & 1u) ? vote : ~vote;
        }
#else
        const uint32_t peers = __match_any_sync(kFullMask, d);
#endif
        const int leader = __ffs(peers) - 1;
        uint32_t pre = 0;
        if (lane == leader && d < (uint32_t)kRadix) {
            pre = s_warp_hist[warp][d];
            s_warp_hist[warp][d] = (uint16_t)(pre + __popc(peers));
        }
        pre = __shfl_sync(kFullMask, pre, leader);
        rank[i] = (uint16_t)(pre + __popc(peers & lt_mask));
        __syncwarp();
    }
    __syncthreads();

    // ---- per digit (kBinsPerThread digits per thread, digit = tid*kBinsPerThread + j so that a
    // thread's digits are adjacent): prefix over warps, tile count, publish, look back
    uint32_t tile_count[kBinsPerThread];
    uint32_t pair_count = 0, gpair = 0;
    uint32_t gcount[kBinsPerThread];
#pragma unroll
    for (int j = 0; j < kBinsPerThread; ++j) {
        const int d = tid * kBinsPerThread + j;
        tile_count[j] = 0;
        gcount[j] = 0;
        if (d < nbins) {
#pragma unroll
            for (int w = 0; w < kSortThreads / 32; ++w) {
                const uint32_t c = s_warp_hist[w][d];
                s_warp_hist[w][d] = (uint16_t)tile_count[j];
                tile_count[j] += c;
            }
            st_relaxed_u32(lookback + (size_t)tile * kRadix + d, (tile == 0 ? kLbIncl : kLbAgg) | tile_count[j]);
            gcount[j] = ghist[d];
        }
        pair_count += tile_count[j];
        gpair += gcount[j];
    }
    uint32_t total;
    uint32_t tile_start = block_exclusive_scan_256<uint32_t>(pair_count, s_scan, total);
    uint32_t gstart = block_exclusive_scan_256<uint32_t>(gpair, s_scan, total);
#pragma unroll
    for (int j = 0; j < kBinsPerThread; ++j) {
        const int d = tid * kBinsPerThread + j;
        if (d < nbins) {
            // Walk back over earlier tiles until one holds an inclusive prefix.  kLbBatch words are
            // requested per round trip: a walk of depth k costs k / kLbBatch L2 latencies, not k
            // (with ~600 tiles in flight the serial walk was the longest chain in the pass).
            uint32_t excl = 0;
            int spins = 0;
            bool done = false;
#if defined(TDFO_SORT_TIMING_NO_LOOKBACK)      // timing experiment only (tools/build_variant.py): results are wrong
            done = true;
#endif
            for (int t = tile - 1; t >= 0 && !done; t -= kLbBatch) {
                uint32_t w[kLbBatch];
#pragma unroll
                for (int q = 0; q < kLbBatch; ++q)
                    w[q] = (t - q >= 0) ? ld_relaxed_u32(lookback + (size_t)(t - q) * kRadix + d) : kLbIncl;
#pragma unroll
                for (int q = 0; q < kLbBatch; ++q) {
                    if (done) break;
                    while ((w[q] >> 30) == 0) {          // that tile has not published yet
                        if (++spins > kScanWatchdog) {
                            if (err_flag) atomicOr(err_flag, TDFO_FLAG_SCAN_TIMEOUT);
                            w[q] = kLbIncl;
                            break;
                        }
                        w[q] = ld_relaxed_u32(lookback + (size_t)(t - q) * kRadix + d);
                    }
                    excl += w[q] & kLbMask;
                    if ((w[q] >> 30) == 2) done = true;
                }
            }
            if (tile > 0)
                st_relaxed_u32(lookback + (size_t)tile * kRadix + d, kLbIncl | ((excl + tile_count[j]) & kLbMask));
            s_tile_start[d] = tile_start;
            s_out_base[d] = gstart + excl - tile_start;
        }
        tile_start += tile_count[j];
        gstart += gcount[j];
    }
    __syncthreads();

    // ---- reorder inside the tile through shared memory, then write runs
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int slot = warp * (kSortItems * 32) + i * 32 + lane;
        if (slot < count) {
            const uint32_t d = (key[i] >> shift) & dmask;
            const uint32_t s = s_tile_start[d] + s_warp_hist[warp][d] + rank[i];
            s_keys[s] = key[i];
            s_vals[s] = val[i];
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < kSortItems; ++i) {
        const int s = i * kSortThreads + tid;
        if (s < count) {
            const uint32_t k = s_keys[s];
            const uint32_t d = (k >> shift) & dmask;
            const uint32_t o = s_out_base[d] + (uint32_t)s;
            keys_out[o] = k;
            vals_out[o] = s_vals[s];
        }
    }
}

// Sorts in place over the two buffer pairs; returns 0 if the result is in (k0,v0),
// 1 if it is in (k1,v1).  Input is in (k0,v0).
// `kin` / `vin` (optional): the unsorted pairs live somewhere else and are only read; the first pass then
// goes from there into (k1, v1) and (k0, v0) is scratch, with the result in the same place as without them.
int radix_sort_impl(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int64_t n, int key_bits,
                    int32_t* err_flag, void* scratch, cudaStream_t st, int* result_buf,
                    const uint32_t* kin, const uint32_t* vin) {
    *result_buf = 0;
    if (n == 0) return TDFO_OK;
    const SortPlan p = make_sort_plan(n, key_bits);
    char* base = static_cast<char*>(scratch);
    TDFO_CUDA(cudaMemsetAsync(base, 0, p.total, st));
    uint32_t* ghist = reinterpret_cast<uint32_t*>(base + p.off_hist);
    unsigned int* tickets = reinterpret_cast<unsigned int*>(base + p.off_ticket);
    uint32_t* lookback = reinterpret_cast<uint32_t*>(base + p.off_lookback);
    HistArgs h{};
    h.passes = p.passes;
    for (int i = 0; i < p.passes; ++i) { h.bits[i] = p.bits[i]; h.shift[i] = p.shift[i]; }
    int64_t hb = (n + kSortThreads * 16 - 1) / (kSortThreads * 16);
    if (hb > kNumSMs * 8) hb = kNumSMs * 8;
    radix_histogram_kernel<<<(unsigned)hb, kSortThreads, 0, st>>>(kin ? kin : k0, n, h, ghist);
    count_launch();
    TDFO_LAUNCH_CHECK("radix_histogram");
    uint32_t *ki = k0, *vi = v0, *ko = k1, *vo = v1;
    for (int i = 0; i < p.passes; ++i) {
        radix_onesweep_kernel<<<(unsigned)p.tiles, kSortThreads, 0, st>>>(
            (i == 0 && kin) ? kin : ki, (i == 0 && vin) ? vin : vi, ko, vo, n, p.shift[i], p.bits[i], ghist + i * kRadix,
            lookback + (size_t)i * p.tiles * kRadix, tickets + i, err_flag);
        count_launch();
        TDFO_LAUNCH_CHECK("radix_onesweep");
        uint32_t* t;
        t = ki; ki = ko; ko = t;
        t = vi; vi = vo; vo = t;
    }
    *result_buf = p.passes & 1;
    return TDFO_OK;
}

int radix_sort_passes_impl(int key_bits) { return (key_bits + kRadixLog - 1) / kRadixLog; }

size_t radix_sort_scratch_bytes_impl(int64_t n, int key_bits) {
    if (n <= 0) return 256;
    return make_sort_plan(n, key_bits).total;
}

}  // namespace tdfo

using namespace tdfo;

extern "C" size_t tdfo_radix_sort_scratch_bytes(int64_t n, int32_t key_bits) {
    if (key_bits < 1) key_bits = 1;
    if (key_bits > 32) key_bits = 32;
    return radix_sort_scratch_bytes_impl(n, key_bits);
}

extern "C" int tdfo_radix_sort_pairs(uint32_t* keys_in, uint32_t* vals_in, uint32_t* keys_out,
                                     uint32_t* vals_out, int64_t n, int32_t key_bits,
                                     int32_t* err_flag, void* scratch, size_t scratch_bytes,
                                     void* stream) {
    TDFO_CHECK_ARG(n >= 0 && n < (1ll << 30), "radix_sort_pairs: n must be in [0, 2^30)");
    TDFO_CHECK_ARG(key_bits >= 1 && key_bits <= 32, "radix_sort_pairs: key_bits must be in [1, 32]");
    if (n == 0) return TDFO_OK;
    TDFO_CHECK_ARG(keys_in && vals_in && keys_out && vals_out, "radix_sort_pairs: null pointer");
    if (!scratch || !aligned(scratch, 256) || scratch_bytes < tdfo_radix_sort_scratch_bytes(n, key_bits)) {
        set_error("radix_sort_pairs: scratch null, misaligned or smaller than %zu bytes",
                  tdfo_radix_sort_scratch_bytes(n, key_bits));
        return TDFO_ERR_SCRATCH;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int where = 0;
    const int rc = radix_sort_impl(keys_in, vals_in, keys_out, vals_out, n, key_bits, err_flag, scratch, st, &where, nullptr, nullptr);
    if (rc) return rc;
    if (where == 0) {   // even number of passes: result sits in the input pair
        TDFO_CUDA(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
        TDFO_CUDA(cudaMemcpyAsync(vals_out, vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, st));
    }
    return TDFO_OK;
}
