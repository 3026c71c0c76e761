// Backward of the embedding lookup with the optimizer fused in:
//   linearise ids -> stable radix sort by row -> run-length encode ->
//   per unique row: ordered fp32 reduction of its gradient rows + in-place update.
// No dense [rows, D] gradient exists at any point.  HBM-bound: per unique row one
// row read + one row write (+ optimizer state); gradient rows are re-read out of
// L2 (sorted order walks the tables one after another, so the live gradient
// slice is B*D*4 bytes per table, far below the 126 MB L2).
#include "common.cuh"

namespace tdfo {

int radix_sort_impl(uint32_t* k0, uint32_t* v0, uint32_t* k1, uint32_t* v1, int64_t n, int key_bits,
                    int32_t* err_flag, void* scratch, cudaStream_t st, int* result_buf,
                    const uint32_t* kin, const uint32_t* vin);
size_t radix_sort_scratch_bytes_impl(int64_t n, int key_bits);
int radix_sort_passes_impl(int key_bits);
int unique_segments_impl(const uint32_t* sorted_keys, const uint32_t* sorted_vals, int64_t n,
                         uint32_t* unique_keys, uint32_t* seg_starts, uint32_t* first_vals,
                         uint32_t* num_unique, int32_t* err_flag, void* scratch, cudaStream_t st,
                         uint32_t* long_hint);

// Weight rows are read and written once per step.  STREAM = false keeps them as ordinary L2 residents
// (ld.cg / plain store); STREAM = true marks both evict-first (ld.cs / st.cs) so that they do not push
// the gradient rows, which ARE re-read, out of L2.  Measured on B200 (profiles/r2_tuning_kernels.txt):
// streaming wins 5-7 % when one table's gradient slice is tens of MB (67 MB: one rank of 8, table-wise)
// and loses ~1 % when the slice is small (8 MB: BASELINE config 2); explicit L2::evict_first policies
// and write-through stores lose everywhere.  dispatch_update picks by slice size.
template <bool STREAM>
__device__ __forceinline__ uint4 upd_load_weight(const uint4* p) {
    return STREAM ? __ldcs(p) : ld_row_128(p);
}
template <bool STREAM>
__device__ __forceinline__ void upd_store_weight(uint4* p, uint4 v) {
    if (STREAM) __stcs(p, v);
    else st_128(p, v);
}
__device__ __forceinline__ uint4 upd_load_grad(const float* p) { return ldg_128(p); }

constexpr int kUpdThreads = 256;    // linearise kernel

// ---------------------------------------------------------------- linearise
__global__ void __launch_bounds__(kUpdThreads)
linearize_kernel(const int64_t* __restrict__ indices, const int64_t* __restrict__ offsets,
                 const int64_t* __restrict__ feat_row_base, const int64_t* __restrict__ feat_rows,
                 int64_t B, int64_t num_bags, int64_t L, uint32_t oob_key, int vals_are_positions,
                 uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint32_t* __restrict__ pos_bag,
                 int32_t* err_flag) {
    const int64_t g = (int64_t)blockIdx.x * (kUpdThreads / 32) + (threadIdx.x >> 5);
    if (g >= num_bags) return;
    const int lane = lane_id();
    const int f = (int)(g / B);
    int64_t s = __ldg(offsets + g), e = __ldg(offsets + g + 1);
    if (e > L) {                  // offsets reach past the ids given: cut the bag, never read beyond indices[L)
        if (lane == 0 && err_flag) atomicOr(err_flag, TDFO_FLAG_OFFSETS_RANGE);
        e = L;
        s = min(s, L);
    }
    const int64_t row_base = __ldg(feat_row_base + f);
    const uint64_t rows = (uint64_t)__ldg(feat_rows + f);
    bool oob = false;
    for (int64_t p = s + lane; p < e; p += 32) {
        const int64_t idx = __ldg(indices + p);
        uint32_t k;
        if ((uint64_t)idx < rows) k = (uint32_t)(row_base + idx);
        else { k = oob_key; oob = true; }
        keys[p] = k;
        vals[p] = vals_are_positions ? (uint32_t)p : (uint32_t)g;
        if (pos_bag) pos_bag[p] = (uint32_t)g;      // weighted pooled backward: payload = position, bag looked up
    }
    if (oob && err_flag) atomicOr(err_flag, TDFO_FLAG_INDEX_OOB);
}

// ---------------------------------------------------------------- update
struct UpdateArgs {
    void* weights;
    float* state0;
    float* state1;
    const float* grad;
    const uint32_t* sorted_vals;
    const uint32_t* unique_keys;
    const uint32_t* seg_starts;
    const uint32_t* first_vals;
    // hot rows (segments longer than kLongSegment), see HotLayout; hot_hdr == nullptr: no hot-row path
    uint32_t* hot_hdr;        // [0] records  [1] work items  [2] partial rows  [3] ticket of the block-per-chunk kernel
    uint4* hot_recs;          // {unique-row index, first partial row, chunks, chunks done}
    uint2* hot_items;         // {record, chunk}
    float* hot_parts;         // partial sums of multi-chunk rows, `row floats` apart
    uint32_t rec_cap, item_cap, part_cap, hot_chunk;
    const uint32_t* num_unique;
    const int64_t* offsets;
    const int64_t* feat_out_col;
    const float* psw;         // per-sample weights by position (pooled modes only) or null
    const uint32_t* pos_bag;  // bag id of every position; set together with psw (payloads are positions then)
    int64_t grad_stride;
    int64_t B;
    uint64_t B_magic;    // ceil(2^64 / B): bag id / B == __umul64hi(id, B_magic) for every 32-bit id (B > 1)
    uint32_t total_rows;
    uint32_t L;
    int32_t D;
    int32_t nchunks;
    int32_t pooling_mode;
    float lr, eps, wd, beta1, beta2, bc1, bc2;   // bc = 1 - beta^step
    float inv_D;
};

// Launch shape of the update kernels: persistent grid of 148 x MINB CTAs of 8 warps (the block-per-chunk kernel and
// most row sizes) or 9 warps (the main kernel for D = 128 rows, see dispatch_update).
constexpr int kUpdCtaThreads = 256;
#ifndef TDFO_UPD_THREADS
#define TDFO_UPD_THREADS 288
#endif
constexpr int kUpdMainThreads = TDFO_UPD_THREADS;    // CTA size of the main kernel for 512-byte rows (see dispatch_update)
constexpr uint32_t kLongSegment = 64;    // occurrences above which a row goes to the block-per-chunk kernel
constexpr int kLongMaxFloats = 8192;      // shared-memory floats for the partial sums of that kernel
constexpr uint32_t kHotChunkMin = 2048;  // occurrences one CTA reduces at most (grows with L, see hot_layout)
constexpr uint32_t kHotMaxRowFloats = 2048;   // widest row in fp32 accumulators (4096 B of bf16)

// Hot-row work area inside the caller's worklist (uint32 words).  A row with more than kLongSegment
// occurrences becomes one record and ceil(len / chunk) work items; rows longer than one chunk also get
// that many partial-sum rows.  Counts: records <= L/65, items <= records + L/chunk, partial rows
// <= 2 L/chunk, so nothing here can overflow for the L the layout was sized for.
struct HotLayout {
    uint32_t chunk, rec_cap, item_cap, part_cap;
    size_t off_recs, off_items, off_parts, words;   // in uint32 words from the worklist start
};

static HotLayout hot_layout(int64_t L) {
    HotLayout h;
    const uint64_t l = (uint64_t)(L > 0 ? L : 0);
    const uint64_t chunk = (l + 2047) / 2048 > kHotChunkMin ? (l + 2047) / 2048 : kHotChunkMin;
    h.chunk = (uint32_t)chunk;
    h.rec_cap = (uint32_t)(l / kLongSegment + 1);
    h.item_cap = (uint32_t)(h.rec_cap + l / chunk + 1);
    h.part_cap = (uint32_t)(2 * (l / chunk) + 2);
    h.off_recs = 16;
    h.off_items = h.off_recs + 4 * (size_t)h.rec_cap;
    h.off_parts = (h.off_items + 2 * (size_t)h.item_cap + 3) / 4 * 4;
    h.words = h.off_parts + (size_t)h.part_cap * kHotMaxRowFloats;
    return h;
}

// Per unique row: key, first / one-past-last sorted position, payload of the first occurrence.
struct RowMeta {
    uint32_t key, s, e, v0;
};

__device__ __forceinline__ RowMeta load_row_meta(const UpdateArgs& a, uint32_t u, uint32_t nu) {
    RowMeta m;
    if (u < nu) {
        m.key = __ldg(a.unique_keys + u);
        m.s = __ldg(a.seg_starts + u);
        m.e = __ldg(a.seg_starts + u + 1);
        m.v0 = __ldg(a.first_vals + u);
    } else {
        m.key = 0xffffffffu;
        m.s = m.e = m.v0 = 0u;
    }
    return m;
}

// Gradient row of one occurrence (payload v) and its scale.
__device__ __forceinline__ const float* grad_row_of(const UpdateArgs& a, uint32_t v, float& scale) {
    scale = 1.f;
    if (a.pooling_mode == TDFO_POOL_NONE) return a.grad + (size_t)v * a.D;
    if (a.psw != nullptr) {
        scale = __ldg(a.psw + v);
        v = __ldg(a.pos_bag + v);
    }
    const uint32_t f = a.B_magic ? (uint32_t)__umul64hi((uint64_t)v, a.B_magic) : v;
    const uint32_t b = v - f * (uint32_t)a.B;
    if (a.pooling_mode == TDFO_POOL_MEAN)
        scale = scale / (float)(__ldg(a.offsets + v + 1) - __ldg(a.offsets + v));
    return a.grad + (size_t)b * a.grad_stride + __ldg(a.feat_out_col + f);
}

// Optimizer step for one row held by an LPR-lane group: g = reduced gradient, w = current row
// (both fp32, CPL chunks per lane); writes the row (and its state) back if `valid`.
template <typename T, int LPR, int CPL, int OPT, bool STREAM = false>
__device__ __forceinline__ void apply_row_update(const UpdateArgs& a, uint32_t row, bool valid, float h_old,
                                                 float (&g)[CPL][Chunk<T>::kElems], float (&w)[CPL][Chunk<T>::kElems],
                                                 int sub, unsigned gmask) {
    constexpr int E = Chunk<T>::kElems;
    uint4* __restrict__ wbase = reinterpret_cast<uint4*>(a.weights);
    if (OPT == TDFO_OPT_SGD) {
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) {
                const float gi = fmaf(a.wd, w[k][i], g[k][i]);
                w[k][i] = fmaf(-a.lr, gi, w[k][i]);
            }
    } else if (OPT == TDFO_OPT_ROWWISE_ADAGRAD) {
        float ss = 0.f;
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) {
                g[k][i] = fmaf(a.wd, w[k][i], g[k][i]);
                if (sub + k * LPR >= a.nchunks) g[k][i] = 0.f;       // lanes past the row end
                ss = fmaf(g[k][i], g[k][i], ss);
            }
#pragma unroll
        for (int o = 1; o < LPR; o <<= 1) ss += __shfl_xor_sync(gmask, ss, o);
        const float h = h_old + ss * a.inv_D;
        // sqrt / divide through the SFU (2-3 ulp, far inside the 1e-5 parity bar)
        const float sq = h > 0.f ? h * rsqrtf(h) : 0.f;
        const float mult = __fdividef(a.lr, sq + a.eps);
        if (valid && sub == 0) a.state0[row] = h;
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) w[k][i] = fmaf(-mult, g[k][i], w[k][i]);
    } else if (valid) {   // ADAM
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = sub + k * LPR;
            if (c < a.nchunks) {
                float* m = a.state0 + (size_t)row * a.D + (size_t)c * E;
                float* v = a.state1 + (size_t)row * a.D + (size_t)c * E;
#pragma unroll
                for (int t = 0; t < E; t += 4) {
                    uint4 mr = ld_128(m + t), vr = ld_128(v + t);
                    float mm[4] = {__uint_as_float(mr.x), __uint_as_float(mr.y), __uint_as_float(mr.z), __uint_as_float(mr.w)};
                    float vv[4] = {__uint_as_float(vr.x), __uint_as_float(vr.y), __uint_as_float(vr.z), __uint_as_float(vr.w)};
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float gi = g[k][t + i];
                        mm[i] = a.beta1 * mm[i] + (1.f - a.beta1) * gi;
                        vv[i] = a.beta2 * vv[i] + (1.f - a.beta2) * gi * gi;
                        const float mhat = mm[i] / a.bc1;
                        const float vhat = vv[i] / a.bc2;
                        w[k][t + i] -= a.lr * (mhat / (sqrtf(vhat) + a.eps) + a.wd * w[k][t + i]);
                    }
                    st_128(m + t, make_uint4(__float_as_uint(mm[0]), __float_as_uint(mm[1]), __float_as_uint(mm[2]), __float_as_uint(mm[3])));
                    st_128(v + t, make_uint4(__float_as_uint(vv[0]), __float_as_uint(vv[1]), __float_as_uint(vv[2]), __float_as_uint(vv[3])));
                }
            }
        }
    }

    if (valid) {
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = sub + k * LPR;
            if (c < a.nchunks) upd_store_weight<STREAM>(wbase + (size_t)row * a.nchunks + c, Chunk<T>::pack(w[k]));
        }
    }
}

// g += sum over occurrences [lo, hi) of one row, in increasing order.  Payloads are fetched one per
// lane of the group (coalesced), gradient rows four at a time, so a long segment is a stream of
// independent loads rather than a chain of dependent ones.
template <typename T, int LPR, int CPL, bool DEEP>
__device__ __forceinline__ void accumulate_occurrences(const UpdateArgs& a, uint32_t lo, uint32_t hi,
                                                       float (&g)[CPL][Chunk<T>::kElems], int sub, int grp,
                                                       unsigned gmask) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int GQ = E / 4;
    // gradient rows in flight: one in the main kernel (register budget), several in the hot-row kernel
#ifndef TDFO_UPD_UB_MAIN
#define TDFO_UPD_UB_MAIN 1
#endif
    constexpr int UB = !DEEP ? TDFO_UPD_UB_MAIN : ((CPL * GQ <= 2) ? 8 : ((CPL * GQ <= 4) ? 4 : ((CPL * GQ <= 8) ? 2 : 1)));
    for (uint32_t i = lo; i < hi; i += LPR) {
        const uint32_t n = min((uint32_t)LPR, hi - i);
        const uint32_t my_v = ((uint32_t)sub < n) ? __ldg(a.sorted_vals + i + sub) : 0u;
        for (uint32_t j = 0; j < n; j += UB) {
            uint4 x[UB][CPL][GQ];
            float sc[UB];
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                const uint32_t v = __shfl_sync(gmask, my_v, grp * LPR + (int)((j + u) % LPR));
                const float* grow = grad_row_of(a, (j + u < n) ? v : 0u, sc[u]);
                if (j + u >= n) sc[u] = 0.f;
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    const int c = min(sub + k * LPR, a.nchunks - 1);
#pragma unroll
                    for (int t = 0; t < GQ; ++t) x[u][k][t] = upd_load_grad(grow + (size_t)c * E + 4 * t);
                }
            }
#pragma unroll
            for (int u = 0; u < UB; ++u) {
                if (j + u < n) {
#pragma unroll
                    for (int k = 0; k < CPL; ++k)
#pragma unroll
                        for (int t = 0; t < GQ; ++t) {
                            g[k][4 * t]     = fmaf(sc[u], __uint_as_float(x[u][k][t].x), g[k][4 * t]);
                            g[k][4 * t + 1] = fmaf(sc[u], __uint_as_float(x[u][k][t].y), g[k][4 * t + 1]);
                            g[k][4 * t + 2] = fmaf(sc[u], __uint_as_float(x[u][k][t].z), g[k][4 * t + 2]);
                            g[k][4 * t + 3] = fmaf(sc[u], __uint_as_float(x[u][k][t].w), g[k][4 * t + 3]);
                        }
                }
            }
        }
    }
}

// Persistent warps; a warp works on RPW*UR consecutive unique rows per step, one row per
// LPR-lane group and unroll slot.  Under load a B200 memory round trip takes ~3 us, and what the
// kernel can keep in flight is set by the register file (shared-memory staging with cp.async was
// measured and is 13-60 % slower, profiles/r2_tuning_kernels.txt):
//   * per step every lane requests its chunks of the weight row, of the first gradient row and
//     the row state before consuming any of them (the unique pass also emits each row's first
//     payload, so finding the first gradient row needs no dependent load); all of these loads are
//     unconditional (rows that do not exist are redirected to row 0 and masked) so ptxas issues
//     them back to back;
//   * few lanes per row (LPR 8 x 4 chunks for a 512 B row): per-row bookkeeping is replicated
//     in every lane of a group, so narrow groups spend a quarter of the issue slots per row;
//   * as many warps per SM as run WITHOUT spilling the row loop: 27 (288 threads x 3 CTAs,
//     72 registers) for D = 128 rows; 30 warps leave 64 registers and lose 15 %.
template <typename T, int LPR, int CPL, int UR, int MINB, int OPT, bool PREFETCH = true, int THREADS = kUpdCtaThreads,
          bool STREAM = false>
__global__ void __launch_bounds__(THREADS, MINB)
sparse_update_kernel(const UpdateArgs a) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int RPW = 32 / LPR;                  // rows processed in parallel by one warp
    constexpr int STEP = RPW * UR;                 // rows per warp step
    constexpr int GQ = E / 4;                      // 16-byte gradient pieces per table chunk
    const int lane = lane_id();
    const int sub = lane % LPR;
    const int grp = lane / LPR;
    // Row groups of one warp run different rows (and different trip counts): group-level
    // shuffles name only the lanes of their own group.
    const unsigned gmask = (LPR == 32) ? kFullMask : (((1u << (LPR & 31)) - 1u) << (grp * LPR));
    const uint32_t nu = *a.num_unique;
    const uint32_t stride = gridDim.x * (THREADS / 32) * STEP;
    uint32_t u0 = (blockIdx.x * (THREADS / 32) + (threadIdx.x >> 5)) * STEP;
    if (u0 >= nu) return;
    uint4* __restrict__ wbase = reinterpret_cast<uint4*>(a.weights);

    RowMeta cur[UR];
#pragma unroll
    for (int q = 0; q < UR; ++q) cur[q] = load_row_meta(a, u0 + q * RPW + grp, nu);

    while (u0 < nu) {
        RowMeta nxt[UR];                            // on their way while this step's rows are processed
        if (PREFETCH) {
#pragma unroll
            for (int q = 0; q < UR; ++q) nxt[q] = load_row_meta(a, u0 + stride + q * RPW + grp, nu);
        }

        bool valid[UR];
        float scale[UR], h_old[UR];
        uint4 wv[UR][CPL];
        uint4 gv[UR][CPL][GQ];
        // ---- issue every request of this step
#pragma unroll
        for (int q = 0; q < UR; ++q) {
            valid[q] = cur[q].key < a.total_rows;      // padding rows and the out-of-range marker fail this
            // A hot row (very long segment) would be reduced serially by this one group: skip it here.
            // (hot_rows_enqueue_kernel has queued it for the block-per-chunk kernel that runs after this one)
            if (a.hot_hdr != nullptr && cur[q].e - cur[q].s > kLongSegment) valid[q] = false;
            const float* grow = grad_row_of(a, valid[q] ? cur[q].v0 : 0u, scale[q]);
            const size_t wrow = (size_t)(valid[q] ? cur[q].key : 0u) * a.nchunks;
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const int c = min(sub + k * LPR, a.nchunks - 1);
                wv[q][k] = upd_load_weight<STREAM>(wbase + wrow + c);
#pragma unroll
                for (int t = 0; t < GQ; ++t) gv[q][k][t] = upd_load_grad(grow + (size_t)c * E + 4 * t);
            }
            h_old[q] = 0.f;
            if (OPT == TDFO_OPT_ROWWISE_ADAGRAD) h_old[q] = __ldcg(a.state0 + (valid[q] ? cur[q].key : 0u));
        }
        // ---- reduce, update, store
#pragma unroll
        for (int q = 0; q < UR; ++q) {
            const uint32_t row = cur[q].key;
            float g[CPL][E], w[CPL][E];
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
#pragma unroll
                for (int t = 0; t < GQ; ++t) {
                    g[k][4 * t]     = scale[q] * __uint_as_float(gv[q][k][t].x);
                    g[k][4 * t + 1] = scale[q] * __uint_as_float(gv[q][k][t].y);
                    g[k][4 * t + 2] = scale[q] * __uint_as_float(gv[q][k][t].z);
                    g[k][4 * t + 3] = scale[q] * __uint_as_float(gv[q][k][t].w);
                }
                Chunk<T>::unpack(wv[q][k], w[k]);
            }
            // further occurrences of the row, in sorted (= original) order
            if (valid[q] && cur[q].e - cur[q].s > 1)
                accumulate_occurrences<T, LPR, CPL, false>(a, cur[q].s + 1, cur[q].e, g, sub, grp, gmask);

            apply_row_update<T, LPR, CPL, OPT, STREAM>(a, row, valid[q], h_old[q], g, w, sub, gmask);
        }
        u0 += stride;
#pragma unroll
        for (int q = 0; q < UR; ++q) cur[q] = PREFETCH ? nxt[q] : load_row_meta(a, u0 + q * RPW + grp, nu);
    }
}

// Hot rows, step 1: one thread per unique row; a row with more than kLongSegment occurrences becomes a
// record plus one work item per chunk of `hot_chunk` occurrences.  (Slots are handed out by atomics, so
// their order varies from run to run; the sums below do not depend on it.)
__global__ void __launch_bounds__(256)
hot_rows_enqueue_kernel(const uint32_t* __restrict__ seg_starts, const uint32_t* __restrict__ num_unique,
                        const uint32_t* __restrict__ unique_keys, uint32_t total_rows, uint32_t* hdr, uint4* recs,
                        uint2* items, uint32_t rec_cap, uint32_t item_cap, uint32_t part_cap, uint32_t chunk,
                        int32_t* err_flag, const uint32_t* long_hint) {
    if (long_hint != nullptr && *long_hint == 0u) return;     // the run-length pass saw no run longer than 32 keys
    const uint32_t nu = *num_unique;
    const int lane = lane_id();
    // whole warps walk the segment starts: one coalesced load per step, the neighbour's start by shuffle
    for (uint32_t u0 = (blockIdx.x * blockDim.x + threadIdx.x) & ~31u; u0 < nu; u0 += gridDim.x * blockDim.x) {
        const uint32_t u = u0 + lane;
        const uint32_t s = (u <= nu) ? __ldg(seg_starts + u) : 0u;
        uint32_t e = __shfl_down_sync(kFullMask, s, 1);
        if (lane == 31) e = (u + 1 <= nu) ? __ldg(seg_starts + u + 1) : 0u;
        if (u >= nu) continue;
        const uint32_t len = e - s;
        if (len <= kLongSegment || __ldg(unique_keys + u) >= total_rows) continue;
        const uint32_t nch = (len + chunk - 1) / chunk;
        const uint32_t slot = atomicAdd(hdr + 0, 1u);
        const uint32_t ibase = atomicAdd(hdr + 1, nch);
        const uint32_t pbase = nch > 1 ? atomicAdd(hdr + 2, nch) : 0u;
        if (slot >= rec_cap || ibase + nch > item_cap || (nch > 1 && pbase + nch > part_cap)) {
            if (err_flag) atomicOr(err_flag, TDFO_FLAG_SCAN_TIMEOUT);      // cannot happen for a correctly sized worklist
            continue;
        }
        recs[slot] = make_uint4(u, pbase, nch, 0u);
        for (uint32_t c = 0; c < nch; ++c) items[ibase + c] = make_uint2(slot, c);
    }
}

// Hot rows, step 2: persistent CTAs take work items (row, chunk) from a ticket.  The G = 256 / LPR groups
// of the CTA each reduce one contiguous share of the chunk's occurrences in increasing order, the partial
// sums meet in shared memory and are added in group order.  A row of one chunk is updated right away; for a
// longer row the chunk sum goes to its partial-sum row, and the CTA that finishes the row's last open chunk
// adds the partial rows in chunk order and applies the optimizer.  Every sum has a fixed order, so the
// result does not depend on timing or on which CTA did what.
template <typename T, int LPR, int CPL, int OPT>
__global__ void __launch_bounds__(kUpdCtaThreads)
sparse_update_long_kernel(const UpdateArgs a) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int G = kUpdCtaThreads / LPR;
    __shared__ float s_part[kLongMaxFloats];
    __shared__ uint32_t s_item;
    __shared__ int s_last;
    const int lane = lane_id();
    const int sub = lane % LPR;
    const int grp_in_warp = lane / LPR;
    const int g_id = threadIdx.x / LPR;
    const unsigned gmask = (LPR == 32) ? kFullMask : (((1u << (LPR & 31)) - 1u) << (grp_in_warp * LPR));
    const uint32_t n_items = min(a.hot_hdr[1], a.item_cap);
    uint4* __restrict__ wbase = reinterpret_cast<uint4*>(a.weights);
    const int rowf = a.nchunks * E;                      // floats per row
    // groups that fit the shared buffer take part in the reduction (all of them for rows <= 1 KB)
    const int g_used = min(G, kLongMaxFloats / rowf);
    while (true) {
        if (threadIdx.x == 0) s_item = atomicAdd(a.hot_hdr + 3, 1u);
        __syncthreads();
        const uint32_t item = s_item;
        if (item >= n_items) break;
        const uint2 it = a.hot_items[item];
        const uint4 rec = a.hot_recs[it.x];
        const uint32_t u = rec.x, pbase = rec.y, nch = rec.z, c = it.y;
        const uint32_t row = __ldg(a.unique_keys + u);
        const uint32_t s0 = __ldg(a.seg_starts + u), e0 = __ldg(a.seg_starts + u + 1);
        const uint32_t lo_c = s0 + c * a.hot_chunk, hi_c = min(e0, lo_c + a.hot_chunk);
        float g[CPL][E];
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) g[k][i] = 0.f;
        if (g_id < g_used) {
            // group g_id reduces the g_id-th contiguous share of the chunk, in order
            const uint32_t share = (hi_c - lo_c + g_used - 1) / g_used;
            const uint32_t lo = min(hi_c, lo_c + (uint32_t)g_id * share), hi = min(hi_c, lo + share);
            accumulate_occurrences<T, LPR, CPL, true>(a, lo, hi, g, sub, grp_in_warp, gmask);
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const int cc = sub + k * LPR;
                if (cc < a.nchunks)
#pragma unroll
                    for (int i = 0; i < E; ++i) s_part[g_id * rowf + cc * E + i] = g[k][i];
            }
        }
        __syncthreads();
        if (g_id == 0) {
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const int cc = sub + k * LPR;
#pragma unroll
                for (int i = 0; i < E; ++i) {
                    float acc = 0.f;
                    if (cc < a.nchunks)
                        for (int gi = 0; gi < g_used; ++gi) acc += s_part[gi * rowf + cc * E + i];
                    g[k][i] = acc;
                }
            }
        }
        bool apply = g_id == 0;
        if (nch > 1) {                                   // uniform over the CTA
            float* prow = a.hot_parts + (size_t)pbase * rowf;
            if (g_id == 0) {
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    const int cc = sub + k * LPR;
                    if (cc < a.nchunks)
#pragma unroll
                        for (int i = 0; i < E; ++i) __stcg(prow + (size_t)c * rowf + cc * E + i, g[k][i]);
                }
                __threadfence();
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                const uint32_t done = atomicAdd(reinterpret_cast<uint32_t*>(a.hot_recs + it.x) + 3, 1u);
                s_last = done + 1 == nch;
            }
            __syncthreads();
            apply = apply && s_last;
            if (apply) {                                 // last chunk of the row to finish: add the partial rows in chunk order
                __threadfence();
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    const int cc = sub + k * LPR;
#pragma unroll
                    for (int i = 0; i < E; ++i) {
                        float acc = 0.f;
                        if (cc < a.nchunks)
                            for (uint32_t pc = 0; pc < nch; ++pc) acc += __ldcg(prow + (size_t)pc * rowf + cc * E + i);
                        g[k][i] = acc;
                    }
                }
            }
        }
        if (apply) {
            float w[CPL][E];
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const int cc = sub + k * LPR;
                Chunk<T>::unpack(ld_row_128(wbase + (size_t)row * a.nchunks + min(cc, a.nchunks - 1)), w[k]);
            }
            const float h_old = (OPT == TDFO_OPT_ROWWISE_ADAGRAD) ? __ldcg(a.state0 + row) : 0.f;
            apply_row_update<T, LPR, CPL, OPT>(a, row, true, h_old, g, w, sub, gmask);
        }
        __syncthreads();
    }
}

// Hot rows (if the caller gave a worklist): the block-per-chunk kernel runs after the main one.
template <typename T, int LPR, int CPL>
static int launch_update_long(const UpdateArgs& a, int opt, cudaStream_t st) {
    if (a.hot_hdr == nullptr) return TDFO_OK;
    const unsigned lb = kNumSMs * 2;
    switch (opt) {
        case TDFO_OPT_SGD:
            sparse_update_long_kernel<T, LPR, CPL, TDFO_OPT_SGD><<<lb, kUpdCtaThreads, 0, st>>>(a); break;
        case TDFO_OPT_ROWWISE_ADAGRAD:
            sparse_update_long_kernel<T, LPR, CPL, TDFO_OPT_ROWWISE_ADAGRAD><<<lb, kUpdCtaThreads, 0, st>>>(a); break;
        default:
            sparse_update_long_kernel<T, LPR, CPL, TDFO_OPT_ADAM><<<lb, kUpdCtaThreads, 0, st>>>(a); break;
    }
    count_launch();
    TDFO_LAUNCH_CHECK("sparse_update_long");
    return TDFO_OK;
}

template <typename T, int LPR, int CPL, int UR, int MINB, bool PREFETCH = true, int THREADS = kUpdCtaThreads, bool STREAM = false>
static int launch_update(const UpdateArgs& a, int opt, int64_t max_unique, cudaStream_t st) {
    constexpr int STEP = (32 / LPR) * UR;
    const int64_t steps = (max_unique + STEP - 1) / STEP;             // one warp step per STEP rows
    int64_t blocks = (steps + THREADS / 32 - 1) / (THREADS / 32);
    const int64_t cap = (int64_t)kNumSMs * MINB;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    switch (opt) {
        case TDFO_OPT_SGD:
            sparse_update_kernel<T, LPR, CPL, UR, MINB, TDFO_OPT_SGD, PREFETCH, THREADS, STREAM><<<(unsigned)blocks, THREADS, 0, st>>>(a); break;
        case TDFO_OPT_ROWWISE_ADAGRAD:
            sparse_update_kernel<T, LPR, CPL, UR, MINB, TDFO_OPT_ROWWISE_ADAGRAD, PREFETCH, THREADS, STREAM><<<(unsigned)blocks, THREADS, 0, st>>>(a); break;
        default:
            sparse_update_kernel<T, LPR, CPL, UR, MINB, TDFO_OPT_ADAM, PREFETCH, THREADS, STREAM><<<(unsigned)blocks, THREADS, 0, st>>>(a); break;
    }
    count_launch();
    TDFO_LAUNCH_CHECK("sparse_update");
    return launch_update_long<T, LPR, CPL>(a, opt, st);
}

// Shapes measured on B200 (tools/tune_update.py, tools/tune_update_multi.py, BASELINE config 2): 8 lanes x 4 chunks
// per 512 B row, one row per group per step; anything that spills, or 32 lanes per row, is 15-60 % slower.
#ifndef TDFO_UPD_PREFETCH
#define TDFO_UPD_PREFETCH 0
#endif
#ifndef TDFO_UPD_MINB
#define TDFO_UPD_MINB 3
#endif
#ifndef TDFO_UPD_STREAM
#define TDFO_UPD_STREAM 2      // weight rows evict-first: 0 never, 1 always, 2 when one feature's gradient slice is large
#endif
constexpr int64_t kStreamSliceBytes = 32ll << 20;

// Shapes measured on B200 for D = 128 rows (tools/tune_update_multi.py, profiles/r2_tuning_kernels.txt): the kernel is
// occupancy-bound up to 27-28 warps per SM and falls off a cliff past that (registers: 30 warps leave 64 per
// thread, which spills), so the CTA is 288 threads x 3 per SM at 72 registers rather than 256 x 3 at 80.
template <typename T>
static int dispatch_update(const UpdateArgs& a, int opt, int64_t max_unique, int64_t grad_slice_bytes, cudaStream_t st) {
    const int n = a.nchunks;
    constexpr bool kPf = TDFO_UPD_PREFETCH != 0;
    constexpr int kMb = TDFO_UPD_MINB;
    constexpr int kTh = kUpdMainThreads;
    const bool stream = TDFO_UPD_STREAM == 2 ? grad_slice_bytes > kStreamSliceBytes : TDFO_UPD_STREAM != 0;
#ifndef TDFO_UPD_LPR512
#define TDFO_UPD_LPR512 8
#endif
    constexpr int kL = TDFO_UPD_LPR512, kC = 32 / TDFO_UPD_LPR512;     // lanes x chunks for a 512-byte row
    constexpr bool kWide = sizeof(T) == 4;          // fp32: 4 chunks per lane; bf16 (8 elems / chunk): 2
    if (kWide) {
        if (n <= 1)   return launch_update<T, 1, 1, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 2)   return launch_update<T, 1, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 4)   return launch_update<T, 1, 4, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 8)   return launch_update<T, 2, 4, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 16)  return launch_update<T, 4, 4, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 32)  return stream ? launch_update<T, kL, kC, 1, kMb, kPf, kTh, true>(a, opt, max_unique, st)
                                    : launch_update<T, kL, kC, 1, kMb, kPf, kTh, false>(a, opt, max_unique, st);
        if (n <= 64)  return launch_update<T, 16, 4, 1, 2, false>(a, opt, max_unique, st);
        if (n <= 128) return launch_update<T, 32, 4, 1, 2, false>(a, opt, max_unique, st);
        if (n <= 256) return launch_update<T, 32, 8, 1, 2, false>(a, opt, max_unique, st);
    } else {
        if (n <= 1)   return launch_update<T, 1, 1, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 2)   return launch_update<T, 1, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 4)   return launch_update<T, 2, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 8)   return launch_update<T, 4, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 16)  return stream ? launch_update<T, 8, 2, 1, kMb, kPf, kTh, true>(a, opt, max_unique, st)
                                    : launch_update<T, 8, 2, 1, kMb, kPf, kTh, false>(a, opt, max_unique, st);
        if (n <= 32)  return launch_update<T, 16, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 64)  return launch_update<T, 32, 2, 1, 3, false>(a, opt, max_unique, st);
        if (n <= 128) return launch_update<T, 32, 4, 1, 2, false>(a, opt, max_unique, st);
        if (n <= 256) return launch_update<T, 32, 8, 1, 2, false>(a, opt, max_unique, st);
    }
    set_error("sparse_update: row of %d bytes exceeds the 4096-byte maximum", n * 16);
    return TDFO_ERR_UNSUPPORTED;
}

static int key_bits_for(int64_t total_rows) {   // bits needed to hold the value total_rows itself
    int b = 1;
    while (b < 32 && (total_rows >> b) != 0) ++b;
    return b;
}

static int check_common(const char* who, int32_t weights_dtype, int64_t total_rows, int32_t D) {
    if (weights_dtype != TDFO_F32 && weights_dtype != TDFO_BF16) {
        set_error("%s: unknown weights_dtype %d", who, weights_dtype);
        return TDFO_ERR_INVALID_ARG;
    }
    const size_t esz = weights_dtype == TDFO_F32 ? 4 : 2;
    if (D <= 0 || (D * esz) % 16 != 0) {
        set_error("%s: row size D*elem = %zu bytes is not a positive multiple of 16", who, (size_t)D * esz);
        return TDFO_ERR_INVALID_ARG;
    }
    if (total_rows < 0 || total_rows >= 0xffffffffll) {
        set_error("%s: total_rows %lld must be < 2^32-1", who, (long long)total_rows);
        return TDFO_ERR_INVALID_ARG;
    }
    return TDFO_OK;
}

static int check_opt(const char* who, const tdfo_optimizer_args* opt, const float* state0, const float* state1) {
    if (!opt) { set_error("%s: optimizer args are null", who); return TDFO_ERR_INVALID_ARG; }
    if (opt->optimizer < TDFO_OPT_SGD || opt->optimizer > TDFO_OPT_ADAM) {
        set_error("%s: unknown optimizer %d", who, opt->optimizer);
        return TDFO_ERR_INVALID_ARG;
    }
    if (opt->optimizer == TDFO_OPT_ROWWISE_ADAGRAD && !state0) {
        set_error("%s: row-wise Adagrad needs state0 (float[total_rows])", who);
        return TDFO_ERR_INVALID_ARG;
    }
    if (opt->optimizer == TDFO_OPT_ADAM && (!state0 || !state1 || opt->step < 1)) {
        set_error("%s: ADAM needs state0, state1 (float[total_rows, D]) and step >= 1", who);
        return TDFO_ERR_INVALID_ARG;
    }
    return TDFO_OK;
}

int sparse_update_impl(void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                       float* state0, float* state1, const float* grad_out, int64_t grad_stride,
                       const uint32_t* sorted_vals, const uint32_t* unique_keys,
                       const uint32_t* seg_starts, const uint32_t* first_vals, const uint32_t* num_unique,
                       const int64_t* offsets, const int64_t* feat_out_col, int32_t F, int64_t B,
                       int64_t L, int32_t pooling_mode, const tdfo_optimizer_args* opt,
                       uint32_t* worklist, int64_t worklist_entries, const float* psw, const uint32_t* pos_bag,
                       int32_t* err_flag, cudaStream_t st, const uint32_t* long_hint = nullptr) {
    if (L == 0) return TDFO_OK;
    const size_t esz = weights_dtype == TDFO_F32 ? 4 : 2;
    UpdateArgs a;
    a.hot_hdr = nullptr; a.hot_recs = nullptr; a.hot_items = nullptr; a.hot_parts = nullptr;
    a.rec_cap = a.item_cap = a.part_cap = 0; a.hot_chunk = kHotChunkMin;
    if (worklist != nullptr) {
        const HotLayout h = hot_layout(L);
        if ((size_t)worklist_entries < h.words || !aligned(worklist, 16)) {
            set_error("sparse_update: the hot-row worklist needs %zu 16-byte-aligned uint32 entries for L = %lld "
                      "(tdfo_hot_row_worklist_entries), got %lld", h.words, (long long)L, (long long)worklist_entries);
            return TDFO_ERR_SCRATCH;
        }
        TDFO_CUDA(cudaMemsetAsync(worklist, 0, 16 * sizeof(uint32_t), st));
        a.hot_hdr = worklist;
        a.hot_recs = reinterpret_cast<uint4*>(worklist + h.off_recs);
        a.hot_items = reinterpret_cast<uint2*>(worklist + h.off_items);
        a.hot_parts = reinterpret_cast<float*>(worklist + h.off_parts);
        a.rec_cap = h.rec_cap; a.item_cap = h.item_cap; a.hot_chunk = h.chunk;
        // partial rows are `row floats` apart: a narrower row leaves room for more of them than the layout assumed
        a.part_cap = (uint32_t)((size_t)h.part_cap * kHotMaxRowFloats / (size_t)D);
        const int64_t mu = L < total_rows + 1 ? L : total_rows + 1;
        int64_t eb = (mu + 255) / 256;
        if (eb > kNumSMs * 8) eb = kNumSMs * 8;
        hot_rows_enqueue_kernel<<<(unsigned)eb, 256, 0, st>>>(seg_starts, num_unique, unique_keys, (uint32_t)total_rows,
                                                             a.hot_hdr, a.hot_recs, a.hot_items, a.rec_cap, a.item_cap,
                                                             a.part_cap, a.hot_chunk, err_flag, long_hint);
        count_launch();
        TDFO_LAUNCH_CHECK("hot_rows_enqueue");
    }
    a.weights = weights; a.state0 = state0; a.state1 = state1; a.grad = grad_out;
    a.sorted_vals = sorted_vals; a.unique_keys = unique_keys; a.seg_starts = seg_starts; a.first_vals = first_vals;
    a.num_unique = num_unique; a.offsets = offsets; a.feat_out_col = feat_out_col;
    a.psw = psw; a.pos_bag = pos_bag;
    a.grad_stride = grad_stride; a.B = B;
    a.B_magic = B > 1 ? (~0ull / (uint64_t)B + 1ull) : 0ull; a.total_rows = (uint32_t)total_rows; a.L = (uint32_t)L; a.D = D;
    a.nchunks = (int32_t)(D * esz / 16); a.pooling_mode = pooling_mode;
    a.inv_D = 1.f / (float)D;
    a.lr = opt->learning_rate; a.eps = opt->eps; a.wd = opt->weight_decay;
    a.beta1 = opt->beta1; a.beta2 = opt->beta2;
    a.bc1 = 1.f; a.bc2 = 1.f;
    if (opt->optimizer == TDFO_OPT_ADAM) {
        a.bc1 = (float)(1.0 - pow((double)opt->beta1, (double)opt->step));
        a.bc2 = (float)(1.0 - pow((double)opt->beta2, (double)opt->step));
    }
    const int64_t max_unique = L < total_rows + 1 ? L : total_rows + 1;
    // bytes of gradient rows one feature's rows draw from (they are re-read out of L2 in sorted order)
    const int64_t slice_rows = pooling_mode == TDFO_POOL_NONE ? L / (F > 0 ? F : 1) : B;
    const int64_t slice = slice_rows * (int64_t)D * 4;
    return weights_dtype == TDFO_F32 ? dispatch_update<float>(a, opt->optimizer, max_unique, slice, st)
                                     : dispatch_update<__nv_bfloat16>(a, opt->optimizer, max_unique, slice, st);
}

static size_t pair_array_bytes(int64_t L) { return align_up((size_t)(L + 1) * sizeof(uint32_t), 256); }

}  // namespace tdfo

using namespace tdfo;

static int linearize_impl(const int64_t* indices, const int64_t* offsets,
                          const int64_t* feat_row_base, const int64_t* feat_rows,
                          int32_t F, int64_t B, int64_t L, int64_t total_rows, int32_t pooling_mode,
                          uint32_t* keys, uint32_t* vals, uint32_t* pos_bag, int32_t* err_flag, void* stream) {
    TDFO_CHECK_ARG(F >= 0 && B >= 0, "linearize_indices: negative size");
    TDFO_CHECK_ARG(total_rows >= 0 && total_rows < 0xffffffffll, "linearize_indices: total_rows must be < 2^32-1");
    TDFO_CHECK_ARG((int64_t)F * B < (1ll << 32), "linearize_indices: F*B must be < 2^32");
    if (F == 0 || B == 0) return TDFO_OK;
    // indices / keys / vals may be null when every bag is empty
    TDFO_CHECK_ARG(offsets && feat_row_base && feat_rows, "linearize_indices: null pointer");
    const int64_t num_bags = (int64_t)F * B;
    const int64_t blocks = (num_bags + kUpdThreads / 32 - 1) / (kUpdThreads / 32);
    linearize_kernel<<<(unsigned)blocks, kUpdThreads, 0, static_cast<cudaStream_t>(stream)>>>(
        indices, offsets, feat_row_base, feat_rows, B, num_bags, L, (uint32_t)total_rows,
        pooling_mode == TDFO_POOL_NONE || pos_bag != nullptr, keys, vals, pos_bag, err_flag);
    count_launch();
    TDFO_LAUNCH_CHECK("linearize_indices");
    return TDFO_OK;
}

extern "C" int tdfo_linearize_indices(const int64_t* indices, const int64_t* offsets,
                                      const int64_t* feat_row_base, const int64_t* feat_rows,
                                      int32_t F, int64_t B, int64_t total_rows, int32_t pooling_mode,
                                      uint32_t* keys, uint32_t* vals, int32_t* err_flag, void* stream) {
    // this entry point is not told how many ids there are: offsets are taken at face value
    return linearize_impl(indices, offsets, feat_row_base, feat_rows, F, B, INT64_MAX, total_rows, pooling_mode, keys,
                          vals, nullptr, err_flag, stream);
}

extern "C" int tdfo_sparse_update(void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                                  float* state0, float* state1, const float* grad_out, int64_t grad_stride,
                                  const uint32_t* sorted_vals, const uint32_t* unique_keys,
                                  const uint32_t* seg_starts, const uint32_t* first_vals, const uint32_t* num_unique,
                                  const int64_t* offsets, const int64_t* feat_out_col, int32_t F,
                                  int64_t B, int64_t L, int32_t pooling_mode,
                                  const tdfo_optimizer_args* opt,
                                  uint32_t* worklist, int64_t worklist_entries, void* stream) {
    int rc = check_common("sparse_update", weights_dtype, total_rows, D);
    if (rc) return rc;
    rc = check_opt("sparse_update", opt, state0, state1);
    if (rc) return rc;
    TDFO_CHECK_ARG(pooling_mode >= TDFO_POOL_SUM && pooling_mode <= TDFO_POOL_NONE, "sparse_update: unknown pooling_mode");
    TDFO_CHECK_ARG(L >= 0 && L < (1ll << 30), "sparse_update: L must be in [0, 2^30)");
    if (L == 0) return TDFO_OK;
    TDFO_CHECK_ARG(weights && grad_out && sorted_vals && unique_keys && seg_starts && first_vals && num_unique,
                   "sparse_update: null pointer");
    TDFO_CHECK_ARG(aligned(weights, 16) && aligned(grad_out, 16), "sparse_update: weights/grad_out not 16-byte aligned");
    if (pooling_mode != TDFO_POOL_NONE) {
        TDFO_CHECK_ARG(feat_out_col != nullptr, "sparse_update: feat_out_col is null");
        TDFO_CHECK_ARG(grad_stride % 4 == 0, "sparse_update: grad_stride not a multiple of 4");
    }
    TDFO_CHECK_ARG(pooling_mode != TDFO_POOL_MEAN || offsets, "sparse_update: MEAN pooling needs offsets");
    return sparse_update_impl(weights, weights_dtype, total_rows, D, state0, state1, grad_out, grad_stride,
                              sorted_vals, unique_keys, seg_starts, first_vals, num_unique, offsets, feat_out_col,
                              F, B, L, pooling_mode, opt, worklist, worklist_entries, nullptr, nullptr, nullptr,
                              static_cast<cudaStream_t>(stream));
}

extern "C" int tdfo_backward_scratch_layout(int64_t L, int64_t total_rows, tdfo_backward_layout* layout) {
    TDFO_CHECK_ARG(layout != nullptr, "backward_scratch_layout: null layout");
    TDFO_CHECK_ARG(L >= 0 && L < (1ll << 30), "backward_scratch_layout: L must be in [0, 2^30)");
    TDFO_CHECK_ARG(total_rows >= 0 && total_rows < 0xffffffffll, "backward_scratch_layout: total_rows must be < 2^32-1");
    const int bits = key_bits_for(total_rows);
    const size_t A = pair_array_bytes(L);
    const int passes = radix_sort_passes_impl(bits);
    const int res = passes & 1;               // buffer pair the sorted result lands in
    layout->key_bits = bits;
    layout->sorted_keys = (res ? 2 : 0) * A;
    layout->sorted_vals = (res ? 3 : 1) * A;
    layout->unique_keys = (res ? 0 : 2) * A;  // the other pair is free after the sort
    layout->seg_starts  = (res ? 1 : 3) * A;
    layout->first_vals  = 4 * A;
    layout->input_keys = 0;
    layout->input_vals = A;
    layout->num_unique = 5 * A;
    layout->worklist = 5 * A + 256;
    layout->worklist_entries = hot_layout(L).words;                // records, work items, partial-sum rows
    const size_t sort_off = layout->worklist + align_up(layout->worklist_entries * sizeof(uint32_t), 256);
    const size_t scan_off = sort_off + radix_sort_scratch_bytes_impl(L, bits);
    layout->total_bytes = scan_off + tdfo_scan_scratch_bytes(L);
    return TDFO_OK;
}

extern "C" size_t tdfo_hot_row_worklist_entries(int64_t L) { return hot_layout(L).words; }

extern "C" size_t tdfo_backward_scratch_bytes(int64_t L, int64_t total_rows) {
    tdfo_backward_layout lay;
    if (tdfo_backward_scratch_layout(L, total_rows, &lay) != TDFO_OK) return 0;
    return lay.total_bytes;
}

// With per-sample weights the sort payload is the id's position, and the bag of every position
// is kept in one more uint32[L] array placed after the unweighted layout.
extern "C" size_t tdfo_backward_scratch_bytes_weighted(int64_t L, int64_t total_rows) {
    const size_t base = tdfo_backward_scratch_bytes(L, total_rows);
    return base ? align_up(base, 256) + pair_array_bytes(L) : 0;
}

// stages: bit 0 = linearise (unless the pairs are already in scratch) + sort + unique, bit 1 = update
static int backward_stages(int stages, const char* who, void* weights, int32_t weights_dtype, int64_t total_rows,
                           int32_t D, float* state0, float* state1, const float* grad_out, int64_t grad_stride,
                           const int64_t* indices, const int64_t* offsets, const float* per_sample_weights,
                           bool weighted, const int64_t* feat_row_base, const int64_t* feat_rows,
                           const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L, int32_t pooling_mode,
                           const tdfo_optimizer_args* opt, int32_t keys_in_scratch, int32_t* err_flag, void* scratch,
                           size_t scratch_bytes, void* stream, const uint32_t* ext_keys = nullptr,
                           const uint32_t* ext_vals = nullptr) {
    tdfo_backward_layout lay;
    int rc = tdfo_backward_scratch_layout(L, total_rows, &lay);
    if (rc) return rc;
    const size_t need = weighted ? tdfo_backward_scratch_bytes_weighted(L, total_rows) : lay.total_bytes;
    if (!scratch || !aligned(scratch, 256) || scratch_bytes < need) {
        set_error("%s: scratch null, misaligned or smaller than %zu bytes", who, need);
        return TDFO_ERR_SCRATCH;
    }
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    char* base = static_cast<char*>(scratch);
    const size_t A = pair_array_bytes(L);
    uint32_t* k0 = reinterpret_cast<uint32_t*>(base);
    uint32_t* v0 = reinterpret_cast<uint32_t*>(base + A);
    uint32_t* k1 = reinterpret_cast<uint32_t*>(base + 2 * A);
    uint32_t* v1 = reinterpret_cast<uint32_t*>(base + 3 * A);
    uint32_t* num_unique = reinterpret_cast<uint32_t*>(base + lay.num_unique);
    const size_t sort_off = lay.worklist + align_up(lay.worklist_entries * sizeof(uint32_t), 256);
    void* sort_scratch = base + sort_off;
    void* scan_scratch = base + sort_off + radix_sort_scratch_bytes_impl(L, lay.key_bits);
    uint32_t* pos_bag = weighted ? reinterpret_cast<uint32_t*>(base + align_up(lay.total_bytes, 256)) : nullptr;
    uint32_t* sk = reinterpret_cast<uint32_t*>(base + lay.sorted_keys);
    uint32_t* sv = reinterpret_cast<uint32_t*>(base + lay.sorted_vals);
    uint32_t* uk = reinterpret_cast<uint32_t*>(base + lay.unique_keys);
    uint32_t* ss = reinterpret_cast<uint32_t*>(base + lay.seg_starts);
    uint32_t* fv = reinterpret_cast<uint32_t*>(base + lay.first_vals);

    if (stages & 1) {
        if (!keys_in_scratch && !ext_keys) {
            rc = linearize_impl(indices, offsets, feat_row_base, feat_rows, F, B, L, total_rows, pooling_mode,
                                k0, v0, pos_bag, err_flag, stream);
            if (rc) return rc;
        }
        int where = 0;
        rc = radix_sort_impl(k0, v0, k1, v1, L, lay.key_bits, err_flag, sort_scratch, st, &where, ext_keys, ext_vals);
        if (rc) return rc;
        if ((where ? k1 : k0) != sk) {
            set_error("%s: internal layout mismatch", who);
            return TDFO_ERR_INVALID_ARG;
        }
        rc = unique_segments_impl(sk, sv, L, uk, ss, fv, num_unique, err_flag, scan_scratch, st, num_unique + 1);
        if (rc) return rc;
    }
    if (stages & 2) {
        rc = sparse_update_impl(weights, weights_dtype, total_rows, D, state0, state1, grad_out, grad_stride,
                                sv, uk, ss, fv, num_unique, offsets, feat_out_col, F, B, L, pooling_mode, opt,
                                reinterpret_cast<uint32_t*>(base + lay.worklist), (int64_t)lay.worklist_entries,
                                per_sample_weights, pos_bag, err_flag, st, num_unique + 1);
        if (rc) return rc;
    }
    return TDFO_OK;
}

extern "C" int tdfo_embedding_backward_prepare(const int64_t* indices, const int64_t* offsets,
                                               const int64_t* feat_row_base, const int64_t* feat_rows,
                                               int32_t F, int64_t B, int64_t L, int64_t total_rows,
                                               int32_t pooling_mode, int32_t weighted, int32_t keys_in_scratch,
                                               int32_t* err_flag, void* scratch, size_t scratch_bytes,
                                               void* stream) {
    TDFO_CHECK_ARG(total_rows >= 0 && total_rows < 0xffffffffll, "embedding_backward_prepare: total_rows must be < 2^32-1");
    TDFO_CHECK_ARG(pooling_mode >= TDFO_POOL_SUM && pooling_mode <= TDFO_POOL_NONE,
                   "embedding_backward_prepare: unknown pooling_mode");
    TDFO_CHECK_ARG(F >= 0 && B >= 0 && L >= 0 && L < (1ll << 30), "embedding_backward_prepare: size out of range");
    TDFO_CHECK_ARG(keys_in_scratch == 0 || keys_in_scratch == 1, "embedding_backward_prepare: keys_in_scratch must be 0 or 1");
    TDFO_CHECK_ARG(!(weighted && (pooling_mode == TDFO_POOL_NONE || keys_in_scratch)),
                   "embedding_backward_prepare: weighted needs a pooled mode and keys_in_scratch = 0");
    if (F == 0 || B == 0 || L == 0) return TDFO_OK;
    TDFO_CHECK_ARG(indices && offsets && feat_row_base && feat_rows, "embedding_backward_prepare: null pointer");
    return backward_stages(1, "embedding_backward_prepare", nullptr, TDFO_F32, total_rows, 4, nullptr, nullptr, nullptr,
                           0, indices, offsets, nullptr, weighted != 0, feat_row_base, feat_rows, nullptr, F, B, L,
                           pooling_mode, nullptr, keys_in_scratch, err_flag, scratch, scratch_bytes, stream);
}

extern "C" int tdfo_embedding_backward_update(void* weights, int32_t weights_dtype, int64_t total_rows,
                                              int32_t D, float* state0, float* state1,
                                              const float* grad_out, int64_t grad_stride,
                                              const int64_t* indices, const int64_t* offsets,
                                              const float* per_sample_weights,
                                              const int64_t* feat_row_base, const int64_t* feat_rows,
                                              const int64_t* feat_out_col, int32_t F, int64_t B,
                                              int64_t L, int32_t pooling_mode,
                                              const tdfo_optimizer_args* opt, int32_t keys_in_scratch,
                                              int32_t* err_flag,
                                              void* scratch, size_t scratch_bytes, void* stream) {
    int rc = check_common("embedding_backward_update", weights_dtype, total_rows, D);
    if (rc) return rc;
    rc = check_opt("embedding_backward_update", opt, state0, state1);
    if (rc) return rc;
    TDFO_CHECK_ARG(pooling_mode >= TDFO_POOL_SUM && pooling_mode <= TDFO_POOL_NONE,
                   "embedding_backward_update: unknown pooling_mode");
    TDFO_CHECK_ARG(F >= 0 && B >= 0 && L >= 0 && L < (1ll << 30), "embedding_backward_update: size out of range");
    TDFO_CHECK_ARG(keys_in_scratch >= 0 && keys_in_scratch <= TDFO_SCRATCH_PREPARED,
                   "embedding_backward_update: keys_in_scratch must be 0, 1 or TDFO_SCRATCH_PREPARED");
    const bool weighted = per_sample_weights != nullptr;
    if (weighted && pooling_mode == TDFO_POOL_NONE) {
        set_error("embedding_backward_update: per-sample weights need a pooled mode (SUM / MEAN)");
        return TDFO_ERR_INVALID_ARG;
    }
    if (weighted && keys_in_scratch == 1) {
        set_error("embedding_backward_update: keys_in_scratch (pairs written by the forward) cannot be combined "
                  "with per-sample weights");
        return TDFO_ERR_INVALID_ARG;
    }
    if (F == 0 || B == 0 || L == 0) return TDFO_OK;
    const bool prepared = keys_in_scratch == TDFO_SCRATCH_PREPARED;
    TDFO_CHECK_ARG(weights && grad_out && offsets && (prepared || (indices && feat_row_base && feat_rows)),
                   "embedding_backward_update: null pointer");
    TDFO_CHECK_ARG(aligned(weights, 16) && aligned(grad_out, 16), "embedding_backward_update: weights/grad_out not 16-byte aligned");
    if (pooling_mode != TDFO_POOL_NONE) {
        TDFO_CHECK_ARG(feat_out_col != nullptr, "embedding_backward_update: feat_out_col is null");
        TDFO_CHECK_ARG(grad_stride % 4 == 0, "embedding_backward_update: grad_stride not a multiple of 4");
    }
    return backward_stages(prepared ? 2 : 3, "embedding_backward_update", weights, weights_dtype, total_rows, D, state0,
                           state1, grad_out, grad_stride, indices, offsets, per_sample_weights, weighted, feat_row_base,
                           feat_rows, feat_out_col, F, B, L, pooling_mode, opt, prepared ? 0 : keys_in_scratch, err_flag,
                           scratch, scratch_bytes, stream);
}

extern "C" int tdfo_embedding_backward_prepare_pairs(const uint32_t* keys, const uint32_t* vals, int64_t L,
                                                     int64_t total_rows, int32_t* err_flag, void* scratch,
                                                     size_t scratch_bytes, void* stream) {
    TDFO_CHECK_ARG(total_rows >= 0 && total_rows < 0xffffffffll, "embedding_backward_prepare_pairs: total_rows must be < 2^32-1");
    TDFO_CHECK_ARG(L >= 0 && L < (1ll << 30), "embedding_backward_prepare_pairs: L must be in [0, 2^30)");
    if (L == 0) return TDFO_OK;
    TDFO_CHECK_ARG(keys && vals, "embedding_backward_prepare_pairs: null pointer");
    return backward_stages(1, "embedding_backward_prepare_pairs", nullptr, TDFO_F32, total_rows, 4, nullptr, nullptr,
                           nullptr, 0, nullptr, nullptr, nullptr, false, nullptr, nullptr, nullptr, 0, 0, L, TDFO_POOL_SUM,
                           nullptr, 0, err_flag, scratch, scratch_bytes, stream, keys, vals);
}
