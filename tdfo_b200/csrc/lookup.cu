// Embedding lookup (forward): offsets-driven gather of table rows with 128-bit
// loads and warp-level segmented pooling.  HBM-bound: the algorithmic traffic is
// L * (D*w + 8) + bags * (D*4 + 8) bytes, all of it streamed exactly once.
//
// Mapping (D = 128 fp32, the BASELINE shape): one warp per bag at a time, one 16-byte
// chunk of the row per lane, so a row is a single fully coalesced 512 B request.
// Smaller rows split the warp into 32/LPR sub-groups that gather different rows
// of the same bag in parallel and are combined with a shuffle tree at the end.
#include "common.cuh"

namespace tdfo {

struct LookupArgs {
    const void* weights;
    const int64_t* indices;
    const int64_t* offsets;
    const float* psw;
    const int64_t* feat_row_base;
    const int64_t* feat_rows;
    const int64_t* feat_out_col;
    float* out;
    const uint64_t* peer_out;   // optional: out base pointer of every source rank (NVLink peer memory)
    int64_t peer_batch;         // samples per source rank when peer_out is set
    const int64_t* peer_pos;    // sequence mode over peer memory: output row of every id at its source rank
    int32_t peer_world;         //   ... and the number of ranks (segment s belongs to source rank s % world)
    uint32_t* sort_keys;   // optional: backward's (row key, payload) pairs, emitted while the ids are in registers
    uint32_t* sort_vals;
    int32_t* err_flag;
    int64_t B;
    int64_t num_bags;     // F * B
    int64_t L;
    int64_t L_shape;      // number of ids the launch shape is chosen for (== L unless the caller knows better)
    int64_t out_stride;
    int32_t F;
    int32_t nchunks;      // 16-byte chunks per row
    int32_t mean;
    uint32_t oob_key;     // key given to out-of-range ids (== total_rows)
};

constexpr int kFwdThreads = 256;
#ifndef TDFO_FWD_THREADS
#define TDFO_FWD_THREADS 288
#endif
#ifndef TDFO_FWD_MINB
#define TDFO_FWD_MINB 3
#endif
// Bag-per-warp pooled kernel: persistent grid of 148 x 3 CTAs of 288 threads = 27 warps per SM at 72 registers.
// Measured on B200 (profiles/r2_tuning_kernels.txt): 24 warps (256 x 3, 80 registers) 6260 GB/s, 27-28 warps
// 6400-6430 GB/s, 30-32 warps (64 registers, spills) 5500 GB/s.
constexpr int kPoolThreads = TDFO_FWD_THREADS;
constexpr int kFwdUnroll = 8;      // independent row requests in flight per sub-group
constexpr int kFwdBlocksPerSM = TDFO_FWD_MINB;

// First chunk (<= 32 ids) of a bag, one id per lane, already turned into a row of the group.
struct BagHead {
    int64_t s, e;
    uint32_t row;   // 0xffffffff = no id in this lane / id out of range
    float w;
    uint32_t flags; // TDFO_FLAG_* bits raised while loading
};

__device__ __forceinline__ BagHead load_bag_head(const LookupArgs& a, int64_t g, int lane) {
    BagHead h;
    h.s = __ldg(a.offsets + g);
    h.e = __ldg(a.offsets + g + 1);
    h.row = 0xffffffffu;
    h.w = 1.f;
    h.flags = 0;
    if (h.e > a.L) {          // offsets reach past the ids we were given (e.g. a receive buffer that overflowed)
        h.flags = TDFO_FLAG_OFFSETS_RANGE;
        h.e = a.L;
        h.s = min(h.s, a.L);
    }
    if (h.s + lane < h.e) {
        const int f = (int)((uint64_t)g / (uint64_t)a.B);
        const int64_t idx = __ldg(a.indices + h.s + lane);
        if ((uint64_t)idx < (uint64_t)__ldg(a.feat_rows + f)) h.row = (uint32_t)(__ldg(a.feat_row_base + f) + idx);
        else h.flags |= TDFO_FLAG_INDEX_OOB;
        if (a.psw) h.w = __ldg(a.psw + h.s + lane);
        if (a.sort_keys) {
            a.sort_keys[h.s + lane] = (h.row == 0xffffffffu) ? a.oob_key : h.row;
            a.sort_vals[h.s + lane] = (uint32_t)g;
        }
    }
    return h;
}

// Persistent warps: warp w handles bags w, w + nwarps, ...  While the rows of bag g are in
// flight, the offsets and ids of the warp's NEXT bag are already being fetched into registers,
// so the only exposed latency per bag is the row gather itself (measured on B200 with
// tools/gather_micro.cu: 6.5 TB/s vs 4.5 TB/s for the same loop without the prefetch, and
// vs 5.3 TB/s for rows staged in shared memory by per-row bulk async copies).
// One batch of UB row requests per sub-group.  Every load is UNCONDITIONAL: lanes without a
// valid row read row 0 / the last chunk instead and are masked when the data is consumed, so
// the compiler can issue all UB*CPL loads back to back and only then start adding.
template <typename T, int LPR, int CPL, int UB, bool WEIGHTED>
__device__ __forceinline__ void gather_batch(const uint4* __restrict__ wbase, int nchunks, int j0, int n,
                                             uint32_t my_row, float my_w, int sub, int grp,
                                             float (&acc)[CPL][Chunk<T>::kElems]) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int RPW = 32 / LPR;
    uint4 v[UB][CPL];
    float w[UB];
    bool ok[UB];
#pragma unroll
    for (int u = 0; u < UB; ++u) {
        const int r = j0 + u * RPW + grp;
        const uint32_t row = __shfl_sync(kFullMask, my_row, r & 31);
        ok[u] = (r < n) && (row != 0xffffffffu);
        w[u] = WEIGHTED ? __shfl_sync(kFullMask, my_w, r & 31) : 1.f;
        const size_t rbase = (size_t)(ok[u] ? row : 0u) * nchunks;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = min(sub + k * LPR, nchunks - 1);
            v[u][k] = ldg_stream_128(wbase + rbase + c);
        }
    }
#pragma unroll
    for (int u = 0; u < UB; ++u)
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            float x[E];
            Chunk<T>::unpack(v[u][k], x);
#pragma unroll
            for (int i = 0; i < E; ++i) acc[k][i] = fmaf(w[u], ok[u] ? x[i] : 0.f, acc[k][i]);
        }
}

// Persistent warps: warp w handles bags w, w + nwarps, ...  While the rows of bag g are in
// flight, the offsets and ids of the warp's NEXT bag are already being fetched into registers,
// so the only exposed latency per bag is the row gather itself (measured on B200 with
// tools/gather_micro.cu: 6.5 TB/s vs 4.5 TB/s for the same loop without the prefetch, and
// vs 5.3 TB/s for rows staged in shared memory by per-row bulk async copies).
template <typename T, int LPR, int CPL, bool WEIGHTED>
__global__ void __launch_bounds__(kPoolThreads, kFwdBlocksPerSM)
pooled_forward_kernel(const LookupArgs a) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int RPW = 32 / LPR;                       // rows gathered in parallel per warp
    constexpr int U = (CPL > 2) ? 2 : (kFwdUnroll / CPL);
    const int lane = lane_id();
    const int sub = lane % LPR;
    const int grp = lane / LPR;
    const int64_t nwarps = (int64_t)gridDim.x * (kPoolThreads / 32);
    int64_t g = (int64_t)blockIdx.x * (kPoolThreads / 32) + (threadIdx.x >> 5);
    if (g >= a.num_bags) return;
    const uint4* __restrict__ wbase = reinterpret_cast<const uint4*>(a.weights);
    uint32_t flags = 0;

    BagHead cur = load_bag_head(a, g, lane);
    while (g < a.num_bags) {
        const int64_t gn = g + nwarps;
        BagHead nxt;
        nxt.s = nxt.e = 0; nxt.row = 0xffffffffu; nxt.w = 1.f; nxt.flags = 0;
        if (gn < a.num_bags) nxt = load_bag_head(a, gn, lane);      // in flight during this bag's gather

        float acc[CPL][E];
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) acc[k][i] = 0.f;
        flags |= cur.flags;
        const int f = (int)((uint64_t)g / (uint64_t)a.B);
        const int64_t b = g - (int64_t)f * a.B;

        uint32_t my_row = cur.row;
        float my_w = cur.w;
        for (int64_t base = cur.s; base < cur.e; base += 32) {
            const int n = (int)min((int64_t)32, cur.e - base);
            if (base != cur.s) {                                     // bags longer than 32 ids
                my_row = 0xffffffffu;
                my_w = 1.f;
                if (lane < n) {
                    const int64_t idx = __ldg(a.indices + base + lane);
                    if ((uint64_t)idx < (uint64_t)__ldg(a.feat_rows + f))
                        my_row = (uint32_t)(__ldg(a.feat_row_base + f) + idx);
                    else flags |= TDFO_FLAG_INDEX_OOB;
                    if (WEIGHTED) my_w = __ldg(a.psw + base + lane);
                    if (a.sort_keys) {
                        a.sort_keys[base + lane] = (my_row == 0xffffffffu) ? a.oob_key : my_row;
                        a.sort_vals[base + lane] = (uint32_t)g;
                    }
                }
            }
            // full batches of U rows per sub-group, then one shorter batch for the tail
            int j0 = 0;
            for (; j0 + RPW * U <= n; j0 += RPW * U)
                gather_batch<T, LPR, CPL, U, WEIGHTED>(wbase, a.nchunks, j0, n, my_row, my_w, sub, grp, acc);
            if (U >= 8 && j0 + RPW * 4 <= n) {
                gather_batch<T, LPR, CPL, (U >= 8 ? 4 : 1), WEIGHTED>(wbase, a.nchunks, j0, n, my_row, my_w, sub, grp, acc);
                j0 += RPW * 4;
            }
            if (U >= 4 && j0 + RPW * 2 <= n) {
                gather_batch<T, LPR, CPL, (U >= 4 ? 2 : 1), WEIGHTED>(wbase, a.nchunks, j0, n, my_row, my_w, sub, grp, acc);
                j0 += RPW * 2;
            }
            for (; j0 < n; j0 += RPW)
                gather_batch<T, LPR, CPL, 1, WEIGHTED>(wbase, a.nchunks, j0, n, my_row, my_w, sub, grp, acc);
        }

        // combine the sub-groups (fixed tree => deterministic)
#pragma unroll
        for (int o = LPR; o < 32; o <<= 1)
#pragma unroll
            for (int k = 0; k < CPL; ++k)
#pragma unroll
                for (int i = 0; i < E; ++i) acc[k][i] += __shfl_xor_sync(kFullMask, acc[k][i], o);

        if (grp == 0) {
            const float scale = (a.mean && cur.e > cur.s) ? 1.f / (float)(cur.e - cur.s) : 1.f;
            // Sharded table-wise output dist fused into the lookup: sample b of the global batch
            // belongs to source rank b / peer_batch, and its pooled row is stored straight into
            // that rank's output buffer over NVLink instead of going through an all-to-all.
            float* obase = a.out;
            int64_t brow = b;
            if (a.peer_out) {
                const int64_t src = b / a.peer_batch;
                brow = b - src * a.peer_batch;
                obase = reinterpret_cast<float*>(__ldg(a.peer_out + src));
            }
            float* orow = obase + brow * a.out_stride + __ldg(a.feat_out_col + f);
#pragma unroll
            for (int k = 0; k < CPL; ++k) {
                const int c = sub + k * LPR;
                if (c < a.nchunks) {
#pragma unroll
                    for (int i = 0; i < E; i += 4) {
                        const uint4 o4 = make_uint4(__float_as_uint(acc[k][i] * scale),
                                                    __float_as_uint(acc[k][i + 1] * scale),
                                                    __float_as_uint(acc[k][i + 2] * scale),
                                                    __float_as_uint(acc[k][i + 3] * scale));
                        st_stream_128(orow + (size_t)c * E + i, o4);
                    }
                }
            }
        }
        g = gn;
        cur = nxt;
    }
    if (flags && a.err_flag) atomicOr(a.err_flag, (int)flags);
}

// Short bags (about one id per bag: the Criteo-style one-hot features of BASELINE config 3).  The bag-at-a-time
// kernel above has a single row in flight per warp there.  Here a warp takes a block of 32 consecutive bags,
// reads their ids as one coalesced stream and gathers the rows kStreamUnroll at a time ACROSS bag boundaries;
// a bag's pooled row is written when the stream moves on to the next bag.  Rows are added in bag order by
// the whole warp (no sub-group tree), so the result equals the sequential sum.
constexpr int kStreamUnroll = 8;
constexpr int kStreamBlocksPerSM = 2;    // 128 registers: the eight rows in flight must not spill

template <typename T, int CPL, bool WEIGHTED>
__global__ void __launch_bounds__(kFwdThreads, kStreamBlocksPerSM)
pooled_forward_stream_kernel(const LookupArgs a) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int U = (CPL > 1) ? (kStreamUnroll / CPL > 1 ? kStreamUnroll / CPL : 1) : kStreamUnroll;
    const int lane = lane_id();
    const int64_t nwarps = (int64_t)gridDim.x * (kFwdThreads / 32);
    const uint4* __restrict__ wbase = reinterpret_cast<const uint4*>(a.weights);
    uint32_t flags = 0;

    auto store_row = [&](int64_t g, float scale, const float (&acc)[CPL][E]) {
        const int f = (int)((uint64_t)g / (uint64_t)a.B);
        const int64_t b = g - (int64_t)f * a.B;
        float* obase = a.out;
        int64_t brow = b;
        if (a.peer_out) {                               // sharded output dist fused in (see pooled_forward_kernel)
            const int64_t src = b / a.peer_batch;
            brow = b - src * a.peer_batch;
            obase = reinterpret_cast<float*>(__ldg(a.peer_out + src));
        }
        float* orow = obase + brow * a.out_stride + __ldg(a.feat_out_col + f);
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = lane + k * 32;
            if (c < a.nchunks) {
#pragma unroll
                for (int i = 0; i < E; i += 4)
                    st_stream_128(orow + (size_t)c * E + i,
                                  make_uint4(__float_as_uint(acc[k][i] * scale), __float_as_uint(acc[k][i + 1] * scale),
                                             __float_as_uint(acc[k][i + 2] * scale), __float_as_uint(acc[k][i + 3] * scale)));
            }
        }
    };

    for (int64_t g0 = ((int64_t)blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5)) * 32; g0 < a.num_bags;
         g0 += nwarps * 32) {
        const int nb = (int)min((int64_t)32, a.num_bags - g0);           // bags in this block
        int64_t my_s = __ldg(a.offsets + g0 + min(lane, nb - 1));
        int64_t my_e = __ldg(a.offsets + g0 + min(lane, nb - 1) + 1);
        if (my_e > a.L) {                                                 // offsets past the ids given: cut there
            flags |= TDFO_FLAG_OFFSETS_RANGE;
            my_e = a.L;
            my_s = min(my_s, a.L);
        }
        if (lane >= nb) my_s = my_e;                                      // lanes past the block: an empty bag at the end
        const int64_t s0 = __shfl_sync(kFullMask, my_s, 0), e0 = __shfl_sync(kFullMask, my_e, nb - 1);
        float acc[CPL][E];
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) acc[k][i] = 0.f;
        // empty bags never show up in the id stream: their rows are zero
        unsigned empty = __ballot_sync(kFullMask, lane < nb && my_e == my_s);
        while (empty) {
            const int eb = __ffs(empty) - 1;
            empty &= empty - 1;
            store_row(g0 + eb, 1.f, acc);
        }
        int cur_bag = -1;
        for (int64_t base = s0; base < e0; base += 32) {
            const int n = (int)min((int64_t)32, e0 - base);
            const int64_t p = base + lane;
            uint32_t my_row = 0xffffffffu;
            int my_bag = 0;
            float my_w = 1.f;
            {   // bag of position p: number of bags of the block that end at or before p (offsets are non-decreasing)
                int lo = 0;
#pragma unroll
                for (int step = 16; step > 0; step >>= 1) {
                    const int64_t e = __shfl_sync(kFullMask, my_e, min(lo + step - 1, 31));
                    if (lo + step <= 32 && e <= p) lo += step;
                }
                my_bag = min(lo, nb - 1);
            }
            if (lane < n) {
                const int64_t g = g0 + my_bag;
                const int f = (int)((uint64_t)g / (uint64_t)a.B);
                const int64_t idx = __ldg(a.indices + p);
                if ((uint64_t)idx < (uint64_t)__ldg(a.feat_rows + f)) my_row = (uint32_t)(__ldg(a.feat_row_base + f) + idx);
                else flags |= TDFO_FLAG_INDEX_OOB;
                if (WEIGHTED) my_w = __ldg(a.psw + p);
                if (a.sort_keys) {
                    a.sort_keys[p] = (my_row == 0xffffffffu) ? a.oob_key : my_row;
                    a.sort_vals[p] = (uint32_t)g;
                }
            }
            for (int j0 = 0; j0 < n; j0 += U) {
                uint4 v[U][CPL];
                int bag[U];
                float w[U];
                bool ok[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {                             // unconditional loads, masked when consumed
                    const int j = min(j0 + u, 31);
                    const uint32_t row = __shfl_sync(kFullMask, my_row, j);
                    bag[u] = __shfl_sync(kFullMask, my_bag, j);
                    w[u] = WEIGHTED ? __shfl_sync(kFullMask, my_w, j) : 1.f;
                    ok[u] = (j0 + u < n) && (row != 0xffffffffu);
                    const size_t rbase = (size_t)(ok[u] ? row : 0u) * a.nchunks;
#pragma unroll
                    for (int k = 0; k < CPL; ++k) v[u][k] = ldg_stream_128(wbase + rbase + min(lane + k * 32, a.nchunks - 1));
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (j0 + u < n) {                                     // uniform over the warp
                        if (bag[u] != cur_bag) {
                            if (cur_bag >= 0) {
                                const int64_t len = __shfl_sync(kFullMask, my_e - my_s, cur_bag);
                                store_row(g0 + cur_bag, a.mean ? 1.f / (float)len : 1.f, acc);
                            }
#pragma unroll
                            for (int k = 0; k < CPL; ++k)
#pragma unroll
                                for (int i = 0; i < E; ++i) acc[k][i] = 0.f;
                            cur_bag = bag[u];
                        }
#pragma unroll
                        for (int k = 0; k < CPL; ++k) {
                            float x[E];
                            Chunk<T>::unpack(v[u][k], x);
#pragma unroll
                            for (int i = 0; i < E; ++i) acc[k][i] = fmaf(w[u], ok[u] ? x[i] : 0.f, acc[k][i]);
                        }
                    }
                }
            }
        }
        if (cur_bag >= 0) {
            const int64_t len = __shfl_sync(kFullMask, my_e - my_s, cur_bag);
            store_row(g0 + cur_bag, a.mean ? 1.f / (float)len : 1.f, acc);
        }
    }
    if (flags && a.err_flag) atomicOr(a.err_flag, (int)flags);
}

// Sequence lookup (EmbeddingCollection): out[p,:] = W[row_base(f(p)) + indices[p], :].
// One LPR-lane group per id; feature of p by binary search over the F+1 feature
// boundaries offsets[f*B] staged in shared memory.
constexpr int kSeqMaxF = 2048;

template <typename T, int LPR, int CPL>
__global__ void __launch_bounds__(kFwdThreads)
sequence_forward_kernel(const LookupArgs a) {
    constexpr int E = Chunk<T>::kElems;
    __shared__ int64_t fbound[kSeqMaxF + 1];
    for (int f = threadIdx.x; f <= a.F; f += blockDim.x) fbound[f] = a.offsets[(int64_t)f * a.B];
    __syncthreads();
    const int sub = threadIdx.x % LPR;
    const int64_t groups_per_block = kFwdThreads / LPR;
    const uint4* __restrict__ wbase = reinterpret_cast<const uint4*>(a.weights);
    bool oob = false;
    // ids of this call: positions [first, last) of `indices` (the whole array unless the offsets are a slice of a
    // larger batch, as in the sharded landing zone); never past the L ids given
    const int64_t first = fbound[0], last = min(a.L, fbound[a.F]);
    for (int64_t p = first + (int64_t)blockIdx.x * groups_per_block + threadIdx.x / LPR; p < last;
         p += (int64_t)gridDim.x * groups_per_block) {
        int lo = 0, hi = a.F;                 // largest f with fbound[f] <= p
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (fbound[mid] <= p) lo = mid; else hi = mid;
        }
        const int64_t idx = __ldg(a.indices + p);
        const bool valid = (uint64_t)idx < (uint64_t)__ldg(a.feat_rows + lo);
        if (!valid) oob = true;
        const size_t row = (size_t)(__ldg(a.feat_row_base + lo) + idx);
        if (a.sort_keys && sub == 0) {
            a.sort_keys[p] = valid ? (uint32_t)row : a.oob_key;
            a.sort_vals[p] = (uint32_t)(p - first);         // payload: the id's gradient row, counted from this call's first id
        }
        float* orow = a.out + (size_t)(p - first) * a.nchunks * E;
        if (a.peer_out)     // sharded output dist fused in: the row goes to its id's place at the rank that sent the id
            orow = reinterpret_cast<float*>(__ldg(a.peer_out + lo % a.peer_world)) +
                   (size_t)__ldg(a.peer_pos + p) * a.nchunks * E;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = sub + k * LPR;
            if (c < a.nchunks) {
                float x[E];
                Chunk<T>::unpack(ldg_stream_128(wbase + (valid ? row : 0) * a.nchunks + c), x);
                if (!valid) {
#pragma unroll
                    for (int i = 0; i < E; ++i) x[i] = 0.f;
                }
#pragma unroll
                for (int i = 0; i < E; i += 4)
                    st_stream_128(orow + (size_t)c * E + i,
                                  make_uint4(__float_as_uint(x[i]), __float_as_uint(x[i + 1]),
                                             __float_as_uint(x[i + 2]), __float_as_uint(x[i + 3])));
            }
        }
    }
    if (oob && a.err_flag) atomicOr(a.err_flag, TDFO_FLAG_INDEX_OOB);
}

template <typename T, int LPR, int CPL>
static int launch_forward(const LookupArgs& a, int pooling_mode, cudaStream_t st) {
    if (pooling_mode == TDFO_POOL_NONE) {
        if (a.L == 0) return TDFO_OK;
        const int64_t gpb = kFwdThreads / LPR;
        int64_t blocks = (a.L + gpb - 1) / gpb;
        const int64_t cap = (int64_t)kNumSMs * 64;       // grid-stride beyond that
        if (blocks > cap) blocks = cap;
        sequence_forward_kernel<T, LPR, CPL><<<(unsigned)blocks, kFwdThreads, 0, st>>>(a);
    } else {
        if (a.num_bags == 0) return TDFO_OK;
        if (a.L_shape < 4 * a.num_bags) {
            // short bags (fewer than 4 ids on average): stream the ids of 32-bag blocks, rows in flight across bags
            constexpr int SCPL = (LPR == 32) ? CPL : 1;
            const int64_t warps = (a.num_bags + 31) / 32;
            int64_t blocks = (warps + kFwdThreads / 32 - 1) / (kFwdThreads / 32);
            if (blocks > (int64_t)kNumSMs * kStreamBlocksPerSM) blocks = (int64_t)kNumSMs * kStreamBlocksPerSM;
            if (a.psw) pooled_forward_stream_kernel<T, SCPL, true><<<(unsigned)blocks, kFwdThreads, 0, st>>>(a);
            else       pooled_forward_stream_kernel<T, SCPL, false><<<(unsigned)blocks, kFwdThreads, 0, st>>>(a);
            count_launch();
            TDFO_LAUNCH_CHECK("embedding_forward");
            return TDFO_OK;
        }
        int64_t blocks = (a.num_bags + kPoolThreads / 32 - 1) / (kPoolThreads / 32);
        if (blocks > (int64_t)kNumSMs * kFwdBlocksPerSM) blocks = (int64_t)kNumSMs * kFwdBlocksPerSM;
        if (a.psw) pooled_forward_kernel<T, LPR, CPL, true><<<(unsigned)blocks, kPoolThreads, 0, st>>>(a);
        else       pooled_forward_kernel<T, LPR, CPL, false><<<(unsigned)blocks, kPoolThreads, 0, st>>>(a);
    }
    count_launch();
    TDFO_LAUNCH_CHECK("embedding_forward");
    return TDFO_OK;
}

template <typename T>
static int dispatch_forward(const LookupArgs& a, int pooling_mode, cudaStream_t st) {
    const int n = a.nchunks;
    if (n <= 1)   return launch_forward<T, 1, 1>(a, pooling_mode, st);
    if (n <= 2)   return launch_forward<T, 2, 1>(a, pooling_mode, st);
    if (n <= 4)   return launch_forward<T, 4, 1>(a, pooling_mode, st);
    if (n <= 8)   return launch_forward<T, 8, 1>(a, pooling_mode, st);
    if (n <= 16)  return launch_forward<T, 16, 1>(a, pooling_mode, st);
    if (n <= 32)  return launch_forward<T, 32, 1>(a, pooling_mode, st);
    if (n <= 64)  return launch_forward<T, 32, 2>(a, pooling_mode, st);
    if (n <= 128) return launch_forward<T, 32, 4>(a, pooling_mode, st);
    if (n <= 256) return launch_forward<T, 32, 8>(a, pooling_mode, st);
    set_error("embedding_forward: row of %d bytes exceeds the 4096-byte maximum", n * 16);
    return TDFO_ERR_UNSUPPORTED;
}

}  // namespace tdfo

using namespace tdfo;

static int embedding_forward_impl(const void* weights, int32_t weights_dtype, int64_t total_rows,
                                  int32_t D, const int64_t* indices, const int64_t* offsets,
                                  const float* per_sample_weights,
                                  const int64_t* feat_row_base, const int64_t* feat_rows,
                                  const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                  int32_t pooling_mode, float* out, int64_t out_stride,
                                  const uint64_t* peer_out, int64_t peer_batch,
                                  uint32_t* sort_keys, uint32_t* sort_vals,
                                  int32_t* err_flag, void* stream,
                                  const int64_t* peer_pos = nullptr, int32_t peer_world = 0, int64_t L_hint = 0) {
    TDFO_CHECK_ARG(weights_dtype == TDFO_F32 || weights_dtype == TDFO_BF16,
                   "embedding_forward: unknown weights_dtype %d", weights_dtype);
    TDFO_CHECK_ARG(pooling_mode >= TDFO_POOL_SUM && pooling_mode <= TDFO_POOL_NONE,
                   "embedding_forward: unknown pooling_mode %d", pooling_mode);
    TDFO_CHECK_ARG(F >= 0 && B >= 0 && L >= 0 && D > 0, "embedding_forward: negative size");
    const size_t esz = weights_dtype == TDFO_F32 ? 4 : 2;
    TDFO_CHECK_ARG((D * esz) % 16 == 0,
                   "embedding_forward: row size D*elem = %zu bytes is not a multiple of 16", D * esz);
    TDFO_CHECK_ARG(total_rows >= 0 && total_rows < 0xffffffffll,
                   "embedding_forward: total_rows %lld must be < 2^32-1", (long long)total_rows);
    if (F == 0 || B == 0 || L == 0 && pooling_mode == TDFO_POOL_NONE) return TDFO_OK;
    TDFO_CHECK_ARG(total_rows >= 1, "embedding_forward: the storage group has no rows");
    TDFO_CHECK_ARG(weights && (indices || L == 0) && offsets && feat_row_base && feat_rows && (out || peer_out),
                   "embedding_forward: null pointer");
    TDFO_CHECK_ARG(aligned(weights, 16) && aligned(out, 16), "embedding_forward: weights/out not 16-byte aligned");
    if (peer_out && pooling_mode == TDFO_POOL_NONE) {
        TDFO_CHECK_ARG(peer_pos != nullptr && peer_world >= 1 && F % peer_world == 0,
                       "embedding_forward_seq_peer: needs positions and a segment count that is a multiple of the world size");
    } else if (peer_out) {
        TDFO_CHECK_ARG(peer_batch > 0 && B % peer_batch == 0,
                       "embedding_forward_peer: batch %lld is not a multiple of the per-rank batch %lld",
                       (long long)B, (long long)peer_batch);
    }
    if (pooling_mode != TDFO_POOL_NONE) {
        TDFO_CHECK_ARG(feat_out_col != nullptr, "embedding_forward: feat_out_col is null");
        TDFO_CHECK_ARG(out_stride % 4 == 0, "embedding_forward: out_stride %lld not a multiple of 4",
                       (long long)out_stride);
    } else {
        TDFO_CHECK_ARG(F <= kSeqMaxF, "embedding_forward: sequence mode supports at most %d features", kSeqMaxF);
    }
    LookupArgs a;
    a.weights = weights; a.indices = indices; a.offsets = offsets; a.psw = per_sample_weights;
    a.feat_row_base = feat_row_base; a.feat_rows = feat_rows; a.feat_out_col = feat_out_col;
    TDFO_CHECK_ARG((sort_keys == nullptr) == (sort_vals == nullptr),
                   "embedding_forward: sort_keys and sort_vals must both be given or both be null");
    TDFO_CHECK_ARG(sort_keys == nullptr || ((int64_t)F * B < (1ll << 32) && L < (1ll << 32)),
                   "embedding_forward: sort keys need F*B and L below 2^32");
    a.peer_pos = peer_pos; a.peer_world = peer_world;
    a.L_shape = (L_hint > 0 && L_hint < L) ? L_hint : L;
    a.out = out; a.peer_out = peer_out; a.peer_batch = peer_batch; a.sort_keys = sort_keys; a.sort_vals = sort_vals; a.oob_key = (uint32_t)total_rows; a.err_flag = err_flag; a.B = B; a.num_bags = (int64_t)F * B; a.L = L;
    a.out_stride = out_stride; a.F = F; a.nchunks = (int32_t)(D * esz / 16);
    a.mean = pooling_mode == TDFO_POOL_MEAN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return weights_dtype == TDFO_F32 ? dispatch_forward<float>(a, pooling_mode, st)
                                     : dispatch_forward<__nv_bfloat16>(a, pooling_mode, st);
}

extern "C" int tdfo_embedding_forward(const void* weights, int32_t weights_dtype, int64_t total_rows,
                                      int32_t D, const int64_t* indices, const int64_t* offsets,
                                      const float* per_sample_weights,
                                      const int64_t* feat_row_base, const int64_t* feat_rows,
                                      const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                      int32_t pooling_mode, float* out, int64_t out_stride,
                                      uint32_t* sort_keys, uint32_t* sort_vals,
                                      int32_t* err_flag, void* stream) {
    return embedding_forward_impl(weights, weights_dtype, total_rows, D, indices, offsets, per_sample_weights,
                                  feat_row_base, feat_rows, feat_out_col, F, B, L, pooling_mode, out, out_stride,
                                  nullptr, 0, sort_keys, sort_vals, err_flag, stream);
}

extern "C" int tdfo_embedding_forward_peer(const void* weights, int32_t weights_dtype, int64_t total_rows,
                                           int32_t D, const int64_t* indices, const int64_t* offsets,
                                           const float* per_sample_weights,
                                           const int64_t* feat_row_base, const int64_t* feat_rows,
                                           const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                           int32_t pooling_mode, const uint64_t* peer_out, int64_t peer_batch,
                                           int64_t out_stride, uint32_t* sort_keys, uint32_t* sort_vals,
                                           int32_t* err_flag, int64_t L_hint, void* stream) {
    if (!peer_out) {
        set_error("embedding_forward_peer: peer_out is null");
        return TDFO_ERR_INVALID_ARG;
    }
    return embedding_forward_impl(weights, weights_dtype, total_rows, D, indices, offsets, per_sample_weights,
                                  feat_row_base, feat_rows, feat_out_col, F, B, L, pooling_mode, nullptr, out_stride,
                                  peer_out, peer_batch, sort_keys, sort_vals, err_flag, stream, nullptr, 0, L_hint);
}

extern "C" int tdfo_embedding_forward_seq_peer(const void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                                               const int64_t* indices, const int64_t* offsets,
                                               const int64_t* seg_row_base, const int64_t* seg_rows, int32_t S,
                                               int64_t B, int64_t L, const int64_t* positions,
                                               const uint64_t* peer_out, int32_t world,
                                               uint32_t* sort_keys, uint32_t* sort_vals, int32_t* err_flag,
                                               void* stream) {
    if (!peer_out || !positions) {
        set_error("embedding_forward_seq_peer: peer_out / positions is null");
        return TDFO_ERR_INVALID_ARG;
    }
    return embedding_forward_impl(weights, weights_dtype, total_rows, D, indices, offsets, nullptr, seg_row_base, seg_rows,
                                  nullptr, S, B, L, TDFO_POOL_NONE, nullptr, D, peer_out, 0, sort_keys, sort_vals,
                                  err_flag, stream, positions, world);
}
