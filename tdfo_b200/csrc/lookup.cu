// Embedding lookup (forward): offsets-driven gather of table rows with 128-bit
// loads and warp-level segmented pooling.  HBM-bound: the algorithmic traffic is
// L * (D*w + 8) + bags * (D*4 + 8) bytes, all of it streamed exactly once.
//
// Mapping (D = 128 fp32, the BASELINE shape): one warp per bag, one 16-byte
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
    int32_t* err_flag;
    int64_t B;
    int64_t num_bags;     // F * B
    int64_t L;
    int64_t out_stride;
    int32_t F;
    int32_t nchunks;      // 16-byte chunks per row
    int32_t mean;
};

constexpr int kFwdThreads = 256;
constexpr int kFwdUnroll = 8;      // independent row requests in flight per sub-group

template <typename T, int LPR, int CPL>
__global__ void __launch_bounds__(kFwdThreads)
pooled_forward_kernel(const LookupArgs a) {
    constexpr int E = Chunk<T>::kElems;
    constexpr int RPW = 32 / LPR;                       // rows gathered in parallel per warp
    constexpr int U = (CPL > 2) ? 2 : (kFwdUnroll / CPL);
    const int64_t g = (int64_t)blockIdx.x * (kFwdThreads / 32) + (threadIdx.x >> 5);
    if (g >= a.num_bags) return;
    const int lane = lane_id();
    const int sub = lane % LPR;
    const int grp = lane / LPR;
    const int f = (int)(g / a.B);
    const int64_t b = g - (int64_t)f * a.B;
    const int64_t s = __ldg(a.offsets + g);
    const int64_t e = __ldg(a.offsets + g + 1);
    const int64_t row_base = __ldg(a.feat_row_base + f);
    const uint64_t rows = (uint64_t)__ldg(a.feat_rows + f);
    const uint4* __restrict__ wbase = reinterpret_cast<const uint4*>(a.weights);
    const bool weighted = a.psw != nullptr;

    float acc[CPL][E];
#pragma unroll
    for (int k = 0; k < CPL; ++k)
#pragma unroll
        for (int i = 0; i < E; ++i) acc[k][i] = 0.f;

    bool oob = false;
    for (int64_t base = s; base < e; base += 32) {
        const int n = (int)min((int64_t)32, e - base);
        uint32_t my_row = 0xffffffffu;
        float my_w = 1.f;
        if (lane < n) {
            const int64_t idx = __ldg(a.indices + base + lane);
            if ((uint64_t)idx < rows) my_row = (uint32_t)(row_base + idx);
            else oob = true;
            if (weighted) my_w = __ldg(a.psw + base + lane);
        }
        for (int j0 = 0; j0 < n; j0 += RPW * U) {
            uint4 v[U][CPL];
            float w[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int r = j0 + u * RPW + grp;
                const uint32_t row = __shfl_sync(kFullMask, my_row, r & 31);
                w[u] = __shfl_sync(kFullMask, my_w, r & 31);
                const bool valid = (r < n) && (row != 0xffffffffu);
                if (!valid) w[u] = 0.f;
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    const int c = sub + k * LPR;
                    if (valid && c < a.nchunks)
                        v[u][k] = ldg_stream_128(wbase + (size_t)row * a.nchunks + c);
                    else
                        v[u][k] = make_uint4(0, 0, 0, 0);
                }
            }
#pragma unroll
            for (int u = 0; u < U; ++u)
#pragma unroll
                for (int k = 0; k < CPL; ++k) {
                    float x[E];
                    Chunk<T>::unpack(v[u][k], x);
#pragma unroll
                    for (int i = 0; i < E; ++i) acc[k][i] = fmaf(w[u], x[i], acc[k][i]);
                }
        }
    }
    if (oob && a.err_flag) atomicOr(a.err_flag, TDFO_FLAG_INDEX_OOB);

    // combine the sub-groups (fixed tree => deterministic)
#pragma unroll
    for (int o = LPR; o < 32; o <<= 1)
#pragma unroll
        for (int k = 0; k < CPL; ++k)
#pragma unroll
            for (int i = 0; i < E; ++i) acc[k][i] += __shfl_xor_sync(kFullMask, acc[k][i], o);

    if (grp == 0) {
        const float scale = (a.mean && e > s) ? 1.f / (float)(e - s) : 1.f;
        float* orow = a.out + b * a.out_stride + __ldg(a.feat_out_col + f);
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
    for (int64_t p = (int64_t)blockIdx.x * groups_per_block + threadIdx.x / LPR; p < a.L;
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
        float* orow = a.out + (size_t)p * a.nchunks * E;
#pragma unroll
        for (int k = 0; k < CPL; ++k) {
            const int c = sub + k * LPR;
            if (c < a.nchunks) {
                float x[E];
                if (valid) {
                    Chunk<T>::unpack(ldg_stream_128(wbase + row * a.nchunks + c), x);
                } else {
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
        const int64_t blocks = (a.num_bags + kFwdThreads / 32 - 1) / (kFwdThreads / 32);
        pooled_forward_kernel<T, LPR, CPL><<<(unsigned)blocks, kFwdThreads, 0, st>>>(a);
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

extern "C" int tdfo_embedding_forward(const void* weights, int32_t weights_dtype, int64_t total_rows,
                                      int32_t D, const int64_t* indices, const int64_t* offsets,
                                      const float* per_sample_weights,
                                      const int64_t* feat_row_base, const int64_t* feat_rows,
                                      const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                      int32_t pooling_mode, float* out, int64_t out_stride,
                                      int32_t* err_flag, void* stream) {
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
    if (F == 0 || B == 0) return TDFO_OK;
    TDFO_CHECK_ARG(weights && (indices || L == 0) && offsets && feat_row_base && feat_rows && out,
                   "embedding_forward: null pointer");
    TDFO_CHECK_ARG(aligned(weights, 16) && aligned(out, 16), "embedding_forward: weights/out not 16-byte aligned");
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
    a.out = out; a.err_flag = err_flag; a.B = B; a.num_bags = (int64_t)F * B; a.L = L;
    a.out_stride = out_stride; a.F = F; a.nchunks = (int32_t)(D * esz / 16);
    a.mean = pooling_mode == TDFO_POOL_MEAN;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    return weights_dtype == TDFO_F32 ? dispatch_forward<float>(a, pooling_mode, st)
                                     : dispatch_forward<__nv_bfloat16>(a, pooling_mode, st);
}
