// KeyedJaggedTensor routing kernels (row-wise bucketise, 2-D permute) and the
// jagged <-> dense copies the reference model calls around the lookup.
// All of it is integer / copy work bounded by HBM; results are bit-exact.
#include "common.cuh"

namespace tdfo {

int complete_cumsum_impl(const int32_t* lengths, int64_t n, int64_t* offsets, void* scratch, cudaStream_t st);

constexpr int kJagThreads = 256;
constexpr int kMaxWorld = 32;

// ---------------------------------------------------------------- bucketise
// Pass 1: per (bucket, f, b) counts.  One warp per bag.
__global__ void __launch_bounds__(kJagThreads)
bucketize_count_kernel(const int64_t* __restrict__ offsets, const int64_t* __restrict__ indices,
                       const int64_t* __restrict__ block_sizes, int64_t B, int64_t num_bags, int world,
                       int32_t* __restrict__ new_lengths, int32_t* err_flag) {
    const int64_t g = (int64_t)blockIdx.x * (kJagThreads / 32) + (threadIdx.x >> 5);
    if (g >= num_bags) return;
    const int lane = lane_id();
    const int f = (int)(g / B);
    const int64_t s = __ldg(offsets + g), e = __ldg(offsets + g + 1);
    const int64_t bs = __ldg(block_sizes + f);
    int my_count = 0;               // lane w counts bucket w
    bool oob = false;
    for (int64_t base = s; base < e; base += 32) {
        int bucket = -1;
        if (base + lane < e) {
            const int64_t idx = __ldg(indices + base + lane);
            int64_t q = idx / bs;
            if (idx < 0 || q >= world) { oob = true; q = world - 1; }
            bucket = (int)q;
        }
        for (int w = 0; w < world; ++w) {
            const unsigned m = __ballot_sync(kFullMask, bucket == w);
            if (lane == w) my_count += __popc(m);
        }
    }
    if (lane < world) new_lengths[(int64_t)lane * num_bags + g] = my_count;
    if (oob && err_flag) atomicOr(err_flag, TDFO_FLAG_INDEX_OOB);
}

// Pass 2: stable scatter into [bucket][f][b] order.
__global__ void __launch_bounds__(kJagThreads)
bucketize_scatter_kernel(const int64_t* __restrict__ offsets, const int64_t* __restrict__ indices,
                         const float* __restrict__ weights, const int64_t* __restrict__ block_sizes,
                         int64_t B, int64_t num_bags, int world,
                         const int64_t* __restrict__ new_offsets, int64_t* __restrict__ new_indices,
                         float* __restrict__ new_weights, int64_t* __restrict__ unbucketize_permute) {
    const int64_t g = (int64_t)blockIdx.x * (kJagThreads / 32) + (threadIdx.x >> 5);
    if (g >= num_bags) return;
    const int lane = lane_id();
    const int f = (int)(g / B);
    const int64_t s = __ldg(offsets + g), e = __ldg(offsets + g + 1);
    const int64_t bs = __ldg(block_sizes + f);
    int64_t my_cursor = (lane < world) ? __ldg(new_offsets + (int64_t)lane * num_bags + g) : 0;
    const unsigned lt = (1u << lane) - 1u;
    for (int64_t base = s; base < e; base += 32) {
        int bucket = -1;
        int64_t idx = 0;
        if (base + lane < e) {
            idx = __ldg(indices + base + lane);
            int64_t q = idx / bs;
            if (idx < 0 || q >= world) q = world - 1;
            bucket = (int)q;
        }
        int64_t pos = 0;
        for (int w = 0; w < world; ++w) {
            const unsigned m = __ballot_sync(kFullMask, bucket == w);
            const int64_t cur = __shfl_sync(kFullMask, my_cursor, w);
            if (bucket == w) pos = cur + __popc(m & lt);
            if (lane == w) my_cursor += __popc(m);
        }
        if (bucket >= 0) {
            new_indices[pos] = idx - (int64_t)bucket * bs;
            if (weights) new_weights[pos] = __ldg(weights + base + lane);
            if (unbucketize_permute) unbucketize_permute[base + lane] = pos;
        }
    }
}

// ---------------------------------------------------------------- permute 2-D
__global__ void permute_lengths_kernel(const int32_t* __restrict__ permute, int64_t B, int64_t n_out,
                                       const int32_t* __restrict__ in_lengths, int32_t* __restrict__ out_lengths) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_out) return;
    const int64_t seg = i / B;
    out_lengths[i] = in_lengths[(int64_t)__ldg(permute + seg) * B + (i - seg * B)];
}

__global__ void __launch_bounds__(kJagThreads)
permute_data_kernel(const int32_t* __restrict__ permute, int64_t B, int64_t n_out,
                    const int64_t* __restrict__ in_offsets, const int64_t* __restrict__ out_offsets,
                    const int64_t* __restrict__ in_indices, const float* __restrict__ in_weights,
                    int64_t* __restrict__ out_indices, float* __restrict__ out_weights) {
    const int64_t g = (int64_t)blockIdx.x * (kJagThreads / 32) + (threadIdx.x >> 5);
    if (g >= n_out) return;
    const int lane = lane_id();
    const int64_t seg = g / B;
    const int64_t src = (int64_t)__ldg(permute + seg) * B + (g - seg * B);
    const int64_t so = __ldg(in_offsets + src);
    const int64_t len = __ldg(in_offsets + src + 1) - so;
    const int64_t dofs = __ldg(out_offsets + g);
    for (int64_t j = lane; j < len; j += 32) {
        out_indices[dofs + j] = __ldg(in_indices + so + j);
        if (in_weights) out_weights[dofs + j] = __ldg(in_weights + so + j);
    }
}

// ---------------------------------------------------------------- jagged <-> dense
// VEC floats per thread-chunk (4 when D % 4 == 0, else 1).
template <int VEC>
__global__ void jagged_2d_to_dense_kernel(const float* __restrict__ values, const int64_t* __restrict__ offsets,
                                          int64_t B, int32_t D, int64_t max_len, float* __restrict__ dense) {
    const int cpr = D / VEC;                                   // chunks per row
    const int64_t total = B * max_len * cpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t r = i / cpr;
        const int c = (int)(i - r * cpr);
        const int64_t b = r / max_len;
        const int64_t t = r - b * max_len;
        const int64_t s = __ldg(offsets + b);
        const int64_t len = __ldg(offsets + b + 1) - s;
        if (VEC == 4) {
            uint4 v = make_uint4(0, 0, 0, 0);
            if (t < len) v = ldg_128(values + (size_t)(s + t) * D + (size_t)c * 4);
            st_128(dense + (size_t)r * D + (size_t)c * 4, v);
        } else {
            dense[(size_t)r * D + c] = (t < len) ? __ldg(values + (size_t)(s + t) * D + c) : 0.f;
        }
    }
}

template <int VEC>
__global__ void dense_to_jagged_kernel(const float* __restrict__ dense, const int64_t* __restrict__ offsets,
                                       int64_t B, int32_t D, int64_t max_len, int64_t L,
                                       float* __restrict__ values) {
    const int cpr = D / VEC;
    const int64_t total = L * cpr;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t p = i / cpr;
        const int c = (int)(i - p * cpr);
        int64_t lo = 0, hi = B;                                 // largest b with offsets[b] <= p
        while (hi - lo > 1) {
            const int64_t mid = (lo + hi) >> 1;
            if (__ldg(offsets + mid) <= p) lo = mid; else hi = mid;
        }
        const int64_t t = p - __ldg(offsets + lo);
        if (VEC == 4) {
            uint4 v = make_uint4(0, 0, 0, 0);
            if (t < max_len) v = ldg_128(dense + ((size_t)lo * max_len + t) * D + (size_t)c * 4);
            st_128(values + (size_t)p * D + (size_t)c * 4, v);
        } else {
            values[(size_t)p * D + c] = (t < max_len) ? __ldg(dense + ((size_t)lo * max_len + t) * D + c) : 0.f;
        }
    }
}

__global__ void jagged_1d_to_padded_dense_kernel(const int64_t* __restrict__ values,
                                                 const int64_t* __restrict__ offsets, int64_t B,
                                                 int64_t desired_length, int64_t padding_value,
                                                 int64_t* __restrict__ dense) {
    const int64_t total = B * desired_length;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / desired_length;
        const int64_t t = i - b * desired_length;
        const int64_t s = __ldg(offsets + b);
        const int64_t len = __ldg(offsets + b + 1) - s;
        dense[i] = (t < len) ? __ldg(values + s + t) : padding_value;
    }
}

// Strided block copy used around the pooled-output exchanges: slot j moves B rows of D floats
// from src + src_off[j] (+ b * src_stride[j]) to dst + dst_off[j] (+ b * dst_stride[j]).
struct CopyItem {
    const uint4* src;
    uint4* dst;
};

__device__ __forceinline__ CopyItem copy_item(const float* src, float* dst, int64_t i, int32_t D4, int32_t nslots,
                                              const int64_t* __restrict__ meta) {
    const int c = (int)(i % D4);
    const int64_t rb = i / D4;
    const int j = (int)(rb % nslots);
    const int64_t b = rb / nslots;
    const int64_t so = __ldg(meta + j), ss = __ldg(meta + nslots + j);
    const int64_t dof = __ldg(meta + 2 * nslots + j), ds = __ldg(meta + 3 * nslots + j);
    CopyItem it;
    it.src = reinterpret_cast<const uint4*>(src + so + b * ss) + c;
    it.dst = reinterpret_cast<uint4*>(dst + dof + b * ds) + c;
    return it;
}

// UNROLL independent 16-byte loads are issued before their stores.  The peer-push form (dst == NULL,
// destinations are NVLink peer addresses) runs with few CTAs and UNROLL 4: it is NVLink-bound, and a
// small grid leaves the SMs to the kernels that overlap it (the sort of the sharded backward).
// MC: the destinations are NVLS multicast addresses; one multimem store lands in the same place of every
// rank's buffer (the switch replicates it), so an all-gather costs each GPU one copy of its data on the wire.
__device__ __forceinline__ void multimem_st_128(void* mc_addr, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};"
                 :: "l"(mc_addr), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)),
                    "f"(__uint_as_float(v.w)) : "memory");
}

template <int UNROLL, bool MC = false>
__global__ void __launch_bounds__(256)
copy_feature_blocks_kernel(const float* __restrict__ src, float* __restrict__ dst, int64_t B,
                           int32_t D4, int32_t nslots, const int64_t* __restrict__ meta) {
    const int64_t total = (int64_t)nslots * B * D4;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += stride * UNROLL) {
        CopyItem it[UNROLL];
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) {
            const int64_t i = i0 + u * stride;
            it[u] = copy_item(src, dst, i < total ? i : i0, D4, nslots, meta);
        }
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = __ldcs(it[u].src);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u)
            if (i0 + u * stride < total) {
                if (MC) multimem_st_128(it[u].dst, v[u]);
                else __stcs(it[u].dst, v[u]);
            }
    }
}

// ---------------------------------------------------------------- row-wise partial sums
// out[b][out_col[k] + 4c ..] = parts[0][b][k][c] + parts[1][b][k][c] + ... in source-rank order.
// W independent 16-byte loads per output chunk; HBM-bound (W reads, 1 write).
__global__ void __launch_bounds__(256)
reduce_partial_sums_kernel(const float* __restrict__ parts, int W, int64_t part_stride, int64_t B, int32_t nslots,
                           int32_t D4, const int64_t* __restrict__ out_col, float* __restrict__ out,
                           int64_t out_stride) {
    const int64_t row4 = (int64_t)nslots * D4;                 // 16-byte chunks per sample in one slot
    const int64_t total = B * row4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t b = i / row4;
        const int64_t kc = i - b * row4;
        const int k = (int)(kc / D4);
        const int c = (int)(kc - (int64_t)k * D4);
        const uint4* src = reinterpret_cast<const uint4*>(parts) + i;
        const int64_t stride4 = part_stride / 4;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        int s = 0;
        for (; s + 4 <= W; s += 4) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldcs(src + (int64_t)(s + u) * stride4);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc[0] += __uint_as_float(v[u].x); acc[1] += __uint_as_float(v[u].y);
                acc[2] += __uint_as_float(v[u].z); acc[3] += __uint_as_float(v[u].w);
            }
        }
        for (; s < W; ++s) {
            const uint4 v = __ldcs(src + (int64_t)s * stride4);
            acc[0] += __uint_as_float(v.x); acc[1] += __uint_as_float(v.y);
            acc[2] += __uint_as_float(v.z); acc[3] += __uint_as_float(v.w);
        }
        float* o = out + b * out_stride + __ldg(out_col + k) + 4 * c;
        st_128(o, make_uint4(__float_as_uint(acc[0]), __float_as_uint(acc[1]), __float_as_uint(acc[2]),
                             __float_as_uint(acc[3])));
    }
}

// ---------------------------------------------------------------- input dist over peer memory
// Slot j copies B lengths from src + src_off[j] to the absolute address dst_addr[j]
// (a peer rank's receive buffer, already in its final [feature][source rank][batch] layout).
__global__ void push_lengths_kernel(const int32_t* __restrict__ src, int64_t B, int32_t nslots,
                                    const int64_t* __restrict__ meta) {
    const int64_t total = (int64_t)nslots * B;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
         i += (int64_t)gridDim.x * blockDim.x) {
        const int j = (int)(i / B);
        const int64_t b = i - (int64_t)j * B;
        int32_t* dst = reinterpret_cast<int32_t*>(__ldg(meta + nslots + j));
        dst[b] = __ldg(src + __ldg(meta + j) + b);
    }
}

// Slot j moves the ids of one whole feature segment (B consecutive bags, contiguous on both
// sides) to the owner rank: the destination offset is read from the owner's offsets array
// (one remote 8-byte load per slot), the ids are then streamed over NVLink with plain stores.
// blockIdx.y = slot; blockIdx.x strides over the segment.
__global__ void __launch_bounds__(256)
push_ids_kernel(const int64_t* __restrict__ values, const int64_t* __restrict__ offsets, int64_t B,
                const int64_t* __restrict__ meta, int32_t nslots, int64_t capacity, int32_t* err_flag) {
    const int j = blockIdx.y;
    const int64_t seg = __ldg(meta + j);
    const int64_t s0 = __ldg(offsets + seg * B), s1 = __ldg(offsets + (seg + 1) * B);
    const int64_t n = s1 - s0;
    if (n <= 0) return;
    const int64_t dst_off = *reinterpret_cast<const volatile int64_t*>(__ldg(meta + nslots + j));
    if (dst_off < 0 || dst_off + n > capacity) {
        if (err_flag && threadIdx.x == 0 && blockIdx.x == 0) atomicOr(err_flag, TDFO_FLAG_RECV_OVERFLOW);
        return;
    }
    int64_t* dst = reinterpret_cast<int64_t*>(__ldg(meta + 2 * nslots + j)) + dst_off;
    const int64_t* src = values + s0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        dst[i] = __ldg(src + i);
}

// int32 -> int64 ids, four per thread (16-byte load, two 16-byte stores)
__global__ void __launch_bounds__(256)
widen_ids_kernel(const int32_t* __restrict__ in, int64_t n, int64_t* __restrict__ out) {
    const int64_t n4 = n / 4;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const int4 v = __ldcs(reinterpret_cast<const int4*>(in) + i);
        longlong2* o = reinterpret_cast<longlong2*>(out) + 2 * i;
        o[0] = make_longlong2((long long)v.x, (long long)v.y);
        o[1] = make_longlong2((long long)v.z, (long long)v.w);
    }
    const int64_t t = n4 * 4 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;      // tail (n % 4 ids)
    if (t < n) out[t] = (int64_t)in[t];
}

// Slot j moves the rows of one whole segment (B consecutive bags of the sender's offsets) to the owner:
// element i of the segment is row rows[positions ? positions[s0 + i] : s0 + i] and lands at row
// (remote offset word + i) of the owner's buffer.  blockIdx.y = slot.
__global__ void __launch_bounds__(256)
push_rows_kernel(const float* __restrict__ rows, const int64_t* __restrict__ positions,
                 const int64_t* __restrict__ offsets, int64_t B, int32_t D4, const int64_t* __restrict__ meta,
                 int32_t nslots, int64_t capacity, int32_t* err_flag) {
    const int j = blockIdx.y;
    const int64_t seg = __ldg(meta + j);
    const int64_t s0 = __ldg(offsets + seg * B), s1 = __ldg(offsets + (seg + 1) * B);
    const int64_t n = s1 - s0;
    if (n <= 0) return;
    const int64_t dst_off = *reinterpret_cast<const volatile int64_t*>(__ldg(meta + nslots + j));
    if (dst_off < 0 || dst_off + n > capacity) {
        if (err_flag && threadIdx.x == 0 && blockIdx.x == 0) atomicOr(err_flag, TDFO_FLAG_RECV_OVERFLOW);
        return;
    }
    uint4* dst = reinterpret_cast<uint4*>(__ldg(meta + 2 * nslots + j)) + dst_off * D4;
    const uint4* src = reinterpret_cast<const uint4*>(rows);
    const int64_t total = n * D4;
    for (int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = t / D4;
        const int c = (int)(t - i * D4);
        const int64_t r = positions ? __ldg(positions + s0 + i) : s0 + i;
        __stcs(dst + t, __ldcs(src + r * D4 + c));
    }
}

// out[permute[i]] = positions_in ? positions_in[i] : i   (the inverse of a permutation, composed with a map)
__global__ void scatter_positions_kernel(const int64_t* __restrict__ permute, const int64_t* __restrict__ positions_in,
                                         int64_t n, const int64_t* __restrict__ n_dev, int64_t* __restrict__ out) {
    if (n_dev) n = min(n, __ldg(n_dev));          // the caller's buffers may be longer than the id count the device knows
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        out[__ldg(permute + i)] = positions_in ? __ldg(positions_in + i) : i;
}

static unsigned grid_for(int64_t items, int threads) {
    int64_t blocks = (items + threads - 1) / threads;
    const int64_t cap = (int64_t)kNumSMs * 32;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (unsigned)blocks;
}

}  // namespace tdfo

using namespace tdfo;

extern "C" size_t tdfo_bucketize_scratch_bytes(int64_t num_bags, int32_t world) {
    if (num_bags < 0 || world < 1) return 256;
    return tdfo_scan_scratch_bytes(num_bags * (int64_t)world);
}

extern "C" int tdfo_kjt_bucketize_row_wise(const int32_t* lengths, const int64_t* offsets,
                                           const int64_t* indices, const float* weights,
                                           const int64_t* block_sizes, int32_t F, int64_t B, int32_t world,
                                           int32_t* new_lengths, int64_t* new_offsets,
                                           int64_t* new_indices, float* new_weights,
                                           int64_t* unbucketize_permute, int32_t* err_flag,
                                           void* scratch, size_t scratch_bytes, void* stream) {
    (void)lengths;
    TDFO_CHECK_ARG(F >= 0 && B >= 0, "bucketize_row_wise: negative size");
    TDFO_CHECK_ARG(world >= 1 && world <= kMaxWorld, "bucketize_row_wise: world must be in [1, %d]", kMaxWorld);
    TDFO_CHECK_ARG(new_offsets != nullptr, "bucketize_row_wise: new_offsets is null");
    TDFO_CHECK_ARG((weights == nullptr) == (new_weights == nullptr),
                   "bucketize_row_wise: weights and new_weights must both be given or both be null");
    const int64_t num_bags = (int64_t)F * B;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    if (num_bags == 0) {
        TDFO_CUDA(cudaMemsetAsync(new_offsets, 0, sizeof(int64_t), st));
        return TDFO_OK;
    }
    // indices / new_indices may be null when every bag is empty
    TDFO_CHECK_ARG(offsets && block_sizes && new_lengths, "bucketize_row_wise: null pointer");
    const size_t need = tdfo_bucketize_scratch_bytes(num_bags, world);
    if (!scratch || !aligned(scratch, 256) || scratch_bytes < need) {
        set_error("bucketize_row_wise: scratch null, misaligned or smaller than %zu bytes", need);
        return TDFO_ERR_SCRATCH;
    }
    const unsigned blocks = (unsigned)((num_bags + kJagThreads / 32 - 1) / (kJagThreads / 32));
    bucketize_count_kernel<<<blocks, kJagThreads, 0, st>>>(offsets, indices, block_sizes, B, num_bags, world,
                                                          new_lengths, err_flag);
    count_launch();
    TDFO_LAUNCH_CHECK("bucketize_count");
    const int rc = complete_cumsum_impl(new_lengths, num_bags * world, new_offsets, scratch, st);
    if (rc) return rc;
    bucketize_scatter_kernel<<<blocks, kJagThreads, 0, st>>>(offsets, indices, weights, block_sizes, B, num_bags,
                                                            world, new_offsets, new_indices, new_weights,
                                                            unbucketize_permute);
    count_launch();
    TDFO_LAUNCH_CHECK("bucketize_scatter");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_permute_2d(const int32_t* permute, int32_t S_in, int32_t S_out, int64_t B,
                                   const int32_t* in_lengths, const int64_t* in_offsets,
                                   const int64_t* in_indices, const float* in_weights,
                                   int32_t* out_lengths, int64_t* out_offsets,
                                   int64_t* out_indices, float* out_weights,
                                   void* scratch, size_t scratch_bytes, void* stream) {
    TDFO_CHECK_ARG(S_in >= 0 && S_out >= 0 && B >= 0, "permute_2d: negative size");
    TDFO_CHECK_ARG(out_offsets != nullptr, "permute_2d: out_offsets is null");
    TDFO_CHECK_ARG((in_weights == nullptr) == (out_weights == nullptr),
                   "permute_2d: in_weights and out_weights must both be given or both be null");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t n_out = (int64_t)S_out * B;
    if (n_out == 0) {
        TDFO_CUDA(cudaMemsetAsync(out_offsets, 0, sizeof(int64_t), st));
        return TDFO_OK;
    }
    TDFO_CHECK_ARG(permute && in_lengths && in_offsets && out_lengths, "permute_2d: null pointer");
    const size_t need = tdfo_scan_scratch_bytes(n_out);
    if (!scratch || !aligned(scratch, 256) || scratch_bytes < need) {
        set_error("permute_2d: scratch null, misaligned or smaller than %zu bytes", need);
        return TDFO_ERR_SCRATCH;
    }
    permute_lengths_kernel<<<(unsigned)((n_out + 255) / 256), 256, 0, st>>>(permute, B, n_out, in_lengths, out_lengths);
    count_launch();
    TDFO_LAUNCH_CHECK("permute_lengths");
    const int rc = complete_cumsum_impl(out_lengths, n_out, out_offsets, scratch, st);
    if (rc) return rc;
    if (in_indices && out_indices) {
        const unsigned blocks = (unsigned)((n_out + kJagThreads / 32 - 1) / (kJagThreads / 32));
        permute_data_kernel<<<blocks, kJagThreads, 0, st>>>(permute, B, n_out, in_offsets, out_offsets,
                                                           in_indices, in_weights, out_indices, out_weights);
        count_launch();
        TDFO_LAUNCH_CHECK("permute_data");
    }
    return TDFO_OK;
}

extern "C" int tdfo_jagged_2d_to_dense(const float* values, const int64_t* offsets, int64_t B, int32_t D,
                                       int64_t max_len, float* dense, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && D > 0 && max_len >= 0, "jagged_2d_to_dense: bad size");
    if (B == 0 || max_len == 0) return TDFO_OK;
    TDFO_CHECK_ARG(offsets && dense, "jagged_2d_to_dense: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool vec = (D % 4 == 0) && aligned(values, 16) && aligned(dense, 16);
    if (vec)
        jagged_2d_to_dense_kernel<4><<<grid_for(B * max_len * (D / 4), 256), 256, 0, st>>>(values, offsets, B, D, max_len, dense);
    else
        jagged_2d_to_dense_kernel<1><<<grid_for(B * max_len * D, 256), 256, 0, st>>>(values, offsets, B, D, max_len, dense);
    count_launch();
    TDFO_LAUNCH_CHECK("jagged_2d_to_dense");
    return TDFO_OK;
}

extern "C" int tdfo_dense_to_jagged(const float* dense, const int64_t* offsets, int64_t B, int32_t D,
                                    int64_t max_len, int64_t L, float* values, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && D > 0 && max_len >= 0 && L >= 0, "dense_to_jagged: bad size");
    if (L == 0) return TDFO_OK;
    TDFO_CHECK_ARG(offsets && values && (dense || max_len == 0) && B > 0, "dense_to_jagged: null pointer");
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const bool vec = (D % 4 == 0) && aligned(values, 16) && aligned(dense, 16);
    if (vec)
        dense_to_jagged_kernel<4><<<grid_for(L * (D / 4), 256), 256, 0, st>>>(dense, offsets, B, D, max_len, L, values);
    else
        dense_to_jagged_kernel<1><<<grid_for(L * D, 256), 256, 0, st>>>(dense, offsets, B, D, max_len, L, values);
    count_launch();
    TDFO_LAUNCH_CHECK("dense_to_jagged");
    return TDFO_OK;
}

extern "C" int tdfo_jagged_1d_to_padded_dense(const int64_t* values, const int64_t* offsets, int64_t B,
                                              int64_t desired_length, int64_t padding_value,
                                              int64_t* dense, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && desired_length >= 0, "jagged_1d_to_padded_dense: bad size");
    if (B == 0 || desired_length == 0) return TDFO_OK;
    TDFO_CHECK_ARG(offsets && dense, "jagged_1d_to_padded_dense: null pointer");
    jagged_1d_to_padded_dense_kernel<<<grid_for(B * desired_length, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        values, offsets, B, desired_length, padding_value, dense);
    count_launch();
    TDFO_LAUNCH_CHECK("jagged_1d_to_padded_dense");
    return TDFO_OK;
}

extern "C" int tdfo_copy_feature_blocks(const float* src, float* dst, int64_t B, int32_t D, int32_t nslots,
                                        const int64_t* meta, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && D > 0 && D % 4 == 0 && nslots >= 0, "copy_feature_blocks: bad size (D must be a multiple of 4)");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(src && meta, "copy_feature_blocks: null pointer");   // dst == NULL: absolute destinations
    TDFO_CHECK_ARG(aligned(src, 16) && aligned(dst, 16), "copy_feature_blocks: buffers not 16-byte aligned");
    const int64_t items = (int64_t)nslots * B * (D / 4);
    if (dst == nullptr) {       // peer push: NVLink-bound, keep the grid small (see the kernel comment)
        unsigned g = grid_for((items + 3) / 4, 256);
        if (g > (unsigned)kNumSMs * 2) g = kNumSMs * 2;
        copy_feature_blocks_kernel<4><<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, B, D / 4, nslots, meta);
    } else {
        copy_feature_blocks_kernel<1><<<grid_for(items, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
            src, dst, B, D / 4, nslots, meta);
    }
    count_launch();
    TDFO_LAUNCH_CHECK("copy_feature_blocks");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_push_lengths(const int32_t* lengths, int64_t B, int32_t nslots, const int64_t* meta,
                                     void* stream) {
    TDFO_CHECK_ARG(B >= 0 && nslots >= 0, "kjt_push_lengths: negative size");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(lengths && meta, "kjt_push_lengths: null pointer");
    push_lengths_kernel<<<grid_for((int64_t)nslots * B, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        lengths, B, nslots, meta);
    count_launch();
    TDFO_LAUNCH_CHECK("kjt_push_lengths");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_push_ids(const int64_t* values, const int64_t* offsets, int64_t B, int32_t nslots,
                                 const int64_t* meta, int64_t capacity, int32_t* err_flag, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && nslots >= 0 && capacity >= 0, "kjt_push_ids: negative size");
    TDFO_CHECK_ARG(nslots <= 65535, "kjt_push_ids: at most 65535 segments per call");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(offsets && meta, "kjt_push_ids: null pointer");
    dim3 grid(64, (unsigned)nslots);          // 64 x 256 threads stream one segment
    push_ids_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(values, offsets, B, meta, nslots, capacity,
                                                                        err_flag);
    count_launch();
    TDFO_LAUNCH_CHECK("kjt_push_ids");
    return TDFO_OK;
}

extern "C" int tdfo_reduce_partial_sums(const float* parts, int32_t W, int64_t part_stride, int64_t B, int32_t nslots,
                                        int32_t D, const int64_t* out_col, float* out, int64_t out_stride,
                                        void* stream) {
    TDFO_CHECK_ARG(W >= 1 && B >= 0 && nslots >= 0 && D > 0 && D % 4 == 0,
                   "reduce_partial_sums: bad size (D must be a multiple of 4)");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(parts && out && out_col, "reduce_partial_sums: null pointer");
    TDFO_CHECK_ARG(aligned(parts, 16) && aligned(out, 16) && part_stride % 4 == 0 && out_stride % 4 == 0,
                   "reduce_partial_sums: buffers / strides not 16-byte aligned");
    TDFO_CHECK_ARG(part_stride >= B * (int64_t)nslots * D, "reduce_partial_sums: part_stride smaller than one slot");
    const int64_t items = B * (int64_t)nslots * (D / 4);
    reduce_partial_sums_kernel<<<grid_for(items, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(
        parts, W, part_stride, B, nslots, D / 4, out_col, out, out_stride);
    count_launch();
    TDFO_LAUNCH_CHECK("reduce_partial_sums");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_widen_ids(const int32_t* in, int64_t n, int64_t* out, void* stream) {
    TDFO_CHECK_ARG(n >= 0, "kjt_widen_ids: n < 0");
    if (n == 0) return TDFO_OK;
    TDFO_CHECK_ARG(in && out, "kjt_widen_ids: null pointer");
    TDFO_CHECK_ARG(aligned(in, 16) && aligned(out, 16), "kjt_widen_ids: buffers not 16-byte aligned");
    widen_ids_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(in, n, out);
    count_launch();
    TDFO_LAUNCH_CHECK("kjt_widen_ids");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_push_rows(const float* rows, const int64_t* positions, const int64_t* offsets, int64_t B,
                                  int32_t D, int32_t nslots, const int64_t* meta, int64_t capacity, int32_t* err_flag,
                                  void* stream) {
    TDFO_CHECK_ARG(B >= 0 && nslots >= 0 && capacity >= 0 && D > 0 && D % 4 == 0, "kjt_push_rows: bad size (D must be a multiple of 4)");
    TDFO_CHECK_ARG(nslots <= 65535, "kjt_push_rows: at most 65535 segments per call");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(offsets && meta, "kjt_push_rows: null pointer");
    TDFO_CHECK_ARG(aligned(rows, 16), "kjt_push_rows: rows not 16-byte aligned");
    dim3 grid(64, (unsigned)nslots);
    push_rows_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(rows, positions, offsets, B, D / 4, meta, nslots,
                                                                         capacity, err_flag);
    count_launch();
    TDFO_LAUNCH_CHECK("kjt_push_rows");
    return TDFO_OK;
}

extern "C" int tdfo_kjt_scatter_positions(const int64_t* permute, const int64_t* positions_in, int64_t n,
                                          const int64_t* n_dev, int64_t* out, void* stream) {
    TDFO_CHECK_ARG(n >= 0, "kjt_scatter_positions: n < 0");
    if (n == 0) return TDFO_OK;
    TDFO_CHECK_ARG(permute && out, "kjt_scatter_positions: null pointer");
    scatter_positions_kernel<<<grid_for(n, 256), 256, 0, static_cast<cudaStream_t>(stream)>>>(permute, positions_in, n, n_dev, out);
    count_launch();
    TDFO_LAUNCH_CHECK("kjt_scatter_positions");
    return TDFO_OK;
}

extern "C" int tdfo_copy_feature_blocks_multicast(const float* src, int64_t B, int32_t D, int32_t nslots,
                                                  const int64_t* meta, void* stream) {
    TDFO_CHECK_ARG(B >= 0 && D > 0 && D % 4 == 0 && nslots >= 0, "copy_feature_blocks_multicast: bad size (D must be a multiple of 4)");
    if (B == 0 || nslots == 0) return TDFO_OK;
    TDFO_CHECK_ARG(src && meta, "copy_feature_blocks_multicast: null pointer");
    TDFO_CHECK_ARG(aligned(src, 16), "copy_feature_blocks_multicast: src not 16-byte aligned");
    const int64_t items = (int64_t)nslots * B * (D / 4);
    unsigned g = grid_for((items + 3) / 4, 256);
    if (g > (unsigned)kNumSMs * 2) g = kNumSMs * 2;
    copy_feature_blocks_kernel<4, true><<<g, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, nullptr, B, D / 4, nslots, meta);
    count_launch();
    TDFO_LAUNCH_CHECK("copy_feature_blocks_multicast");
    return TDFO_OK;
}
