/*
 * tdfo_oracle.c — CPU restatement of the sparse-embedding hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under tdfo_b200/ may import, link or call
 * this file; only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs use it, and there only as the checker / the CPU arm.
 *
 * PARITY STATUS: unpinned by the reference.  massquantity/tdfo ships no tests and
 * no golden vectors, and the arithmetic of this path lives in un-vendored
 * third-party packages (torchrec==0.4.0 pinned at torchrec/README.md:23;
 * fbgemm_gpu unpinned, imported at torchrec/train.py:13) that cannot be
 * installed here.  What pins this file instead (tests/golden/make_golden.py):
 *   - torch.nn.EmbeddingBag / nn.Embedding forward and autograd (the modules an
 *     unsharded torchrec EmbeddingBagCollection / EmbeddingCollection wraps),
 *   - FBGEMM's own CPU reference `fbgemm::rowwise_sparse_adagrad_ref`, exported
 *     by libtorch_cpu.so,
 *   - torch.optim.SGD / AdamW for the fused SGD / ADAM update,
 *   - hand-computed known-answer cases for the routing (bucketise, permute).
 *
 * Every function names the reference call site (relative to /root/reference)
 * whose behaviour it restates; [UPSTREAM] marks semantics of the third-party
 * implementation behind that call site (SURVEY.md §8a rows a9–a15).
 *
 * Plain C99, sequential fp32 arithmetic in a fixed order (compile with
 * -ffp-contract=off).  OpenMP only parallelises over independent bags / rows,
 * never inside a reduction, so results do not depend on the thread count.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif


#define ORC_F32 0
#define ORC_BF16 1
#define ORC_POOL_SUM 0
#define ORC_POOL_MEAN 1
#define ORC_POOL_NONE 2
#define ORC_OPT_SGD 0
#define ORC_OPT_ROWWISE_ADAGRAD 1
#define ORC_OPT_ADAM 2

static inline float bf16_to_f32(uint16_t h) {
    uint32_t u = (uint32_t)h << 16;
    float f;
    memcpy(&f, &u, 4);
    return f;
}

static inline uint16_t f32_to_bf16_rne(float f) {
    uint32_t u;
    memcpy(&u, &f, 4);
    if ((u & 0x7fffffffu) > 0x7f800000u) return (uint16_t)((u >> 16) | 0x40u); /* NaN */
    u += 0x7fffu + ((u >> 16) & 1u);
    return (uint16_t)(u >> 16);
}

static inline float load_w(const void* w, int dtype, int64_t i) {
    return dtype == ORC_F32 ? ((const float*)w)[i] : bf16_to_f32(((const uint16_t*)w)[i]);
}

static inline void store_w(void* w, int dtype, int64_t i, float v) {
    if (dtype == ORC_F32) ((float*)w)[i] = v;
    else ((uint16_t*)w)[i] = f32_to_bf16_rne(v);
}

/* KeyedJaggedTensor.from_lengths_sync -> offsets (torchrec/train.py:38-40).
 * [UPSTREAM] complete cumulative sum: offsets[0] = 0, offsets[i+1] = sum(lengths[0..i]). */
void orc_complete_cumsum(const int32_t* lengths, int64_t n, int64_t* offsets) {
    offsets[0] = 0;
    for (int64_t i = 0; i < n; ++i) offsets[i + 1] = offsets[i] + lengths[i];
}

/* Row-wise input dist of the sharded collection built by DMP at torchrec/train.py:241-247.
 * [UPSTREAM] block_bucketize_sparse_features(bucketize_pos=False): bucket = id / block_size,
 * new id = id - bucket*block_size, output grouped [bucket][feature][batch], original order
 * kept inside each group; unbucketize_permute[p] = slot of original id p.
 * Returns 1 if any id fell outside [0, world*block_size) (it is then sent to the last bucket). */
int orc_bucketize_row_wise(const int64_t* offsets, const int64_t* indices, const float* weights,
                           const int64_t* block_sizes, int32_t F, int64_t B, int32_t world,
                           int32_t* new_lengths, int64_t* new_offsets, int64_t* new_indices,
                           float* new_weights, int64_t* unbucketize_permute) {
    const int64_t bags = (int64_t)F * B;
    int oob = 0;
    memset(new_lengths, 0, sizeof(int32_t) * (size_t)(bags * world));
    for (int64_t g = 0; g < bags; ++g) {
        const int64_t bs = block_sizes[g / B];
        for (int64_t p = offsets[g]; p < offsets[g + 1]; ++p) {
            int64_t q = indices[p] / bs;
            if (indices[p] < 0 || q >= world) { q = world - 1; oob = 1; }
            new_lengths[q * bags + g] += 1;
        }
    }
    orc_complete_cumsum(new_lengths, bags * world, new_offsets);
    int64_t* cursor = (int64_t*)malloc(sizeof(int64_t) * (size_t)(bags * world + 1));
    memcpy(cursor, new_offsets, sizeof(int64_t) * (size_t)(bags * world + 1));
    for (int64_t g = 0; g < bags; ++g) {
        const int64_t bs = block_sizes[g / B];
        for (int64_t p = offsets[g]; p < offsets[g + 1]; ++p) {
            int64_t q = indices[p] / bs;
            if (indices[p] < 0 || q >= world) q = world - 1;
            const int64_t slot = cursor[q * bags + g]++;
            new_indices[slot] = indices[p] - q * bs;
            if (weights) new_weights[slot] = weights[p];
            if (unbucketize_permute) unbucketize_permute[p] = slot;
        }
    }
    free(cursor);
    return oob;
}

/* [UPSTREAM] permute_2D_sparse_data inside the same input dist (torchrec/train.py:98 -> DMP
 * forward), also KeyedJaggedTensor.permute: output segment i is input segment permute[i],
 * each segment being B consecutive bags. */
void orc_permute_2d(const int32_t* permute, int32_t S_out, int64_t B, const int32_t* in_lengths,
                    const int64_t* in_offsets, const int64_t* in_indices, const float* in_weights,
                    int32_t* out_lengths, int64_t* out_offsets, int64_t* out_indices,
                    float* out_weights) {
    for (int64_t i = 0; i < (int64_t)S_out * B; ++i)
        out_lengths[i] = in_lengths[(int64_t)permute[i / B] * B + i % B];
    orc_complete_cumsum(out_lengths, (int64_t)S_out * B, out_offsets);
    if (!in_indices) return;
    for (int64_t i = 0; i < (int64_t)S_out * B; ++i) {
        const int64_t src = (int64_t)permute[i / B] * B + i % B;
        for (int64_t j = 0; j < out_lengths[i]; ++j) {
            out_indices[out_offsets[i] + j] = in_indices[in_offsets[src] + j];
            if (in_weights) out_weights[out_offsets[i] + j] = in_weights[in_offsets[src] + j];
        }
    }
}

/* `self.embed_collection(id_list_features)` torchrec/models.py:164 (POOL_NONE) and the
 * EmbeddingBagCollection of the BASELINE configs (SUM / MEAN).
 * [UPSTREAM] unsharded torchrec modules wrap nn.Embedding / nn.EmbeddingBag per table:
 * out[b, col_f : col_f + D] = sum_j w_j * W_f[idx_j, :], MEAN divides by the bag length,
 * empty bag -> zeros; sequence mode returns one row per id.
 * Ids outside the table are skipped and reported (return value 1). */
int orc_embedding_forward(const void* weights, int32_t dtype, int32_t D, const int64_t* indices,
                          const int64_t* offsets, const float* psw, const int64_t* feat_row_base,
                          const int64_t* feat_rows, const int64_t* feat_out_col, int32_t F, int64_t B,
                          int64_t L, int32_t pooling_mode, float* out, int64_t out_stride) {
    int oob = 0;
    (void)L;
    if (pooling_mode == ORC_POOL_NONE) {
#pragma omp parallel for schedule(static) reduction(| : oob)
        for (int64_t g = 0; g < (int64_t)F * B; ++g) {
            const int f = (int)(g / B);
            for (int64_t p = offsets[g]; p < offsets[g + 1]; ++p) {
                const int64_t idx = indices[p];
                float* o = out + p * D;
                if (idx < 0 || idx >= feat_rows[f]) {
                    oob |= 1;
                    for (int d = 0; d < D; ++d) o[d] = 0.f;
                    continue;
                }
                const int64_t row = feat_row_base[f] + idx;
                for (int d = 0; d < D; ++d) o[d] = load_w(weights, dtype, row * D + d);
            }
        }
        return oob;
    }
#pragma omp parallel for schedule(static) reduction(| : oob)
    for (int64_t g = 0; g < (int64_t)F * B; ++g) {
        const int f = (int)(g / B);
        const int64_t b = g % B;
        float* o = out + b * out_stride + feat_out_col[f];
        for (int d = 0; d < D; ++d) o[d] = 0.f;
        const int64_t s = offsets[g], e = offsets[g + 1];
        for (int64_t p = s; p < e; ++p) {
            const int64_t idx = indices[p];
            if (idx < 0 || idx >= feat_rows[f]) { oob |= 1; continue; }
            const int64_t row = feat_row_base[f] + idx;
            const float w = psw ? psw[p] : 1.f;
            for (int d = 0; d < D; ++d) o[d] = fmaf(w, load_w(weights, dtype, row * D + d), o[d]);
        }
        if (pooling_mode == ORC_POOL_MEAN && e > s) {
            const float inv = 1.f / (float)(e - s);
            for (int d = 0; d < D; ++d) o[d] *= inv;
        }
    }
    return oob;
}

/* Backward prologue reached from loss.backward() at torchrec/train.py:104.
 * [UPSTREAM] linear row = table row offset + id; payload = bag id (pooled) or position
 * (sequence).  Out-of-range ids get key total_rows. Returns 1 if any id was out of range. */
int orc_linearize_indices(const int64_t* indices, const int64_t* offsets, const int64_t* feat_row_base,
                          const int64_t* feat_rows, int32_t F, int64_t B, int64_t total_rows,
                          int32_t pooling_mode, uint32_t* keys, uint32_t* vals) {
    int oob = 0;
#pragma omp parallel for schedule(static) reduction(| : oob)
    for (int64_t g = 0; g < (int64_t)F * B; ++g) {
        const int f = (int)(g / B);
        for (int64_t p = offsets[g]; p < offsets[g + 1]; ++p) {
            const int64_t idx = indices[p];
            if (idx < 0 || idx >= feat_rows[f]) { keys[p] = (uint32_t)total_rows; oob |= 1; }
            else keys[p] = (uint32_t)(feat_row_base[f] + idx);
            vals[p] = pooling_mode == ORC_POOL_NONE ? (uint32_t)p : (uint32_t)g;
        }
    }
    return oob;
}

/* [UPSTREAM] cub::DeviceRadixSort::SortPairs in the TBE backward (torchrec/train.py:104):
 * a stable sort of (row, payload) by row.  Restated as a stable LSD counting sort with two
 * 16-bit digits; any stable sort gives the same permutation. */
void orc_sort_pairs_stable(const uint32_t* keys_in, const uint32_t* vals_in, int64_t n,
                           uint32_t* keys_out, uint32_t* vals_out) {
    /* Each pass: the input is cut into T contiguous chunks (one per thread); chunk c counts its digits,
     * the exclusive prefix over (digit, chunk) gives every chunk its own output cursor per digit, and the
     * chunks scatter independently.  Chunks are taken in input order inside a digit, so the pass is stable
     * and the result is the same for any T. */
    uint32_t* tk = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    uint32_t* tv = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(n > 0 ? n : 1));
    int T = 1;
#ifdef _OPENMP
    T = omp_get_max_threads();
#endif
    if (n < 65536 || T < 1) T = 1;
    int64_t* count = (int64_t*)malloc(sizeof(int64_t) * 65536 * (size_t)T);
    for (int pass = 0; pass < 2; ++pass) {
        const uint32_t* sk = pass == 0 ? keys_in : tk;
        const uint32_t* sv = pass == 0 ? vals_in : tv;
        uint32_t* dk = pass == 0 ? tk : keys_out;
        uint32_t* dv = pass == 0 ? tv : vals_out;
        const int shift = 16 * pass;
        memset(count, 0, sizeof(int64_t) * 65536 * (size_t)T);
#pragma omp parallel for schedule(static, 1) num_threads(T)
        for (int c = 0; c < T; ++c) {
            int64_t* cnt = count + (size_t)c * 65536;
            const int64_t lo = n * c / T, hi = n * (c + 1) / T;
            for (int64_t i = lo; i < hi; ++i) cnt[(sk[i] >> shift) & 0xffffu]++;
        }
        int64_t run = 0;
        for (int d = 0; d < 65536; ++d)
            for (int c = 0; c < T; ++c) {
                const int64_t x = count[(size_t)c * 65536 + d];
                count[(size_t)c * 65536 + d] = run;
                run += x;
            }
#pragma omp parallel for schedule(static, 1) num_threads(T)
        for (int c = 0; c < T; ++c) {
            int64_t* cnt = count + (size_t)c * 65536;
            const int64_t lo = n * c / T, hi = n * (c + 1) / T;
            for (int64_t i = lo; i < hi; ++i) {
                const int64_t slot = cnt[(sk[i] >> shift) & 0xffffu]++;
                dk[slot] = sk[i];
                dv[slot] = sv[i];
            }
        }
    }
    free(tk); free(tv); free(count);
}

/* [UPSTREAM] cub::DeviceRunLengthEncode in the TBE backward: unique rows and where each
 * run starts; seg_starts[num_unique] = n.  Returns num_unique. */
int64_t orc_unique_segments(const uint32_t* sorted_keys, int64_t n, uint32_t* unique_keys,
                            uint32_t* seg_starts) {
    /* chunks count their run heads, an exclusive prefix places them, chunks write independently */
    int T = 1;
#ifdef _OPENMP
    T = omp_get_max_threads();
#endif
    if (n < 65536 || T < 1) T = 1;
    int64_t* heads = (int64_t*)calloc((size_t)T + 1, sizeof(int64_t));
#pragma omp parallel for schedule(static, 1) num_threads(T)
    for (int c = 0; c < T; ++c) {
        const int64_t lo = n * c / T, hi = n * (c + 1) / T;
        int64_t h = 0;
        for (int64_t i = lo; i < hi; ++i) h += (i == 0 || sorted_keys[i] != sorted_keys[i - 1]);
        heads[c + 1] = h;
    }
    for (int c = 0; c < T; ++c) heads[c + 1] += heads[c];
#pragma omp parallel for schedule(static, 1) num_threads(T)
    for (int c = 0; c < T; ++c) {
        const int64_t lo = n * c / T, hi = n * (c + 1) / T;
        int64_t u = heads[c];
        for (int64_t i = lo; i < hi; ++i) {
            if (i == 0 || sorted_keys[i] != sorted_keys[i - 1]) {
                unique_keys[u] = sorted_keys[i];
                seg_starts[u] = (uint32_t)i;
                ++u;
            }
        }
    }
    const int64_t u = heads[T];
    free(heads);
    seg_starts[u] = (uint32_t)n;
    return u;
}

typedef struct orc_optimizer_args {
    int32_t optimizer;
    float learning_rate, eps, weight_decay, beta1, beta2;
    int64_t step;
} orc_optimizer_args;

/* Fused backward + optimizer reached from loss.backward() (torchrec/train.py:104), with the
 * hyper-parameters of torchrec/train.py:236-240.  [UPSTREAM] fbgemm TBE "exact" optimizers:
 * the gradient of a row is the ordered sum over its occurrences (MEAN: each scaled by 1/len),
 * then per touched row only
 *   SGD              g += wd*w;  w -= lr * g
 *   ROWWISE_ADAGRAD  g += wd*w;  h += mean_d(g_d^2);  w -= lr / (sqrt(h) + eps) * g
 *   ADAM             m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
 *                    w -= lr * ( (m/(1-b1^t)) / (sqrt(v/(1-b2^t)) + eps) + wd*w )
 * Untouched rows and their state are left alone.  bf16 tables: fp32 math, RNE store. */
void orc_sparse_update(void* weights, int32_t dtype, int64_t total_rows, int32_t D, float* state0,
                       float* state1, const float* grad_out, int64_t grad_stride,
                       const uint32_t* sorted_vals, const uint32_t* unique_keys,
                       const uint32_t* seg_starts, int64_t num_unique, const int64_t* offsets,
                       const int64_t* feat_out_col, int64_t B, int32_t pooling_mode,
                       const orc_optimizer_args* opt) {
    const float lr = opt->learning_rate, eps = opt->eps, wd = opt->weight_decay;
    const float b1 = opt->beta1, b2 = opt->beta2;
    float bc1 = 1.f, bc2 = 1.f;
    if (opt->optimizer == ORC_OPT_ADAM) {
        bc1 = (float)(1.0 - pow((double)b1, (double)opt->step));
        bc2 = (float)(1.0 - pow((double)b2, (double)opt->step));
    }
#pragma omp parallel
    {
        float* g = (float*)malloc(sizeof(float) * (size_t)D);
#pragma omp for schedule(static)
        for (int64_t u = 0; u < num_unique; ++u) {
            const int64_t row = unique_keys[u];
            if (row >= total_rows) continue;
            for (int d = 0; d < D; ++d) g[d] = 0.f;
            for (uint32_t i = seg_starts[u]; i < seg_starts[u + 1]; ++i) {
                const uint32_t v = sorted_vals[i];
                const float* gr;
                float scale = 1.f;
                if (pooling_mode == ORC_POOL_NONE) gr = grad_out + (int64_t)v * D;
                else {
                    const int64_t f = v / B, b = v % B;
                    gr = grad_out + b * grad_stride + feat_out_col[f];
                    if (pooling_mode == ORC_POOL_MEAN) scale = 1.f / (float)(offsets[v + 1] - offsets[v]);
                }
                for (int d = 0; d < D; ++d) g[d] = fmaf(scale, gr[d], g[d]);
            }
            if (opt->optimizer == ORC_OPT_SGD) {
                for (int d = 0; d < D; ++d) {
                    const float w = load_w(weights, dtype, row * D + d);
                    store_w(weights, dtype, row * D + d, w - lr * (g[d] + wd * w));
                }
            } else if (opt->optimizer == ORC_OPT_ROWWISE_ADAGRAD) {
                float ss = 0.f;
                for (int d = 0; d < D; ++d) {
                    g[d] += wd * load_w(weights, dtype, row * D + d);
                    ss += g[d] * g[d];
                }
                const float h = state0[row] + ss / (float)D;
                state0[row] = h;
                const float mult = lr / (sqrtf(h) + eps);
                for (int d = 0; d < D; ++d) {
                    const float w = load_w(weights, dtype, row * D + d);
                    store_w(weights, dtype, row * D + d, w - mult * g[d]);
                }
            } else {
                for (int d = 0; d < D; ++d) {
                    float* m = state0 + row * D + d;
                    float* v = state1 + row * D + d;
                    const float w = load_w(weights, dtype, row * D + d);
                    *m = b1 * *m + (1.f - b1) * g[d];
                    *v = b2 * *v + (1.f - b2) * g[d] * g[d];
                    const float mhat = *m / bc1, vhat = *v / bc2;
                    store_w(weights, dtype, row * D + d, w - lr * (mhat / (sqrtf(vhat) + eps) + wd * w));
                }
            }
        }
        free(g);
    }
}

/* The whole backward: linearise -> stable sort -> unique -> update.  Scratch is malloc'd here.
 * If the out_* pointers are non-NULL they receive the intermediate results (arrays of L, L,
 * L, L+1 entries and the unique count) for the bit-exact routing checks.
 * Returns 1 if any id was out of range. */
int orc_embedding_backward_update(void* weights, int32_t dtype, int64_t total_rows, int32_t D,
                                  float* state0, float* state1, const float* grad_out,
                                  int64_t grad_stride, const int64_t* indices, const int64_t* offsets,
                                  const int64_t* feat_row_base, const int64_t* feat_rows,
                                  const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                  int32_t pooling_mode, const orc_optimizer_args* opt,
                                  uint32_t* out_sorted_keys, uint32_t* out_sorted_vals,
                                  uint32_t* out_unique_keys, uint32_t* out_seg_starts,
                                  int64_t* out_num_unique) {
    const size_t n = (size_t)(L > 0 ? L : 1);
    uint32_t* k = (uint32_t*)malloc(4 * n);
    uint32_t* v = (uint32_t*)malloc(4 * n);
    uint32_t* sk = out_sorted_keys ? out_sorted_keys : (uint32_t*)malloc(4 * n);
    uint32_t* sv = out_sorted_vals ? out_sorted_vals : (uint32_t*)malloc(4 * n);
    uint32_t* uk = out_unique_keys ? out_unique_keys : (uint32_t*)malloc(4 * n);
    uint32_t* ss = out_seg_starts ? out_seg_starts : (uint32_t*)malloc(4 * (n + 1));
    const int oob = orc_linearize_indices(indices, offsets, feat_row_base, feat_rows, F, B, total_rows,
                                          pooling_mode, k, v);
    orc_sort_pairs_stable(k, v, L, sk, sv);
    const int64_t nu = orc_unique_segments(sk, L, uk, ss);
    if (out_num_unique) *out_num_unique = nu;
    orc_sparse_update(weights, dtype, total_rows, D, state0, state1, grad_out, grad_stride, sv, uk, ss, nu,
                      offsets, feat_out_col, B, pooling_mode, opt);
    free(k); free(v);
    if (!out_sorted_keys) free(sk);
    if (!out_sorted_vals) free(sv);
    if (!out_unique_keys) free(uk);
    if (!out_seg_starts) free(ss);
    return oob;
}

/* Weighted pooled lookup backward (SURVEY 8 a14 "x per-sample weight"; KJT weights are the optional
 * field of a9 -- the reference's own call sites never set them).  [UPSTREAM] nn.EmbeddingBag(mode=
 * "sum", per_sample_weights=w): d out / d W[idx_p] = w_p * grad_out[bag(p)].  Restated on its own:
 * the sort payload is the position p, bag(p) comes from a per-position table, the optimizer step
 * reuses orc_sparse_update by materialising the scaled rows as a sequence-mode gradient [L, D].
 * The weights receive no gradient.  Returns 1 if any id was out of range. */
int orc_embedding_backward_update_weighted(void* weights, int32_t dtype, int64_t total_rows, int32_t D,
                                           float* state0, float* state1, const float* grad_out,
                                           int64_t grad_stride, const int64_t* indices,
                                           const int64_t* offsets, const float* psw,
                                           const int64_t* feat_row_base, const int64_t* feat_rows,
                                           const int64_t* feat_out_col, int32_t F, int64_t B, int64_t L,
                                           int32_t pooling_mode, const orc_optimizer_args* opt,
                                           uint32_t* out_sorted_vals) {
    const size_t n = (size_t)(L > 0 ? L : 1);
    uint32_t* k = (uint32_t*)malloc(4 * n);
    uint32_t* v = (uint32_t*)malloc(4 * n);
    uint32_t* sk = (uint32_t*)malloc(4 * n);
    uint32_t* sv = out_sorted_vals ? out_sorted_vals : (uint32_t*)malloc(4 * n);
    uint32_t* uk = (uint32_t*)malloc(4 * n);
    uint32_t* ss = (uint32_t*)malloc(4 * (n + 1));
    float* rows = (float*)malloc(sizeof(float) * n * (size_t)D);
    /* payload = position: linearise in sequence mode */
    const int oob = orc_linearize_indices(indices, offsets, feat_row_base, feat_rows, F, B, total_rows,
                                          ORC_POOL_NONE, k, v);
    for (int64_t g = 0; g < (int64_t)F * B; ++g) {
        const int64_t f = g / B, b = g % B;
        const int64_t s = offsets[g], e = offsets[g + 1];
        const float* gr = grad_out + b * grad_stride + feat_out_col[f];
        for (int64_t p = s; p < e; ++p) {
            float scale = psw[p];
            if (pooling_mode == ORC_POOL_MEAN) scale = scale / (float)(e - s);
            for (int d = 0; d < D; ++d) rows[p * D + d] = scale * gr[d];
        }
    }
    orc_sort_pairs_stable(k, v, L, sk, sv);
    const int64_t nu = orc_unique_segments(sk, L, uk, ss);
    orc_sparse_update(weights, dtype, total_rows, D, state0, state1, rows, D, sv, uk, ss, nu,
                      offsets, feat_out_col, B, ORC_POOL_NONE, opt);
    free(k); free(v); free(sk); free(uk); free(ss); free(rows);
    if (!out_sorted_vals) free(sv);
    return oob;
}

/* torch.ops.fbgemm.jagged_2d_to_dense at torchrec/models.py:168-172.
 * [UPSTREAM] dense[b, t, :] = values[offsets[b] + t, :] for t < min(len_b, max_len), else 0. */
void orc_jagged_2d_to_dense(const float* values, const int64_t* offsets, int64_t B, int32_t D,
                            int64_t max_len, float* dense) {
    for (int64_t b = 0; b < B; ++b) {
        const int64_t len = offsets[b + 1] - offsets[b];
        for (int64_t t = 0; t < max_len; ++t)
            for (int d = 0; d < D; ++d)
                dense[(b * max_len + t) * D + d] = t < len ? values[(offsets[b] + t) * D + d] : 0.f;
    }
}

/* Autograd of the above (loss.backward(), torchrec/train.py:104); truncated rows get zero. */
void orc_dense_to_jagged(const float* dense, const int64_t* offsets, int64_t B, int32_t D,
                         int64_t max_len, float* values) {
    for (int64_t b = 0; b < B; ++b) {
        const int64_t len = offsets[b + 1] - offsets[b];
        for (int64_t t = 0; t < len; ++t)
            for (int d = 0; d < D; ++d)
                values[(offsets[b] + t) * D + d] = t < max_len ? dense[(b * max_len + t) * D + d] : 0.f;
    }
}

/* JaggedTensor.to_padded_dense(desired_length, padding_value) at torchrec/models.py:210-212. */
void orc_jagged_1d_to_padded_dense(const int64_t* values, const int64_t* offsets, int64_t B,
                                   int64_t desired_length, int64_t padding_value, int64_t* dense) {
    for (int64_t b = 0; b < B; ++b) {
        const int64_t len = offsets[b + 1] - offsets[b];
        for (int64_t t = 0; t < desired_length; ++t)
            dense[b * desired_length + t] = t < len ? values[offsets[b] + t] : padding_value;
    }
}

/* Thread control for the timed CPU baseline (bench.py): launchers such as torchrun export
 * OMP_NUM_THREADS=1, which would silently turn the all-cores baseline into a one-core one. */
int orc_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
    return omp_get_max_threads();
#else
    (void)n;
    return 1;
#endif
}
