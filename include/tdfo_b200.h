/*
 * tdfo_b200.h — C ABI of the B200-native sparse-embedding hot path.
 *
 * This is the drop-in boundary for the one path massquantity/tdfo delegates to
 * torchrec / fbgemm_gpu: KeyedJaggedTensor plumbing, the (sharded) embedding
 * lookup, and the sort/unique backward with the optimizer fused in.  The
 * reference has no C FFI of its own: it reaches this path through
 * `torch.ops.fbgemm.*` and the torchrec Python modules.  Every entry point
 * below names the reference call site (path relative to /root/reference) whose
 * third-party implementation it replaces.
 *
 * Conventions
 *  - Plain pointers and sizes only.  Unless a parameter is marked HOST, every
 *    pointer is a device pointer on the current CUDA device.
 *  - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default
 *    stream).  All work is enqueued on it; nothing synchronises the host.
 *  - No entry point allocates or frees device memory.  Kernels that need
 *    workspace take a caller-owned `scratch` pointer; the matching
 *    `*_scratch_bytes` function (pure host arithmetic) says how large it must
 *    be.  Scratch must be 256-byte aligned.  The one exception is
 *    tdfo_kjt_complete_cumsum_static, which uses a fixed workspace inside the
 *    library (its address is looked up on the first call per device); it is
 *    limited to TDFO_CUMSUM_STATIC_MAX_N lengths and its calls on one device
 *    must not overlap (see its declaration).
 *  - Return value: TDFO_OK (0) or a negative TDFO_ERR_* code; the message is
 *    available from tdfo_last_error() (thread-local).
 *  - ids (`indices`) are int64, `lengths` are int32, `offsets` are int64 and
 *    are *complete* cumulative sums (`offsets[0] == 0`, `n + 1` entries).
 *  - Out-of-range ids never touch memory: the lookup is skipped (treated as a
 *    zero row) and bit TDFO_FLAG_INDEX_OOB is OR-ed into `*err_flag` (a device
 *    word the caller zeroes and reads back lazily).  `err_flag` may be NULL.
 *    Likewise a bag whose offsets reach past the `L` ids passed in is cut at L
 *    (bit TDFO_FLAG_OFFSETS_RANGE): the lookup never reads beyond `indices[L)`.
 *  - Row size `D * sizeof(element)` must be a multiple of 16 bytes (128-bit
 *    vector loads); tables must be 16-byte aligned.
 */
#ifndef TDFO_B200_H
#define TDFO_B200_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TDFO_ABI_VERSION 1

/* return codes */
#define TDFO_OK                 0
#define TDFO_ERR_INVALID_ARG   (-1)
#define TDFO_ERR_CUDA          (-2)
#define TDFO_ERR_UNSUPPORTED   (-3)
#define TDFO_ERR_SCRATCH       (-4)

/* element types of tables / outputs */
#define TDFO_F32   0
#define TDFO_BF16  1

/* pooling modes (torchrec PoolingType; NONE = EmbeddingCollection, one output row per id) */
#define TDFO_POOL_SUM   0
#define TDFO_POOL_MEAN  1
#define TDFO_POOL_NONE  2

/* fused optimizers (fbgemm_gpu EmbOptimType as selected at torchrec/train.py:191,195) */
#define TDFO_OPT_SGD              0   /* EmbOptimType.SGD  (torchrec/train.py:195) */
#define TDFO_OPT_ROWWISE_ADAGRAD  1   /* EmbOptimType.EXACT_ROWWISE_ADAGRAD (north star) */
#define TDFO_OPT_ADAM             2   /* EmbOptimType.ADAM (torchrec/train.py:191) */

/* bits of *err_flag */
#define TDFO_FLAG_INDEX_OOB     1   /* an id was < 0 or >= rows of its table */
#define TDFO_FLAG_SCAN_TIMEOUT  2   /* internal look-back watchdog fired (bug guard) */
#define TDFO_FLAG_RECV_OVERFLOW 4   /* a peer's id receive buffer was too small for this batch */
#define TDFO_FLAG_OFFSETS_RANGE 8   /* an offset pointed past the L ids given: the bag was cut at L */

/* ------------------------------------------------------------------ misc */

/* ABI version of the loaded library (== TDFO_ABI_VERSION it was built with). */
int tdfo_abi_version(void);

/* Thread-local message of the last failing call ("" if none). */
const char* tdfo_last_error(void);

/* Kernels launched by this library since it was loaded (host-side counter). */
unsigned long long tdfo_launch_count(void);

/* ------------------------------------------------------------------ KJT */

/*
 * lengths[n] -> offsets[n+1], complete cumulative sum.
 * Replaces: KeyedJaggedTensor.from_lengths_sync(...) / .offsets()
 *   (torchrec/train.py:38-40; upstream fbgemm asynchronous_complete_cumsum).
 * scratch: tdfo_scan_scratch_bytes(n).
 */
size_t tdfo_scan_scratch_bytes(int64_t n);
int tdfo_kjt_complete_cumsum(const int32_t* lengths, int64_t n, int64_t* offsets,
                             void* scratch, size_t scratch_bytes, void* stream);

/*
 * The same lengths -> offsets scan for callers that cannot hand over device scratch because
 * they cannot allocate any: a plug-in running inside a framework that has already claimed
 * nearly all of HBM for its own pool (XLA and TensorFlow do this by default, and those are
 * the runtimes behind the reference's jax-flax and tensorflow2 trainers), where a cudaMalloc
 * from the plug-in fails or eats into the host framework's budget; or a serving loop that
 * must not hit the device-wide synchronisation cudaMalloc / cudaFree imply.  This entry point
 * has no counterpart in the reference.
 *
 * The workspace is a fixed array inside the library image (about 16 KB per GPU that loads the
 * library, held until process exit), which is why the input is limited to
 * TDFO_CUMSUM_STATIC_MAX_N lengths; larger n returns TDFO_ERR_SCRATCH and the message names
 * the limit.  The call performs no cudaMalloc / cudaFree and calls nothing that allocates;
 * `lengths` and `offsets` are still the caller's buffers.
 *
 *  - The first call on each device looks up the array's device address
 *    (cudaGetSymbolAddress, with that device current).  Make that first call before starting
 *    any stream capture on the device.
 *  - All calls on a device share the one array, and it still holds the previous call's
 *    contents; each call re-zeroes the words it needs on `stream` before its kernel runs.
 *    Two calls whose GPU work can overlap (different streams without an ordering between them,
 *    or several host threads) race on it: order them, or use tdfo_kjt_complete_cumsum with
 *    separate scratch buffers.
 * Results are identical to tdfo_kjt_complete_cumsum.
 */
#define TDFO_CUMSUM_STATIC_MAX_N (1 << 22)   /* 4 194 304 lengths, e.g. 256 features x batch 16 384 */
int tdfo_kjt_complete_cumsum_static(const int32_t* lengths, int64_t n, int64_t* offsets, void* stream);

/*
 * ids that fit 32 bits can cross PCIe as int32 (half the bytes of the per-step host->device copy, the
 * end-to-end bound on 8 GPUs): this widens them to the int64 the lookups take, the first time the
 * device touches them.  out[i] = in[i] (sign-extended), i < n.
 * Replaces: nothing in the reference, whose DataLoader hands int64 ids to `_to_kjt`
 *   (torchrec/train.py:33-41, 95-97); it is the device half of a narrower host format.
 */
int tdfo_kjt_widen_ids(const int32_t* in, int64_t n, int64_t* out, void* stream);

/*
 * Row-wise input dist, step 1: bucketise ids by destination rank.
 * For feature f with `block_sizes[f]` rows per rank, id x goes to bucket
 * x / block_sizes[f] and becomes x - bucket*block_sizes[f].  Output is laid
 * out [bucket][feature][batch], stable inside each (bucket, f, b).
 *   new_lengths  int32[world * F * B]
 *   new_offsets  int64[world * F * B + 1]   (complete cumsum of new_lengths)
 *   new_indices  int64[L]
 *   new_weights  float[L] or NULL (iff weights == NULL)
 *   unbucketize_permute int64[L] or NULL: position of original id p in new_indices
 *   (needed by the sequence / EmbeddingCollection output dist).
 * Replaces: the input_dist of the sharded EmbeddingCollection reached from
 *   DMP(...) at torchrec/train.py:241-247 (upstream
 *   fbgemm block_bucketize_sparse_features); the only explicit row partition in
 *   the reference is MinSizePartitioner at tensorflow2/train_ps.py:55-61.
 * Ids outside [0, world*block_size) set TDFO_FLAG_INDEX_OOB and go to the last bucket.
 * scratch: tdfo_bucketize_scratch_bytes(F*B, world).
 */
size_t tdfo_bucketize_scratch_bytes(int64_t num_bags, int32_t world);
int tdfo_kjt_bucketize_row_wise(const int32_t* lengths, const int64_t* offsets,
                                const int64_t* indices, const float* weights,
                                const int64_t* block_sizes /* [F] */,
                                int32_t F, int64_t B, int32_t world,
                                int32_t* new_lengths, int64_t* new_offsets,
                                int64_t* new_indices, float* new_weights,
                                int64_t* unbucketize_permute,
                                int32_t* err_flag,
                                void* scratch, size_t scratch_bytes, void* stream);

/*
 * Permute the (segment x batch) rows of a jagged 2-D layout:
 *   out segment i = in segment permute[i]; every segment has B bags.
 *   in_lengths int32[S_in*B], in_offsets int64[S_in*B+1]; out_* sized for S_out.
 * Used after the input all-to-all to turn [src_rank][feature][batch] into
 * [feature][src_rank][batch], and by KeyedJaggedTensor.permute().
 * Replaces: upstream fbgemm permute_2D_sparse_data inside the same input_dist
 *   (torchrec/train.py:98 -> DMP forward).
 * scratch: tdfo_scan_scratch_bytes(S_out*B).
 */
int tdfo_kjt_permute_2d(const int32_t* permute /* [S_out] */, int32_t S_in, int32_t S_out, int64_t B,
                        const int32_t* in_lengths, const int64_t* in_offsets,
                        const int64_t* in_indices, const float* in_weights,
                        int32_t* out_lengths, int64_t* out_offsets,
                        int64_t* out_indices, float* out_weights,
                        void* scratch, size_t scratch_bytes, void* stream);

/*
 * Input dist over NVLink peer memory (no all-to-all, no host round trip for split sizes).
 * Every feature segment (the B bags of one feature, or of one (bucket, feature) after
 * tdfo_kjt_bucketize_row_wise) is contiguous at the sender and, in the owner's
 * [feature][source rank][batch] layout, contiguous at the receiver.
 *
 * tdfo_kjt_push_lengths: slot j copies B lengths, lengths[src_off[j] .. +B) -> *(int32*)dst_addr[j].
 *   meta = device int64[2*nslots] = src_off (elements) | dst_addr (absolute, peer memory).
 * (barrier; the owner runs tdfo_kjt_complete_cumsum over what it received; barrier)
 * tdfo_kjt_push_ids: slot j sends segment seg[j] = values[offsets[seg*B] .. offsets[(seg+1)*B])
 *   to ids_base[j] + *(int64*)off_addr[j], where off_addr[j] points at the owner's offset word for
 *   this (feature, source rank) (read remotely), ids_base[j] at its id buffer.
 *   meta = device int64[3*nslots] = seg | off_addr | ids_base.  A segment that would run past
 *   `capacity` ids is dropped and TDFO_FLAG_RECV_OVERFLOW is set.
 * Replaces: KJTAllToAll (two all_to_all_single + permute_2D_sparse_data) in the input_dist of the
 *   sharded module DMP builds (torchrec/train.py:241-247, 98).
 */
int tdfo_kjt_push_lengths(const int32_t* lengths, int64_t B, int32_t nslots, const int64_t* meta, void* stream);
/*
 * The same transport for rows of D floats, used by the sharded EmbeddingCollection's backward: element i of
 * segment seg[j] is rows[positions ? positions[s0 + i] : s0 + i] (s0 = offsets[seg*B]) and lands at row
 * *(int64*)off_addr[j] + i of the float buffer at rows_base[j].  meta = seg | off_addr | rows_base.
 * Replaces: the gradient half of SequenceEmbeddingsAllToAll in the sharded module DMP builds for the
 *   reference's item table (torchrec/models.py:158-164, torchrec/train.py:241-247, 104).
 */
int tdfo_kjt_push_rows(const float* rows, const int64_t* positions, const int64_t* offsets, int64_t B, int32_t D,
                       int32_t nslots, const int64_t* meta, int64_t capacity, int32_t* err_flag, void* stream);
/*
 * out[permute[i]] = positions_in ? positions_in[i] : i, i < min(n, *n_dev) (n_dev: optional device word, for
 * buffers sized to an upper bound of an id count only the device knows): turns tdfo_kjt_bucketize_row_wise's
 * unbucketize_permute (original position -> bucketised position) into "original position of every
 * bucketised id", which travels with the ids so the owner knows where each looked-up row belongs.
 */
int tdfo_kjt_scatter_positions(const int64_t* permute, const int64_t* positions_in, int64_t n, const int64_t* n_dev,
                               int64_t* out, void* stream);
int tdfo_kjt_push_ids(const int64_t* values, const int64_t* offsets, int64_t B, int32_t nslots,
                      const int64_t* meta, int64_t capacity, int32_t* err_flag, void* stream);

/* ------------------------------------------------------------------ lookup */

/*
 * Embedding lookup over one storage group (all tables of one dim/dtype live in
 * one [total_rows, D] buffer).  KJT layout: bag g = f*B + b, ids of bag g are
 * indices[offsets[g] .. offsets[g+1]).
 *   feat_row_base[f]  first row of feature f's table inside `weights`
 *   feat_rows[f]      number of rows of that table (bounds check)
 *   feat_out_col[f]   first output column of feature f (elements)
 * Pooled (SUM/MEAN):  out is float[B, out_stride], bag (f,b) writes D floats at
 *   out[b*out_stride + feat_out_col[f]].  MEAN divides by the bag length, an
 *   empty bag yields zeros.  per_sample_weights (float[L]) may be NULL.
 * Sequence (NONE):    out is float[L, D], out[p,:] = row of indices[p]
 *   (out_stride, feat_out_col ignored).
 * Accumulation is fp32 in bag order (deterministic).
 * sort_keys / sort_vals (uint32[L], both or neither): when given, the kernel also
 *   writes the (row key, payload) pair of every id exactly as tdfo_linearize_indices
 *   would, while the ids are in registers.  Pointing them at the input_keys /
 *   input_vals slots of the backward scratch (tdfo_backward_scratch_layout) lets
 *   tdfo_embedding_backward_update(keys_in_scratch = 1) skip its linearise stage.
 * Replaces: `self.embed_collection(id_list_features)` torchrec/models.py:164
 *   (EmbeddingCollection, upstream fbgemm TBE forward "nobag"), and the
 *   EmbeddingBagCollection the BASELINE configs name (TBE pooled forward);
 *   jax-flax/models.py:78-88 and tensorflow2/models.py:48-58 are the P=1 case.
 */
int tdfo_embedding_forward(const void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                           const int64_t* indices, const int64_t* offsets,
                           const float* per_sample_weights,
                           const int64_t* feat_row_base, const int64_t* feat_rows,
                           const int64_t* feat_out_col,
                           int32_t F, int64_t B, int64_t L /* number of ids == offsets[F*B] */,
                           int32_t pooling_mode,
                           float* out, int64_t out_stride,
                           uint32_t* sort_keys, uint32_t* sort_vals,
                           int32_t* err_flag, void* stream);

/*
 * Pooled lookup with the table-wise output dist fused in: the B bags of every local feature
 * cover the GLOBAL batch (world * peer_batch samples, source-rank major), and the pooled row of
 * sample b is stored straight into the output buffer of its source rank b / peer_batch:
 *   peer_out[src] + (b % peer_batch) * out_stride + feat_out_col[f]
 * peer_out is a device array of `B / peer_batch` base addresses (as uint64), one per source rank,
 * pointing into NVLink peer memory (e.g. a torch symmetric-memory buffer); the caller puts a
 * barrier between this call and the readers.  Everything else as tdfo_embedding_forward.
 * `indices` is normally a receive buffer whose fill the host does not know yet: pass its capacity as L
 * (bags are cut there) and, as L_hint, how many ids to expect (0 = unknown); the hint only selects the
 * launch shape (short bags are streamed in blocks of 32 bags).
 * Replaces: TBE forward + PooledEmbeddingsAllToAll of the sharded module DMP builds
 *   (torchrec/train.py:241-247, 98).
 */
int tdfo_embedding_forward_peer(const void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                                const int64_t* indices, const int64_t* offsets,
                                const float* per_sample_weights,
                                const int64_t* feat_row_base, const int64_t* feat_rows,
                                const int64_t* feat_out_col,
                                int32_t F, int64_t B, int64_t L, int32_t pooling_mode,
                                const uint64_t* peer_out, int64_t peer_batch, int64_t out_stride,
                                uint32_t* sort_keys, uint32_t* sort_vals,
                                int32_t* err_flag, int64_t L_hint, void* stream);

/*
 * Sequence lookup (EmbeddingCollection, one row per id) with the output dist fused in, on the owner of the
 * rows.  `offsets` describes S = (local features x world) segments of B bags each, in the landing-zone order
 * [feature][source rank][batch]; `indices` is the landing zone, read at positions [offsets[0], offsets[S*B])
 * (never past L); seg_row_base / seg_rows give the table slice per SEGMENT (a feature's values repeated world
 * times).  The row of landed id p is stored at row positions[p] of the output buffer of source rank
 * (segment of p) % world: peer_out[src] + positions[p] * D floats, i.e. at the place the id had in the
 * sender's own KeyedJaggedTensor (positions travel with the ids: tdfo_kjt_push_ids of a position array).
 * sort_keys / sort_vals (optional): the backward's pairs, payload = p - offsets[0].
 * Replaces: TBE "nobag" forward + SequenceEmbeddingsAllToAll (+ the un-bucketise permute for row-wise
 *   tables) of the sharded EmbeddingCollection DMP builds for the reference's item table
 *   (torchrec/models.py:158-164, torchrec/train.py:241-247, 98).
 */
int tdfo_embedding_forward_seq_peer(const void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                                    const int64_t* indices, const int64_t* offsets,
                                    const int64_t* seg_row_base, const int64_t* seg_rows, int32_t S,
                                    int64_t B, int64_t L, const int64_t* positions,
                                    const uint64_t* peer_out, int32_t world,
                                    uint32_t* sort_keys, uint32_t* sort_vals, int32_t* err_flag, void* stream);

/* ------------------------------------------------------------------ backward */

/*
 * Stage 1 of the backward: linearise (feature, id) to a row of the storage
 * group.  keys[p] = feat_row_base[f] + indices[p] (uint32; total_rows < 2^32),
 * vals[p] = bag id g = f*B + b (uint32), or the position p itself when
 * pooling_mode is TDFO_POOL_NONE.  Out-of-range ids get key `total_rows`
 * (sorted last, skipped by the update) and set TDFO_FLAG_INDEX_OOB.
 * Replaces: the linearize_cache_indices prologue of the fbgemm TBE backward
 *   reached from loss.backward() at torchrec/train.py:104.
 */
int tdfo_linearize_indices(const int64_t* indices, const int64_t* offsets,
                           const int64_t* feat_row_base, const int64_t* feat_rows,
                           int32_t F, int64_t B, int64_t total_rows, int32_t pooling_mode,
                           uint32_t* keys, uint32_t* vals,
                           int32_t* err_flag, void* stream);

/*
 * Stage 2: stable LSD radix sort of n (key, val) uint32 pairs on the low
 * `key_bits` bits of the key (1..32).  Result lands in keys_out / vals_out;
 * keys_in / vals_in are clobbered (used as the ping-pong buffer).
 * n < 2^30.  Bit-exact with a stable CPU sort.
 * Replaces: cub::DeviceRadixSort::SortPairs in the upstream TBE backward
 *   (torchrec/train.py:104).
 * scratch: tdfo_radix_sort_scratch_bytes(n, key_bits).
 */
size_t tdfo_radix_sort_scratch_bytes(int64_t n, int32_t key_bits);
int tdfo_radix_sort_pairs(uint32_t* keys_in, uint32_t* vals_in,
                          uint32_t* keys_out, uint32_t* vals_out,
                          int64_t n, int32_t key_bits,
                          int32_t* err_flag,
                          void* scratch, size_t scratch_bytes, void* stream);

/*
 * Stage 3: run-length encode sorted keys.
 *   unique_keys uint32[n]     (first *num_unique entries valid)
 *   seg_starts  uint32[n + 1] (seg_starts[u] = first sorted position of
 *                              unique row u; seg_starts[*num_unique] = n)
 *   first_vals  uint32[n]     optional (needs sorted_vals): payload of the first occurrence
 *                             of unique row u, i.e. sorted_vals[seg_starts[u]] -- lets the
 *                             update kernel find a row's first gradient without a dependent load
 *   num_unique  uint32[1]     device scalar
 * Replaces: cub::DeviceRunLengthEncode in the upstream TBE backward.
 * scratch: tdfo_scan_scratch_bytes(n).
 */
int tdfo_unique_segments(const uint32_t* sorted_keys, const uint32_t* sorted_vals, int64_t n,
                         uint32_t* unique_keys, uint32_t* seg_starts, uint32_t* first_vals,
                         uint32_t* num_unique,
                         int32_t* err_flag,
                         void* scratch, size_t scratch_bytes, void* stream);

/* Hyper-parameters of the fused optimizer (HOST struct, passed by pointer). */
typedef struct tdfo_optimizer_args {
    int32_t optimizer;      /* TDFO_OPT_* */
    float   learning_rate;
    float   eps;
    float   weight_decay;   /* L2: g += weight_decay * w  (ADAM: decoupled into the step, as fbgemm) */
    float   beta1, beta2;   /* ADAM */
    int64_t step;           /* ADAM bias correction, 1-based iteration count */
} tdfo_optimizer_args;

/*
 * Stage 4: per unique row, reduce the gradient rows of all its occurrences in
 * sorted (= original) order in fp32 and apply the optimizer in place.  No dense
 * [rows, D] gradient is ever formed.
 *   grad_out   float[B, grad_stride]; bag (f,b) reads D floats at
 *              grad_out[b*grad_stride + feat_out_col[f]]           (pooled)
 *              float[L, D], occurrence p reads grad_out[p,:]        (POOL_NONE;
 *              sorted_vals then hold positions p, see tdfo_embedding_backward_update)
 *   offsets    int64[F*B+1] (MEAN scaling; may be NULL for SUM / NONE)
 *   state0     ROWWISE_ADAGRAD: float[total_rows] sum of mean squared grads
 *              ADAM: float[total_rows, D] first moment;  SGD: NULL
 *   state1     ADAM: float[total_rows, D] second moment; else NULL
 * worklist (optional, uint32[worklist_entries], 16-byte aligned, entries >=
 *   tdfo_hot_row_worklist_entries(L)): rows with more than 64 occurrences ("hot" rows of a skewed id
 *   distribution) are skipped by the main kernel; a small kernel queues them here as (row, chunk) work
 *   items and a third kernel puts a whole thread block on each chunk of at most max(2048, L/2048)
 *   occurrences, so the hottest row is reduced by many blocks at once.  Shares of a chunk, chunks of a
 *   row and their partial sums are combined in a fixed order: results are reproducible; the summation
 *   order differs from the strictly sequential one.  Without a worklist every row is reduced by one
 *   lane group.
 * ROWWISE_ADAGRAD: h += mean_d(g_d^2); w -= lr / (sqrt(h) + eps) * g.
 * SGD: w -= lr * g.  bf16 tables: fp32 math, round-to-nearest-even on store.
 * Replaces: upstream split_embedding_backward_codegen_{sgd,rowwise_adagrad,adam}
 *   reached from loss.backward() (torchrec/train.py:104); optimizer.step()
 *   (torchrec/train.py:105) is then a no-op for these rows.
 */
int tdfo_sparse_update(void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                       float* state0, float* state1,
                       const float* grad_out, int64_t grad_stride,
                       const uint32_t* sorted_vals,
                       const uint32_t* unique_keys, const uint32_t* seg_starts,
                       const uint32_t* first_vals, const uint32_t* num_unique,
                       const int64_t* offsets, const int64_t* feat_out_col,
                       int32_t F, int64_t B, int64_t L, int32_t pooling_mode,
                       const tdfo_optimizer_args* opt /* HOST */,
                       uint32_t* worklist, int64_t worklist_entries,
                       void* stream);

/*
 * The whole backward in one call: linearise -> sort -> unique -> update.
 * Same arguments as the stages above.  On return (stream order) scratch holds,
 * at the offsets reported by tdfo_backward_scratch_layout, the sorted keys,
 * sorted vals, unique keys, segment starts and the unique count, so a caller
 * (or a parity test) can read them.
 * keys_in_scratch == 1: the caller already had tdfo_embedding_forward write the
 * linearised pairs for these same indices / offsets into the input_keys /
 * input_vals slots of `scratch`; the linearise stage is skipped.
 * keys_in_scratch == TDFO_SCRATCH_PREPARED: tdfo_embedding_backward_prepare already ran on this
 * scratch; only the update kernel runs (indices / feat_row_base / feat_rows may be NULL).
 * per_sample_weights (float[L], pooled modes only, may be NULL): the gradient row
 * of occurrence p is scaled by per_sample_weights[p] (and by 1/len for MEAN), the
 * backward of the weighted pooled lookup.  The weights themselves
 * receive no gradient.  With weights the sort payload is the id's position, the
 * scratch must be tdfo_backward_scratch_bytes_weighted(L, total_rows) bytes, the
 * sorted_vals read back are positions, and keys_in_scratch must be 0.
 * scratch: tdfo_backward_scratch_bytes(L, total_rows) otherwise.
 */
typedef struct tdfo_backward_layout {
    size_t input_keys, input_vals;   /* where linearised pairs go in (clobbered by the sort) */
    size_t sorted_keys, sorted_vals, unique_keys, seg_starts, first_vals, num_unique; /* byte offsets */
    size_t worklist, worklist_entries;   /* hot-row worklist for tdfo_sparse_update: byte offset, uint32 entries */
    size_t total_bytes;
    int32_t key_bits;
} tdfo_backward_layout;

size_t tdfo_backward_scratch_bytes(int64_t L, int64_t total_rows);
size_t tdfo_backward_scratch_bytes_weighted(int64_t L, int64_t total_rows);

/*
 * The id-only half of the backward (linearise -> sort -> unique) on its own, so a caller can run
 * it on `stream` while the gradient is still being produced or exchanged (the sharded path
 * overlaps it with the NVLink gradient push): it needs the ids, not grad_out.  Follow with
 * tdfo_embedding_backward_update(..., keys_in_scratch = TDFO_SCRATCH_PREPARED, same scratch, same
 * indices / offsets / L / total_rows / pooling_mode), which then runs only the update kernel.
 * weighted != 0 selects the weighted layout (scratch of tdfo_backward_scratch_bytes_weighted).
 * keys_in_scratch: 0, or 1 if tdfo_embedding_forward already wrote the pairs (see above).
 */
#define TDFO_SCRATCH_PREPARED 2
/*
 * The same from pairs that already exist: `keys` / `vals` (uint32[L], read only, anywhere in device memory)
 * are the (row key, payload) pairs tdfo_embedding_forward / _peer wrote through its sort_keys / sort_vals
 * arguments.  Used by the sharded path, where the forward runs before the host knows how many ids the rank
 * received: it writes the pairs into a capacity-sized buffer, and the backward sorts the first L of them.
 */
int tdfo_embedding_backward_prepare_pairs(const uint32_t* keys, const uint32_t* vals, int64_t L, int64_t total_rows,
                                          int32_t* err_flag, void* scratch, size_t scratch_bytes, void* stream);
int tdfo_embedding_backward_prepare(const int64_t* indices, const int64_t* offsets,
                                    const int64_t* feat_row_base, const int64_t* feat_rows,
                                    int32_t F, int64_t B, int64_t L, int64_t total_rows,
                                    int32_t pooling_mode, int32_t weighted, int32_t keys_in_scratch,
                                    int32_t* err_flag,
                                    void* scratch, size_t scratch_bytes, void* stream);
int tdfo_backward_scratch_layout(int64_t L, int64_t total_rows, tdfo_backward_layout* layout /* HOST */);
/* uint32 entries the hot-row worklist of tdfo_sparse_update needs for L ids (pure host arithmetic). */
size_t tdfo_hot_row_worklist_entries(int64_t L);
int tdfo_embedding_backward_update(void* weights, int32_t weights_dtype, int64_t total_rows, int32_t D,
                                   float* state0, float* state1,
                                   const float* grad_out, int64_t grad_stride,
                                   const int64_t* indices, const int64_t* offsets,
                                   const float* per_sample_weights,
                                   const int64_t* feat_row_base, const int64_t* feat_rows,
                                   const int64_t* feat_out_col,
                                   int32_t F, int64_t B, int64_t L, int32_t pooling_mode,
                                   const tdfo_optimizer_args* opt /* HOST */,
                                   int32_t keys_in_scratch,
                                   int32_t* err_flag,
                                   void* scratch, size_t scratch_bytes, void* stream);

/*
 * Output dist helper: move pooled rows between the per-rank exchange layout
 * ([rank][batch][features of that rank * D]) and the KeyedTensor layout ([batch][all features * D]).
 * Slot j copies B rows of D floats: src[src_off[j] + b*src_stride[j] ..] -> dst[dst_off[j] + b*dst_stride[j] ..].
 * meta = device int64[4*nslots] = src_off | src_stride | dst_off | dst_stride (elements; multiples of 4).
 * With dst == NULL the dst_off entries are absolute addresses divided by 4, so different slots can
 * target different buffers, including NVLink peer memory (gradient rows pushed to the table owners).
 * Replaces: the column permute after PooledEmbeddingsAllToAll in the sharded module DMP builds
 *   (torchrec/train.py:241-247), and its gradient.
 */
int tdfo_copy_feature_blocks(const float* src, float* dst, int64_t B, int32_t D, int32_t nslots,
                             const int64_t* meta, void* stream);

/*
 * tdfo_copy_feature_blocks with absolute destinations that are NVLS MULTICAST addresses (dst_off = address / 4
 * of a multicast mapping of a buffer every rank holds): each 16-byte chunk is written with one multimem store
 * and the NVSwitch delivers it to every rank's copy.  Used for the row-wise backward, where every rank needs
 * the gradient rows of every sample: an all-gather that costs each GPU one copy of its own rows on its link
 * instead of world - 1.  Stores only; nothing is loaded through the multicast mapping.
 * Replaces: the all-gather behind the backward of PooledEmbeddingsReduceScatter (torchrec/train.py:241-247, 104).
 */
int tdfo_copy_feature_blocks_multicast(const float* src, int64_t B, int32_t D, int32_t nslots,
                                       const int64_t* meta, void* stream);

/*
 * Row-wise output dist, second half: every rank s holds the partial pooled sums of its row block for
 * all samples; the owners' lookups (tdfo_embedding_forward_peer) store them into slot s of the sample
 * owner's buffer `parts` = float[W][B][nslots * D], and this kernel adds the W slots in rank order
 * (s = 0, 1, ..., W-1: a reduce-scatter with a fixed, reproducible summation order) into the
 * KeyedTensor layout: out[b * out_stride + out_col[k] + d] = sum_s parts[s][b][k * D + d].
 * out_col = device int64[nslots] (multiples of 4); `part_stride` = floats between two source slots.
 * Replaces: PooledEmbeddingsReduceScatter of the row-wise sharded module DMP builds
 *   (torchrec/train.py:241-247; the reference's only explicit row partition is
 *   tensorflow2/train_ps.py:55-61).
 */
int tdfo_reduce_partial_sums(const float* parts, int32_t W, int64_t part_stride, int64_t B, int32_t nslots,
                             int32_t D, const int64_t* out_col, float* out, int64_t out_stride, void* stream);

/* ------------------------------------------------------------------ jagged */

/*
 * values float[L, D], offsets int64[B+1] -> dense float[B, max_len, D];
 * zero padded, rows beyond max_len dropped.
 * Replaces: torch.ops.fbgemm.jagged_2d_to_dense at torchrec/models.py:168-172.
 */
int tdfo_jagged_2d_to_dense(const float* values, const int64_t* offsets,
                            int64_t B, int32_t D, int64_t max_len,
                            float* dense, void* stream);

/*
 * Gradient of the above: dense float[B, max_len, D] -> values float[L, D];
 * rows that were dropped (position >= max_len) get zeros.
 * Replaces: autograd of jagged_2d_to_dense (upstream fbgemm dense_to_jagged),
 *   reached from loss.backward() at torchrec/train.py:104.
 */
int tdfo_dense_to_jagged(const float* dense, const int64_t* offsets,
                         int64_t B, int32_t D, int64_t max_len, int64_t L,
                         float* values, void* stream);

/*
 * values int64[L], offsets int64[B+1] -> dense int64[B, desired_length] padded
 * with padding_value.
 * Replaces: JaggedTensor.to_padded_dense at torchrec/models.py:210-212.
 */
int tdfo_jagged_1d_to_padded_dense(const int64_t* values, const int64_t* offsets,
                                   int64_t B, int64_t desired_length, int64_t padding_value,
                                   int64_t* dense, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* TDFO_B200_H */
