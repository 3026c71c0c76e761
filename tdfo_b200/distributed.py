"""Sharded embedding path: table-wise / row-wise / mixed sharding of an EmbeddingBagCollection
over the GPUs of one box, one process per GPU.

This is the part of the reference that `DistributedModelParallel(module, device, sharders=[...])`
(torchrec/train.py:241-247) delegates to torchrec's sharded modules (SURVEY.md §8 a10, a11, a13,
§8e).  Per step and rank:

  input dist   TW features are permuted into owner order, RW features are bucketised by row
               block (`tdfo_kjt_bucketize_row_wise`); lengths, then ids, are exchanged with two
               all-to-alls; the received [src][feature][batch] layout is permuted to
               [feature][src][batch] (`tdfo_kjt_permute_2d`) so each local feature sees the
               global batch W*B.
  lookup       `tdfo_embedding_forward` over the local storage (whole TW tables, row blocks of RW
               tables).
  output dist  TW: all-to-all of pooled rows back to the rank that owns the samples;
               RW: partial sums are exchanged and added in source-rank order (a reduce-scatter
               with a fixed, reproducible order).
  backward     the reverse exchanges (all-to-all of gradients for TW, all-gather for RW), then the
               fused sort/unique/update kernels on the local storage.

torch.distributed is the plumbing (NCCL over NVLink on the GPU box, gloo in the CPU tests).  The
compute steps go through a small `KernelOps` table whose only implementation in this package is
the CUDA one; the CPU tests in `tests/` inject their own checker-backed table to exercise the
routing logic without a GPU.
"""
from __future__ import annotations

import contextlib
import enum
import os
import zlib
from dataclasses import dataclass
from types import SimpleNamespace
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed as dist
import torch.nn as nn

from . import _lib
from .modules import EmbeddingBagConfig, PoolingType
from .optim import EmbOptimType, FusedOptimizer
from .sparse import KeyedJaggedTensor, KeyedTensor


class ShardingType(enum.Enum):
    TABLE_WISE = "table_wise"
    ROW_WISE = "row_wise"
    COLUMN_WISE = "column_wise"


@dataclass
class ParameterSharding:
    sharding_type: ShardingType
    rank: Optional[int] = None          # owner for TABLE_WISE
    ranks: Optional[List[int]] = None   # owners of the column shards for COLUMN_WISE (shard j -> ranks[j])


def table_wise(rank: int) -> ParameterSharding:
    return ParameterSharding(ShardingType.TABLE_WISE, rank)


def row_wise() -> ParameterSharding:
    return ParameterSharding(ShardingType.ROW_WISE)


def column_wise(ranks: Sequence[int]) -> ParameterSharding:
    """Split the embedding dimension into len(ranks) equal column shards; shard j lives on ranks[j].
    Every shard sees all ids of the table and returns its own columns of the pooled row, so unlike
    row-wise there are no partial sums to reduce: it balances a table count that does not divide
    the world size at table-wise cost."""
    return ParameterSharding(ShardingType.COLUMN_WISE, None, [int(r) for r in ranks])


def balanced_plan(tables: Sequence[EmbeddingBagConfig], world: int, min_shard_bytes: int = 64
                  ) -> Dict[str, ParameterSharding]:
    """Equal tables: floor(T / world) * world of them table-wise round-robin, the remainder R split
    column-wise into world / gcd(R, world) shards each so every rank holds the same number of columns.
    Falls back to table-wise for the remainder when a shard's row would drop below `min_shard_bytes`."""
    import math
    T = len(tables)
    n_tw = (T // world) * world
    plan = {t.name: table_wise(i % world) for i, t in enumerate(tables[:n_tw])}
    rest = list(tables[n_tw:])
    if rest:
        k = world // math.gcd(len(rest), world)
        esz = 4 if tables[0].data_type.value == "FP32" else 2
        ok = all(t.embedding_dim % k == 0 and (t.embedding_dim // k) * esz >= min_shard_bytes
                 and ((t.embedding_dim // k) * esz) % 16 == 0 for t in rest)
        for i, t in enumerate(rest):
            plan[t.name] = column_wise([(i * k + j) % world for j in range(k)]) if ok and k > 1 else table_wise(i % world)
    return plan


def balanced_table_wise_plan(tables: Sequence[EmbeddingBagConfig], world: int) -> Dict[str, ParameterSharding]:
    """Greedy size-balanced table -> rank assignment (largest table first onto the lightest rank)."""
    load = [0] * world
    plan: Dict[str, ParameterSharding] = {}
    for t in sorted(tables, key=lambda t: -t.num_embeddings * t.embedding_dim):
        r = min(range(world), key=lambda i: (load[i], i))
        plan[t.name] = table_wise(r)
        load[r] += t.num_embeddings * t.embedding_dim
    return plan


class CudaKernelOps:
    """The compute steps of the sharded path, on the sm_100a kernels (the only shipped table)."""

    def __init__(self):
        from . import ops
        self._ops = ops

    def make_group(self, weights, feat_row_base, feat_rows, feat_out_col, out_stride, state0, state1, err_flag):
        return self._ops.TableGroup(weights=weights, feat_row_base=feat_row_base, feat_rows=feat_rows,
                                    feat_out_col=feat_out_col, out_stride=out_stride, state0=state0,
                                    state1=state1, err_flag=err_flag)

    def complete_cumsum(self, lengths):
        return self._ops.complete_cumsum(lengths)

    def bucketize_row_wise(self, lengths, offsets, indices, block_sizes, F, B, world, err_flag):
        nl, no, ni, _, _ = self._ops.bucketize_row_wise(lengths, offsets, indices, None, block_sizes, F, B, world,
                                                        False, err_flag)
        return nl, no, ni

    def bucketize_row_wise_with_permute(self, lengths, offsets, indices, block_sizes, F, B, world, err_flag):
        nl, no, ni, _, perm = self._ops.bucketize_row_wise(lengths, offsets, indices, None, block_sizes, F, B, world,
                                                           True, err_flag)
        return nl, no, ni, perm

    def permute_2d(self, permute, B, lengths, offsets, indices, out_total):
        ol, oo, oi, _ = self._ops.permute_2d(permute, B, lengths, offsets, indices, None, out_total=out_total)
        return ol, oo, oi

    def embedding_forward(self, group, indices, offsets, B, pooling_mode, out, backward_scratch=None):
        return self._ops.embedding_forward(group, indices, offsets, B, pooling_mode, out=out,
                                           backward_scratch=backward_scratch)

    def embedding_backward_update(self, group, grad_out, indices, offsets, B, pooling_mode, opt, scratch=None,
                                  keys_in_scratch=False):
        self._ops.embedding_backward_update(group, grad_out, indices, offsets, B, pooling_mode, opt, scratch=scratch,
                                            keys_in_scratch=keys_in_scratch)

    def make_peer_exchange(self, *args, **kwargs):
        return _PeerExchange(*args, **kwargs)

    def embedding_backward_prepare(self, group, indices, offsets, B, pooling_mode, scratch, keys_in_scratch=False):
        self._ops.embedding_backward_prepare(group, indices, offsets, B, pooling_mode, scratch,
                                             keys_in_scratch=keys_in_scratch)

    def backward_scratch_bytes(self, L, total_rows):
        return self._ops.backward_layout(L, total_rows).total_bytes

    def embedding_forward_peer(self, group, indices, offsets, B, pooling_mode, peer_out, peer_batch,
                               backward_scratch=None, sort_pairs=None, expected_ids=0):
        self._ops.embedding_forward_peer(group, indices, offsets, B, pooling_mode, peer_out, peer_batch,
                                         backward_scratch=backward_scratch, sort_pairs=sort_pairs,
                                         expected_ids=expected_ids)

    def embedding_backward_prepare_pairs(self, group, keys, vals, L, scratch):
        self._ops.embedding_backward_prepare_pairs(group, keys, vals, L, scratch)

    def embedding_forward_seq_peer(self, group, indices, offsets, B, positions, peer_out, world, sort_pairs=None):
        self._ops.embedding_forward_seq_peer(group, indices, offsets, B, positions, peer_out, world, sort_pairs=sort_pairs)

    def push_rows(self, rows, positions, offsets, B, meta, capacity, err_flag):
        self._ops.push_rows(rows, positions, offsets, B, meta, capacity, err_flag)

    def scatter_positions(self, permute, positions_in, n, n_dev=None):
        return self._ops.scatter_positions(permute, positions_in, n, n_dev)

    def copy_feature_blocks(self, src, dst, B, D, meta):
        self._ops.copy_feature_blocks(src, dst, B, D, meta)

    def push_lengths(self, lengths, B, meta):
        self._ops.push_lengths(lengths, B, meta)

    def push_ids(self, values, offsets, B, meta, capacity, err_flag):
        self._ops.push_ids(values, offsets, B, meta, capacity, err_flag)

    def complete_cumsum_into(self, lengths, out):
        return self._ops.complete_cumsum(lengths, out=out)

    def reduce_partial_sums(self, parts, W, B, nslots, D, out_col, out, out_stride):
        self._ops.reduce_partial_sums(parts, W, B, nslots, D, out_col, out, out_stride)

    def copy_feature_blocks_multicast(self, src, B, D, meta):
        self._ops.copy_feature_blocks_multicast(src, B, D, meta)


class _Shard:
    """Local storage of one sharding flavour (all TW tables this rank owns, or all RW row blocks)."""

    def __init__(self):
        self.tables: List[EmbeddingBagConfig] = []
        self.local_rows: List[int] = []
        self.row0: List[int] = []
        self.features: List[int] = []        # global feature indices served, in local order
        self.feat_table: List[int] = []
        self.total_rows = 0

    def add(self, table: EmbeddingBagConfig, rows: int, feature_ids: List[int]):
        self.row0.append(self.total_rows)
        self.tables.append(table)
        self.local_rows.append(rows)
        self.total_rows += rows
        for f in feature_ids:
            self.features.append(f)
            self.feat_table.append(len(self.tables) - 1)


class _PeerExchange:
    """Symmetric (NVLink peer-mapped) buffers of one sharded module for a given local batch size:
    two rotating output buffers [B, F_total*D] that the owners of table-wise features store pooled
    rows into, and one gradient landing buffer per rank that the sample owners push gradient rows
    into.  Barriers are the symmetric-memory signal-pad barriers, enqueued on the current stream."""

    def __init__(self, pg, device, B: int, F_total: int, D: int, n_tw: List[int], n_rw: int, ids_capacity: int,
                 n_cw: Optional[List[int]] = None, Dc: int = 0, seq_cap: int = 0):
        import torch.distributed._symmetric_memory as symm
        self.seq_cap = int(seq_cap)
        W = len(n_tw)
        group = pg if pg is not None else dist.group.WORLD
        self.B = B
        # input dist landing zones, already in the [feature][source rank][batch] layout
        n_cw = n_cw if n_cw is not None else [0] * W
        fl_max = max(a + b for a, b in zip(n_tw, n_cw)) + n_rw
        self.capacity = int(ids_capacity)
        self.len_x = symm.empty(max(fl_max * W * B, 4), dtype=torch.int32, device=device)
        self.off_x = symm.empty(fl_max * W * B + 4, dtype=torch.int64, device=device)
        self.ids_x = symm.empty(max(self.capacity, 4), dtype=torch.int64, device=device)
        self.len_ptrs = list(symm.rendezvous(self.len_x, group).buffer_ptrs)
        self.off_ptrs = list(symm.rendezvous(self.off_x, group).buffer_ptrs)
        self.ids_ptrs = list(symm.rendezvous(self.ids_x, group).buffer_ptrs)
        self.counts_host = torch.zeros(3, dtype=torch.int64).pin_memory()     # ids: end of TW, end of CW, total
        self.counts_event = torch.cuda.Event()
        self.pos_x, self.pos_ptrs = None, []
        if self.seq_cap:                           # sequence module: every landed id comes with its place at the sender
            self.pos_x = symm.empty(max(self.capacity, 4), dtype=torch.int64, device=device)
            self.pos_ptrs = list(symm.rendezvous(self.pos_x, group).buffer_ptrs)
        # Landing zones of the input dist.  Zone 0 is this object's own buffers above; a second one is added
        # only when input dists are prefetched (ensure_second_zone).
        self._zone_args = (symm, group, device, fl_max * W * B)
        self.zones = [self]
        self.out = [symm.empty(B * F_total * D, dtype=torch.float32, device=device) for _ in range(2)]
        self.out_hdl = [symm.rendezvous(t_, group) for t_ in self.out]
        self.out_ptrs = [torch.tensor(list(h.buffer_ptrs), dtype=torch.int64, device=device) for h in self.out_hdl]
        self.grad = symm.empty(max(W * B * max(n_tw) * D, 4), dtype=torch.float32, device=device)
        self.grad_hdl = symm.rendezvous(self.grad, group)
        self.grad_ptrs = list(self.grad_hdl.buffer_ptrs)
        self.grad_cw, self.grad_cw_ptrs = None, []
        if max(n_cw) > 0:                          # landing buffer for the gradient columns of column shards
            self.grad_cw = symm.empty(max(W * B * max(n_cw) * Dc, 4), dtype=torch.float32, device=device)
            self.grad_cw_ptrs = list(symm.rendezvous(self.grad_cw, group).buffer_ptrs)
        # row-wise tables: slot s of `rw_part` receives rank s's partial pooled sums for MY samples
        # ([W][B][n_rw*D]); `grad_rw` receives every rank's gradient rows of the row-wise features ([W*B][n_rw*D])
        self.rw_part, self.rw_part_ptrs, self.grad_rw, self.grad_rw_ptrs, self.grad_rw_mc = None, [], None, [], 0
        if n_rw > 0:
            self.rw_part = symm.empty(W * B * n_rw * D, dtype=torch.float32, device=device)
            self.rw_part_ptrs = list(symm.rendezvous(self.rw_part, group).buffer_ptrs)
            self.grad_rw = symm.empty(W * B * n_rw * D, dtype=torch.float32, device=device)
            hdl = symm.rendezvous(self.grad_rw, group)
            self.grad_rw_ptrs = list(hdl.buffer_ptrs)
            # NVLS multicast address of the same buffer (0 when the fabric has none): one store reaches every rank
            self.grad_rw_mc = 0
            if os.environ.get("TDFO_NO_MULTICAST") != "1":
                try:
                    self.grad_rw_mc = int(hdl.multicast_ptr or 0)
                except Exception:
                    self.grad_rw_mc = 0
        # sequence module: rotating row buffers [seq_cap, D] the owners store looked-up rows into (at the ids' own
        # positions), and the landing buffer for the gradient rows of the ids this rank owns
        self.seq_out, self.seq_out_ptrs, self.grad_seq, self.grad_seq_ptrs = [], [], None, []
        if self.seq_cap:
            self.seq_out = [symm.empty(max(self.seq_cap * D, 4), dtype=torch.float32, device=device) for _ in range(2)]
            self.seq_out_ptrs = [torch.tensor(list(symm.rendezvous(t_, group).buffer_ptrs), dtype=torch.int64, device=device)
                                 for t_ in self.seq_out]
            self.grad_seq = symm.empty(max(self.capacity * D, 4), dtype=torch.float32, device=device)
            self.grad_seq_ptrs = list(symm.rendezvous(self.grad_seq, group).buffer_ptrs)
        self.step = 0

    def _make_zone(self):
        """A further landing zone: lengths, offsets and ids in the owner's layout, plus the id counts the
        backward needs on the host (same fields as zone 0 = this object)."""
        symm, group, device, n_bags = self._zone_args
        z = SimpleNamespace()
        z.len_x = symm.empty(max(n_bags, 4), dtype=torch.int32, device=device)
        z.off_x = symm.empty(n_bags + 4, dtype=torch.int64, device=device)
        z.ids_x = symm.empty(max(self.capacity, 4), dtype=torch.int64, device=device)
        z.len_ptrs = list(symm.rendezvous(z.len_x, group).buffer_ptrs)
        z.off_ptrs = list(symm.rendezvous(z.off_x, group).buffer_ptrs)
        z.ids_ptrs = list(symm.rendezvous(z.ids_x, group).buffer_ptrs)
        z.counts_host = torch.zeros(3, dtype=torch.int64).pin_memory()     # ids: end of TW, end of CW, total
        z.counts_event = torch.cuda.Event()
        z.pos_x, z.pos_ptrs = None, []
        if self.seq_cap:
            z.pos_x = symm.empty(max(self.capacity, 4), dtype=torch.int64, device=device)
            z.pos_ptrs = list(symm.rendezvous(z.pos_x, group).buffer_ptrs)
        return z

    def ensure_second_zone(self):
        """COLLECTIVE, first call only: double-buffer the landing zones so that the input dist of the next
        batch can run while this batch's lookup / backward still read theirs."""
        if len(self.zones) < 2:
            self.zones.append(self._make_zone())

    def barrier(self, channel: int = 0):
        """Channel 0: the training stream; channel 1: the stream prefetched input dists run on (two streams
        must not share a signal pad)."""
        self.grad_hdl.barrier(channel=channel)


class _ShardedLookupFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, module, values, lengths, B, dist_ctx=None):
        out, saved = module._forward_impl(values, lengths, B, dist_ctx)
        ctx.module, ctx.saved, ctx.B = module, saved, B
        return out

    @staticmethod
    def backward(ctx, grad_out):
        ctx.module._backward_impl(grad_out.contiguous(), ctx.saved, ctx.B)
        return None, None, None, None, None, None


class ShardedEmbeddingBagCollection(nn.Module):
    """`EmbeddingBagCollection` whose tables are spread over the ranks of `process_group`.

    `plan` maps table name -> `table_wise(rank)` / `row_wise()`.  Every rank feeds its own local
    batch (same batch size B on every rank) and gets back the pooled embeddings of ITS samples,
    exactly what the unsharded module would return for them.
    All tables must share embedding_dim, data_type and pooling in this version.
    """

    def __init__(self, tables: List[EmbeddingBagConfig], plan: Dict[str, ParameterSharding],
                 device: torch.device, process_group: Optional[dist.ProcessGroup] = None,
                 kernel_ops=None, init_seed: int = 0, init_on_device: bool = False,
                 peer_memory: Optional[bool] = None, max_ids_per_rank: Optional[int] = None):
        super().__init__()
        self._pg = process_group
        self._world = dist.get_world_size(process_group)
        self._rank = dist.get_rank(process_group)
        self._ops = kernel_ops if kernel_ops is not None else CudaKernelOps()
        self._configs = list(tables)
        W, r = self._world, self._rank
        dims = {t.embedding_dim for t in tables}
        kinds = {(t.data_type, getattr(t, "pooling", PoolingType.SUM)) for t in tables}
        if len(dims) != 1 or len(kinds) != 1:
            raise ValueError("sharded tables must share embedding_dim, data_type and pooling in this version")
        self._D = tables[0].embedding_dim
        self._pooling = _lib.POOL_MEAN if getattr(tables[0], "pooling", PoolingType.SUM) == PoolingType.MEAN else _lib.POOL_SUM
        if self._pooling == _lib.POOL_MEAN and any(plan[t.name].sharding_type == ShardingType.ROW_WISE for t in tables):
            raise ValueError("row-wise sharding supports SUM pooling only (a bag's length is split across ranks)")
        dtype = torch.float32 if tables[0].data_type.value == "FP32" else torch.bfloat16
        device = torch.device(device)

        self._feature_names: List[str] = [f for t in tables for f in t.feature_names]
        F_total = len(self._feature_names)
        feat_ids: Dict[str, List[int]] = {}
        k = 0
        for t in tables:
            feat_ids[t.name] = list(range(k, k + len(t.feature_names)))
            k += len(t.feature_names)

        # who serves which feature
        self._tw_feats: List[List[int]] = [[] for _ in range(W)]   # per owner rank, module order
        self._cw_slots: List[List[tuple]] = [[] for _ in range(W)] # per owner rank: (feature, column shard)
        self._rw_feats: List[int] = []
        self._rw_block: List[int] = []                             # block size per RW feature
        self._tw, self._cw, self._rw = _Shard(), _Shard(), _Shard()
        self._cw_local_shard: List[int] = []                       # column shard index of each local CW table
        self._Dc = 0                                               # columns per column-wise shard
        for t in tables:
            ps = plan[t.name]
            if ps.sharding_type == ShardingType.COLUMN_WISE:
                k = len(ps.ranks or [])
                if k < 1 or len(set(ps.ranks)) != k or not all(0 <= x < W for x in ps.ranks):
                    raise ValueError(f"table {t.name}: column-wise owners {ps.ranks} must be distinct ranks in [0, {W})")
                if self._D % k:
                    raise ValueError(f"table {t.name}: embedding_dim {self._D} does not split into {k} column shards")
                dc = self._D // k
                if self._Dc not in (0, dc):
                    raise ValueError("all column-wise tables must use the same number of shards in this version")
                if (dc * (4 if dtype == torch.float32 else 2)) % 16:
                    raise ValueError(f"table {t.name}: a column shard of {dc} elements is not a multiple of 16 bytes")
                self._Dc = dc
                for j, owner in enumerate(ps.ranks):
                    self._cw_slots[owner] += [(f, j) for f in feat_ids[t.name]]
                    if owner == r:
                        self._cw.add(t, t.num_embeddings, feat_ids[t.name])
                        self._cw_local_shard.append(j)
            elif ps.sharding_type == ShardingType.TABLE_WISE:
                if ps.rank is None or not 0 <= ps.rank < W:
                    raise ValueError(f"table {t.name}: table-wise owner rank {ps.rank} outside [0, {W})")
                self._tw_feats[ps.rank] += feat_ids[t.name]
                if ps.rank == r:
                    self._tw.add(t, t.num_embeddings, feat_ids[t.name])
            else:
                bs = (t.num_embeddings + W - 1) // W
                self._rw_feats += feat_ids[t.name]
                self._rw_block += [bs] * len(t.feature_names)
                self._rw.add(t, max(0, min(bs, t.num_embeddings - r * bs)), feat_ids[t.name])
        self._F_total = F_total
        self._n_tw = [len(x) for x in self._tw_feats]
        self._n_cw = [len(x) for x in self._cw_slots]
        self._n_rw = len(self._rw_feats)
        self._F_local = self._n_tw[r] + self._n_cw[r] + self._n_rw      # local order: TW, CW, RW features

        # storage: identical initial values to an unsharded module built with the same seed would
        # need a shared generator; each table is drawn from its own seeded stream instead.
        for name, sh in (("tw", self._tw), ("cw", self._cw), ("rw", self._rw)):
            width = (self._Dc or 4) if name == "cw" else self._D
            w = torch.empty((max(sh.total_rows, 1), width), dtype=dtype, device=device)
            for ti, (t, r0, rows) in enumerate(zip(sh.tables, sh.row0, sh.local_rows)):
                if init_on_device:          # large tables: draw the local rows in place (not rank-consistent)
                    w[r0:r0 + rows].uniform_(t.get_weight_init_min(), t.get_weight_init_max())
                    continue
                full = self.initial_table(t, init_seed)
                if name == "cw":
                    c0 = self._cw_local_shard[ti] * self._Dc
                    w[r0:r0 + rows].copy_(full[:, c0:c0 + self._Dc].to(dtype))
                    continue
                lo = 0 if sh is self._tw else r * ((t.num_embeddings + W - 1) // W)
                w[r0:r0 + rows].copy_(full[lo:lo + rows].to(dtype))
            p = nn.Parameter(w, requires_grad=False)
            self.register_parameter(f"_{name}_weight", p)
            n_f = len(sh.features)
            self.register_buffer(f"_{name}_row_base", torch.tensor([sh.row0[ti] for ti in sh.feat_table] or [0],
                                                                    dtype=torch.int64, device=device)[:max(n_f, 0)],
                                 persistent=False)
            self.register_buffer(f"_{name}_rows", torch.tensor([sh.local_rows[ti] for ti in sh.feat_table] or [0],
                                                                dtype=torch.int64, device=device)[:max(n_f, 0)],
                                 persistent=False)
            self.register_buffer(f"_{name}_out_col", (torch.arange(max(n_f, 1), device=device) * width)[:n_f],
                                 persistent=False)
        self.register_buffer("_rw_block_sizes", torch.tensor(self._rw_block or [1], dtype=torch.int64, device=device),
                             persistent=False)
        self.register_buffer("_err_flag", torch.zeros(1, dtype=torch.int32, device=device), persistent=False)

        # static permutes
        # features in destination order: for every owner its TW features, then the features of its column shards
        tw_order = [f for d in range(W) for f in self._tw_feats[d] + [f_ for (f_, _) in self._cw_slots[d]]]
        self._send_mult = max([tw_order.count(f) for f in set(tw_order)] or [1])
        self.register_buffer("_tw_send_perm", torch.tensor(tw_order or [0], dtype=torch.int32, device=device),
                             persistent=False)
        self.register_buffer("_rw_select_perm", torch.tensor(self._rw_feats or [0], dtype=torch.int32, device=device),
                             persistent=False)
        Fl = self._F_local
        recv_perm = [s * Fl + f for f in range(Fl) for s in range(W)]       # [src][f][b] -> [f][src][b]
        self.register_buffer("_recv_perm", torch.tensor(recv_perm or [0], dtype=torch.int32, device=device),
                             persistent=False)
        # strided-copy tables between the exchange layout [rank d][B][n_tw[d]*D] and the output
        # [B][F_total*D]; slots = TW features in owner order
        D_ = self._D
        so, ss, do, ds = [], [], [], []
        self._tw_block_off: List[int] = []          # element offset of rank d's block for batch size 1
        acc = 0
        for d in range(W):
            self._tw_block_off.append(acc)
            acc += self._n_tw[d] * D_
        self._tw_cols_total = acc
        self._unpack_slots = [(d, j, f) for d in range(W) for j, f in enumerate(self._tw_feats[d])]
        self._cw_block_off: List[int] = []
        acc = 0
        for d in range(W):
            self._cw_block_off.append(acc)
            acc += self._n_cw[d] * self._Dc
        self._cw_cols_total = acc
        self._cw_unpack_slots = [(d, jj, f, j) for d in range(W) for jj, (f, j) in enumerate(self._cw_slots[d])]
        self._anchor: Optional[torch.Tensor] = None
        self._fused: Optional[FusedOptimizer] = None
        self._meta_cache: Dict[tuple, torch.Tensor] = {}
        self._scratch: Dict[str, torch.Tensor] = {}
        self._emit_keys = False                      # set per forward: training with the CUDA kernel table
        self._fwd_token = 0
        self._peer_by_B: Dict[int, "_PeerExchange"] = {}
        self._prefetched = None                      # (features, ctx, event) of prefetch_input_dist
        self._track_positions = False                # sequence output: remember where every sent id came from
        self._routing: Dict[str, object] = {}
        # Table-wise output dist / gradient return through NVLink peer memory instead of NCCL
        # all-to-alls (needs the CUDA kernels and symmetric memory; on by default when available).
        if peer_memory is None:
            peer_memory = isinstance(self._ops, CudaKernelOps) and W > 1 and device.type == "cuda"
        self._use_peer = bool(peer_memory)
        self._max_ids_per_rank = max_ids_per_rank
        self._peer: Optional[_PeerExchange] = None
        self.register_buffer("_tw_global_col", torch.tensor([f * self._D for f in self._tw.features] or [0],
                                                            dtype=torch.int64, device=device)[:len(self._tw.features)],
                             persistent=False)
        cw_cols = [f * self._D + self._cw_local_shard[ti] * self._Dc
                   for f, ti in zip(self._cw.features, self._cw.feat_table)]
        self.register_buffer("_cw_global_col", torch.tensor(cw_cols or [0], dtype=torch.int64,
                                                            device=device)[:len(cw_cols)], persistent=False)
        self.register_buffer("_rw_global_col", torch.tensor([f * self._D for f in self._rw_feats] or [0],
                                                            dtype=torch.int64, device=device)[:self._n_rw],
                             persistent=False)
        self.set_fused_optimizer(EmbOptimType.EXACT_SGD, learning_rate=0.01)

    def _tw_meta(self, B: int, to_output: bool, device, cls: str = "tw") -> torch.Tensor:
        """Copy table for batch size B: exchange layout <-> KeyedTensor layout (cached per B).
        cls "tw": whole-row slots of width D; "cw": column-shard slots of width Dc."""
        key = (B, to_output, str(device), cls)
        m = self._meta_cache.get(key)
        if m is None:
            D_, Ft, Dc = self._D, self._F_total, self._Dc
            ex_off, ex_stride, kt_off, kt_stride = [], [], [], []
            if cls == "tw":
                for (d, j, f) in self._unpack_slots:
                    ex_off.append(self._tw_block_off[d] * B + j * D_)
                    ex_stride.append(self._n_tw[d] * D_)
                    kt_off.append(f * D_)
                    kt_stride.append(Ft * D_)
            else:
                for (d, jj, f, j) in self._cw_unpack_slots:
                    ex_off.append(self._cw_block_off[d] * B + jj * Dc)
                    ex_stride.append(self._n_cw[d] * Dc)
                    kt_off.append(f * D_ + j * Dc)
                    kt_stride.append(Ft * D_)
            rows = [ex_off, ex_stride, kt_off, kt_stride] if to_output else [kt_off, kt_stride, ex_off, ex_stride]
            m = torch.tensor(rows, dtype=torch.int64, device=device).reshape(-1)
            self._meta_cache[key] = m
        return m

    def _peer_grad_meta(self, B: int, device, cls: str = "tw") -> torch.Tensor:
        """Copy table: my gradient columns -> rows [rank*B, rank*B + B) of each owner's landing buffer."""
        key = (B, "peer_grad", str(device), cls)
        m = self._meta_cache.get(key)
        if m is None:
            D_, Ft, r, Dc = self._D, self._F_total, self._rank, self._Dc
            so, ss, do, ds = [], [], [], []
            if cls == "tw":
                for (d, j, f) in self._unpack_slots:
                    n = self._n_tw[d]
                    so.append(f * D_)
                    ss.append(Ft * D_)
                    do.append(self._peer_by_B[B].grad_ptrs[d] // 4 + r * B * n * D_ + j * D_)     # absolute address / 4
                    ds.append(n * D_)
            elif cls == "rw_mc":
                # the same all-gather through the NVLS multicast mapping of the landing buffer: one slot per feature
                n = self._n_rw
                for k_, f in enumerate(self._rw_feats):
                    so.append(f * D_)
                    ss.append(Ft * D_)
                    do.append(self._peer_by_B[B].grad_rw_mc // 4 + r * B * n * D_ + k_ * D_)
                    ds.append(n * D_)
            elif cls == "rw":
                # every rank needs the gradient of every sample for the row-wise features (an all-gather,
                # written by the sample owners straight into the other ranks' landing buffers)
                n = self._n_rw
                for d in range(self._world):
                    for k_, f in enumerate(self._rw_feats):
                        so.append(f * D_)
                        ss.append(Ft * D_)
                        do.append(self._peer_by_B[B].grad_rw_ptrs[d] // 4 + r * B * n * D_ + k_ * D_)
                        ds.append(n * D_)
            else:
                for (d, jj, f, j) in self._cw_unpack_slots:
                    n = self._n_cw[d]
                    so.append(f * D_ + j * Dc)
                    ss.append(Ft * D_)
                    do.append(self._peer_by_B[B].grad_cw_ptrs[d] // 4 + r * B * n * Dc + jj * Dc)
                    ds.append(n * Dc)
            m = torch.tensor([so, ss, do, ds], dtype=torch.int64, device=device).reshape(-1)
            self._meta_cache[key] = m
        return m

    # ------------------------------------------------------------------ optimizer
    def set_fused_optimizer(self, optimizer=EmbOptimType.EXACT_SGD, learning_rate=0.01, eps=1e-8, weight_decay=0.0,
                            beta1=0.9, beta2=0.999) -> FusedOptimizer:
        params = {"_tw_weight": self._tw_weight, "_cw_weight": self._cw_weight, "_rw_weight": self._rw_weight}
        self._fused = FusedOptimizer(params, optimizer, learning_rate, eps, weight_decay, beta1, beta2)
        for p in params.values():
            p._in_backward_optimizers = [self._fused]
        return self._fused

    @property
    def fused_optimizer(self) -> FusedOptimizer:
        return self._fused

    def _state_for(self, name: str, rows: int):
        code = self._fused.optimizer.code()
        w = getattr(self, f"_{name}_weight")
        st = self._fused.state
        k0, k1 = f"_{name}_weight.state0", f"_{name}_weight.state1"
        if code == _lib.OPT_SGD:
            return None, None
        if code == _lib.OPT_ROWWISE_ADAGRAD:
            if k0 not in st:
                st[k0] = torch.zeros(w.shape[0], dtype=torch.float32, device=w.device)
            return st[k0], None
        for k in (k0, k1):
            if k not in st:
                st[k] = torch.zeros(tuple(w.shape), dtype=torch.float32, device=w.device)
        return st[k0], st[k1]

    def _group(self, name: str, with_state: bool):
        sh = {"tw": self._tw, "cw": self._cw, "rw": self._rw}[name]
        s0, s1 = self._state_for(name, sh.total_rows) if with_state else (None, None)
        width = self._Dc if name == "cw" else self._D
        return self._ops.make_group(getattr(self, f"_{name}_weight").data, getattr(self, f"_{name}_row_base"),
                                    getattr(self, f"_{name}_rows"), getattr(self, f"_{name}_out_col"),
                                    len(sh.features) * width, s0, s1, self._err_flag)

    # ------------------------------------------------------------------ collectives (plumbing)
    def _a2a(self, send: torch.Tensor, send_splits: Optional[List[int]], recv_splits: Optional[List[int]],
             recv_shape_tail=()) -> torch.Tensor:
        n = sum(recv_splits) if recv_splits is not None else send.shape[0]
        recv = torch.empty((n, *recv_shape_tail), dtype=send.dtype, device=send.device)
        dist.all_to_all_single(recv, send.contiguous(), recv_splits, send_splits, group=self._pg)
        return recv

    # ------------------------------------------------------------------ forward
    def _forward_impl(self, values, lengths, B, ctx=None):
        return self._pooled_lookup(ctx if ctx is not None else self._input_dist(values, lengths, B), B)

    def _input_dist(self, values, lengths, B, channel: int = 0):
        """Route ids to the ranks that own their rows; returns the local KJT in
        [feature][source rank][batch] layout (ids `lv`, offsets `lo`).  `channel` 1: called from
        prefetch_input_dist on its own stream."""
        W, r, D, ops = self._world, self._rank, self._D, self._ops
        dev = values.device
        offsets = ops.complete_cumsum(lengths)
        track = self._track_positions
        pos_parts: List[List[tuple]] = [[] for _ in range(W)]
        Fl, n_rw = self._F_local, self._n_rw

        n_tw_all = sum(self._n_tw) + sum(self._n_cw)      # slots served whole-table or column-wise
        n_tw_r, n_cw_r = self._n_tw[r], self._n_cw[r]
        n_own = n_tw_r + n_cw_r                            # local features ahead of the row-wise ones
        peer = None
        if self._use_peer:
            # ---- input dist through NVLink peer memory: lengths, then ids, are stored by the senders
            # directly where the owner's lookup reads them; no all-to-all, no split sizes on the host
            if self._peer is None or self._peer.B != B:
                # one exchange (symmetric buffers + rendezvous, collective to create) per batch size, kept: a
                # train / eval alternation or a short last batch switches between them without reallocating
                self._peer = self._peer_by_B.get(B)
            if self._peer is None:
                cap = self._max_ids_per_rank or int(1.5 * W * max(values.numel(), 1)) + 1024
                # symmetric buffers must have one size on every rank: agree on the largest proposal
                cap_t = torch.tensor([cap], dtype=torch.int64, device=dev)
                dist.all_reduce(cap_t, op=dist.ReduceOp.MAX, group=self._pg)
                cap = int(cap_t.item())
                kw = {}
                if track:      # sequence module: also agree on the largest local id count (rows of the output buffers)
                    sc = torch.tensor([int(1.25 * values.numel()) + 256], dtype=torch.int64, device=dev)
                    dist.all_reduce(sc, op=dist.ReduceOp.MAX, group=self._pg)
                    kw["seq_cap"] = max(int(sc.item()), self._max_local_ids or 0)
                self._peer = self._peer_by_B[B] = ops.make_peer_exchange(
                    self._pg, dev, B, self._F_total, D, self._n_tw, n_rw, cap, n_cw=self._n_cw, Dc=self._Dc, **kw)
                self._peer.dist_count = 0
            peer = self._peer
            # landing zone: always zone 0 unless input dists are being prefetched, then alternating
            slot = peer.dist_count % len(peer.zones)
            peer.dist_count += 1
            z = peer.zones[slot]
            peer.barrier(channel)          # everyone is done reading this zone (and last step's outputs)
            if n_tw_all:
                ops.push_lengths(lengths, B, self._peer_meta("len_tw", B, dev, slot))
            seq = None
            if track:
                seq = dict(offsets=offsets, bo=None, bpos=None, arange=self._arange(values.numel(), dev))
            if n_rw:
                rl, ro, rv = ops.permute_2d(self._rw_select_perm[:n_rw], B, lengths, offsets, values, values.numel())
                if track:
                    # where in MY KeyedJaggedTensor every bucketised id came from (travels with the ids)
                    _, _, rpos = ops.permute_2d(self._rw_select_perm[:n_rw], B, lengths, offsets, seq["arange"],
                                                values.numel())
                    bl, bo, bv, unb = ops.bucketize_row_wise_with_permute(rl, ro, rv, self._rw_block_sizes[:n_rw], n_rw,
                                                                          B, W, self._err_flag)
                    seq["bo"] = bo
                    seq["bpos"] = ops.scatter_positions(unb, rpos, values.numel(), ro[n_rw * B:])
                else:
                    bl, bo, bv = ops.bucketize_row_wise(rl, ro, rv, self._rw_block_sizes[:n_rw], n_rw, B, W,
                                                        self._err_flag)
                ops.push_lengths(bl, B, self._peer_meta("len_rw", B, dev, slot))
            peer.barrier(channel)          # lengths have landed
            lo_ = z.off_x[: Fl * W * B + 1]
            ops.complete_cumsum_into(z.len_x[: Fl * W * B], lo_)
            z.counts_host.copy_(torch.stack([lo_[n_tw_r * W * B], lo_[n_own * W * B], lo_[Fl * W * B]]),
                                non_blocking=True)
            z.counts_event.record()
            peer.barrier(channel)          # every owner's offsets are readable
            if n_tw_all:
                ops.push_ids(values, offsets, B, self._peer_meta("ids_tw", B, dev, slot), peer.capacity, self._err_flag)
            if n_rw:
                ops.push_ids(bv, bo, B, self._peer_meta("ids_rw", B, dev, slot), peer.capacity, self._err_flag)
            if track:                      # the ids' positions, through the same transport into the zone's pos_x
                if n_tw_all:
                    ops.push_ids(seq["arange"], offsets, B, self._peer_meta("pos_tw", B, dev, slot), peer.capacity,
                                 self._err_flag)
                if n_rw:
                    ops.push_ids(seq["bpos"], bo, B, self._peer_meta("pos_rw", B, dev, slot), peer.capacity, self._err_flag)
            peer.barrier(channel)          # ids have landed
            lv = z.ids_x
            tw_total = cw_total = None     # known on the host only later (counts_event)
        else:
            # ---- input dist: per-destination blocks  [tw features of d | rw bucket d]
            len_parts: List[List[torch.Tensor]] = [[] for _ in range(W)]
            val_parts: List[List[torch.Tensor]] = [[] for _ in range(W)]
            if n_tw_all:
                # a column-wise feature is sent to each of its shard owners: size for the repeats
                tl, to, tv = ops.permute_2d(self._tw_send_perm[:n_tw_all], B, lengths, offsets, values,
                                            values.numel() * self._send_mult)
                f0 = 0
                bounds = to[:: max(B, 1)] if B else to
                for d in range(W):
                    f1 = f0 + self._n_tw[d] + self._n_cw[d]
                    len_parts[d].append(tl[f0 * B:f1 * B])
                    val_parts[d].append((tv, bounds, f0, f1))
                    f0 = f1
                if track:       # where in the local KJT each sent id came from
                    arange = torch.arange(values.numel(), dtype=torch.int64, device=dev)
                    _, _, tpos = ops.permute_2d(self._tw_send_perm[:n_tw_all], B, lengths, offsets, arange,
                                                values.numel())
                    for d in range(W):
                        pos_parts[d].append(tpos)
            if n_rw:
                rl, ro, rv = ops.permute_2d(self._rw_select_perm[:n_rw], B, lengths, offsets, values, values.numel())
                if track:
                    arange = torch.arange(values.numel(), dtype=torch.int64, device=dev)
                    _, _, rpos = ops.permute_2d(self._rw_select_perm[:n_rw], B, lengths, offsets, arange,
                                                values.numel())
                    bl, bo, bv, unb = ops.bucketize_row_wise_with_permute(rl, ro, rv, self._rw_block_sizes[:n_rw],
                                                                          n_rw, B, W, self._err_flag)
                    n_rw_ids = int(ro[n_rw * B].item())        # buffers may be larger than the id count
                    bpos = torch.empty(max(n_rw_ids, 1), dtype=torch.int64, device=dev)
                    bpos[unb[:n_rw_ids]] = rpos[:n_rw_ids]     # slot in the bucketised order -> original position
                    for d in range(W):
                        pos_parts[d].append(bpos)
                else:
                    bl, bo, bv = ops.bucketize_row_wise(rl, ro, rv, self._rw_block_sizes[:n_rw], n_rw, B, W,
                                                        self._err_flag)
                bbounds = bo[:: max(n_rw * B, 1)]
                for d in range(W):
                    len_parts[d].append(bl[d * n_rw * B:(d + 1) * n_rw * B])
                    val_parts[d].append((bv, bbounds, d, d + 1))
            send_lengths = torch.cat([p for d in range(W) for p in len_parts[d]])
            # value split sizes need host ints: one small device->host read per step
            bound_list = []
            for d in range(W):
                for (_, bnd, i0, i1) in val_parts[d]:
                    bound_list += [bnd[i0], bnd[i1]]
            b_host = torch.stack(bound_list).tolist() if bound_list else []
            send_vals, send_pos, send_splits, k = [], [], [0] * W, 0
            for d in range(W):
                for j, (buf, _, _, _) in enumerate(val_parts[d]):
                    lo, hi = b_host[k], b_host[k + 1]
                    k += 2
                    send_vals.append(buf[lo:hi])
                    if track:
                        send_pos.append(pos_parts[d][j][lo:hi])
                    send_splits[d] += hi - lo
            send_values = torch.cat(send_vals) if send_vals else values[:0]

            len_send_splits = [(self._n_tw[d] + self._n_cw[d] + n_rw) * B for d in range(W)]
            recv_lengths = self._a2a(send_lengths, len_send_splits, [Fl * B] * W)
            if Fl * B:
                rl3 = recv_lengths.view(W, Fl, B)
                # per-source id counts and the id counts of the TW / TW+CW parts: one device->host read
                host = torch.cat([rl3.sum(dim=(1, 2), dtype=torch.int64),
                                  rl3[:, :n_tw_r].sum(dtype=torch.int64).view(1),
                                  rl3[:, :n_own].sum(dtype=torch.int64).view(1)]).tolist()
                recv_splits, tw_total, cw_total = host[:W], host[W], host[W + 1]
            else:
                recv_splits, tw_total, cw_total = [0] * W, 0, 0
            recv_values = self._a2a(send_values, send_splits, recv_splits)

            # ---- [src][f][b] -> [f][src][b]
            if Fl:
                recv_offsets = ops.complete_cumsum(recv_lengths)
                ll, lo_, lv = ops.permute_2d(self._recv_perm[:Fl * W], B, recv_lengths, recv_offsets, recv_values,
                                             int(sum(recv_splits)))
                if track:
                    n_recv = int(sum(recv_splits))
                    _, _, lv_pos = ops.permute_2d(self._recv_perm[:Fl * W], B, recv_lengths, recv_offsets,
                                                  torch.arange(n_recv, dtype=torch.int64, device=dev), n_recv)
                    self._routing = dict(send_pos=torch.cat(send_pos) if send_pos else values[:0],
                                         send_splits=send_splits, recv_splits=recv_splits, lv_pos=lv_pos[:n_recv])
            else:
                ll = recv_lengths
                lo_ = torch.zeros(1, dtype=torch.int64, device=dev)
                lv = recv_values
        return dict(lv=lv, lo=lo_, tw_total=tw_total, cw_total=cw_total, n_tw_r=n_tw_r, n_cw_r=n_cw_r, peer=peer,
                    dev=dev, slot=slot if peer is not None else 0, seq=seq if peer is not None else None,
                    mean_len=values.numel() / max(lengths.numel(), 1))

    def _pooled_lookup(self, c, B):
        W, r, D, ops = self._world, self._rank, self._D, self._ops
        lv, lo_, tw_total, n_tw_r, peer, dev = c["lv"], c["lo"], c["tw_total"], c["n_tw_r"], c["peer"], c["dev"]
        cw_total, n_cw_r = c["cw_total"], c["n_cw_r"]
        n_own = n_tw_r + n_cw_r
        Fl, n_rw, n_tw_all, n_cw_all = self._F_local, self._n_rw, sum(self._n_tw), sum(self._n_cw)
        GB = W * B                                            # global batch seen by each local feature

        # ---- lookup on the local storage
        self._fwd_token += 1
        if peer is not None:
            k = peer.step % 2
            peer.step += 1
            out = peer.out[k].view(B, self._F_total * D)     # valid until the forward after next
            # Training: the lookups also write the backward's (row key, bag) pairs, at the ids' own positions in
            # a capacity-sized buffer (how many ids arrived is not known on the host yet), so the backward can
            # go straight to the sort.
            pairs = None
            if self._emit_keys and hasattr(ops, "embedding_backward_prepare_pairs"):
                pairs = getattr(peer, "sort_pairs", None)
                if pairs is None:
                    pairs = peer.sort_pairs = (torch.empty(max(peer.capacity, 4), dtype=torch.int32, device=dev),
                                               torch.empty(max(peer.capacity, 4), dtype=torch.int32, device=dev))
            # how many ids a shard should expect (the landing zone's fill is not known on the host yet): the local
            # batch's mean bag length times the shard's bags; it only picks the lookup's launch shape
            mean_len = c.get("mean_len", 0.0)
            hint = lambda n_feat: int(mean_len * n_feat * GB) + 1
            if n_tw_r:
                # pooled rows go straight into the output buffer of the rank that owns the sample
                grp = self._group("tw", False)
                grp.feat_out_col = self._tw_global_col        # columns of the final [B, F_total*D] layout
                grp.out_stride = self._F_total * D
                ops.embedding_forward_peer(grp, lv, lo_[: n_tw_r * GB + 1], GB, self._pooling, peer.out_ptrs[k], B,
                                           sort_pairs=pairs, expected_ids=hint(n_tw_r))
            if n_cw_r:
                # column shards: the same, Dc columns per pooled row
                grp = self._group("cw", False)
                grp.feat_out_col = self._cw_global_col
                grp.out_stride = self._F_total * D
                ops.embedding_forward_peer(grp, lv, lo_[n_tw_r * GB: n_own * GB + 1], GB, self._pooling,
                                           peer.out_ptrs[k], B, sort_pairs=pairs, expected_ids=hint(n_cw_r))
            if n_rw:
                # row-wise: my block's partial sums of sample (src, b) go into slot `r` of rank src's buffer
                key = (B, "rw_part_out", str(dev))
                slot_ptrs = self._meta_cache.get(key)
                if slot_ptrs is None:
                    slot_ptrs = self._meta_cache[key] = torch.tensor(
                        [p + 4 * r * B * n_rw * D for p in peer.rw_part_ptrs], dtype=torch.int64, device=dev)
                ops.embedding_forward_peer(self._group("rw", False), lv, lo_[n_own * GB: Fl * GB + 1], GB, _lib.POOL_SUM,
                                           slot_ptrs, B, sort_pairs=pairs, expected_ids=hint(n_rw) // W + 1)
            peer.barrier()                                    # all pooled rows / partial sums have landed
            if n_rw:
                # add the W slots in rank order into the row-wise columns of the output (fixed order: reproducible)
                ops.reduce_partial_sums(peer.rw_part, W, B, n_rw, D, self._rw_global_col, out, self._F_total * D)
            return out, dict(peer=True, px=peer, lo=lo_, n_tw_r=n_tw_r, n_cw_r=n_cw_r, GB=GB, token=self._fwd_token,
                             fused=False, slot=c["slot"], pairs=pairs)

        out = torch.empty((B, self._F_total * D), dtype=torch.float32, device=dev)
        out3 = out.view(B, self._F_total, D)
        if n_tw_all:
            if n_tw_r:
                tw_out = torch.empty((GB, n_tw_r * D), dtype=torch.float32, device=dev)
                ops.embedding_forward(self._group("tw", False), lv[:tw_total], lo_[: n_tw_r * GB + 1], GB,
                                      self._pooling, tw_out, self._bw_scratch("tw", tw_total, dev))
            else:
                tw_out = torch.empty((0,), dtype=torch.float32, device=dev)
            # output dist TW: rank d returns B rows of n_tw[d]*D floats to every source rank
            recv_flat = self._a2a(tw_out.reshape(-1), [B * n_tw_r * D] * W, [B * self._n_tw[d] * D for d in range(W)])
            ops.copy_feature_blocks(recv_flat, out, B, D, self._tw_meta(B, True, dev))
        saved_cw = None
        if n_cw_all:
            Dc = self._Dc
            if n_cw_r:
                cw_vals = lv[tw_total:cw_total]
                cw_off = lo_[n_tw_r * GB: n_own * GB + 1] - tw_total if n_tw_r else lo_[: n_own * GB + 1]
                cw_out = torch.empty((GB, n_cw_r * Dc), dtype=torch.float32, device=dev)
                ops.embedding_forward(self._group("cw", False), cw_vals, cw_off, GB, self._pooling, cw_out,
                                      self._bw_scratch("cw", cw_vals.numel(), dev))
                saved_cw = (cw_vals, cw_off)
            else:
                cw_out = torch.empty((0,), dtype=torch.float32, device=dev)
            recv_flat = self._a2a(cw_out.reshape(-1), [B * n_cw_r * Dc] * W, [B * self._n_cw[d] * Dc for d in range(W)])
            ops.copy_feature_blocks(recv_flat, out, B, Dc, self._tw_meta(B, True, dev, "cw"))
        saved_rw = None
        if n_rw:
            rw_off = lo_[n_own * GB:] - cw_total if n_own else lo_
            rw_vals = lv[cw_total:]
            rw_out = torch.empty((GB, n_rw * D), dtype=torch.float32, device=dev)
            ops.embedding_forward(self._group("rw", False), rw_vals, rw_off, GB, _lib.POOL_SUM, rw_out,
                                  self._bw_scratch("rw", rw_vals.numel(), dev))
            out3[:, self._rw_feats, :] = self._exchange_partial_sums(rw_out, B, n_rw)
            saved_rw = (rw_vals, rw_off)
        # the TW part of the local ids only: the backward sorts exactly the ids it is given
        saved = dict(peer=False, lv=lv[:tw_total], lo=lo_, n_tw_r=n_tw_r, n_cw_r=n_cw_r, GB=GB, cw=saved_cw, rw=saved_rw,
                     fused=self._emit_keys, token=self._fwd_token)
        return out, saved

    def _exchange_partial_sums(self, rw_out: torch.Tensor, B: int, n_rw: int) -> torch.Tensor:
        """Row-wise output dist: every rank holds partial sums for all W*B samples; exchange them and
        add in source-rank order (a reduce-scatter with a fixed, reproducible order)."""
        W, D = self._world, self._D
        recv_flat = self._a2a(rw_out.reshape(-1), [B * n_rw * D] * W, [B * n_rw * D] * W)
        parts = recv_flat.view(W, B, n_rw, D)
        acc = parts[0].clone()
        for s_ in range(1, W):
            acc += parts[s_]
        return acc

    def _peer_meta(self, kind: str, B: int, device, slot: int = 0) -> torch.Tensor:
        """Push tables for the peer-memory input dist (cached per batch size and landing zone)."""
        key = (B, kind, str(device), slot)
        m = self._meta_cache.get(key)
        if m is not None:
            return m
        W, r, n_rw, p = self._world, self._rank, self._n_rw, self._peer_by_B[B].zones[slot]
        rows: List[List[int]] = []
        if kind in ("len_tw", "ids_tw", "pos_tw", "gseq_tw"):
            # whole-table slots, then the column-shard slots of each owner (a column-wise feature is
            # pushed to every owner of one of its shards)
            slots = [(d, j, f) for d in range(W)
                     for j, f in enumerate(self._tw_feats[d] + [f_ for (f_, _) in self._cw_slots[d]])]
            seg = [f for (_, _, f) in slots]                               # segment = original feature
            word = [(j * W + r) * B for (_, j, _) in slots]                # (feature j, source r, batch 0) on owner d
        else:
            slots = [(d, self._n_tw[d] + self._n_cw[d] + k_, None) for d in range(W) for k_ in range(n_rw)]
            seg = [d * n_rw + k_ for d in range(W) for k_ in range(n_rw)]  # segment of the bucketised layout
            word = [(j * W + r) * B for (_, j, _) in slots]
        dsts = [d for (d, _, _) in slots]
        if kind.startswith("len"):
            rows = [[sg * B for sg in seg], [p.len_ptrs[d] + 4 * w_ for d, w_ in zip(dsts, word)]]
        else:
            base = (p.ids_ptrs if kind.startswith("ids") else p.pos_ptrs if kind.startswith("pos")
                    else self._peer_by_B[B].grad_seq_ptrs)
            rows = [seg, [p.off_ptrs[d] + 8 * w_ for d, w_ in zip(dsts, word)], [base[d] for d in dsts]]
        m = torch.tensor(rows, dtype=torch.int64, device=device).reshape(-1)
        self._meta_cache[key] = m
        return m

    _max_local_ids: Optional[int] = None

    def _arange(self, n: int, device) -> torch.Tensor:
        """0..n-1 on the device, grown geometrically and kept (positions of the local ids)."""
        buf = getattr(self, "_arange_buf", None)
        if buf is None or buf.numel() < n or buf.device != device:
            buf = self._arange_buf = torch.arange(max(int(1.5 * n), 1024), dtype=torch.int64, device=device)
        return buf[:n]

    def _bw_scratch(self, name: str, L: int, device):
        """Backward workspace of a shard; the forward writes the sort pairs into it when training."""
        if not self._emit_keys or L == 0:
            return None
        from . import ops as _o
        sh = {"tw": self._tw, "cw": self._cw, "rw": self._rw}[name]
        need = _o.backward_layout(L, max(sh.total_rows, 1)).total_bytes
        buf = self._scratch.get(name)
        if buf is None or buf.numel() < need or buf.device != device:
            buf = torch.empty(int(need * 1.1) + 256, dtype=torch.uint8, device=device)
            self._scratch[name] = buf
        return buf

    # ------------------------------------------------------------------ backward
    def _side_stream(self, device):
        st = getattr(self, "_side", None)
        if st is None or st.device != device:
            st = self._side = torch.cuda.Stream(device=device, priority=-1)
        return st

    def _prepared_scratch(self, name: str, L: int, device):
        """Workspace the id-only half of the backward sorts into (one per shard kind: they are in flight together)."""
        sh = {"tw": self._tw, "cw": self._cw, "rw": self._rw}[name]
        need = self._ops.backward_scratch_bytes(L, max(sh.total_rows, 1))
        buf = self._scratch.get("prep_" + name)
        if buf is None or buf.numel() < need or buf.device != device:
            buf = torch.empty(int(need * 1.1) + 256, dtype=torch.uint8, device=device)
            self._scratch["prep_" + name] = buf
        return buf

    def _backward_impl(self, grad_out, saved, B):
        W, D, Dc, ops = self._world, self._D, self._Dc, self._ops
        stale = saved["token"] != self._fwd_token          # another forward ran since
        self._fused.iteration += 1
        opt = self._fused.args()
        GB, n_tw_r, n_cw_r, n_rw = saved["GB"], saved["n_tw_r"], saved["n_cw_r"], self._n_rw
        n_own, Fl = n_tw_r + n_cw_r, self._F_local
        n_tw_all, n_cw_all = sum(self._n_tw), sum(self._n_cw)
        if saved["peer"]:
            if stale:
                raise RuntimeError("tdfo_b200: with peer-memory exchange, backward must run before the next forward "
                                   "(the received ids live in the shared landing buffers)")
            px = saved["px"]
            z = px.zones[saved["slot"]]
            z.counts_event.synchronize()                    # counts were copied during the forward: no stall
            tw_total, cw_total, total = (int(x) for x in z.counts_host.tolist())
            if total > px.capacity:
                # senders dropped what did not fit (TDFO_FLAG_RECV_OVERFLOW on their side) and the lookup cut the
                # bags at the buffer's end: nothing was read out of bounds, but this step's results are incomplete
                raise RuntimeError(f"tdfo_b200: rank {self._rank} received {total} ids this step, its exchange buffer "
                                   f"holds {px.capacity}; construct the module with a larger max_ids_per_rank")
            lo_ = saved["lo"]
            lv_tw, lo_tw = z.ids_x[:tw_total], lo_[: n_tw_r * GB + 1]
            cw = (z.ids_x[tw_total:cw_total], lo_[n_tw_r * GB: n_own * GB + 1] - tw_total) if n_cw_r else None
            rw = (z.ids_x[cw_total:total], lo_[n_own * GB: Fl * GB + 1] - cw_total) if n_rw else None
            fused = False
        else:
            lv_tw, lo_tw, cw, rw = saved["lv"], saved["lo"][: n_tw_r * GB + 1], saved["cw"], saved["rw"]
            fused = saved["fused"] and not stale
        recv = recv_cw = recv_rw = None
        prepared = {}
        if saved["peer"]:
            # Push every gradient row (column slice) into the landing buffer of its owner over NVLink, on a
            # side stream: the push is NVLink-bound and the id-only half of the backward (linearise, sort,
            # unique: latency-bound, needs no gradient) runs on the main stream meanwhile.
            todo = [(name, ids_, off_) for name, have, ids_, off_ in
                    (("tw", n_tw_r, lv_tw, lo_tw), ("cw", n_cw_r) + (cw if cw else (None, None)),
                     ("rw", n_rw) + (rw if rw else (None, None)))
                    if have and ids_.numel() > 0]
            # (Measured at N=8: linearising before the push starts does not help, 3.28 vs 3.24 ms/step: every
            # kernel that runs next to the NVLink stores slows down, the streaming linearise pass most.)
            for name, ids_, off_ in todo:
                prepared[name] = self._prepared_scratch(name, ids_.numel(), grad_out.device)
            on_gpu = grad_out.is_cuda                       # (the CPU tests drive this branch with a fake exchange)
            if on_gpu:
                main = torch.cuda.current_stream(grad_out.device)
                side = self._side_stream(grad_out.device)
                side.wait_stream(main)                      # grad_out has been produced
                grad_out.record_stream(side)
            with (torch.cuda.stream(side) if on_gpu else contextlib.nullcontext()):
                if n_tw_all:
                    ops.copy_feature_blocks(grad_out, None, B, D, self._peer_grad_meta(B, grad_out.device))
                    recv = px.grad[: GB * n_tw_r * D]
                if n_cw_all:
                    ops.copy_feature_blocks(grad_out, None, B, Dc, self._peer_grad_meta(B, grad_out.device, "cw"))
                    recv_cw = px.grad_cw[: GB * n_cw_r * Dc]
                if n_rw:
                    if getattr(px, "grad_rw_mc", 0) and hasattr(ops, "copy_feature_blocks_multicast"):
                        ops.copy_feature_blocks_multicast(grad_out, B, D, self._peer_grad_meta(B, grad_out.device, "rw_mc"))
                    else:
                        ops.copy_feature_blocks(grad_out, None, B, D, self._peer_grad_meta(B, grad_out.device, "rw"))
                    recv_rw = px.grad_rw.view(GB, n_rw * D)
                if n_tw_all or n_cw_all or n_rw:
                    px.barrier()
                if on_gpu:
                    landed = torch.cuda.Event()
                    landed.record(side)
            starts = {"tw": 0, "cw": tw_total, "rw": cw_total}      # where each shard's ids start in the landing zone
            for name, ids_, off_ in todo:
                if saved.get("pairs") is not None:
                    s0 = starts[name]
                    ops.embedding_backward_prepare_pairs(self._group(name, False), saved["pairs"][0][s0:s0 + ids_.numel()],
                                                         saved["pairs"][1][s0:s0 + ids_.numel()], ids_.numel(),
                                                         prepared[name])
                else:
                    ops.embedding_backward_prepare(self._group(name, False), ids_, off_, GB,
                                                   _lib.POOL_SUM if name == "rw" else self._pooling, prepared[name])
            if on_gpu:
                main.wait_event(landed)
        else:
            # the owner of a TW feature / column shard needs the gradient of every rank's samples
            if n_tw_all:
                send = torch.empty(B * self._tw_cols_total, dtype=torch.float32, device=grad_out.device)
                ops.copy_feature_blocks(grad_out, send, B, D, self._tw_meta(B, False, grad_out.device))
                recv = self._a2a(send, [B * self._n_tw[d] * D for d in range(W)], [B * n_tw_r * D] * W)
            if n_cw_all:
                send = torch.empty(B * self._cw_cols_total, dtype=torch.float32, device=grad_out.device)
                ops.copy_feature_blocks(grad_out, send, B, Dc, self._tw_meta(B, False, grad_out.device, "cw"))
                recv_cw = self._a2a(send, [B * self._n_cw[d] * Dc for d in range(W)], [B * n_cw_r * Dc] * W)
        if n_tw_r:
            sc = prepared.get("tw", self._scratch.get("tw") if fused else None)
            ops.embedding_backward_update(self._group("tw", True), recv.view(GB, n_tw_r * D), lv_tw, lo_tw, GB,
                                          self._pooling, opt, scratch=sc,
                                          keys_in_scratch="prepared" if "tw" in prepared else
                                          (sc is not None and lv_tw.numel() > 0))
        if n_cw_r:
            cw_vals, cw_off = cw
            sc = prepared.get("cw", self._scratch.get("cw") if fused else None)
            ops.embedding_backward_update(self._group("cw", True), recv_cw.view(GB, n_cw_r * Dc), cw_vals, cw_off, GB,
                                          self._pooling, opt, scratch=sc,
                                          keys_in_scratch="prepared" if "cw" in prepared else
                                          (sc is not None and cw_vals.numel() > 0))
        if n_rw:
            rw_vals, rw_off = rw
            if saved["peer"]:
                full = recv_rw
                sc = prepared.get("rw")
                kis = "prepared" if "rw" in prepared else False
            else:
                g3 = grad_out.view(B, self._F_total, D)
                mine = g3[:, self._rw_feats, :].reshape(B, n_rw * D).contiguous()
                full = torch.empty((GB, n_rw * D), dtype=torch.float32, device=grad_out.device)
                dist.all_gather_into_tensor(full, mine, group=self._pg)
                sc = self._scratch.get("rw") if fused else None
                kis = sc is not None and rw_vals.numel() > 0
            ops.embedding_backward_update(self._group("rw", True), full, rw_vals, rw_off, GB, _lib.POOL_SUM, opt,
                                          scratch=sc, keys_in_scratch=kis)

    bounds_check_mode = None      # set to a modules.BoundsCheckMode to change what check_errors() does with bad ids

    def check_errors(self) -> None:
        """Synchronising check of this rank's device error word: ids outside their table (IndexError, unless
        bounds_check_mode is IGNORE), a receive buffer that was too small for a batch, or offsets past the
        received ids (RuntimeError).  The sharded kernels never read out of bounds in any of these cases; the
        word says the step's results are incomplete."""
        from .modules import BoundsCheckMode
        from . import ops as _o
        try:
            _o.check_error_flag(self._err_flag, "the sharded lookup")
        except IndexError:
            if self.bounds_check_mode != BoundsCheckMode.IGNORE:
                raise

    def _in_module_order(self, features: KeyedJaggedTensor) -> KeyedJaggedTensor:
        if list(features.keys()) != self._feature_names:
            keys = list(features.keys())
            features = features.permute([keys.index(k) for k in self._feature_names])
        return features

    def prefetch_input_dist(self, features: KeyedJaggedTensor) -> None:
        """COLLECTIVE.  Start the input dist (lengths / ids exchange) of a batch that a later
        `forward(features)` -- the same KeyedJaggedTensor object -- will consume, so that it runs next to
        the backward of the batch in flight (the sparse-dist stage of torchrec's train
        pipelines [UPSTREAM]):

            out = module(kjt_i); module.prefetch_input_dist(kjt_next); loss(out).backward()

        On CUDA it is enqueued on its own stream (ordered after everything already enqueued on the current
        one); with the peer-memory exchange the landing zones are double-buffered from the first call on
        and its barriers use their own signal-pad channel.  Every rank must call it for the same steps.
        Status: checked on CPU ranks (gloo all-to-all path and the fake peer memory of tests/), not yet
        timed on GPUs."""
        if self._track_positions:
            raise NotImplementedError("prefetch_input_dist is not available for ShardedEmbeddingCollection")
        if self._prefetched is not None:
            raise RuntimeError("tdfo_b200: one input dist is already prefetched; run its forward first")
        feats = self._in_module_order(features)
        values, lengths, B = feats.values(), feats.lengths(), feats.stride()
        if self._use_peer and B in self._peer_by_B:
            self._peer_by_B[B].ensure_second_zone()
        ev = None
        if values.is_cuda:
            main = torch.cuda.current_stream(values.device)
            side = getattr(self, "_dist_stream", None)
            if side is None or side.device != values.device:
                side = self._dist_stream = torch.cuda.Stream(device=values.device)
            side.wait_stream(main)
            with torch.cuda.stream(side):
                ctx = self._input_dist(values, lengths, B, channel=1)
                ev = torch.cuda.Event()
                ev.record(side)
        else:
            ctx = self._input_dist(values, lengths, B, channel=1)
        self._prefetched = (features, ctx, ev)

    def _take_prefetched(self, features: KeyedJaggedTensor):
        if self._prefetched is None:
            return None
        src, ctx, ev = self._prefetched
        if src is not features:
            raise RuntimeError("tdfo_b200: forward() got another batch than the one whose input dist was prefetched")
        self._prefetched = None
        if ev is not None:
            cur = torch.cuda.current_stream(ctx["dev"])
            cur.wait_event(ev)
            for v in ctx.values():                       # tensors produced on the prefetch stream
                if isinstance(v, torch.Tensor) and v.is_cuda:
                    v.record_stream(cur)
        return ctx

    def forward(self, features: KeyedJaggedTensor) -> KeyedTensor:
        dist_ctx = self._take_prefetched(features)
        features = self._in_module_order(features)
        values, lengths, B = features.values(), features.lengths(), features.stride()
        if self.training and torch.is_grad_enabled():
            if self._anchor is None or self._anchor.device != values.device:
                self._anchor = torch.zeros(1, device=values.device, requires_grad=True)
            self._emit_keys = isinstance(self._ops, CudaKernelOps)
            out = _ShardedLookupFn.apply(self._anchor, self, values, lengths, B, dist_ctx)
        else:
            self._emit_keys = False
            out, _ = self._forward_impl(values, lengths, B, dist_ctx)
        return KeyedTensor(self._feature_names, [self._D] * self._F_total, out)

    # ------------------------------------------------------------------ checkpoints
    # `torch.save(model.state_dict(), ...)` at torchrec/train.py:172-176 is called on every rank.
    # state_dict() therefore carries, under the same per-table keys as the unsharded module
    # (`embedding_bags.<table>.weight` / `embeddings.<table>.weight`), THIS rank's rows of every table
    # (an empty [0, D] tensor for a table-wise table owned elsewhere) plus `_shard_metadata` saying
    # which global rows they are.  load_state_dict() takes either such local shards or full tables
    # (e.g. the state_dict of an unsharded module, or full_state_dict()), so a checkpoint can be
    # resharded onto another plan / world size.
    _state_prefix = "embedding_bags"

    def _table_shards(self):
        """(table, storage name, first local row, local rows, first global row, first global column,
        columns) per local shard."""
        W, r = self._world, self._rank
        for shn, sh in (("tw", self._tw), ("cw", self._cw), ("rw", self._rw)):
            for ti, (t, r0, rows) in enumerate(zip(sh.tables, sh.row0, sh.local_rows)):
                g0 = r * ((t.num_embeddings + W - 1) // W) if shn == "rw" else 0
                c0, cols = (self._cw_local_shard[ti] * self._Dc, self._Dc) if shn == "cw" else (0, self._D)
                yield t, shn, r0, rows, g0, c0, cols

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        local = {t.name: (getattr(self, f"_{shn}_weight")[r0:r0 + rows], g0, c0)
                 for t, shn, r0, rows, g0, c0, _ in self._table_shards()}
        ref = self._tw_weight
        meta = {"world": self._world, "rank": self._rank, "tables": {}}
        for t in self._configs:
            v, g0, c0 = local.get(t.name, (ref.new_empty((0, self._D)), 0, 0))
            destination[f"{prefix}{self._state_prefix}.{t.name}.weight"] = v if keep_vars else v.detach()
            meta["tables"][t.name] = {"row_offset": g0, "rows": int(v.shape[0]), "num_embeddings": t.num_embeddings}
            if v.shape[1] != self._D:
                meta["tables"][t.name].update(col_offset=c0, cols=int(v.shape[1]))
        destination[f"{prefix}_shard_metadata"] = meta

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        # Shards are only interchangeable with the rank / world / plan they were cut for: a row block of
        # rank 0 has the same shape as rank 1's, so shape alone cannot tell them apart.  When the checkpoint
        # carries `_shard_metadata` (every state_dict() of a sharded module does), a local shard is accepted
        # only if the metadata says it is THIS rank's slice; full tables need no metadata.
        meta = state_dict.get(f"{prefix}_shard_metadata")
        for t, shn, r0, rows, g0, c0, cols in self._table_shards():
            key = f"{prefix}{self._state_prefix}.{t.name}.weight"
            if key not in state_dict:
                if strict:
                    missing_keys.append(key)
                continue
            src = state_dict[key]
            is_full = tuple(src.shape) == (t.num_embeddings, self._D)
            if is_full and (rows, cols) != (t.num_embeddings, self._D):
                src = src[g0:g0 + rows, c0:c0 + cols]         # a full table: take this rank's rows / columns
            elif not is_full and isinstance(meta, dict):
                tm = meta.get("tables", {}).get(t.name)
                want = {"world": self._world, "rank": self._rank, "row_offset": g0, "col_offset": c0,
                        "num_embeddings": t.num_embeddings}
                have = None if tm is None else {"world": meta.get("world"), "rank": meta.get("rank"),
                                                "row_offset": tm.get("row_offset"), "col_offset": tm.get("col_offset", 0),
                                                "num_embeddings": tm.get("num_embeddings")}
                if have != want:
                    error_msgs.append(f"shard mismatch for {key}: the checkpoint holds {have}, this rank needs {want} "
                                      "(per-rank checkpoints need per-rank files; or save full_state_dict() once)")
                    continue
            if tuple(src.shape) != (rows, cols):
                error_msgs.append(f"size mismatch for {key}: {tuple(src.shape)} is neither this rank's shard "
                                  f"{(rows, cols)} nor the full table {(t.num_embeddings, self._D)}")
                continue
            with torch.no_grad():
                getattr(self, f"_{shn}_weight")[r0:r0 + rows].copy_(src)

    def full_state_dict(self, to_cpu: bool = True) -> Dict[str, torch.Tensor]:
        """COLLECTIVE: every table whole, on every rank, under the unsharded module's keys."""
        W = self._world
        local = {t.name: getattr(self, f"_{shn}_weight")[r0:r0 + rows]
                 for t, shn, r0, rows, _, _, _ in self._table_shards()}
        owner = {f: d for d in range(W) for f in self._tw_feats[d]}
        cw_owner = {(f, j): d for d in range(W) for (f, j) in self._cw_slots[d]}
        out: Dict[str, torch.Tensor] = {}
        ref = self._tw_weight
        grank = (lambda d: dist.get_global_rank(self._pg, d)) if self._pg is not None else (lambda d: d)
        k = 0
        for t in self._configs:
            f0 = k
            k += len(t.feature_names)
            if f0 in owner:                                   # table-wise: the owner broadcasts
                src = owner[f0]
                full = local[t.name].contiguous() if src == self._rank else ref.new_empty((t.num_embeddings, self._D))
                dist.broadcast(full, src=grank(src), group=self._pg)
            elif (f0, 0) in cw_owner:                         # column-wise: every shard owner broadcasts its columns
                parts = []
                for j in range(self._D // self._Dc):
                    src = cw_owner[(f0, j)]
                    part = local[t.name].contiguous() if src == self._rank else \
                        ref.new_empty((t.num_embeddings, self._Dc))
                    dist.broadcast(part, src=grank(src), group=self._pg)
                    parts.append(part)
                full = torch.cat(parts, dim=1)
            else:                                             # row-wise: gather the (padded) blocks
                bs = (t.num_embeddings + W - 1) // W
                mine = ref.new_zeros((bs, self._D))
                mine[:local[t.name].shape[0]].copy_(local[t.name])
                parts = [torch.empty_like(mine) for _ in range(W)]
                dist.all_gather(parts, mine, group=self._pg)
                full = torch.cat(parts)[:t.num_embeddings]
            out[f"{self._state_prefix}.{t.name}.weight"] = full.cpu() if to_cpu else full.clone()
        return out

    # ------------------------------------------------------------------ inspection helpers
    def local_table(self, name: str) -> torch.Tensor:
        """This rank's rows of table `name` (whole table for TW on the owner, the row block for RW)."""
        for shn, sh in (("tw", self._tw), ("cw", self._cw), ("rw", self._rw)):
            for t, r0, rows in zip(sh.tables, sh.row0, sh.local_rows):
                if t.name == name:
                    return getattr(self, f"_{shn}_weight")[r0:r0 + rows]
        return torch.empty((0, self._D))

    @staticmethod
    def initial_table(table: EmbeddingBagConfig, init_seed: int = 0) -> torch.Tensor:
        """The full initial value of `table` as every rank draws it (for tests / checkpoints)."""
        g = torch.Generator(device="cpu").manual_seed(init_seed * 1_000_003 + zlib.crc32(table.name.encode()) % 1_000_003)
        return torch.empty((table.num_embeddings, table.embedding_dim), dtype=torch.float32).uniform_(
            table.get_weight_init_min(), table.get_weight_init_max(), generator=g)


# --------------------------------------------------------------------------- sequence (EmbeddingCollection)
class _ShardedSequenceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, module, values, lengths, B):
        rows, saved = module._seq_forward(values, lengths, B)
        ctx.module, ctx.saved, ctx.B = module, saved, B
        return rows

    @staticmethod
    def backward(ctx, grad_rows):
        ctx.module._seq_backward(grad_rows.contiguous(), ctx.saved, ctx.B)
        return None, None, None, None, None


class ShardedEmbeddingCollection(ShardedEmbeddingBagCollection):
    """`EmbeddingCollection` (one embedding row per id, torchrec/models.py:158-164) whose tables are
    spread over the ranks: the module DMP builds for the reference's Bert4Rec item table when it runs
    on more than one device (torchrec/train.py:241-247).

    Input dist as for the pooled module; the owners look the rows up (`TDFO_POOL_NONE`), put them
    back into the order the ids arrived in and return them with one all-to-all of `[n, D]` rows
    (split sizes = the id counts of the input dist, SURVEY.md §8 a13 "sequence"); the sender drops
    each row at the position its id had in the local KeyedJaggedTensor (for row-wise tables this is
    the un-bucketise permute).  Backward sends the gradient rows the same way round and runs the
    fused sort / unique / update on the owner.
    """

    _state_prefix = "embeddings"

    def __init__(self, tables, plan: Dict[str, ParameterSharding], device: torch.device,
                 process_group: Optional[dist.ProcessGroup] = None, kernel_ops=None, init_seed: int = 0,
                 init_on_device: bool = False, peer_memory: Optional[bool] = None,
                 max_ids_per_rank: Optional[int] = None, max_local_ids: Optional[int] = None):
        if any(ps.sharding_type == ShardingType.COLUMN_WISE for ps in plan.values()):
            raise ValueError("ShardedEmbeddingCollection supports table-wise and row-wise sharding (no column-wise)")
        super().__init__(tables, plan, device, process_group, kernel_ops, init_seed, init_on_device,
                         peer_memory=peer_memory, max_ids_per_rank=max_ids_per_rank)
        self._track_positions = True
        self._max_local_ids = max_local_ids        # rows of the output buffers (default: 1.25 x the first batch)
        W = self._world
        for name in ("tw", "rw"):                   # table slice per landing-zone SEGMENT (feature x source rank)
            self.register_buffer(f"_{name}_seg_row_base", getattr(self, f"_{name}_row_base").repeat_interleave(W),
                                 persistent=False)
            self.register_buffer(f"_{name}_seg_rows", getattr(self, f"_{name}_rows").repeat_interleave(W),
                                 persistent=False)

    def _seg_group(self, name: str):
        grp = self._group(name, False)
        grp.feat_row_base, grp.feat_rows = getattr(self, f"_{name}_seg_row_base"), getattr(self, f"_{name}_seg_rows")
        return grp

    # ---- over NVLink peer memory: the owners' lookups store every row at its id's own position in the sender's
    # output buffer; the gradient rows travel back the same way (tdfo_kjt_push_rows); no collective library call,
    # no index ops on the step
    def _seq_forward_peer(self, values, lengths, B):
        W, D, ops, r = self._world, self._D, self._ops, self._rank
        c = self._input_dist(values, lengths, B)
        peer, lo_, dev = c["peer"], c["lo"], c["dev"]
        z = peer.zones[c["slot"]]
        n_tw_r, n_rw, GB = c["n_tw_r"], self._n_rw, W * B
        L_local = values.numel()
        if L_local > peer.seq_cap:
            raise RuntimeError(f"tdfo_b200: this batch has {L_local} ids, the output buffers hold {peer.seq_cap}; "
                               "construct the module with a larger max_local_ids")
        k = peer.step % 2
        peer.step += 1
        self._fwd_token += 1
        pairs = None
        if self._emit_keys and hasattr(ops, "embedding_backward_prepare_pairs"):
            pairs = getattr(peer, "sort_pairs", None)
            if pairs is None:
                pairs = peer.sort_pairs = (torch.empty(max(peer.capacity, 4), dtype=torch.int32, device=dev),
                                           torch.empty(max(peer.capacity, 4), dtype=torch.int32, device=dev))
        if n_tw_r:
            ops.embedding_forward_seq_peer(self._seg_group("tw"), z.ids_x, lo_[: n_tw_r * GB + 1], B, z.pos_x,
                                           peer.seq_out_ptrs[k], W, sort_pairs=pairs)
        if n_rw:
            ops.embedding_forward_seq_peer(self._seg_group("rw"), z.ids_x, lo_[n_tw_r * GB: (n_tw_r + n_rw) * GB + 1], B,
                                           z.pos_x, peer.seq_out_ptrs[k], W, sort_pairs=pairs)
        peer.barrier()                                   # every row of my ids has landed
        out = peer.seq_out[k][: L_local * D].view(L_local, D)     # valid until the forward after next
        saved = dict(px=peer, slot=c["slot"], lo=lo_, n_tw_r=n_tw_r, GB=GB, pairs=pairs, seq=c["seq"],
                     token=self._fwd_token)
        return out, saved

    def _seq_backward_peer(self, grad_rows, saved, B):
        W, D, ops = self._world, self._D, self._ops
        if saved["token"] != self._fwd_token:
            raise RuntimeError("tdfo_b200: with peer-memory exchange, backward must run before the next forward")
        self._fused.iteration += 1
        opt = self._fused.args()
        px, lo_, n_tw_r, GB, seq = saved["px"], saved["lo"], saved["n_tw_r"], saved["GB"], saved["seq"]
        n_rw, dev = self._n_rw, grad_rows.device
        z = px.zones[saved["slot"]]
        z.counts_event.synchronize()
        tw_total, _, total = (int(x) for x in z.counts_host.tolist())
        if total > px.capacity:
            raise RuntimeError(f"tdfo_b200: rank {self._rank} received {total} ids this step, its exchange buffer "
                               f"holds {px.capacity}; construct the module with a larger max_ids_per_rank")
        # gradient rows to the owners of their ids, landing in the order the ids landed
        if sum(self._n_tw):
            ops.push_rows(grad_rows, None, seq["offsets"], B, self._peer_meta("gseq_tw", B, dev, saved["slot"]),
                          px.capacity, self._err_flag)
        if n_rw:
            ops.push_rows(grad_rows, seq["bpos"], seq["bo"], B, self._peer_meta("gseq_rw", B, dev, saved["slot"]),
                          px.capacity, self._err_flag)
        shards = [(name, s0, n, off) for name, have, s0, n, off in
                  (("tw", n_tw_r, 0, tw_total, lo_[: n_tw_r * GB + 1]),
                   ("rw", n_rw, tw_total, total - tw_total, lo_[n_tw_r * GB: (n_tw_r + n_rw) * GB + 1]))
                  if have and n > 0]
        prepared = {}
        for name, s0, n, off in shards:                  # the id-only half of the backward runs while the rows travel
            if saved["pairs"] is not None:
                prepared[name] = self._prepared_scratch(name, n, dev)
                ops.embedding_backward_prepare_pairs(self._group(name, False), saved["pairs"][0][s0:s0 + n],
                                                     saved["pairs"][1][s0:s0 + n], n, prepared[name])
        px.barrier()                                     # gradient rows have landed
        for name, s0, n, off in shards:
            grad = px.grad_seq[s0 * D:(s0 + n) * D].view(n, D)
            if name in prepared:
                ops.embedding_backward_update(self._group(name, True), grad, z.ids_x[s0:s0 + n], off, GB, _lib.POOL_NONE,
                                              opt, scratch=prepared[name], keys_in_scratch="prepared")
            else:
                ops.embedding_backward_update(self._group(name, True), grad, z.ids_x[s0:s0 + n], off - s0 if s0 else off,
                                              GB, _lib.POOL_NONE, opt)

    def _seq_forward(self, values, lengths, B):
        if self._use_peer:
            return self._seq_forward_peer(values, lengths, B)
        W, D, ops = self._world, self._D, self._ops
        c = self._input_dist(values, lengths, B)
        lv, lo_, tw_total, n_tw_r, dev = c["lv"], c["lo"], c["tw_total"], c["n_tw_r"], c["dev"]
        GB, n_rw = W * B, self._n_rw
        n_recv = lv.numel()
        rows_local = torch.empty((n_recv, D), dtype=torch.float32, device=dev)
        if n_tw_r and tw_total:
            ops.embedding_forward(self._group("tw", False), lv[:tw_total], lo_[: n_tw_r * GB + 1], GB, _lib.POOL_NONE,
                                  rows_local[:tw_total])
        rw = None
        if n_rw and n_recv > tw_total:
            rw = (lv[tw_total:], lo_[n_tw_r * GB:] - tw_total if n_tw_r else lo_)
            ops.embedding_forward(self._group("rw", False), rw[0], rw[1], GB, _lib.POOL_NONE, rows_local[tw_total:])
        routing = dict(self._routing)
        # back into arrival order, home with one all-to-all, then to the ids' original positions
        back = torch.empty_like(rows_local)
        back[routing["lv_pos"]] = rows_local
        home = self._a2a(back, routing["recv_splits"], routing["send_splits"], (D,))
        out = torch.zeros((values.numel(), D), dtype=torch.float32, device=dev)
        out[routing["send_pos"]] = home
        self._fwd_token += 1
        saved = dict(lv=lv, lo=lo_, tw_total=tw_total, n_tw_r=n_tw_r, GB=GB, rw=rw, routing=routing)
        return out, saved

    def _seq_backward(self, grad_rows, saved, B):
        if "px" in saved:
            return self._seq_backward_peer(grad_rows, saved, B)
        W, D, ops = self._world, self._D, self._ops
        self._fused.iteration += 1
        opt = self._fused.args()
        routing, tw_total, n_tw_r, GB = saved["routing"], saved["tw_total"], saved["n_tw_r"], saved["GB"]
        g_send = grad_rows[routing["send_pos"]]
        g_recv = self._a2a(g_send, routing["send_splits"], routing["recv_splits"], (D,))
        g_local = g_recv[routing["lv_pos"]]                       # [feature][source rank][batch] order
        if n_tw_r and tw_total:
            ops.embedding_backward_update(self._group("tw", True), g_local[:tw_total].contiguous(),
                                          saved["lv"][:tw_total], saved["lo"][: n_tw_r * GB + 1], GB, _lib.POOL_NONE, opt)
        if saved["rw"] is not None:
            rw_vals, rw_off = saved["rw"]
            ops.embedding_backward_update(self._group("rw", True), g_local[tw_total:].contiguous(), rw_vals, rw_off, GB,
                                          _lib.POOL_NONE, opt)

    def forward(self, features: KeyedJaggedTensor):
        from .sparse import JaggedTensor
        if list(features.keys()) != self._feature_names:
            keys = list(features.keys())
            features = features.permute([keys.index(k) for k in self._feature_names])
        values, lengths, B = features.values(), features.lengths(), features.stride()
        if self.training and torch.is_grad_enabled():
            if self._anchor is None or self._anchor.device != values.device:
                self._anchor = torch.zeros(1, device=values.device, requires_grad=True)
            self._emit_keys = isinstance(self._ops, CudaKernelOps)
            rows = _ShardedSequenceFn.apply(self._anchor, self, values, lengths, B)
        else:
            self._emit_keys = False
            rows, _ = self._seq_forward(values, lengths, B)
        names = self._feature_names
        offsets = features.offsets()
        if len(names) == 1:
            return {names[0]: JaggedTensor(rows, None, lengths, offsets)}
        bounds = features.offset_per_key()
        out = {}
        for i, k in enumerate(names):
            s0, e0 = bounds[i], bounds[i + 1]
            off = offsets[i * B:(i + 1) * B + 1]
            out[k] = JaggedTensor(rows[s0:e0], None, lengths[i * B:(i + 1) * B], off - s0 if s0 else off)
        return out


# --------------------------------------------------------------------------- DMP-style wrapper
class _Sharder:
    """`EmbeddingCollectionSharder(fused_params)` / `EmbeddingBagCollectionSharder(fused_params)`:
    carries the fused-optimizer settings exactly as the reference passes them
    (torchrec/train.py:236-245): {"optimizer": EmbOptimType, "learning_rate": float,
    "weight_decay": float, optional "eps", "beta1", "beta2"}."""

    module_kind: str = ""

    def __init__(self, fused_params: Optional[Dict[str, object]] = None):
        self.fused_params = dict(fused_params or {})

    def optimizer_kwargs(self) -> Dict[str, object]:
        fp = self.fused_params
        kw: Dict[str, object] = {}
        if "optimizer" in fp:
            kw["optimizer"] = fp["optimizer"]
        for src, dst in (("learning_rate", "learning_rate"), ("weight_decay", "weight_decay"), ("eps", "eps"),
                         ("beta1", "beta1"), ("beta2", "beta2")):
            if src in fp:
                kw[dst] = fp[src]
        return kw


class EmbeddingCollectionSharder(_Sharder):
    module_kind = "EmbeddingCollection"


class EmbeddingBagCollectionSharder(_Sharder):
    module_kind = "EmbeddingBagCollection"


class DistributedModelParallel(nn.Module):
    """`DMP(module=model, device=device, sharders=[...])` as at torchrec/train.py:241-247.

    * every EmbeddingBagCollection / EmbeddingCollection inside `module` gets the sharder's fused
      optimizer (the update then happens inside backward; `fused_optimizer` is the handle the
      reference puts into `CombinedOptimizer`);
    * with more than one rank, EmbeddingBagCollections are replaced by
      `ShardedEmbeddingBagCollection` (plan: `plan[table_name]`, default size-balanced table-wise)
      / EmbeddingCollections by `ShardedEmbeddingCollection`, and the remaining (dense)
      parameters are wrapped in DistributedDataParallel.
    """

    def __init__(self, module: nn.Module, device: Optional[torch.device] = None,
                 sharders: Optional[List[_Sharder]] = None, plan: Optional[Dict[str, ParameterSharding]] = None,
                 init_data_parallel: bool = True, process_group: Optional[dist.ProcessGroup] = None,
                 kernel_ops=None):
        super().__init__()
        from .modules import EmbeddingBagCollection, EmbeddingCollection
        from .optim import CombinedFusedOptimizer
        device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        by_kind = {s.module_kind: s for s in (sharders or [])}
        fused: List[FusedOptimizer] = []

        def adopt(old: nn.Module, new: nn.Module):
            """The sharded module starts from the tables the unsharded one holds (every rank built the same
            model, e.g. from one seed or one checkpoint), each rank keeping its rows / columns.  Modules
            built on the meta device have nothing to hand over."""
            sd = old.state_dict()
            if sd and not any(v.is_meta for v in sd.values()):
                new.load_state_dict(sd)

        def convert(parent: nn.Module):
            for name, child in list(parent.named_children()):
                if isinstance(child, EmbeddingBagCollection):
                    kw = by_kind["EmbeddingBagCollection"].optimizer_kwargs() if "EmbeddingBagCollection" in by_kind else {}
                    if world > 1:
                        tables = child.embedding_bag_configs()
                        p = plan or balanced_table_wise_plan(tables, world)
                        new = ShardedEmbeddingBagCollection(tables, p, device, process_group, kernel_ops=kernel_ops)
                        adopt(child, new)
                        setattr(parent, name, new)
                        child = new
                    else:
                        child.to(device)
                    fused.append(child.set_fused_optimizer(**kw) if kw else child.fused_optimizer)
                elif isinstance(child, EmbeddingCollection):
                    kw = by_kind["EmbeddingCollection"].optimizer_kwargs() if "EmbeddingCollection" in by_kind else {}
                    if world > 1:
                        tables = child.embedding_configs()
                        p = plan or balanced_table_wise_plan(tables, world)
                        new = ShardedEmbeddingCollection(tables, p, device, process_group, kernel_ops=kernel_ops)
                        adopt(child, new)
                        setattr(parent, name, new)
                        child = new
                    else:
                        child.to(device)
                    fused.append(child.set_fused_optimizer(**kw) if kw else child.fused_optimizer)
                else:
                    convert(child)

        holder = nn.ModuleDict({"m": module})
        convert(holder)
        module = holder["m"]
        module.to(device)
        self._fused_optimizer = CombinedFusedOptimizer(fused)
        dense = [p for p in module.parameters() if p.requires_grad]
        if world > 1 and init_data_parallel and dense:
            from torch.nn.parallel import DistributedDataParallel as DDP
            # DDP broadcasts every parameter and buffer of the module from rank 0 when it is built and keeps
            # buffers in sync afterwards: the shards (different on every rank, updated inside backward) and
            # the routing buffers of the sharded modules must stay out of that
            ignore = []
            for mod_name, mod in module.named_modules():
                if isinstance(mod, ShardedEmbeddingBagCollection):
                    pre = mod_name + "." if mod_name else ""
                    ignore += [pre + n for n, _ in mod.named_parameters(recurse=False)]
                    ignore += [pre + n for n, _ in mod.named_buffers(recurse=False)]
            DDP._set_params_and_buffers_to_ignore_for_model(module, ignore)
            module = DDP(module, device_ids=[device.index] if device.type == "cuda" else None,
                         process_group=process_group)
        self.module = module

    @property
    def fused_optimizer(self):
        return self._fused_optimizer

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)
