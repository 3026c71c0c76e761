"""Tensor-level wrappers over the C ABI (``include/tdfo_b200.h``).

Torch is plumbing here: it owns device memory and streams; every op below is one
call into ``libtdfo_b200.so`` with raw pointers on the current CUDA stream.
CPU tensors are rejected — there is no fallback path.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass, field
from typing import Optional

import torch

from . import _lib
from ._lib import (BF16, F32, OPT_ADAM, OPT_ROWWISE_ADAGRAD, OPT_SGD, POOL_MEAN, POOL_NONE, POOL_SUM,
                   BackwardLayout, OptimizerArgs, check)

__all__ = [
    "TableGroup", "complete_cumsum", "embedding_forward", "embedding_backward_update", "linearize_indices",
    "radix_sort_pairs", "unique_segments", "sparse_update", "jagged_2d_to_dense", "dense_to_jagged",
    "jagged_1d_to_padded_dense", "bucketize_row_wise", "permute_2d", "launch_count", "check_error_flag",
]


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cuda(t: torch.Tensor, dtype: torch.dtype, name: str) -> torch.Tensor:
    if not t.is_cuda:
        raise RuntimeError(f"tdfo_b200: `{name}` must be a CUDA tensor (there is no CPU path), got {t.device}")
    if t.dtype != dtype:
        raise TypeError(f"tdfo_b200: `{name}` must be {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


_scratch: dict = {}


def _get_scratch(device: torch.device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) workspace, grown geometrically; never shrinks."""
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes * 1.25), 1 << 16), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def launch_count() -> int:
    return int(_lib.load().tdfo_launch_count())


def _wdtype(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise TypeError(f"tdfo_b200: tables must be float32 or bfloat16, got {t.dtype}")


@dataclass
class TableGroup:
    """All tables of one (dim, dtype) in one ``[total_rows, D]`` buffer plus per-feature metadata.

    ``feat_*`` are device int64 tensors of length F (features in KJT key order):
    first row / row count of the feature's table in ``weights`` and the first output column.
    """
    weights: torch.Tensor
    feat_row_base: torch.Tensor
    feat_rows: torch.Tensor
    feat_out_col: torch.Tensor
    out_stride: int
    state0: Optional[torch.Tensor] = None
    state1: Optional[torch.Tensor] = None
    err_flag: torch.Tensor = field(default=None)  # int32[1]

    def __post_init__(self):
        if self.err_flag is None:
            self.err_flag = torch.zeros(1, dtype=torch.int32, device=self.weights.device)

    @property
    def D(self) -> int:
        return self.weights.shape[1]

    @property
    def F(self) -> int:
        return self.feat_row_base.numel()

    @property
    def total_rows(self) -> int:
        return self.weights.shape[0]


def check_error_flag(flag: torch.Tensor, what: str = "embedding lookup") -> None:
    """Synchronising read of a device error word; raises like fbgemm's BoundsCheckMode.FATAL."""
    v = int(flag.item())
    if v & _lib.FLAG_SCAN_TIMEOUT:
        raise RuntimeError(f"tdfo_b200: internal scan watchdog fired during {what}")
    if v & _lib.FLAG_RECV_OVERFLOW:
        flag.zero_()
        raise RuntimeError(f"tdfo_b200: a rank received more ids than its exchange buffer holds during {what}; "
                           "raise max_ids_per_rank")
    if v & _lib.FLAG_OFFSETS_RANGE:
        flag.zero_()
        raise RuntimeError(f"tdfo_b200: offsets pointed past the ids given during {what} (the bags were cut there)")
    if v & _lib.FLAG_INDEX_OOB:
        flag.zero_()
        raise IndexError(f"tdfo_b200: an id was out of range of its table during {what}")


# --------------------------------------------------------------------------- KJT
def complete_cumsum(lengths: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lengths = _cuda(lengths, torch.int32, "lengths")
    n = lengths.numel()
    offsets = out if out is not None else torch.empty(n + 1, dtype=torch.int64, device=lengths.device)
    if offsets.numel() < n + 1 or offsets.dtype != torch.int64 or not offsets.is_contiguous():
        raise ValueError("complete_cumsum: `out` must be a contiguous int64 tensor with n+1 entries")
    lib = _lib.load()
    nb = lib.tdfo_scan_scratch_bytes(n)
    scratch = _get_scratch(lengths.device, nb)
    check(lib.tdfo_kjt_complete_cumsum(_ptr(lengths), n, _ptr(offsets), _ptr(scratch), scratch.numel(), _stream()))
    return offsets


def widen_ids(ids32: torch.Tensor) -> torch.Tensor:
    """int32 ids (the narrow host->device format) -> the int64 ids the lookups take."""
    ids32 = _cuda(ids32, torch.int32, "ids")
    out = torch.empty(ids32.numel(), dtype=torch.int64, device=ids32.device)
    check(_lib.load().tdfo_kjt_widen_ids(_ptr(ids32), ids32.numel(), _ptr(out), _stream()))
    return out


def bucketize_row_wise(lengths, offsets, indices, weights, block_sizes, F, B, world, want_permute=False,
                       err_flag=None):
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    block_sizes = _cuda(block_sizes, torch.int64, "block_sizes")
    if weights is not None:
        weights = _cuda(weights, torch.float32, "weights")
    dev = indices.device
    bags, L = F * B, indices.numel()
    new_lengths = torch.empty(bags * world, dtype=torch.int32, device=dev)
    new_offsets = torch.empty(bags * world + 1, dtype=torch.int64, device=dev)
    new_indices = torch.empty(L, dtype=torch.int64, device=dev)
    new_weights = None if weights is None else torch.empty(L, dtype=torch.float32, device=dev)
    perm = torch.empty(L, dtype=torch.int64, device=dev) if want_permute else None
    lib = _lib.load()
    scratch = _get_scratch(dev, lib.tdfo_bucketize_scratch_bytes(bags, world))
    check(lib.tdfo_kjt_bucketize_row_wise(_ptr(lengths), _ptr(offsets), _ptr(indices), _ptr(weights),
                                          _ptr(block_sizes), F, B, world, _ptr(new_lengths), _ptr(new_offsets),
                                          _ptr(new_indices), _ptr(new_weights), _ptr(perm), _ptr(err_flag),
                                          _ptr(scratch), scratch.numel(), _stream()))
    return new_lengths, new_offsets, new_indices, new_weights, perm


def push_lengths(lengths: torch.Tensor, B: int, meta: torch.Tensor) -> None:
    lengths = _cuda(lengths, torch.int32, "lengths")
    check(_lib.load().tdfo_kjt_push_lengths(_ptr(lengths), B, meta.numel() // 2, _ptr(meta), _stream()))


def push_ids(values: torch.Tensor, offsets: torch.Tensor, B: int, meta: torch.Tensor, capacity: int, err_flag) -> None:
    values = _cuda(values, torch.int64, "values")
    offsets = _cuda(offsets, torch.int64, "offsets")
    check(_lib.load().tdfo_kjt_push_ids(_ptr(values), _ptr(offsets), B, meta.numel() // 3, _ptr(meta), capacity,
                                        _ptr(err_flag), _stream()))


def push_rows(rows: torch.Tensor, positions: Optional[torch.Tensor], offsets: torch.Tensor, B: int, meta: torch.Tensor,
              capacity: int, err_flag) -> None:
    """Rows of ``rows`` ([n, D] float32), picked through ``positions`` (or in place order), to the owners' row buffers."""
    rows = _cuda(rows, torch.float32, "rows")
    offsets = _cuda(offsets, torch.int64, "offsets")
    if positions is not None:
        positions = _cuda(positions, torch.int64, "positions")
    check(_lib.load().tdfo_kjt_push_rows(_ptr(rows), _ptr(positions), _ptr(offsets), B, rows.shape[1], meta.numel() // 3,
                                         _ptr(meta), capacity, _ptr(err_flag), _stream()))


def scatter_positions(permute: torch.Tensor, positions_in: Optional[torch.Tensor], n: int,
                      n_dev: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out[permute[i]] = positions_in[i]`` (or ``i``) for i < min(n, n_dev[0])."""
    permute = _cuda(permute, torch.int64, "permute")
    if positions_in is not None:
        positions_in = _cuda(positions_in, torch.int64, "positions_in")
    out = torch.empty(max(n, 1), dtype=torch.int64, device=permute.device)
    check(_lib.load().tdfo_kjt_scatter_positions(_ptr(permute), _ptr(positions_in), n, _ptr(n_dev), _ptr(out), _stream()))
    return out


def embedding_forward_seq_peer(group: TableGroup, indices: torch.Tensor, offsets: torch.Tensor, B: int,
                               positions: torch.Tensor, peer_out: torch.Tensor, world: int,
                               sort_pairs: Optional[tuple] = None) -> None:
    """Sequence lookup on the owner with the rows stored at the ids' own positions in their source ranks' buffers.
    ``group.feat_row_base`` / ``feat_rows`` are per SEGMENT (local feature x source rank)."""
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    positions = _cuda(positions, torch.int64, "positions")
    peer_out = _cuda(peer_out, torch.int64, "peer_out")
    w = group.weights
    S = group.feat_row_base.numel()
    if offsets.numel() != S * B + 1:
        raise ValueError(f"offsets has {offsets.numel()} entries, expected S*B+1 = {S * B + 1}")
    sk, sv = (None, None) if sort_pairs is None else (_ptr(sort_pairs[0]), _ptr(sort_pairs[1]))
    check(_lib.load().tdfo_embedding_forward_seq_peer(
        _ptr(w), _wdtype(w), group.total_rows, group.D, _ptr(indices), _ptr(offsets), _ptr(group.feat_row_base),
        _ptr(group.feat_rows), S, B, indices.numel(), _ptr(positions), _ptr(peer_out), world, sk, sv,
        _ptr(group.err_flag), _stream()))


def permute_2d(permute, B, in_lengths, in_offsets, in_indices, in_weights=None, out_total: Optional[int] = None):
    """``out_total``: number of ids the output holds, if the caller knows it (saves a sync);
    otherwise it is read back from the computed offsets (one D2H sync)."""
    permute = _cuda(permute, torch.int32, "permute")
    in_lengths = _cuda(in_lengths, torch.int32, "in_lengths")
    in_offsets = _cuda(in_offsets, torch.int64, "in_offsets")
    in_indices = _cuda(in_indices, torch.int64, "in_indices")
    dev = in_lengths.device
    S_out = permute.numel()
    S_in = in_lengths.numel() // B if B else 0
    out_lengths = torch.empty(S_out * B, dtype=torch.int32, device=dev)
    out_offsets = torch.empty(S_out * B + 1, dtype=torch.int64, device=dev)
    lib = _lib.load()
    scratch = _get_scratch(dev, lib.tdfo_scan_scratch_bytes(S_out * B))
    if out_total is None:
        # lengths first, then size the data buffers from the last offset
        check(lib.tdfo_kjt_permute_2d(_ptr(permute), S_in, S_out, B, _ptr(in_lengths), _ptr(in_offsets), None, None,
                                      _ptr(out_lengths), _ptr(out_offsets), None, None, _ptr(scratch),
                                      scratch.numel(), _stream()))
        out_total = int(out_offsets[-1].item())
    out_indices = torch.empty(out_total, dtype=torch.int64, device=dev)
    out_weights = None
    if in_weights is not None:
        in_weights = _cuda(in_weights, torch.float32, "in_weights")
        out_weights = torch.empty(out_total, dtype=torch.float32, device=dev)
    check(lib.tdfo_kjt_permute_2d(_ptr(permute), S_in, S_out, B, _ptr(in_lengths), _ptr(in_offsets),
                                  _ptr(in_indices), _ptr(in_weights), _ptr(out_lengths), _ptr(out_offsets),
                                  _ptr(out_indices), _ptr(out_weights), _ptr(scratch), scratch.numel(), _stream()))
    return out_lengths, out_offsets, out_indices, out_weights


# --------------------------------------------------------------------------- lookup
def embedding_forward(group: TableGroup, indices: torch.Tensor, offsets: torch.Tensor, B: int, pooling_mode: int,
                      per_sample_weights: Optional[torch.Tensor] = None,
                      out: Optional[torch.Tensor] = None,
                      backward_scratch: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``backward_scratch``: a buffer of ``backward_layout(L, rows).total_bytes`` bytes; when given, the
    kernel also writes the linearised (row key, payload) pairs the backward sorts into it, so
    ``embedding_backward_update(..., keys_in_scratch=True)`` can skip its linearise stage."""
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    if per_sample_weights is not None:
        per_sample_weights = _cuda(per_sample_weights, torch.float32, "per_sample_weights")
    w = group.weights
    if not w.is_cuda:
        raise RuntimeError("tdfo_b200: embedding tables must live on a CUDA device (there is no CPU path)")
    L = indices.numel()
    if offsets.numel() != group.F * B + 1:
        raise ValueError(f"offsets has {offsets.numel()} entries, expected F*B+1 = {group.F * B + 1}")
    if out is None:
        if pooling_mode == POOL_NONE:
            out = torch.empty((L, group.D), dtype=torch.float32, device=w.device)
        else:
            out = torch.empty((B, group.out_stride), dtype=torch.float32, device=w.device)
    sk = sv = None
    if backward_scratch is not None and L > 0:
        lay = backward_layout(L, group.total_rows)
        if backward_scratch.numel() < lay.total_bytes:
            raise ValueError(f"backward_scratch has {backward_scratch.numel()} bytes, needs {lay.total_bytes}")
        sk = ctypes.c_void_p(backward_scratch.data_ptr() + lay.input_keys)
        sv = ctypes.c_void_p(backward_scratch.data_ptr() + lay.input_vals)
    check(_lib.load().tdfo_embedding_forward(
        _ptr(w), _wdtype(w), group.total_rows, group.D, _ptr(indices), _ptr(offsets), _ptr(per_sample_weights),
        _ptr(group.feat_row_base), _ptr(group.feat_rows), _ptr(group.feat_out_col), group.F, B, L, pooling_mode,
        _ptr(out), group.out_stride, sk, sv, _ptr(group.err_flag), _stream()))
    return out


def embedding_forward_peer(group: TableGroup, indices: torch.Tensor, offsets: torch.Tensor, B: int, pooling_mode: int,
                           peer_out: torch.Tensor, peer_batch: int,
                           backward_scratch: Optional[torch.Tensor] = None,
                           sort_pairs: Optional[tuple] = None, expected_ids: int = 0) -> None:
    """Pooled lookup whose rows are stored into the output buffers of the samples' source ranks
    (``peer_out``: int64 device tensor of base addresses, one per source rank)."""
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    peer_out = _cuda(peer_out, torch.int64, "peer_out")
    w = group.weights
    L = indices.numel()
    sk = sv = None
    if backward_scratch is not None and L > 0:
        lay = backward_layout(L, group.total_rows)
        if backward_scratch.numel() < lay.total_bytes:
            raise ValueError(f"backward_scratch has {backward_scratch.numel()} bytes, needs {lay.total_bytes}")
        sk = ctypes.c_void_p(backward_scratch.data_ptr() + lay.input_keys)
        sv = ctypes.c_void_p(backward_scratch.data_ptr() + lay.input_vals)
    if sort_pairs is not None:          # (keys, vals) int32 buffers of at least `indices.numel()` entries
        sk, sv = _ptr(sort_pairs[0]), _ptr(sort_pairs[1])
    check(_lib.load().tdfo_embedding_forward_peer(
        _ptr(w), _wdtype(w), group.total_rows, group.D, _ptr(indices), _ptr(offsets), None,
        _ptr(group.feat_row_base), _ptr(group.feat_rows), _ptr(group.feat_out_col), group.F, B, L, pooling_mode,
        _ptr(peer_out), peer_batch, group.out_stride, sk, sv, _ptr(group.err_flag), int(expected_ids), _stream()))


# --------------------------------------------------------------------------- backward
def make_optimizer_args(optimizer: int, learning_rate: float, eps: float = 1e-8, weight_decay: float = 0.0,
                        beta1: float = 0.9, beta2: float = 0.999, step: int = 1) -> OptimizerArgs:
    return OptimizerArgs(optimizer, learning_rate, eps, weight_decay, beta1, beta2, step)


def backward_layout(L: int, total_rows: int) -> BackwardLayout:
    lay = BackwardLayout()
    check(_lib.load().tdfo_backward_scratch_layout(L, total_rows, ctypes.byref(lay)))
    return lay


def embedding_backward_update(group: TableGroup, grad_out: torch.Tensor, indices: torch.Tensor,
                              offsets: torch.Tensor, B: int, pooling_mode: int, opt: OptimizerArgs,
                              scratch: Optional[torch.Tensor] = None, keys_in_scratch=False,
                              per_sample_weights: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Fused backward + optimizer, in place on ``group.weights`` / state.  Returns the scratch buffer
    (sorted keys / unique rows are readable at ``backward_layout`` offsets).  With
    ``per_sample_weights`` (pooled modes) each occurrence's gradient row is scaled by its weight.
    ``keys_in_scratch``: False / True (pairs written by the forward) / ``"prepared"`` (``embedding_backward_prepare``
    already sorted and run-length encoded into ``scratch``: only the update kernel runs)."""
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    grad_out = _cuda(grad_out, torch.float32, "grad_out")
    w = group.weights
    L = indices.numel()
    lib = _lib.load()
    if per_sample_weights is not None:
        per_sample_weights = _cuda(per_sample_weights, torch.float32, "per_sample_weights")
        if per_sample_weights.numel() != L:
            raise ValueError(f"per_sample_weights has {per_sample_weights.numel()} entries for {L} ids")
        nb = lib.tdfo_backward_scratch_bytes_weighted(L, group.total_rows)
    else:
        nb = lib.tdfo_backward_scratch_bytes(L, group.total_rows)
    if scratch is None or (per_sample_weights is not None and scratch.numel() < nb):
        scratch = _get_scratch(w.device, nb)
    grad_stride = group.D if pooling_mode == POOL_NONE else grad_out.shape[-1]
    check(lib.tdfo_embedding_backward_update(
        _ptr(w), _wdtype(w), group.total_rows, group.D, _ptr(group.state0), _ptr(group.state1),
        _ptr(grad_out), grad_stride, _ptr(indices), _ptr(offsets), _ptr(per_sample_weights),
        _ptr(group.feat_row_base), _ptr(group.feat_rows), _ptr(group.feat_out_col), group.F, B, L, pooling_mode,
        ctypes.byref(opt), _lib.SCRATCH_PREPARED if keys_in_scratch == "prepared" else (1 if keys_in_scratch else 0),
        _ptr(group.err_flag), _ptr(scratch), scratch.numel(), _stream()))
    return scratch


def embedding_backward_prepare(group: TableGroup, indices: torch.Tensor, offsets: torch.Tensor, B: int,
                               pooling_mode: int, scratch: torch.Tensor, keys_in_scratch: bool = False,
                               weighted: bool = False) -> None:
    """The id-only half of the backward (linearise, sort, unique) into ``scratch``; enqueue it on any stream
    while the gradient is still on its way, then call ``embedding_backward_update(..., scratch=scratch,
    keys_in_scratch="prepared")``."""
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    check(_lib.load().tdfo_embedding_backward_prepare(
        _ptr(indices), _ptr(offsets), _ptr(group.feat_row_base), _ptr(group.feat_rows), group.F, B, indices.numel(),
        group.total_rows, pooling_mode, 1 if weighted else 0, 1 if keys_in_scratch else 0, _ptr(group.err_flag),
        _ptr(scratch), scratch.numel(), _stream()))


def embedding_backward_prepare_pairs(group: TableGroup, keys: torch.Tensor, vals: torch.Tensor, L: int,
                                     scratch: torch.Tensor) -> None:
    """`embedding_backward_prepare` from (row key, payload) pairs the forward already wrote (int32 tensors holding
    uint32 bit patterns; only the first ``L`` entries are sorted)."""
    keys = _cuda(keys, torch.int32, "keys")
    vals = _cuda(vals, torch.int32, "vals")
    if keys.numel() < L or vals.numel() < L:
        raise ValueError(f"pair buffers hold {keys.numel()} / {vals.numel()} entries, fewer than L = {L}")
    check(_lib.load().tdfo_embedding_backward_prepare_pairs(_ptr(keys), _ptr(vals), L, group.total_rows,
                                                            _ptr(group.err_flag), _ptr(scratch), scratch.numel(), _stream()))


def linearize_indices(group: TableGroup, indices, offsets, B, pooling_mode):
    indices = _cuda(indices, torch.int64, "indices")
    offsets = _cuda(offsets, torch.int64, "offsets")
    L = indices.numel()
    keys = torch.empty(L, dtype=torch.int32, device=indices.device)   # uint32 bit patterns
    vals = torch.empty(L, dtype=torch.int32, device=indices.device)
    check(_lib.load().tdfo_linearize_indices(_ptr(indices), _ptr(offsets), _ptr(group.feat_row_base),
                                             _ptr(group.feat_rows), group.F, B, group.total_rows, pooling_mode,
                                             _ptr(keys), _ptr(vals), _ptr(group.err_flag), _stream()))
    return keys, vals


def radix_sort_pairs(keys: torch.Tensor, vals: torch.Tensor, key_bits: int, err_flag=None):
    """Stable sort of uint32 pairs held as int32 tensors; inputs are clobbered."""
    keys = _cuda(keys, torch.int32, "keys")
    vals = _cuda(vals, torch.int32, "vals")
    n = keys.numel()
    ko, vo = torch.empty_like(keys), torch.empty_like(vals)
    lib = _lib.load()
    scratch = _get_scratch(keys.device, lib.tdfo_radix_sort_scratch_bytes(n, key_bits))
    check(lib.tdfo_radix_sort_pairs(_ptr(keys), _ptr(vals), _ptr(ko), _ptr(vo), n, key_bits, _ptr(err_flag),
                                    _ptr(scratch), scratch.numel(), _stream()))
    return ko, vo


def unique_segments(sorted_keys: torch.Tensor, err_flag=None, sorted_vals: Optional[torch.Tensor] = None):
    """Returns (unique_keys, seg_starts, num_unique[, first_vals if ``sorted_vals`` is given])."""
    sorted_keys = _cuda(sorted_keys, torch.int32, "sorted_keys")
    n = sorted_keys.numel()
    dev = sorted_keys.device
    uk = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    ss = torch.empty(n + 1, dtype=torch.int32, device=dev)
    nu = torch.empty(1, dtype=torch.int32, device=dev)
    lib = _lib.load()
    scratch = _get_scratch(dev, lib.tdfo_scan_scratch_bytes(n))
    fv = None if sorted_vals is None else torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    check(lib.tdfo_unique_segments(_ptr(sorted_keys), _ptr(sorted_vals), n, _ptr(uk), _ptr(ss), _ptr(fv), _ptr(nu),
                                   _ptr(err_flag), _ptr(scratch), scratch.numel(), _stream()))
    return (uk, ss, nu) if sorted_vals is None else (uk, ss, nu, fv)


def sparse_update(group: TableGroup, grad_out, sorted_vals, unique_keys, seg_starts, first_vals, num_unique, offsets,
                  B, L, pooling_mode, opt: OptimizerArgs, worklist: Optional[torch.Tensor] = None):
    """``worklist``: optional int32 tensor of ``hot_row_worklist_entries(L)`` entries for the hot-row kernels."""
    grad_out = _cuda(grad_out, torch.float32, "grad_out")
    w = group.weights
    grad_stride = group.D if pooling_mode == POOL_NONE else grad_out.shape[-1]
    check(_lib.load().tdfo_sparse_update(
        _ptr(w), _wdtype(w), group.total_rows, group.D, _ptr(group.state0), _ptr(group.state1), _ptr(grad_out),
        grad_stride, _ptr(sorted_vals), _ptr(unique_keys), _ptr(seg_starts), _ptr(first_vals), _ptr(num_unique),
        _ptr(offsets),
        _ptr(group.feat_out_col), group.F, B, L, pooling_mode, ctypes.byref(opt), _ptr(worklist),
        0 if worklist is None else worklist.numel(), _stream()))


def hot_row_worklist_entries(L: int) -> int:
    return int(_lib.load().tdfo_hot_row_worklist_entries(L))


def copy_feature_blocks(src: torch.Tensor, dst: torch.Tensor, B: int, D: int, meta: torch.Tensor) -> None:
    """meta: int64[4*nslots] on the device = src_off | src_stride | dst_off | dst_stride (elements)."""
    src = _cuda(src, torch.float32, "src")
    if dst is not None and not (dst.is_cuda and dst.dtype == torch.float32 and dst.is_contiguous()):
        raise TypeError("dst must be a contiguous float32 CUDA tensor (or None for absolute destinations)")
    check(_lib.load().tdfo_copy_feature_blocks(_ptr(src), _ptr(dst), B, D, meta.numel() // 4, _ptr(meta), _stream()))


def copy_feature_blocks_multicast(src: torch.Tensor, B: int, D: int, meta: torch.Tensor) -> None:
    """`copy_feature_blocks` whose absolute destinations are NVLS multicast addresses (one store reaches every rank)."""
    src = _cuda(src, torch.float32, "src")
    check(_lib.load().tdfo_copy_feature_blocks_multicast(_ptr(src), B, D, meta.numel() // 4, _ptr(meta), _stream()))


def reduce_partial_sums(parts: torch.Tensor, W: int, B: int, nslots: int, D: int, out_col: torch.Tensor,
                        out: torch.Tensor, out_stride: int) -> None:
    """Row-wise output dist: ``out[b, out_col[k] + d] = sum_s parts[s, b, k*D + d]`` with s in rank order
    (``parts``: float32 [W, B, nslots*D], the slots the owners' lookups stored their partial sums into)."""
    parts = _cuda(parts, torch.float32, "parts")
    out_col = _cuda(out_col, torch.int64, "out_col")
    if not (out.is_cuda and out.dtype == torch.float32 and out.is_contiguous()):
        raise TypeError("out must be a contiguous float32 CUDA tensor")
    check(_lib.load().tdfo_reduce_partial_sums(_ptr(parts), W, B * nslots * D, B, nslots, D, _ptr(out_col), _ptr(out),
                                               out_stride, _stream()))


# --------------------------------------------------------------------------- jagged
def _jagged_2d_to_dense_raw(values, offsets, max_len):
    values = _cuda(values, torch.float32, "values")
    offsets = _cuda(offsets, torch.int64, "offsets")
    B, D = offsets.numel() - 1, values.shape[1]
    dense = torch.empty((B, max_len, D), dtype=torch.float32, device=values.device)
    check(_lib.load().tdfo_jagged_2d_to_dense(_ptr(values), _ptr(offsets), B, D, max_len, _ptr(dense), _stream()))
    return dense


def dense_to_jagged(dense, offsets, total_L: int):
    dense = _cuda(dense, torch.float32, "dense")
    offsets = _cuda(offsets, torch.int64, "offsets")
    B, max_len, D = dense.shape
    values = torch.empty((total_L, D), dtype=torch.float32, device=dense.device)
    check(_lib.load().tdfo_dense_to_jagged(_ptr(dense), _ptr(offsets), B, D, max_len, total_L, _ptr(values), _stream()))
    return values


class _Jagged2DToDense(torch.autograd.Function):
    @staticmethod
    def forward(ctx, values, offsets, max_len):
        ctx.save_for_backward(offsets)
        ctx.total_L = values.shape[0]
        return _jagged_2d_to_dense_raw(values, offsets, max_len)

    @staticmethod
    def backward(ctx, grad_dense):
        (offsets,) = ctx.saved_tensors
        return dense_to_jagged(grad_dense.contiguous(), offsets, ctx.total_L), None, None


def jagged_2d_to_dense(values: torch.Tensor, offsets: torch.Tensor, max_sequence_length: int) -> torch.Tensor:
    """Drop-in for ``torch.ops.fbgemm.jagged_2d_to_dense`` (torchrec/models.py:168-172), differentiable."""
    return _Jagged2DToDense.apply(values, offsets, int(max_sequence_length))


def jagged_1d_to_padded_dense(values, offsets, desired_length: int, padding_value: int):
    values = _cuda(values, torch.int64, "values")
    offsets = _cuda(offsets, torch.int64, "offsets")
    B = offsets.numel() - 1
    dense = torch.empty((B, desired_length), dtype=torch.int64, device=values.device)
    check(_lib.load().tdfo_jagged_1d_to_padded_dense(_ptr(values), _ptr(offsets), B, desired_length,
                                                     int(padding_value), _ptr(dense), _stream()))
    return dense
