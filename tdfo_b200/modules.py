"""``EmbeddingBagCollection`` / ``EmbeddingCollection`` with the torchrec call shapes the
reference uses (torchrec/models.py:150-164), backed by the sm_100a kernels.

Storage: every group of tables with the same (dim, dtype, pooling) lives in ONE
``[total_rows, D]`` parameter so a step is one forward launch and one backward
pipeline per group, not per table.  The optimizer is fused into the backward
(no dense gradient is ever produced), so these parameters never receive ``.grad``
and are tagged ``_in_backward_optimizers`` for ``in_backward_optimizer_filter``.
"""
from __future__ import annotations

import enum
import math
from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from . import _lib, ops
from .optim import EmbOptimType, FusedOptimizer
from .sparse import JaggedTensor, KeyedJaggedTensor, KeyedTensor


class PoolingType(enum.Enum):
    SUM = "SUM"
    MEAN = "MEAN"
    NONE = "NONE"


class BoundsCheckMode(enum.Enum):
    """What happens to an id outside its table (SURVEY.md §8b "Errors"; [UPSTREAM] fbgemm's
    BoundsCheckMode).  In every mode the kernels skip such an id (it contributes nothing to the
    lookup and no row is updated) and raise the device error word; the modes differ in who reads it:
    FATAL    the forward synchronises and raises IndexError at once (debugging; costs a sync per step)
    WARNING  nothing is read until `check_errors()`, which raises IndexError (default)
    IGNORE   `check_errors()` clears the word silently"""
    FATAL = "FATAL"
    WARNING = "WARNING"
    IGNORE = "IGNORE"


class DataType(enum.Enum):
    FP32 = "FP32"
    BF16 = "BF16"


_TORCH_DTYPE = {DataType.FP32: torch.float32, DataType.BF16: torch.bfloat16}
_POOL_CODE = {PoolingType.SUM: _lib.POOL_SUM, PoolingType.MEAN: _lib.POOL_MEAN, PoolingType.NONE: _lib.POOL_NONE}


@dataclass
class BaseEmbeddingConfig:
    num_embeddings: int
    embedding_dim: int
    name: str = ""
    feature_names: List[str] = field(default_factory=list)
    weight_init_max: Optional[float] = None
    weight_init_min: Optional[float] = None
    data_type: DataType = DataType.FP32

    def get_weight_init_max(self) -> float:
        return math.sqrt(1.0 / self.num_embeddings) if self.weight_init_max is None else self.weight_init_max

    def get_weight_init_min(self) -> float:
        return -math.sqrt(1.0 / self.num_embeddings) if self.weight_init_min is None else self.weight_init_min


@dataclass
class EmbeddingBagConfig(BaseEmbeddingConfig):
    pooling: PoolingType = PoolingType.SUM


@dataclass
class EmbeddingConfig(BaseEmbeddingConfig):
    """Sequence embedding table — torchrec/models.py:150-157."""


class _Group:
    """Tables of one (dim, dtype, pooling) and the KJT feature range they serve."""

    def __init__(self, index: int, D: int, dtype: torch.dtype, pooling: int):
        self.index, self.D, self.dtype, self.pooling = index, D, dtype, pooling
        self.tables: List[BaseEmbeddingConfig] = []
        self.table_row0: List[int] = []
        self.total_rows = 0
        self.features: List[str] = []
        self.feat_table: List[int] = []       # index into self.tables
        self.f0 = 0                            # first feature of the group in kernel (KJT) order

    @property
    def param_name(self) -> str:
        return f"_group{self.index}_weight"


class _LookupFn(torch.autograd.Function):
    """Autograd edge between the lookup output and the fused backward/update."""

    @staticmethod
    def forward(ctx, anchor, module, values, offsets, B, psw):
        ctx.module, ctx.B = module, B
        ctx.save_for_backward(values, offsets)
        ctx.psw = psw                      # constants of the step: the weights get no gradient
        out = module._forward_impl(values, offsets, B, psw, emit_keys=psw is None)
        ctx.keys_token = module._keys_token
        return out

    @staticmethod
    def backward(ctx, grad_out):
        values, offsets = ctx.saved_tensors
        ctx.module._backward_impl(grad_out.contiguous(), values, offsets, ctx.B, ctx.keys_token, ctx.psw)
        return None, None, None, None, None, None


class _DenseLookupFn(torch.autograd.Function):
    """Dense-gradient mode (`use_dense_gradients()`): the tables are ordinary trainable parameters and
    backward returns a dense [rows, D] gradient per table group, as the unsharded torchrec modules do
    in the reference's DDP branch (torchrec/train.py:255-260).  The gradient is built
    by the same sort / unique / segmented-reduce kernels: "update" a zero buffer with SGD at lr = -1."""

    @staticmethod
    def forward(ctx, module, values, offsets, B, psw, *weights):
        ctx.module, ctx.B, ctx.psw = module, B, psw
        ctx.save_for_backward(values, offsets)
        return module._forward_impl(values, offsets, B, psw)

    @staticmethod
    def backward(ctx, grad_out):
        values, offsets = ctx.saved_tensors
        grads = ctx.module._dense_backward(grad_out.contiguous(), values, offsets, ctx.B, ctx.psw)
        return (None, None, None, None, None) + tuple(grads)


class _FusedEmbeddingBase(nn.Module):
    _state_prefix = "embedding_bags"

    def __init__(self, tables: List[BaseEmbeddingConfig], device: Optional[torch.device], sequence: bool):
        super().__init__()
        if not tables:
            raise ValueError("at least one table is required")
        self._sequence = sequence
        self._configs = list(tables)
        names = [t.name for t in tables]
        if len(set(names)) != len(names):
            raise ValueError(f"duplicate table names: {names}")
        device = torch.device(device) if device is not None else torch.device("cpu")

        # module (output) order: tables in order, each with its features
        self._feature_names: List[str] = [f for t in tables for f in t.feature_names]
        if len(set(self._feature_names)) != len(self._feature_names):
            raise ValueError("a feature name is served by more than one table")
        self._feature_dims: List[int] = [t.embedding_dim for t in tables for _ in t.feature_names]
        out_col, acc = {}, 0
        for f, d in zip(self._feature_names, self._feature_dims):
            out_col[f] = acc
            acc += d
        self._total_D = acc

        groups: Dict[tuple, _Group] = {}
        for t in tables:
            pooling = _lib.POOL_NONE if sequence else _POOL_CODE[getattr(t, "pooling", PoolingType.SUM)]
            if not sequence and pooling == _lib.POOL_NONE:
                raise ValueError("EmbeddingBagCollection tables need SUM or MEAN pooling")
            key = (t.embedding_dim, _TORCH_DTYPE[t.data_type], pooling)
            g = groups.get(key)
            if g is None:
                g = groups[key] = _Group(len(groups), *key)
            g.table_row0.append(g.total_rows)
            g.tables.append(t)
            g.total_rows += t.num_embeddings
            for f in t.feature_names:
                g.features.append(f)
                g.feat_table.append(len(g.tables) - 1)
        self._groups = list(groups.values())
        if sequence and len(self._groups) != 1:
            raise ValueError("EmbeddingCollection tables must share embedding_dim and data_type")

        # kernel order of features = groups back to back
        self._kernel_order: List[str] = []
        for g in self._groups:
            g.f0 = len(self._kernel_order)
            self._kernel_order += g.features

        for g in self._groups:
            esz = 4 if g.dtype == torch.float32 else 2
            if (g.D * esz) % 16:
                raise ValueError(f"embedding_dim {g.D} x {esz} B must be a multiple of 16 bytes")
            w = torch.empty((g.total_rows, g.D), dtype=g.dtype, device=device)
            for t, r0 in zip(g.tables, g.table_row0):
                w[r0:r0 + t.num_embeddings].uniform_(t.get_weight_init_min(), t.get_weight_init_max())
            p = nn.Parameter(w, requires_grad=False)
            self.register_parameter(g.param_name, p)
            self.register_buffer(f"_group{g.index}_row_base", torch.tensor(
                [g.table_row0[ti] for ti in g.feat_table], dtype=torch.int64, device=device), persistent=False)
            self.register_buffer(f"_group{g.index}_rows", torch.tensor(
                [g.tables[ti].num_embeddings for ti in g.feat_table], dtype=torch.int64, device=device), persistent=False)
            self.register_buffer(f"_group{g.index}_out_col", torch.tensor(
                [out_col[f] for f in g.features], dtype=torch.int64, device=device), persistent=False)
        self.register_buffer("_err_flag", torch.zeros(1, dtype=torch.int32, device=device), persistent=False)
        self._anchor: Optional[torch.Tensor] = None
        self._fused: Optional[FusedOptimizer] = None
        # Backward workspace.  In training the forward kernel already writes the (row, bag) pairs the
        # backward sorts into it; `_keys_token` says which forward call they belong to.
        self._bw_scratch: Optional[torch.Tensor] = None
        self._keys_token = 0
        self.set_fused_optimizer(EmbOptimType.EXACT_SGD, learning_rate=0.01)

    # ------------------------------------------------------------------ optimizer
    def set_fused_optimizer(self, optimizer: EmbOptimType = EmbOptimType.EXACT_SGD, learning_rate: float = 0.01,
                            eps: float = 1e-8, weight_decay: float = 0.0, beta1: float = 0.9,
                            beta2: float = 0.999) -> FusedOptimizer:
        """Choose the optimizer fused into backward (the ``fused_params`` of torchrec/train.py:236-240)."""
        params = {g.param_name: getattr(self, g.param_name) for g in self._groups}
        self._fused = FusedOptimizer(params, optimizer, learning_rate, eps, weight_decay, beta1, beta2)
        for p in params.values():
            p._in_backward_optimizers = [self._fused]
        return self._fused

    @property
    def fused_optimizer(self) -> FusedOptimizer:
        return self._fused

    def use_dense_gradients(self) -> "_FusedEmbeddingBase":
        """Leave the fused mode: tables become trainable parameters with dense `.grad`, for callers that
        wrap the model in DDP and hand `model.parameters()` to a dense optimizer (the reference's
        non-model-parallel branch, torchrec/train.py:255-260).  O(rows x D) per step, like the reference."""
        for g in self._groups:
            if g.dtype != torch.float32:
                raise ValueError("dense gradients need fp32 tables")
            p = getattr(self, g.param_name)
            p.requires_grad_(True)
            if hasattr(p, "_in_backward_optimizers"):
                del p._in_backward_optimizers
        self._fused = None
        self._dense = True
        return self

    def _dense_backward(self, grad_out, values, offsets, B, psw):
        minus_one = ops.make_optimizer_args(_lib.OPT_SGD, -1.0)             # w <- 0 - (-1) * g
        grads = []
        for g in self._groups:
            F = len(g.features)
            dense = torch.zeros((g.total_rows, g.D), dtype=torch.float32, device=values.device)
            grp = ops.TableGroup(weights=dense, feat_row_base=getattr(self, f"_group{g.index}_row_base"),
                                 feat_rows=getattr(self, f"_group{g.index}_rows"),
                                 feat_out_col=getattr(self, f"_group{g.index}_out_col"),
                                 out_stride=g.D if self._sequence else self._total_D, err_flag=self._err_flag)
            off = offsets[g.f0 * B:(g.f0 + F) * B + 1]
            if len(self._groups) == 1:
                ops.embedding_backward_update(grp, grad_out, values, off, B, g.pooling, minus_one,
                                              per_sample_weights=psw)
            else:
                lo, hi = self._group_bounds(offsets, g, B)
                ops.embedding_backward_update(grp, grad_out, values[lo:hi], off - lo, B, g.pooling, minus_one,
                                              per_sample_weights=None if psw is None else psw[lo:hi])
            grads.append(dense)
        return grads

    def _state_for(self, g: _Group):
        """Optimizer state tensors of a group, allocated on first use next to the weights."""
        code = self._fused.optimizer.code()
        w = getattr(self, g.param_name)
        st = self._fused.state
        k0, k1 = f"{g.param_name}.state0", f"{g.param_name}.state1"
        if code == _lib.OPT_SGD:
            return None, None
        if code == _lib.OPT_ROWWISE_ADAGRAD:
            if k0 not in st or st[k0].device != w.device:
                st[k0] = torch.zeros(g.total_rows, dtype=torch.float32, device=w.device)
            return st[k0], None
        for k in (k0, k1):
            if k not in st or st[k].device != w.device:
                st[k] = torch.zeros((g.total_rows, g.D), dtype=torch.float32, device=w.device)
        return st[k0], st[k1]

    def _table_group(self, g: _Group, with_state: bool) -> ops.TableGroup:
        s0, s1 = self._state_for(g) if with_state else (None, None)
        return ops.TableGroup(
            weights=getattr(self, g.param_name).data,
            feat_row_base=getattr(self, f"_group{g.index}_row_base"),
            feat_rows=getattr(self, f"_group{g.index}_rows"),
            feat_out_col=getattr(self, f"_group{g.index}_out_col"),
            out_stride=g.D if self._sequence else self._total_D,
            state0=s0, state1=s1, err_flag=self._err_flag)

    # ------------------------------------------------------------------ compute
    def _scratch_for(self, L: int, total_rows: int, device) -> torch.Tensor:
        need = ops.backward_layout(L, total_rows).total_bytes
        if self._bw_scratch is None or self._bw_scratch.numel() < need or self._bw_scratch.device != device:
            self._bw_scratch = torch.empty(int(need * 1.1) + 256, dtype=torch.uint8, device=device)
        return self._bw_scratch

    def _forward_impl(self, values, offsets, B, psw, emit_keys: bool = False):
        dev = values.device
        scratch = None
        if emit_keys and len(self._groups) == 1 and values.numel() > 0:
            scratch = self._scratch_for(values.numel(), self._groups[0].total_rows, dev)
            self._keys_token += 1
        if self._sequence:
            g = self._groups[0]
            return ops.embedding_forward(self._table_group(g, False), values, offsets, B, _lib.POOL_NONE,
                                         backward_scratch=scratch)
        out = torch.empty((B, self._total_D), dtype=torch.float32, device=dev)
        for g in self._groups:
            F = len(g.features)
            off = offsets[g.f0 * B:(g.f0 + F) * B + 1]
            ops.embedding_forward(self._table_group(g, False), values, off, B, g.pooling, psw, out=out,
                                  backward_scratch=scratch)
        return out

    def _backward_impl(self, grad_out, values, offsets, B, keys_token: int = -1, psw=None):
        self._fused.iteration += 1
        opt = self._fused.args()
        for g in self._groups:
            F = len(g.features)
            off = offsets[g.f0 * B:(g.f0 + F) * B + 1]
            # group-local ids only: the kernels index `values` by absolute offsets, and
            # L bounds the sort, so pass the slice this group's offsets cover
            if len(self._groups) == 1:
                scratch = self._scratch_for(values.numel(), g.total_rows, values.device)
                # the pairs in `scratch` are still this call's unless another forward ran in between
                fused = keys_token == self._keys_token and values.numel() > 0 and psw is None
                ops.embedding_backward_update(self._table_group(g, True), grad_out, values, off, B, g.pooling, opt,
                                              scratch=scratch, keys_in_scratch=fused, per_sample_weights=psw)
                self._keys_token += 1          # the sort consumed them
            else:
                lo, hi = self._group_bounds(offsets, g, B)
                ops.embedding_backward_update(self._table_group(g, True), grad_out, values[lo:hi], off - lo, B,
                                              g.pooling, opt,
                                              per_sample_weights=None if psw is None else psw[lo:hi])

    @staticmethod
    def _group_bounds(offsets, g: _Group, B: int):
        ends = offsets[[g.f0 * B, (g.f0 + len(g.features)) * B]].tolist()
        return ends[0], ends[1]

    def _in_kernel_order(self, kjt: KeyedJaggedTensor) -> KeyedJaggedTensor:
        if list(kjt.keys()) != self._kernel_order:
            keys = list(kjt.keys())
            try:
                kjt = kjt.permute([keys.index(k) for k in self._kernel_order])
            except ValueError as e:
                raise KeyError(f"KeyedJaggedTensor keys {keys} do not cover features {self._kernel_order}") from e
        return kjt

    def _lookup(self, kjt: KeyedJaggedTensor) -> torch.Tensor:
        """``kjt`` must already be in kernel feature order."""
        values, offsets, B = kjt.values(), kjt.offsets(), kjt.stride()
        if not values.is_cuda or not getattr(self, self._groups[0].param_name).is_cuda:
            raise RuntimeError("tdfo_b200 embedding modules run on CUDA only (no CPU path): "
                               "move the module and the KeyedJaggedTensor to a cuda device")
        psw = kjt.weights_or_none() if getattr(self, "_is_weighted", False) else None
        if getattr(self, "_dense", False) and torch.is_grad_enabled():
            out = _DenseLookupFn.apply(self, values, offsets, B, psw, *[getattr(self, g.param_name) for g in self._groups])
        elif self.training and torch.is_grad_enabled():
            if self._anchor is None or self._anchor.device != values.device:
                self._anchor = torch.zeros(1, device=values.device, requires_grad=True)
            out = _LookupFn.apply(self._anchor, self, values, offsets, B, psw)
        else:
            out = self._forward_impl(values, offsets, B, psw)
        if self.bounds_check_mode == BoundsCheckMode.FATAL:
            ops.check_error_flag(self._err_flag)
        return out

    bounds_check_mode: BoundsCheckMode = BoundsCheckMode.WARNING

    def check_errors(self) -> None:
        """Synchronising check of the device error word (out-of-range ids)."""
        try:
            ops.check_error_flag(self._err_flag)
        except IndexError:
            if self.bounds_check_mode != BoundsCheckMode.IGNORE:
                raise

    # ------------------------------------------------------------------ state_dict: one entry per table
    def _table_slices(self):
        for g in self._groups:
            w = getattr(self, g.param_name)
            for t, r0 in zip(g.tables, g.table_row0):
                yield t.name, w, r0, r0 + t.num_embeddings

    def _save_to_state_dict(self, destination, prefix, keep_vars):
        for name, w, r0, r1 in self._table_slices():
            v = w[r0:r1]
            destination[f"{prefix}{self._state_prefix}.{name}.weight"] = v if keep_vars else v.detach()

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys,
                              error_msgs):
        for name, w, r0, r1 in self._table_slices():
            key = f"{prefix}{self._state_prefix}.{name}.weight"
            if key not in state_dict:
                if strict:
                    missing_keys.append(key)
                continue
            src = state_dict[key]
            if tuple(src.shape) != (r1 - r0, w.shape[1]):
                error_msgs.append(f"size mismatch for {key}: {tuple(src.shape)} vs {(r1 - r0, w.shape[1])}")
                continue
            with torch.no_grad():
                w[r0:r1].copy_(src)

    def embedding_configs(self) -> List[BaseEmbeddingConfig]:
        return self._configs


class EmbeddingBagCollection(_FusedEmbeddingBase):
    """Pooled lookup: ``forward(KeyedJaggedTensor) -> KeyedTensor [B, sum(D_f)]``."""
    _state_prefix = "embedding_bags"

    def __init__(self, tables: List[EmbeddingBagConfig], is_weighted: bool = False,
                 device: Optional[torch.device] = None):
        super().__init__(tables, device, sequence=False)
        self._is_weighted = is_weighted

    def embedding_bag_configs(self) -> List[EmbeddingBagConfig]:
        return self._configs

    def is_weighted(self) -> bool:
        return self._is_weighted

    def forward(self, features: KeyedJaggedTensor) -> KeyedTensor:
        return KeyedTensor(self._feature_names, self._feature_dims, self._lookup(self._in_kernel_order(features)))


class EmbeddingCollection(_FusedEmbeddingBase):
    """Sequence lookup: ``forward(KeyedJaggedTensor) -> Dict[feature, JaggedTensor]`` with one
    embedding row per id — torchrec/models.py:158-164."""
    _state_prefix = "embeddings"

    def __init__(self, tables: List[EmbeddingConfig], device: Optional[torch.device] = None):
        super().__init__(tables, device, sequence=True)

    def embedding_dim(self) -> int:
        return self._groups[0].D

    def forward(self, features: KeyedJaggedTensor) -> Dict[str, JaggedTensor]:
        features = self._in_kernel_order(features)
        rows = self._lookup(features)                       # [L, D] in kernel feature order
        order = self._kernel_order
        B = features.stride()
        out: Dict[str, JaggedTensor] = {}
        if len(order) == 1:
            out[order[0]] = JaggedTensor(rows, None, features.lengths(), features.offsets())
            return out
        bounds = features.offset_per_key()
        lengths, offsets = features.lengths(), features.offsets()
        for i, k in enumerate(order):
            s, e = bounds[i], bounds[i + 1]
            off = offsets[i * B:(i + 1) * B + 1]
            out[k] = JaggedTensor(rows[s:e], None, lengths[i * B:(i + 1) * B], off - s if s else off)
        return out
