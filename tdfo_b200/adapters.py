"""Framework-neutral face of the lookup for the reference's jax-flax and tensorflow2 variants
(SURVEY.md §8f N3).

Their two-tower model does seven independent gathers per step, `nn.Embed(...)(ids)` at
jax-flax/models.py:78-88 and `keras.layers.Embedding(...)(ids)` at tensorflow2/models.py:48-58, each
`ids int32[B] -> [B, D]`, then all-reduces DENSE table gradients (`jax.lax.pmean(grads)`,
jax-flax/train_dp.py:63) and runs AdamW over every row (`apply_gradients`, train_dp.py:65;
tensorflow2/train_dp.py:72).  `DLPackEmbeddingTables` serves the same call shape from the B200
kernels:

* `lookup({name: ids})` does all the gathers in ONE kernel launch (a KeyedJaggedTensor with one id
  per bag) and returns `{name: [B, D]}`;
* `apply_gradients({name: ids}, {name: d loss / d rows})` is the sparse replacement for
  pmean + dense optimizer: the fused sort / unique / update touches only the looked-up rows.  With a
  process group the (ids, gradient rows) of every replica are all-gathered first, so replicated
  tables stay identical on every rank, and `average=True` reproduces pmean's 1 / world scaling.

Arrays cross the framework boundary through DLPack: inputs may be torch tensors, any object with
`__dlpack__` (jax.Array) or a raw DLPack capsule (`tf.experimental.dlpack.to_dlpack(t)`); outputs are
torch CUDA tensors, which jax / tf import zero-copy with `from_dlpack`.  TensorFlow's embedding
gradients arrive as `IndexedSlices`: `apply_indexed_slices` takes their (indices, values) as they are.
The row partition of tensorflow2/train_ps.py:55-61 (`MinSizePartitioner` over axis 0) is the `row_wise()`
plan of `ShardedEmbeddingBagCollection`.  Neither jax nor tensorflow is
installed on this box, so only the torch side of the protocol is exercised by the tests.

Semantics note: the fused ADAM is "lazy" (rows and moments of ids that do not occur in a step are
left alone), exactly as torchrec's fused optimizers are, whereas optax / keras AdamW decay every
row every step.  For rows that occur the arithmetic is AdamW with decoupled weight decay.
"""
from __future__ import annotations

from typing import Any, Dict, List, Mapping, Optional, Sequence, Tuple, Union

import torch
import torch.distributed as dist

from .modules import EmbeddingBagCollection, EmbeddingBagConfig, PoolingType
from .optim import EmbOptimType
from . import ops


def _as_tensor(x: Any, device: torch.device) -> torch.Tensor:
    if not isinstance(x, torch.Tensor):
        if type(x).__name__ == "PyCapsule":
            # a raw DLPack capsule: what `tf.experimental.dlpack.to_dlpack(tensor)` returns (TF 2.12 tensors
            # have no __dlpack__ method; tensorflow2/README.md:22 pins that version)
            x = torch.utils.dlpack.from_dlpack(x)
        elif hasattr(x, "__dlpack__"):
            x = torch.from_dlpack(x)
        else:
            raise TypeError(f"expected a torch tensor, a DLPack producer or a DLPack capsule, got {type(x).__name__}")
    if x.device != device:
        raise RuntimeError(f"array lives on {x.device}, the tables on {device}: hand over device arrays "
                           "(there is no host path)")
    return x


class DLPackEmbeddingTables:
    """`tables`: `{name: (num_embeddings, embedding_dim)}` (insertion order = output order)."""

    def __init__(self, tables: Mapping[str, Tuple[int, int]], device: Union[str, torch.device],
                 optimizer: EmbOptimType = EmbOptimType.ADAM, learning_rate: float = 1e-3, eps: float = 1e-8,
                 weight_decay: float = 0.0, beta1: float = 0.9, beta2: float = 0.999):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DLPackEmbeddingTables runs on CUDA only (no CPU path)")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.names: List[str] = list(tables)
        self.dims: List[int] = [int(tables[n][1]) for n in self.names]
        cfgs = [EmbeddingBagConfig(num_embeddings=int(tables[n][0]), embedding_dim=int(tables[n][1]), name=n,
                                   feature_names=[n], pooling=PoolingType.SUM) for n in self.names]
        self._ebc = EmbeddingBagCollection(cfgs, device=self.device)
        self._ebc.set_fused_optimizer(optimizer, learning_rate, eps, weight_decay, beta1, beta2)
        # kernel feature order (tables grouped by row size) vs the caller's order
        self._kernel_order: List[str] = list(self._ebc._kernel_order)
        self._offsets: Dict[int, torch.Tensor] = {}

    # ------------------------------------------------------------------ plumbing
    @property
    def module(self) -> EmbeddingBagCollection:
        return self._ebc

    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self._ebc.state_dict()

    def load_state_dict(self, sd: Mapping[str, torch.Tensor]) -> None:
        self._ebc.load_state_dict(dict(sd))

    def _pack_ids(self, ids: Mapping[str, Any]) -> Tuple[torch.Tensor, int]:
        cols = []
        for n in self._kernel_order:
            if n not in ids:
                raise KeyError(f"no ids for table {n!r}")
            v = _as_tensor(ids[n], self.device).reshape(-1)
            if v.dtype not in (torch.int32, torch.int64):
                raise TypeError(f"ids for {n!r} must be int32 or int64, got {v.dtype}")
            cols.append(v.to(torch.int64))
        B = cols[0].numel()
        if any(c.numel() != B for c in cols):
            raise ValueError("every table must be given the same number of ids")
        return torch.cat(cols), B

    def _unit_offsets(self, n: int) -> torch.Tensor:
        off = self._offsets.get(n)
        if off is None:
            off = self._offsets[n] = torch.arange(n + 1, dtype=torch.int64, device=self.device)
        return off

    # ------------------------------------------------------------------ the two calls
    @torch.no_grad()
    def lookup(self, ids: Mapping[str, Any]) -> Dict[str, torch.Tensor]:
        """`{name: ids [B]}` -> `{name: rows [B, D]}` (views of one [B, sum D] buffer; DLPack-exportable)."""
        values, B = self._pack_ids(ids)
        out = self._ebc._forward_impl(values, self._unit_offsets(len(self.names) * B), B, None)
        res, col = {}, 0
        for n, d in zip(self.names, self.dims):
            res[n] = out[:, col:col + d]
            col += d
        return res

    @torch.no_grad()
    def apply_gradients(self, ids: Mapping[str, Any], grads: Mapping[str, Any],
                        process_group: Optional[dist.ProcessGroup] = None, average: bool = True,
                        replicated: bool = False) -> None:
        """One fused optimizer step on the rows named by `ids` with `grads[name]` = d loss / d rows [B, D].
        `replicated=True`: tables are data-parallel replicas; gather every rank's (ids, rows) so all
        replicas apply the same update (`average` divides by the world size like pmean)."""
        values, B = self._pack_ids(ids)
        F = len(self.names)
        g = torch.cat([_as_tensor(grads[n], self.device).reshape(B, d).to(torch.float32)
                       for n, d in zip(self.names, self.dims)], dim=1).contiguous()
        if replicated:
            W = dist.get_world_size(process_group)
            if average:
                g = g / W
            all_v = torch.empty((W, F, B), dtype=torch.int64, device=self.device)
            dist.all_gather_into_tensor(all_v, values.view(F, B), group=process_group)
            all_g = torch.empty((W * B, g.shape[1]), dtype=torch.float32, device=self.device)
            dist.all_gather_into_tensor(all_g, g, group=process_group)
            values = all_v.permute(1, 0, 2).reshape(-1)          # [f][rank][b]: sample index = rank * B + b
            g, B = all_g, W * B
        self._ebc._backward_impl(g, values, self._unit_offsets(F * B), B)

    @torch.no_grad()
    def apply_indexed_slices(self, slices: Mapping[str, Any]) -> None:
        """One fused optimizer step from (indices, values) pairs, the shape TensorFlow hands an optimizer for an
        embedding variable: `tape.gradient` of `Embedding(...)(ids)` is a `tf.IndexedSlices(values [n, D],
        indices [n])` with duplicate indices left in and a different n per table (tensorflow2/models.py:48-58,
        train_dp.py:71-72).  `slices[name]` is such an object (anything with `.indices` and `.values`) or an
        `(indices, values)` tuple; the arrays are torch tensors, DLPack producers or DLPack capsules.  Tables not
        named are left alone.  Duplicates are summed in order of appearance, as the dense scatter-add would."""
        ebc = self._ebc
        ebc._fused.iteration += 1
        opt = ebc._fused.args()
        for g in ebc._groups:
            ids_parts, val_parts, counts = [], [], []
            for n in g.features:
                item = slices.get(n)
                if item is None:
                    counts.append(0)
                    continue
                idx, val = (item.indices, item.values) if hasattr(item, "indices") else item
                idx = _as_tensor(idx, self.device).reshape(-1)
                val = _as_tensor(val, self.device).reshape(-1, g.D).to(torch.float32)
                if idx.dtype not in (torch.int32, torch.int64) or idx.numel() != val.shape[0]:
                    raise ValueError(f"slices for {n!r}: need integer indices [n] and values [n, {g.D}]")
                ids_parts.append(idx.to(torch.int64))
                val_parts.append(val)
                counts.append(idx.numel())
            if not ids_parts:
                continue
            # one bag per table holding all of its slices; sequence-mode backward: gradient row p belongs to id p
            offsets = torch.tensor([0] + counts, dtype=torch.int64).cumsum(0).to(self.device)
            ops.embedding_backward_update(ebc._table_group(g, True), torch.cat(val_parts).contiguous(),
                                          torch.cat(ids_parts), offsets, 1, ops.POOL_NONE, opt)

    def check_errors(self) -> None:
        self._ebc.check_errors()
