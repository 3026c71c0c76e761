"""Jagged containers with the torchrec names the reference uses.

``KeyedJaggedTensor`` / ``JaggedTensor`` / ``KeyedTensor`` mirror the surface called at
torchrec/train.py:38-40 and torchrec/models.py:163-178,208-212 of the reference.  They are
containers: they hold tensors on whatever device they were given (the reference builds the
KJT on the host and calls ``.to(device)``); all arithmetic on the lookup path happens in the
CUDA kernels behind ``tdfo_b200.ops``.

Layout: ``lengths`` is feature-major, ``lengths[f*B + b]``; ``offsets`` is the
complete cumulative sum (int64, ``F*B + 1`` entries); ``values`` is int64.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops


def _complete_cumsum(lengths: torch.Tensor) -> torch.Tensor:
    if lengths.is_cuda:
        return ops.complete_cumsum(lengths.to(torch.int32))
    # host-side container bookkeeping (the reference also builds its KJT on the host)
    out = torch.zeros(lengths.numel() + 1, dtype=torch.int64)
    torch.cumsum(lengths.to(torch.int64), 0, out=out[1:])
    return out


class JaggedTensor:
    """values [L] or [L, D] + lengths [B] / offsets [B+1]."""

    def __init__(self, values: torch.Tensor, weights: Optional[torch.Tensor] = None,
                 lengths: Optional[torch.Tensor] = None, offsets: Optional[torch.Tensor] = None):
        if lengths is None and offsets is None:
            raise ValueError("JaggedTensor needs lengths or offsets")
        self._values, self._weights, self._lengths, self._offsets = values, weights, lengths, offsets

    def values(self) -> torch.Tensor:
        return self._values

    def weights(self) -> torch.Tensor:
        if self._weights is None:
            raise ValueError("this JaggedTensor has no weights")
        return self._weights

    def weights_or_none(self) -> Optional[torch.Tensor]:
        return self._weights

    def lengths(self) -> torch.Tensor:
        if self._lengths is None:
            self._lengths = (self._offsets[1:] - self._offsets[:-1]).to(torch.int32)
        return self._lengths

    def offsets(self) -> torch.Tensor:
        if self._offsets is None:
            self._offsets = _complete_cumsum(self._lengths)
        return self._offsets

    def device(self) -> torch.device:
        return self._values.device

    def to(self, device, non_blocking: bool = False) -> "JaggedTensor":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        return JaggedTensor(mv(self._values), mv(self._weights), mv(self._lengths), mv(self._offsets))

    def record_stream(self, stream) -> None:
        for t in (self._values, self._weights, self._lengths, self._offsets):
            if t is not None and t.is_cuda:
                t.record_stream(stream)

    def to_padded_dense(self, desired_length: Optional[int] = None, padding_value=0) -> torch.Tensor:
        """[B, desired_length] (1-D values) — torchrec/models.py:210-212."""
        if desired_length is None:
            desired_length = int(self.lengths().max().item()) if self.lengths().numel() else 0
        v = self._values
        if v.dim() == 1 and v.dtype == torch.int64:
            return ops.jagged_1d_to_padded_dense(v, self.offsets(), int(desired_length), int(padding_value))
        if v.dim() == 2 and v.dtype == torch.float32:
            if padding_value != 0:
                raise NotImplementedError("2-D to_padded_dense pads with zeros only")
            return ops.jagged_2d_to_dense(v, self.offsets(), int(desired_length))
        raise TypeError(f"to_padded_dense supports int64 [L] or float32 [L, D] values, got {v.dtype} {tuple(v.shape)}")

    def to_dense(self) -> List[torch.Tensor]:
        off = self.offsets().tolist()
        return [self._values[off[i]:off[i + 1]] for i in range(len(off) - 1)]

    def __repr__(self) -> str:
        return f"JaggedTensor(values={tuple(self._values.shape)}, B={self.lengths().numel()})"


class KeyedJaggedTensor:
    """Feature-major jagged batch: keys [F], values [L], lengths [F*B]."""

    def __init__(self, keys: List[str], values: torch.Tensor, weights: Optional[torch.Tensor] = None,
                 lengths: Optional[torch.Tensor] = None, offsets: Optional[torch.Tensor] = None,
                 stride: Optional[int] = None, length_per_key: Optional[List[int]] = None):
        if lengths is None and offsets is None:
            raise ValueError("KeyedJaggedTensor needs lengths or offsets")
        self._keys = list(keys)
        self._values, self._weights, self._lengths, self._offsets = values, weights, lengths, offsets
        n = lengths.numel() if lengths is not None else offsets.numel() - 1
        F = len(self._keys)
        if stride is None:
            stride = n // F if F else 0
        if F and stride * F != n:
            raise ValueError(f"lengths has {n} entries, not a multiple of {F} keys")
        self._stride = stride
        self._length_per_key = length_per_key
        self._offset_per_key: Optional[List[int]] = None

    # ---- constructors used by the reference (torchrec/train.py:38-40)
    @staticmethod
    def from_lengths_sync(keys: List[str], values: torch.Tensor, lengths: torch.Tensor,
                          weights: Optional[torch.Tensor] = None) -> "KeyedJaggedTensor":
        return KeyedJaggedTensor(keys, values, weights, lengths=lengths.to(torch.int32))

    @staticmethod
    def from_offsets_sync(keys: List[str], values: torch.Tensor, offsets: torch.Tensor,
                          weights: Optional[torch.Tensor] = None) -> "KeyedJaggedTensor":
        return KeyedJaggedTensor(keys, values, weights, offsets=offsets.to(torch.int64))

    # ---- accessors
    def keys(self) -> List[str]:
        return self._keys

    def values(self) -> torch.Tensor:
        """int64 ids.  A batch built from int32 ids (half the host->device bytes; fine whenever every table
        has fewer than 2^31 rows) is widened here, once, by `tdfo_kjt_widen_ids` on the device."""
        if self._values.dtype == torch.int32:
            self._values = ops.widen_ids(self._values) if self._values.is_cuda else self._values.to(torch.int64)
        return self._values

    def weights(self) -> torch.Tensor:
        if self._weights is None:
            raise ValueError("this KeyedJaggedTensor has no weights")
        return self._weights

    def weights_or_none(self) -> Optional[torch.Tensor]:
        return self._weights

    def lengths(self) -> torch.Tensor:
        if self._lengths is None:
            self._lengths = (self._offsets[1:] - self._offsets[:-1]).to(torch.int32)
        return self._lengths

    def offsets(self) -> torch.Tensor:
        if self._offsets is None:
            self._offsets = _complete_cumsum(self._lengths)
        return self._offsets

    def stride(self) -> int:
        return self._stride

    def device(self) -> torch.device:
        return self._values.device

    def offset_per_key(self) -> List[int]:
        """Host list of F+1 value offsets (one device->host read, cached)."""
        if self._offset_per_key is None:
            F, B = len(self._keys), self._stride
            if F == 1:
                self._offset_per_key = [0, int(self._values.shape[0])]
            else:
                self._offset_per_key = self.offsets()[:: max(B, 1)][: F + 1].tolist() if B else [0] * (F + 1)
        return self._offset_per_key

    def length_per_key(self) -> List[int]:
        if self._length_per_key is None:
            o = self.offset_per_key()
            self._length_per_key = [o[i + 1] - o[i] for i in range(len(o) - 1)]
        return self._length_per_key

    # ---- movement
    def to(self, device, non_blocking: bool = False) -> "KeyedJaggedTensor":
        mv = lambda t: None if t is None else t.to(device, non_blocking=non_blocking)
        out = KeyedJaggedTensor(self._keys, mv(self._values), mv(self._weights), mv(self._lengths),
                                mv(self._offsets), self._stride, self._length_per_key)
        out._offset_per_key = self._offset_per_key
        return out

    def pin_memory(self) -> "KeyedJaggedTensor":
        pm = lambda t: None if t is None else t.pin_memory()
        return KeyedJaggedTensor(self._keys, pm(self._values), pm(self._weights), pm(self._lengths),
                                 pm(self._offsets), self._stride, self._length_per_key)

    def record_stream(self, stream) -> None:
        """Tell the caching allocator that `stream` uses this batch (it was filled on a copy stream)."""
        for t in (self._values, self._weights, self._lengths, self._offsets):
            if t is not None and t.is_cuda:
                t.record_stream(stream)

    # ---- views
    def __getitem__(self, key: str) -> JaggedTensor:
        f = self._keys.index(key)
        B = self._stride
        o = self.offset_per_key()
        s, e = o[f], o[f + 1]
        offsets = self.offsets()[f * B:(f + 1) * B + 1]
        if s:
            offsets = offsets - s
        w = None if self._weights is None else self._weights[s:e]
        return JaggedTensor(self.values()[s:e], w, self.lengths()[f * B:(f + 1) * B], offsets)

    def to_dict(self) -> Dict[str, JaggedTensor]:
        return {k: self[k] for k in self._keys}

    def permute(self, indices: List[int]) -> "KeyedJaggedTensor":
        """New KJT whose i-th key is ``keys[indices[i]]`` (keys may repeat)."""
        if not self._values.is_cuda:
            raise RuntimeError("KeyedJaggedTensor.permute runs on the GPU (tdfo_kjt_permute_2d); move the KJT first")
        perm = torch.tensor(indices, dtype=torch.int32, device=self._values.device)
        lpk = self.length_per_key() if self._length_per_key is not None else None
        total = sum(lpk[i] for i in indices) if lpk is not None else None
        ol, oo, oi, ow = ops.permute_2d(perm, self._stride, self.lengths(), self.offsets(), self.values(),
                                        self._weights, out_total=total)
        return KeyedJaggedTensor([self._keys[i] for i in indices], oi, ow, ol, oo, self._stride,
                                 None if lpk is None else [lpk[i] for i in indices])

    def split(self, segments: List[int]) -> List["KeyedJaggedTensor"]:
        out, f0 = [], 0
        B = self._stride
        o = self.offset_per_key()
        for seg in segments:
            f1 = f0 + seg
            s, e = o[f0], o[f1]
            w = None if self._weights is None else self._weights[s:e]
            out.append(KeyedJaggedTensor(self._keys[f0:f1], self.values()[s:e], w,
                                         lengths=self.lengths()[f0 * B:f1 * B], stride=B))
            f0 = f1
        return out

    def __repr__(self) -> str:
        return f"KeyedJaggedTensor(keys={self._keys}, stride={self._stride}, values={tuple(self._values.shape)})"


class KeyedTensor:
    """Pooled output: values [B, sum(D_f)], key f owns columns [offset_f, offset_f + D_f)."""

    def __init__(self, keys: List[str], length_per_key: List[int], values: torch.Tensor, key_dim: int = 1):
        self._keys, self._length_per_key, self._values, self._key_dim = list(keys), list(length_per_key), values, key_dim
        self._offsets = [0]
        for n in self._length_per_key:
            self._offsets.append(self._offsets[-1] + n)

    def keys(self) -> List[str]:
        return self._keys

    def values(self) -> torch.Tensor:
        return self._values

    def key_dim(self) -> int:
        return self._key_dim

    def length_per_key(self) -> List[int]:
        return self._length_per_key

    def offset_per_key(self) -> List[int]:
        return self._offsets

    def __getitem__(self, key: str) -> torch.Tensor:
        i = self._keys.index(key)
        return self._values.narrow(self._key_dim, self._offsets[i], self._length_per_key[i])

    def to_dict(self) -> Dict[str, torch.Tensor]:
        return {k: self[k] for k in self._keys}

    def __repr__(self) -> str:
        return f"KeyedTensor(keys={self._keys}, values={tuple(self._values.shape)})"
