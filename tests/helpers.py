"""Shared builders for the parity tests (seeded jagged inputs in numpy)."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np


@dataclass
class Case:
    F: int
    B: int
    D: int
    rows: np.ndarray            # [T] rows per table (feature f -> table f)
    row_base: np.ndarray        # [F]
    out_col: np.ndarray         # [F]
    lengths: np.ndarray         # int32 [F*B]
    offsets: np.ndarray         # int64 [F*B+1]
    indices: np.ndarray         # int64 [L]
    weights: np.ndarray         # float32 [total_rows, D]

    @property
    def L(self) -> int:
        return int(self.indices.size)

    @property
    def total_rows(self) -> int:
        return int(self.rows.sum())


def make_case(seed: int, F: int, B: int, D: int, rows: Optional[List[int]] = None, max_len: int = 6,
              p_empty: float = 0.15, long_bag: int = 0, zipf: bool = False) -> Case:
    rng = np.random.default_rng(seed)
    if rows is None:
        rows = [int(x) for x in rng.integers(3, 200, size=F)]
    rows = np.asarray(rows, dtype=np.int64)
    lengths = rng.integers(0, max_len + 1, size=F * B).astype(np.int32)
    lengths[rng.random(F * B) < p_empty] = 0
    if long_bag and F * B > 2:
        lengths[rng.integers(0, F * B)] = long_bag
        lengths[-1] = long_bag // 2 + 1
    offsets = np.concatenate([[0], np.cumsum(lengths, dtype=np.int64)]).astype(np.int64)
    indices = np.zeros(int(offsets[-1]), dtype=np.int64)
    for f in range(F):
        s, e = int(offsets[f * B]), int(offsets[(f + 1) * B])
        if zipf:
            indices[s:e] = np.minimum(rng.zipf(1.2, size=e - s) - 1, rows[f] - 1)
        else:
            indices[s:e] = rng.integers(0, rows[f], size=e - s)
    row_base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
    out_col = (np.arange(F) * D).astype(np.int64)
    weights = rng.uniform(-1, 1, size=(int(rows.sum()), D)).astype(np.float32)
    return Case(F, B, D, rows, row_base, out_col, lengths, offsets, indices, weights)


from oracle import fp64_backward_update, within_budget  # noqa: E402,F401  (the float64 yardstick lives with the oracle)
