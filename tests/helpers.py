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


def fp64_backward_update(w, D, state0, state1, grad, indices, offsets, row_base, rows, out_col, F, B, mode,
                         optimizer, lr, eps=1e-8, wd=0.0, beta1=0.9, beta2=0.999, step=1, touched_only=True):
    """float64 restatement of the fused backward + optimizer, independent of both the
    C oracle and the kernels: np.add.at over occurrences, then the optimizer formulas.  It is the error
    budget's yardstick: a summation order cannot matter at fp64 for the sizes the tests use.
    `mode`: 0 sum, 1 mean, 2 none (grad is [L, D]).  Returns (w64, state0_64, state1_64, touched_mask, scale):
    `scale[row, d]` = |w_old| + (step multiplier) * sum over occurrences of |gradient term|, the magnitude of
    what enters each updated element.  An fp32 result is "within rtol" of this yardstick when
    |got - w64| <= rtol * scale: the forward-error form for a sum (relative to the sum of magnitudes, not
    to a result that cancellation can bring arbitrarily close to zero)."""
    w64 = np.asarray(w, dtype=np.float64).copy()
    total = w64.shape[0]
    L = int(indices.size)
    bag_of = np.repeat(np.arange(F * B, dtype=np.int64), np.diff(offsets).astype(np.int64))
    f_of = bag_of // B
    valid = (indices >= 0) & (indices < np.asarray(rows, dtype=np.int64)[f_of])
    key = np.asarray(row_base, dtype=np.int64)[f_of] + indices
    g = np.zeros((total, D), dtype=np.float64)
    if mode == 2:
        src = np.asarray(grad, dtype=np.float64)
    else:
        b_of = bag_of - f_of * B
        cols = np.asarray(out_col, dtype=np.int64)[f_of][:, None] + np.arange(D)[None, :]
        src = np.asarray(grad, dtype=np.float64)[b_of[:, None], cols]
        if mode == 1:
            src = src / np.diff(offsets).astype(np.float64)[bag_of][:, None]
    np.add.at(g, key[valid], src[valid])
    gabs = np.zeros((total, D), dtype=np.float64)
    np.add.at(gabs, key[valid], np.abs(src[valid]))
    scale = np.abs(w64).copy()
    touched = np.zeros(total, dtype=bool)
    touched[key[valid]] = True
    s0 = None if state0 is None else np.asarray(state0, dtype=np.float64).copy()
    s1 = None if state1 is None else np.asarray(state1, dtype=np.float64).copy()
    t = touched
    if optimizer == 0:                                   # SGD (L2 decay folded into the gradient)
        gg = g[t] + wd * w64[t]
        scale[t] += lr * gabs[t]
        w64[t] -= lr * gg
    elif optimizer == 1:                                 # row-wise Adagrad
        gg = g[t] + wd * w64[t]
        s0[t] += (gg * gg).sum(axis=1) / D
        scale[t] += (lr / (np.sqrt(s0[t]) + eps))[:, None] * gabs[t]
        w64[t] -= (lr / (np.sqrt(s0[t]) + eps))[:, None] * gg
    else:                                                # ADAM, decoupled decay, touched rows only
        gg = g[t]
        s0[t] = beta1 * s0[t] + (1 - beta1) * gg
        s1[t] = beta2 * s1[t] + (1 - beta2) * gg * gg
        mhat = s0[t] / (1 - beta1 ** step)
        vhat = s1[t] / (1 - beta2 ** step)
        scale[t] += lr
        w64[t] -= lr * (mhat / (np.sqrt(vhat) + eps) + wd * w64[t])
    return w64, s0, s1, touched, scale


def within_budget(got, want64, scale, rtol):
    """max over elements of |got - want64| / (rtol * scale); <= 1 means inside the budget."""
    return float((np.abs(np.asarray(got, dtype=np.float64) - want64) / (rtol * np.maximum(scale, 1e-30))).max())
