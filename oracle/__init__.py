"""numpy front-end of the CPU oracle (``oracle/tdfo_oracle.c``).

TEST INFRASTRUCTURE ONLY — see the header of ``tdfo_oracle.c``.  Imported by
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs;
never by ``tdfo_b200``.  Parity status: unpinned by the reference (it has no
tests); pinned against torch.nn.EmbeddingBag / FBGEMM's CPU reference by
``tests/golden/make_golden.py``.
"""
from __future__ import annotations

import ctypes
import subprocess
from ctypes import POINTER, c_float, c_int32, c_int64, c_void_p
from dataclasses import dataclass
from pathlib import Path

import numpy as np

_DIR = Path(__file__).resolve().parent
_LIB_PATH = _DIR / "libtdfo_oracle.so"

F32, BF16 = 0, 1
POOL_SUM, POOL_MEAN, POOL_NONE = 0, 1, 2
OPT_SGD, OPT_ROWWISE_ADAGRAD, OPT_ADAM = 0, 1, 2


def build(force: bool = False) -> Path:
    src = _DIR / "tdfo_oracle.c"
    if force or not _LIB_PATH.exists() or _LIB_PATH.stat().st_mtime < src.stat().st_mtime:
        r = subprocess.run(["make", "-C", str(_DIR), "-B" if force else "-s", "libtdfo_oracle.so"],
                           capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"oracle build failed:\n{r.stdout}\n{r.stderr}")
    return _LIB_PATH


class _OptArgs(ctypes.Structure):
    _fields_ = [("optimizer", c_int32), ("learning_rate", c_float), ("eps", c_float),
                ("weight_decay", c_float), ("beta1", c_float), ("beta2", c_float), ("step", c_int64)]


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(str(_LIB_PATH))
        _lib.orc_unique_segments.restype = c_int64
        for name in ("orc_bucketize_row_wise", "orc_embedding_forward", "orc_linearize_indices", "orc_embedding_backward_update_weighted",
                     "orc_embedding_backward_update"):
            getattr(_lib, name).restype = ctypes.c_int
    return _lib


def set_num_threads(n: int = 0) -> int:
    """Use `n` OpenMP threads (0: leave as is) in the oracle; returns the thread count in effect."""
    f = lib().orc_set_num_threads
    f.restype = ctypes.c_int
    return int(f(ctypes.c_int(n)))


def _p(a):
    return None if a is None else a.ctypes.data_as(c_void_p)


def _c(a, dtype):
    return None if a is None else np.ascontiguousarray(a, dtype=dtype)


@dataclass
class OptArgs:
    optimizer: int = OPT_SGD
    learning_rate: float = 0.01
    eps: float = 1e-8
    weight_decay: float = 0.0
    beta1: float = 0.9
    beta2: float = 0.999
    step: int = 1

    def c(self) -> _OptArgs:
        return _OptArgs(self.optimizer, self.learning_rate, self.eps, self.weight_decay,
                        self.beta1, self.beta2, self.step)


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit patterns (uint16), round to nearest even."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


def bf16_to_f32(h: np.ndarray) -> np.ndarray:
    return (h.astype(np.uint32) << 16).view(np.float32)


def complete_cumsum(lengths: np.ndarray) -> np.ndarray:
    lengths = _c(lengths, np.int32)
    out = np.empty(lengths.size + 1, dtype=np.int64)
    lib().orc_complete_cumsum(_p(lengths), c_int64(lengths.size), _p(out))
    return out


def bucketize_row_wise(offsets, indices, weights, block_sizes, F, B, world, want_permute=True):
    offsets, indices = _c(offsets, np.int64), _c(indices, np.int64)
    weights, block_sizes = _c(weights, np.float32), _c(block_sizes, np.int64)
    bags, L = F * B, indices.size
    new_lengths = np.zeros(bags * world, dtype=np.int32)
    new_offsets = np.zeros(bags * world + 1, dtype=np.int64)
    new_indices = np.zeros(L, dtype=np.int64)
    new_weights = None if weights is None else np.zeros(L, dtype=np.float32)
    perm = np.zeros(L, dtype=np.int64) if want_permute else None
    oob = lib().orc_bucketize_row_wise(_p(offsets), _p(indices), _p(weights), _p(block_sizes), c_int32(F),
                                       c_int64(B), c_int32(world), _p(new_lengths), _p(new_offsets),
                                       _p(new_indices), _p(new_weights), _p(perm))
    return new_lengths, new_offsets, new_indices, new_weights, perm, bool(oob)


def permute_2d(permute, B, in_lengths, in_offsets, in_indices, in_weights=None):
    permute, in_lengths = _c(permute, np.int32), _c(in_lengths, np.int32)
    in_offsets, in_indices = _c(in_offsets, np.int64), _c(in_indices, np.int64)
    in_weights = _c(in_weights, np.float32)
    S_out = permute.size
    out_lengths = np.zeros(S_out * B, dtype=np.int32)
    out_offsets = np.zeros(S_out * B + 1, dtype=np.int64)
    total = int(in_lengths.reshape(-1, B)[permute].sum()) if B and S_out else 0
    out_indices = np.zeros(total, dtype=np.int64)
    out_weights = None if in_weights is None else np.zeros(total, dtype=np.float32)
    lib().orc_permute_2d(_p(permute), c_int32(S_out), c_int64(B), _p(in_lengths), _p(in_offsets),
                         _p(in_indices), _p(in_weights), _p(out_lengths), _p(out_offsets),
                         _p(out_indices), _p(out_weights))
    return out_lengths, out_offsets, out_indices, out_weights


def _wdtype(weights: np.ndarray) -> int:
    if weights.dtype == np.float32:
        return F32
    if weights.dtype == np.uint16:
        return BF16
    raise TypeError("oracle weights must be float32 or uint16 (bf16 bit patterns)")


def embedding_forward(weights, D, indices, offsets, psw, feat_row_base, feat_rows, feat_out_col, F, B,
                      pooling_mode, out_stride=None):
    """weights: float32[rows, D] or uint16[rows, D] (bf16 bits).  Returns (out fp32, oob)."""
    assert weights.flags.c_contiguous
    indices, offsets = _c(indices, np.int64), _c(offsets, np.int64)
    psw = _c(psw, np.float32)
    frb, fr = _c(feat_row_base, np.int64), _c(feat_rows, np.int64)
    foc = _c(feat_out_col if feat_out_col is not None else np.zeros(F), np.int64)
    L = indices.size
    if pooling_mode == POOL_NONE:
        out = np.zeros((L, D), dtype=np.float32)
        out_stride = D
    else:
        if out_stride is None:
            out_stride = F * D
        out = np.zeros((B, out_stride), dtype=np.float32)
    oob = lib().orc_embedding_forward(_p(weights), c_int32(_wdtype(weights)), c_int32(D), _p(indices),
                                      _p(offsets), _p(psw), _p(frb), _p(fr), _p(foc), c_int32(F), c_int64(B),
                                      c_int64(L), c_int32(pooling_mode), _p(out), c_int64(out_stride))
    return out, bool(oob)


def linearize_indices(indices, offsets, feat_row_base, feat_rows, F, B, total_rows, pooling_mode):
    indices, offsets = _c(indices, np.int64), _c(offsets, np.int64)
    frb, fr = _c(feat_row_base, np.int64), _c(feat_rows, np.int64)
    keys = np.zeros(indices.size, dtype=np.uint32)
    vals = np.zeros(indices.size, dtype=np.uint32)
    oob = lib().orc_linearize_indices(_p(indices), _p(offsets), _p(frb), _p(fr), c_int32(F), c_int64(B),
                                      c_int64(total_rows), c_int32(pooling_mode), _p(keys), _p(vals))
    return keys, vals, bool(oob)


def sort_pairs_stable(keys, vals):
    keys, vals = _c(keys, np.uint32), _c(vals, np.uint32)
    ko, vo = np.zeros_like(keys), np.zeros_like(vals)
    lib().orc_sort_pairs_stable(_p(keys), _p(vals), c_int64(keys.size), _p(ko), _p(vo))
    return ko, vo


def unique_segments(sorted_keys):
    sorted_keys = _c(sorted_keys, np.uint32)
    n = sorted_keys.size
    uk = np.zeros(max(n, 1), dtype=np.uint32)
    ss = np.zeros(n + 1, dtype=np.uint32)
    nu = lib().orc_unique_segments(_p(sorted_keys), c_int64(n), _p(uk), _p(ss))
    return uk[:nu].copy(), ss[: nu + 1].copy(), int(nu)


def embedding_backward_update(weights, D, state0, state1, grad_out, indices, offsets, feat_row_base,
                              feat_rows, feat_out_col, F, B, pooling_mode, opt: OptArgs,
                              grad_stride=None, return_routing=False):
    """In-place on ``weights`` / ``state0`` / ``state1`` (numpy arrays).  Returns dict."""
    assert weights.flags.c_contiguous and weights.flags.writeable
    indices, offsets = _c(indices, np.int64), _c(offsets, np.int64)
    grad_out = _c(grad_out, np.float32)
    frb, fr = _c(feat_row_base, np.int64), _c(feat_rows, np.int64)
    foc = _c(feat_out_col if feat_out_col is not None else np.zeros(F), np.int64)
    total_rows = weights.shape[0]
    L = indices.size
    if grad_stride is None:
        grad_stride = D if pooling_mode == POOL_NONE else grad_out.shape[-1]
    n = max(L, 1)
    sk = np.zeros(n, np.uint32); sv = np.zeros(n, np.uint32)
    uk = np.zeros(n, np.uint32); ss = np.zeros(n + 1, np.uint32)
    nu = c_int64(0)
    c_opt = opt.c()
    oob = lib().orc_embedding_backward_update(
        _p(weights), c_int32(_wdtype(weights)), c_int64(total_rows), c_int32(D), _p(state0), _p(state1),
        _p(grad_out), c_int64(grad_stride), _p(indices), _p(offsets), _p(frb), _p(fr), _p(foc),
        c_int32(F), c_int64(B), c_int64(L), c_int32(pooling_mode), ctypes.byref(c_opt),
        _p(sk), _p(sv), _p(uk), _p(ss), ctypes.byref(nu))
    res = {"oob": bool(oob), "num_unique": int(nu.value)}
    if return_routing:
        res.update(sorted_keys=sk[:L], sorted_vals=sv[:L], unique_keys=uk[: nu.value],
                   seg_starts=ss[: nu.value + 1])
    return res


def embedding_backward_update_weighted(weights, D, state0, state1, grad_out, indices, offsets, psw,
                                       feat_row_base, feat_rows, feat_out_col, F, B, pooling_mode, opt: OptArgs):
    """Weighted pooled backward, in place.  Returns dict(oob, sorted_vals = positions in sorted order)."""
    assert weights.flags.c_contiguous and weights.flags.writeable and pooling_mode != POOL_NONE
    indices, offsets = _c(indices, np.int64), _c(offsets, np.int64)
    grad_out, psw = _c(grad_out, np.float32), _c(psw, np.float32)
    frb, fr, foc = _c(feat_row_base, np.int64), _c(feat_rows, np.int64), _c(feat_out_col, np.int64)
    L = indices.size
    assert psw.size == L
    sv = np.zeros(max(L, 1), np.uint32)
    c_opt = opt.c()
    oob = lib().orc_embedding_backward_update_weighted(
        _p(weights), c_int32(_wdtype(weights)), c_int64(weights.shape[0]), c_int32(D), _p(state0), _p(state1),
        _p(grad_out), c_int64(grad_out.shape[-1]), _p(indices), _p(offsets), _p(psw), _p(frb), _p(fr), _p(foc),
        c_int32(F), c_int64(B), c_int64(L), c_int32(pooling_mode), ctypes.byref(c_opt), _p(sv))
    return {"oob": bool(oob), "sorted_vals": sv[:L]}


def jagged_2d_to_dense(values, offsets, max_len):
    values, offsets = _c(values, np.float32), _c(offsets, np.int64)
    B, D = offsets.size - 1, values.shape[1]
    dense = np.zeros((B, max_len, D), dtype=np.float32)
    lib().orc_jagged_2d_to_dense(_p(values), _p(offsets), c_int64(B), c_int32(D), c_int64(max_len), _p(dense))
    return dense


def dense_to_jagged(dense, offsets):
    dense, offsets = _c(dense, np.float32), _c(offsets, np.int64)
    B, max_len, D = dense.shape
    values = np.zeros((int(offsets[-1]), D), dtype=np.float32)
    lib().orc_dense_to_jagged(_p(dense), _p(offsets), c_int64(B), c_int32(D), c_int64(max_len), _p(values))
    return values


def jagged_1d_to_padded_dense(values, offsets, desired_length, padding_value):
    values, offsets = _c(values, np.int64), _c(offsets, np.int64)
    B = offsets.size - 1
    dense = np.zeros((B, desired_length), dtype=np.int64)
    lib().orc_jagged_1d_to_padded_dense(_p(values), _p(offsets), c_int64(B), c_int64(desired_length),
                                        c_int64(padding_value), _p(dense))
    return dense


# --------------------------------------------------------------------------- float64 yardstick
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
