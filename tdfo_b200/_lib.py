"""ctypes binding of ``libtdfo_b200.so`` — the C ABI declared in ``include/tdfo_b200.h``.

There is no CPU path and no fallback: if the library is missing this module
raises, and every op refuses non-CUDA tensors.
"""
from __future__ import annotations

import ctypes
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_size_t, c_ulonglong, c_void_p
from pathlib import Path

import os

# TDFO_B200_LIB overrides the in-tree library (deployments that install it elsewhere, A/B builds)
LIB_PATH = Path(os.environ.get("TDFO_B200_LIB") or Path(__file__).resolve().parent / "libtdfo_b200.so")

# constants of include/tdfo_b200.h
F32, BF16 = 0, 1
POOL_SUM, POOL_MEAN, POOL_NONE = 0, 1, 2
OPT_SGD, OPT_ROWWISE_ADAGRAD, OPT_ADAM = 0, 1, 2
FLAG_INDEX_OOB, FLAG_SCAN_TIMEOUT, FLAG_RECV_OVERFLOW, FLAG_OFFSETS_RANGE = 1, 2, 4, 8
ABI_VERSION = 1
SCRATCH_PREPARED = 2
CUMSUM_STATIC_MAX_N = 1 << 22


class OptimizerArgs(Structure):
    _fields_ = [("optimizer", c_int32), ("learning_rate", c_float), ("eps", c_float),
                ("weight_decay", c_float), ("beta1", c_float), ("beta2", c_float), ("step", c_int64)]


class BackwardLayout(Structure):
    _fields_ = [("input_keys", c_size_t), ("input_vals", c_size_t), ("sorted_keys", c_size_t), ("sorted_vals", c_size_t), ("unique_keys", c_size_t),
                ("seg_starts", c_size_t), ("first_vals", c_size_t), ("num_unique", c_size_t),
                ("worklist", c_size_t), ("worklist_entries", c_size_t), ("total_bytes", c_size_t),
                ("key_bits", c_int32)]


P = c_void_p
_SIGNATURES = {
    "tdfo_abi_version": (c_int, []),
    "tdfo_last_error": (c_char_p, []),
    "tdfo_launch_count": (c_ulonglong, []),
    "tdfo_scan_scratch_bytes": (c_size_t, [c_int64]),
    "tdfo_kjt_complete_cumsum": (c_int, [P, c_int64, P, P, c_size_t, P]),
    "tdfo_kjt_complete_cumsum_static": (c_int, [P, c_int64, P, P]),
    "tdfo_bucketize_scratch_bytes": (c_size_t, [c_int64, c_int32]),
    "tdfo_kjt_bucketize_row_wise": (c_int, [P, P, P, P, P, c_int32, c_int64, c_int32, P, P, P, P, P, P, P, c_size_t, P]),
    "tdfo_kjt_permute_2d": (c_int, [P, c_int32, c_int32, c_int64, P, P, P, P, P, P, P, P, P, c_size_t, P]),
    "tdfo_kjt_push_lengths": (c_int, [P, c_int64, c_int32, P, P]),
    "tdfo_kjt_push_ids": (c_int, [P, P, c_int64, c_int32, P, c_int64, P, P]),
    "tdfo_embedding_forward": (c_int, [P, c_int32, c_int64, c_int32, P, P, P, P, P, P, c_int32, c_int64, c_int64,
                                       c_int32, P, c_int64, P, P, P, P]),
    "tdfo_embedding_forward_peer": (c_int, [P, c_int32, c_int64, c_int32, P, P, P, P, P, P, c_int32, c_int64, c_int64,
                                            c_int32, P, c_int64, c_int64, P, P, P, c_int64, P]),
    "tdfo_linearize_indices": (c_int, [P, P, P, P, c_int32, c_int64, c_int64, c_int32, P, P, P, P]),
    "tdfo_radix_sort_scratch_bytes": (c_size_t, [c_int64, c_int32]),
    "tdfo_radix_sort_pairs": (c_int, [P, P, P, P, c_int64, c_int32, P, P, c_size_t, P]),
    "tdfo_unique_segments": (c_int, [P, P, c_int64, P, P, P, P, P, P, c_size_t, P]),
    "tdfo_sparse_update": (c_int, [P, c_int32, c_int64, c_int32, P, P, P, c_int64, P, P, P, P, P, P, P, c_int32,
                                   c_int64, c_int64, c_int32, POINTER(OptimizerArgs), P, c_int64, P]),
    "tdfo_backward_scratch_bytes": (c_size_t, [c_int64, c_int64]),
    "tdfo_hot_row_worklist_entries": (c_size_t, [c_int64]),
    "tdfo_backward_scratch_bytes_weighted": (c_size_t, [c_int64, c_int64]),
    "tdfo_backward_scratch_layout": (c_int, [c_int64, c_int64, POINTER(BackwardLayout)]),
    "tdfo_embedding_backward_update": (c_int, [P, c_int32, c_int64, c_int32, P, P, P, c_int64, P, P, P, P, P, P,
                                               c_int32, c_int64, c_int64, c_int32, POINTER(OptimizerArgs), c_int32,
                                               P, P, c_size_t, P]),
    "tdfo_embedding_backward_prepare": (c_int, [P, P, P, P, c_int32, c_int64, c_int64, c_int64, c_int32, c_int32,
                                                c_int32, P, P, c_size_t, P]),
    "tdfo_kjt_widen_ids": (c_int, [P, c_int64, P, P]),
    "tdfo_kjt_push_rows": (c_int, [P, P, P, c_int64, c_int32, c_int32, P, c_int64, P, P]),
    "tdfo_kjt_scatter_positions": (c_int, [P, P, c_int64, P, P, P]),
    "tdfo_embedding_forward_seq_peer": (c_int, [P, c_int32, c_int64, c_int32, P, P, P, P, c_int32, c_int64, c_int64, P,
                                                P, c_int32, P, P, P, P]),
    "tdfo_embedding_backward_prepare_pairs": (c_int, [P, P, c_int64, c_int64, P, P, c_size_t, P]),
    "tdfo_copy_feature_blocks": (c_int, [P, P, c_int64, c_int32, c_int32, P, P]),
    "tdfo_copy_feature_blocks_multicast": (c_int, [P, c_int64, c_int32, c_int32, P, P]),
    "tdfo_reduce_partial_sums": (c_int, [P, c_int32, c_int64, c_int64, c_int32, c_int32, P, P, c_int64, P]),
    "tdfo_jagged_2d_to_dense": (c_int, [P, P, c_int64, c_int32, c_int64, P, P]),
    "tdfo_dense_to_jagged": (c_int, [P, P, c_int64, c_int32, c_int64, c_int64, P, P]),
    "tdfo_jagged_1d_to_padded_dense": (c_int, [P, P, c_int64, c_int64, c_int64, P, P]),
}

_lib = None


class TdfoError(RuntimeError):
    """A C-ABI call returned a TDFO_ERR_* code."""


def load() -> ctypes.CDLL:
    """Load the library once and attach the prototypes.  Fails loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m tdfo_b200.build` "
            "(tdfo_b200 has no CPU or PyTorch fallback)")
    lib = ctypes.CDLL(str(LIB_PATH))
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header / library mismatch
        fn.restype = res
        fn.argtypes = args
    if lib.tdfo_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libtdfo_b200.so has ABI {lib.tdfo_abi_version()}, bindings expect {ABI_VERSION}")
    _lib = lib
    return lib


def exported_symbols() -> list[str]:
    return sorted(_SIGNATURES)


def check(rc: int) -> None:
    if rc != 0:
        msg = load().tdfo_last_error().decode(errors="replace")
        raise TdfoError(f"tdfo_b200 error {rc}: {msg}")
