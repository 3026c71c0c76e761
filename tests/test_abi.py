"""CPU-side checks of the drop-in boundary: the C-ABI library loads without a GPU, exports
every function include/tdfo_b200.h declares, and its host-only helpers behave."""
import ctypes
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "tdfo_b200.h"


def declared_functions():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(tdfo_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from tdfo_b200 import build, _lib
    build.build_library()
    return _lib.load()


def test_header_declares_the_path(lib):
    names = declared_functions()
    assert len(names) >= 18
    for must in ("tdfo_embedding_forward", "tdfo_embedding_backward_update", "tdfo_radix_sort_pairs",
                 "tdfo_unique_segments", "tdfo_kjt_bucketize_row_wise", "tdfo_jagged_2d_to_dense"):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    raw = ctypes.CDLL(str(ROOT / "tdfo_b200" / "libtdfo_b200.so"))
    missing = [n for n in declared_functions() if not hasattr(raw, n)]
    assert not missing, f"declared in the header but not exported: {missing}"


def test_bindings_cover_the_header(lib):
    from tdfo_b200 import _lib
    assert set(declared_functions()) <= set(_lib.exported_symbols())


def test_host_only_helpers(lib):
    from tdfo_b200 import _lib
    assert lib.tdfo_abi_version() == _lib.ABI_VERSION
    assert lib.tdfo_scan_scratch_bytes(0) % 256 == 0
    assert lib.tdfo_scan_scratch_bytes(10_000_000) >= 10_000_000 // 2048 * 8
    lay = _lib.BackwardLayout()
    assert lib.tdfo_backward_scratch_layout(1000, 26_000_000, ctypes.byref(lay)) == 0
    assert lay.key_bits == 25
    offs = sorted([lay.sorted_keys, lay.sorted_vals, lay.unique_keys, lay.seg_starts])
    assert len(set(offs)) == 4 and all(o % 256 == 0 for o in offs) and lay.total_bytes > lay.num_unique
    assert lib.tdfo_backward_scratch_layout(1 << 31, 10, ctypes.byref(lay)) != 0      # L too large
    assert b"2^30" in lib.tdfo_last_error()


def test_argument_errors_without_touching_the_gpu(lib):
    # rejected by host-side validation before any CUDA call
    rc = lib.tdfo_embedding_forward(None, 7, 10, 128, None, None, None, None, None, None, 1, 1, 0, 0, None, 128, None, None,
                                    None, None)
    assert rc == -1 and b"weights_dtype" in lib.tdfo_last_error()
    rc = lib.tdfo_radix_sort_pairs(None, None, None, None, 10, 40, None, None, 0, None)
    assert rc == -1 and b"key_bits" in lib.tdfo_last_error()


def test_no_cpu_path():
    import torch
    import tdfo_b200 as t
    ebc = t.EmbeddingBagCollection([t.EmbeddingBagConfig(num_embeddings=10, embedding_dim=8, name="a", feature_names=["fa"])])
    kjt = t.KeyedJaggedTensor.from_lengths_sync(["fa"], torch.arange(3), torch.tensor([1, 2]))
    with pytest.raises(RuntimeError, match="CUDA only"):
        ebc(kjt)


def test_product_never_imports_the_oracle():
    for p in (ROOT / "tdfo_b200").rglob("*.py"):
        assert "oracle" not in p.read_text(), f"{p} mentions the oracle"
    for p in (ROOT / "tdfo_b200" / "csrc").glob("*"):
        assert "oracle" not in p.read_text(), f"{p} mentions the oracle"


def test_new_entry_points_validate_on_the_host(lib):
    """Weighted scratch size, prepare / PREPARED flag: argument checks run before any CUDA call."""
    from tdfo_b200 import _lib
    base = lib.tdfo_backward_scratch_bytes(1000, 5000)
    wtd = lib.tdfo_backward_scratch_bytes_weighted(1000, 5000)
    assert wtd >= base + 1000 * 4 and wtd % 256 == 0
    assert lib.tdfo_backward_scratch_bytes_weighted(1 << 31, 10) == 0                # L too large
    # prepare: bad pooling mode, bad flag, weighted sequence mode
    P = None
    rc = lib.tdfo_embedding_backward_prepare(P, P, P, P, 1, 1, 10, 100, 9, 0, 0, P, P, 0, P)
    assert rc == -1 and b"pooling_mode" in lib.tdfo_last_error()
    rc = lib.tdfo_embedding_backward_prepare(P, P, P, P, 1, 1, 10, 100, _lib.POOL_SUM, 0, 2, P, P, 0, P)
    assert rc == -1 and b"keys_in_scratch" in lib.tdfo_last_error()
    rc = lib.tdfo_embedding_backward_prepare(P, P, P, P, 1, 1, 10, 100, _lib.POOL_NONE, 1, 0, P, P, 0, P)
    assert rc == -1 and b"weighted" in lib.tdfo_last_error()
    assert lib.tdfo_embedding_backward_prepare(P, P, P, P, 0, 0, 0, 100, _lib.POOL_SUM, 0, 0, P, P, 0, P) == 0   # nothing to do
    rc = lib.tdfo_embedding_backward_prepare(P, P, P, P, 1, 1, 10, 100, _lib.POOL_SUM, 0, 0, P, P, 0, P)
    assert rc == -1 and b"null pointer" in lib.tdfo_last_error()
    # update: keys_in_scratch out of range
    opt = _lib.OptimizerArgs(_lib.OPT_SGD, 0.1, 1e-8, 0.0, 0.9, 0.999, 1)
    rc = lib.tdfo_embedding_backward_update(P, _lib.F32, 100, 16, P, P, P, 16, P, P, P, P, P, P, 1, 1, 10, _lib.POOL_SUM,
                                            ctypes.byref(opt), 3, P, P, 0, P)
    assert rc == -1 and b"keys_in_scratch" in lib.tdfo_last_error()
    assert _lib.SCRATCH_PREPARED == 2
