"""tdfo_kjt_complete_cumsum_static: the lengths -> offsets scan for callers that cannot allocate.

Some users of this library run as a plug-in inside a framework that has already taken almost
all of HBM for its own pool (XLA / TensorFlow preallocation), or inside a serving loop where the
device-wide synchronisation of cudaMalloc is not acceptable.  They cannot provide the scratch
buffer the regular entry point wants, so this variant keeps its workspace in a small fixed array
inside the library.  The test drives it through the C ABI and checks it against the regular
entry point, bit for bit, on two back-to-back calls: the second call finds the first call's
look-back words still in the shared array, so it only passes if they are cleared properly.

Collected last on purpose (zz): a run that stops at the first failure has then already run
every other test.  It also has to pass on its own in a fresh process, where its first call is
the one that looks up the array's device address.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from tdfo_b200 import _lib
    return _lib.load()


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _reference_offsets(lib, lengths):
    """The existing entry point, with caller-provided scratch."""
    n = lengths.numel()
    out = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    nbytes = lib.tdfo_scan_scratch_bytes(n)
    scratch = torch.empty(max(nbytes, 256), dtype=torch.uint8, device="cuda")
    rc = lib.tdfo_kjt_complete_cumsum(_ptr(lengths), n, _ptr(out), _ptr(scratch), scratch.numel(), _stream())
    assert rc == 0, lib.tdfo_last_error()
    return out


def _static_offsets(lib, lengths):
    n = lengths.numel()
    out = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
    rc = lib.tdfo_kjt_complete_cumsum_static(_ptr(lengths), n, _ptr(out), _stream())
    assert rc == 0, lib.tdfo_last_error()
    return out


def test_static_cumsum_matches_regular_entry_point_twice_in_a_row(lib):
    gen = torch.Generator(device="cuda").manual_seed(7)
    # first call: many tiles, so plenty of look-back words are left behind in the shared array
    first = torch.randint(0, 40, (26 * 16384,), dtype=torch.int32, device="cuda", generator=gen)
    # second call right after it: fewer tiles and different values
    second = torch.randint(0, 500, (100_003,), dtype=torch.int32, device="cuda", generator=gen)
    got_first = _static_offsets(lib, first)
    got_second = _static_offsets(lib, second)
    torch.cuda.synchronize()
    assert torch.equal(got_first, _reference_offsets(lib, first))        # integer scan: bit-exact
    assert torch.equal(got_second, _reference_offsets(lib, second))
    assert int(got_second[-1]) == int(second.sum(dtype=torch.int64))
    # and against the CPU oracle's scan, so the static variant is checked by something other than its sibling
    import oracle as orc
    np.testing.assert_array_equal(got_first.cpu().numpy(), orc.complete_cumsum(first.cpu().numpy()))
    np.testing.assert_array_equal(got_second.cpu().numpy(), orc.complete_cumsum(second.cpu().numpy()))
    # and once more with the first input, now after a smaller call
    assert torch.equal(_static_offsets(lib, first), got_first)


def test_static_cumsum_edges_and_limit(lib):
    from tdfo_b200 import _lib
    empty = torch.empty(0, dtype=torch.int32, device="cuda")
    assert _static_offsets(lib, empty).tolist() == [0]
    one = torch.tensor([5], dtype=torch.int32, device="cuda")
    assert _static_offsets(lib, one).tolist() == [0, 5]
    # exactly at the limit is accepted ...
    big = torch.ones(_lib.CUMSUM_STATIC_MAX_N, dtype=torch.int32, device="cuda")
    out = _static_offsets(lib, big)
    assert int(out[-1]) == _lib.CUMSUM_STATIC_MAX_N and torch.equal(out, _reference_offsets(lib, big))
    # ... one more is refused before anything is launched, and the message names the limit
    too_big = torch.ones(_lib.CUMSUM_STATIC_MAX_N + 1, dtype=torch.int32, device="cuda")
    dst = torch.empty(too_big.numel() + 1, dtype=torch.int64, device="cuda")
    rc = lib.tdfo_kjt_complete_cumsum_static(_ptr(too_big), too_big.numel(), _ptr(dst), _stream())
    assert rc == -4
    msg = lib.tdfo_last_error().decode()
    assert str(_lib.CUMSUM_STATIC_MAX_N) in msg and "exceeds" in msg
