"""Property tests (hypothesis) of the oracle's routing stages against plain-Python restatements of the
published semantics ([UPSTREAM] fbgemm `block_bucketize_sparse_features`, `permute_2D_sparse_data`,
`jagged_2d_to_dense`, `asynchronous_complete_cumsum`): empty bags, empty
features, repeated permute entries, tables smaller than the world, truncation.  The oracle is what the
CUDA kernels are compared with bit for bit, so it gets an independent check of its own."""
import numpy as np
from hypothesis import given, settings, strategies as st

import oracle as orc

SET = settings(max_examples=60, deadline=None)


@st.composite
def jagged(draw, max_F=4, max_B=5, max_len=4, max_rows=12):
    F, B = draw(st.integers(1, max_F)), draw(st.integers(1, max_B))
    rows = [draw(st.integers(1, max_rows)) for _ in range(F)]
    lengths = np.array([draw(st.integers(0, max_len)) for _ in range(F * B)], dtype=np.int32)
    ids = []
    for f in range(F):
        n = int(lengths[f * B:(f + 1) * B].sum())
        ids += [draw(st.integers(0, rows[f] - 1)) for _ in range(n)]
    return F, B, rows, lengths, np.array(ids, dtype=np.int64)


@SET
@given(st.lists(st.integers(0, 7), min_size=0, max_size=40))
def test_complete_cumsum(xs):
    got = orc.complete_cumsum(np.array(xs, dtype=np.int32))
    assert got.tolist() == [sum(xs[:i]) for i in range(len(xs) + 1)]


@SET
@given(jagged(), st.integers(1, 4), st.data())
def test_bucketize_row_wise(case, world, data):
    F, B, rows, lengths, ids = case
    block = [(r + world - 1) // world for r in rows]
    weights = np.arange(ids.size, dtype=np.float32)
    off = orc.complete_cumsum(lengths)
    nl, no, ni, nw, perm, oob = orc.bucketize_row_wise(off, ids, weights, block, F, B, world)
    assert not oob
    # restatement: output bag (bucket, f, b) holds the ids of bag (f, b) whose row block is `bucket`,
    # in their original order, rebased to the block; unbucketize_permute[p] = where id p went
    want_len, want_idx, want_w, want_perm = [], [], [], [None] * ids.size
    for bucket in range(world):
        for f in range(F):
            for b in range(B):
                g = f * B + b
                n = 0
                for p in range(int(off[g]), int(off[g + 1])):
                    if ids[p] // block[f] == bucket:
                        want_perm[p] = len(want_idx)
                        want_idx.append(int(ids[p]) - bucket * block[f])
                        want_w.append(float(weights[p]))
                        n += 1
                want_len.append(n)
    assert nl.tolist() == want_len and ni.tolist() == want_idx and nw.tolist() == want_w
    assert perm.tolist() == want_perm and no.tolist() == orc.complete_cumsum(np.array(want_len, np.int32)).tolist()


@SET
@given(jagged(), st.data())
def test_permute_2d(case, data):
    F, B, rows, lengths, ids = case
    permute = data.draw(st.lists(st.integers(0, F - 1), min_size=0, max_size=6))      # repeats allowed
    off = orc.complete_cumsum(lengths)
    ol, oo, oi, _ = orc.permute_2d(np.array(permute, np.int32), B, lengths, off, ids)
    want_len, want_idx = [], []
    for f in permute:
        for b in range(B):
            g = f * B + b
            want_len.append(int(lengths[g]))
            want_idx += ids[int(off[g]):int(off[g + 1])].tolist()
    assert ol.tolist() == want_len and oi[:len(want_idx)].tolist() == want_idx


@SET
@given(st.lists(st.integers(0, 6), min_size=1, max_size=6), st.integers(0, 7), st.integers(1, 3))
def test_jagged_dense_roundtrip(lens, max_len, D):
    off = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
    L = int(off[-1])
    values = (np.arange(L * D, dtype=np.float32) + 1).reshape(L, D)
    dense = orc.jagged_2d_to_dense(values, off, max_len)
    assert dense.shape == (len(lens), max_len, D)
    for b, n in enumerate(lens):
        for t_ in range(max_len):
            want = values[off[b] + t_] if t_ < n else np.zeros(D, np.float32)
            assert dense[b, t_].tolist() == want.tolist()
    back = orc.dense_to_jagged(dense, off)                       # rows cut off by max_len come back as zeros
    for b, n in enumerate(lens):
        for t_ in range(n):
            want = values[off[b] + t_] if t_ < max_len else np.zeros(D, np.float32)
            assert back[off[b] + t_].tolist() == want.tolist()
    ids = np.arange(L, dtype=np.int64) + 5
    padded = orc.jagged_1d_to_padded_dense(ids, off, max_len, -1)
    for b, n in enumerate(lens):
        assert padded[b].tolist() == [int(ids[off[b] + t_]) if t_ < n else -1 for t_ in range(max_len)]


@SET
@given(jagged(max_rows=6))
def test_linearize_sort_unique_pipeline(case):
    F, B, rows, lengths, ids = case
    off = orc.complete_cumsum(lengths)
    base = np.concatenate([[0], np.cumsum(rows)[:-1]]).astype(np.int64)
    total = int(sum(rows))
    for mode in (orc.POOL_SUM, orc.POOL_NONE):
        keys, vals, oob = orc.linearize_indices(ids, off, base, rows, F, B, total, mode)
        assert not oob
        want_keys, want_vals = [], []
        for f in range(F):
            for b in range(B):
                g = f * B + b
                for p in range(int(off[g]), int(off[g + 1])):
                    want_keys.append(int(base[f] + ids[p]))
                    want_vals.append(p if mode == orc.POOL_NONE else g)
        assert keys.tolist() == want_keys and vals.tolist() == want_vals
        sk, sv = orc.sort_pairs_stable(keys, vals)
        order = sorted(range(len(want_keys)), key=lambda i: want_keys[i])          # Python's sort is stable
        assert sk.tolist() == [want_keys[i] for i in order] and sv.tolist() == [want_vals[i] for i in order]
        uk, ss, nu = orc.unique_segments(sk)
        uniq = sorted(set(want_keys))
        assert nu == len(uniq) and uk[:nu].tolist() == uniq
        assert ss[:nu + 1].tolist() == [sk.tolist().index(k) for k in uniq] + [len(want_keys)]


def test_parallel_sort_and_unique_equal_numpy_for_any_thread_count():
    """Above 65536 pairs the oracle's sort / run-length stages run on all OpenMP threads (chunked counting
    sort, chunked head count): same permutation as numpy's stable argsort whatever the thread count."""
    rng = np.random.default_rng(5)
    n = 300_007
    keys = rng.integers(0, 70_000, size=n).astype(np.uint32)          # many duplicates, both 16-bit digits in use
    keys[rng.integers(0, n, size=50)] = np.uint32(0xfffffff0)
    vals = np.arange(n, dtype=np.uint32)
    order = np.argsort(keys, kind="stable")
    heads = np.flatnonzero(np.concatenate([[True], keys[order][1:] != keys[order][:-1]]))
    before = orc.set_num_threads(0)
    try:
        for threads in (1, 3, 8):
            orc.set_num_threads(threads)
            ko, vo = orc.sort_pairs_stable(keys, vals)
            np.testing.assert_array_equal(ko, keys[order])
            np.testing.assert_array_equal(vo, vals[order])
            uk, ss, nu = orc.unique_segments(ko)
            assert nu == heads.size
            np.testing.assert_array_equal(uk, keys[order][heads])
            np.testing.assert_array_equal(ss, np.concatenate([heads, [n]]).astype(np.uint32))
    finally:
        orc.set_num_threads(before)
