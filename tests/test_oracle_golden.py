"""The CPU oracle against the committed known answers (tests/golden/make_golden.py):
torch.nn.EmbeddingBag / nn.Embedding, torch SGD / AdamW, FBGEMM's rowwise_sparse_adagrad_ref
and the hand-computed routing cases.  This is what pins the oracle; the reference itself
ships no vectors for this path."""
import numpy as np

import oracle as orc

RTOL, ATOL = 1e-5, 1e-6   # north star: 1e-5 relative for fp32


def test_forward_matches_embedding_bag(golden):
    g = golden("fwd")
    base, rows, col = g.table_meta()
    for mode, psw, want in ((orc.POOL_SUM, None, g.out_sum), (orc.POOL_MEAN, None, g.out_mean),
                            (orc.POOL_SUM, g.psw, g.out_wsum)):
        out, oob = orc.embedding_forward(g.weights, g.D, g.indices, g.offsets, psw, base, rows, col,
                                         g.F, g.B, mode)
        assert not oob
        np.testing.assert_allclose(out, want, rtol=RTOL, atol=ATOL)
    out, oob = orc.embedding_forward(g.weights, g.D, g.indices, g.offsets, None, base, rows, None,
                                     g.F, g.B, orc.POOL_NONE)
    assert not oob
    np.testing.assert_array_equal(out, g.out_seq)        # pure gather: bit-exact


def test_empty_bags_are_zero(golden):
    g = golden("fwd")
    base, rows, col = g.table_meta()
    out, _ = orc.embedding_forward(g.weights, g.D, g.indices, g.offsets, None, base, rows, col, g.F, g.B,
                                   orc.POOL_MEAN)
    lengths = g.lengths.reshape(g.F, g.B)
    for f in range(g.F):
        for b in range(g.B):
            if lengths[f, b] == 0:
                assert not out[b, f * g.D:(f + 1) * g.D].any()


def test_sgd_matches_torch(golden):
    g = golden("sgd")
    base, rows, col = g.table_meta()
    for mode, want in ((orc.POOL_SUM, g.weights_after_sum), (orc.POOL_MEAN, g.weights_after_mean)):
        w = g.weights.copy()
        orc.embedding_backward_update(w, g.D, None, None, g.grad_out, g.indices, g.offsets, base, rows, col,
                                      g.F, g.B, mode, orc.OptArgs(orc.OPT_SGD, learning_rate=g.lr))
        np.testing.assert_allclose(w, want, rtol=RTOL, atol=ATOL)


def test_rowwise_adagrad_matches_fbgemm_ref(golden):
    g = golden("rwa")
    base, rows, col = g.table_meta()
    for tag in ("nowd", "wd"):
        w, h = g.weights.copy(), g.h0.copy()
        opt = orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, learning_rate=g.lr, eps=g.eps,
                          weight_decay=getattr(g, f"wd_{tag}"))
        orc.embedding_backward_update(w, g.D, h, None, g.grad_out, g.indices, g.offsets, base, rows, col,
                                      g.F, g.B, orc.POOL_SUM, opt)
        np.testing.assert_allclose(w, getattr(g, f"weights_after_{tag}"), rtol=RTOL, atol=ATOL)
        np.testing.assert_allclose(h, getattr(g, f"h_after_{tag}"), rtol=RTOL, atol=ATOL)


def test_adam_matches_torch_adamw(golden):
    g = golden("adam")
    base, rows, col = g.table_meta()
    for tag in ("nowd", "wd"):
        w = g.weights.copy()
        m, v = np.zeros_like(w), np.zeros_like(w)
        for s in range(g.steps):
            opt = orc.OptArgs(orc.OPT_ADAM, learning_rate=g.lr, eps=g.eps, weight_decay=getattr(g, f"wd_{tag}"),
                              beta1=g.beta1, beta2=g.beta2, step=s + 1)
            orc.embedding_backward_update(w, g.D, m, v, getattr(g, f"grad_out{s}"), getattr(g, f"indices{s}"),
                                          getattr(g, f"offsets{s}"), base, rows, col, g.F, g.B, orc.POOL_SUM, opt)
            np.testing.assert_allclose(w, getattr(g, f"weights_after_{tag}{s}"), rtol=RTOL, atol=ATOL)


def test_bucketize_kat(golden):
    g = golden("bucketize")
    offsets = orc.complete_cumsum(g.lengths)
    nl, no, ni, nw, perm, oob = orc.bucketize_row_wise(offsets, g.indices, g.weights, g.block_sizes,
                                                       g.F, g.B, g.world)
    assert not oob
    np.testing.assert_array_equal(nl, g.new_lengths)
    np.testing.assert_array_equal(no, orc.complete_cumsum(g.new_lengths))
    np.testing.assert_array_equal(ni, g.new_indices)
    np.testing.assert_array_equal(nw, g.new_weights)
    np.testing.assert_array_equal(perm, g.unbucketize_permute)


def test_permute_kat(golden):
    g = golden("permute")
    ol, oo, oi, _ = orc.permute_2d(g.permute, g.B, g.in_lengths, orc.complete_cumsum(g.in_lengths), g.in_indices)
    np.testing.assert_array_equal(ol, g.out_lengths)
    np.testing.assert_array_equal(oi, g.out_indices)
    np.testing.assert_array_equal(oo, orc.complete_cumsum(g.out_lengths))


def test_jagged_dense_kat(golden):
    g = golden("jagged")
    np.testing.assert_array_equal(orc.jagged_2d_to_dense(g.values, g.offsets, g.max_len), g.dense)
    np.testing.assert_array_equal(orc.dense_to_jagged(g.dense, g.offsets), g.back)
    np.testing.assert_array_equal(orc.jagged_1d_to_padded_dense(g.ids, g.offsets, g.desired, g.pad), g.padded)


def test_sort_unique_against_numpy():
    rng = np.random.default_rng(1)
    keys = rng.integers(0, 50, size=1000).astype(np.uint32)
    vals = np.arange(1000, dtype=np.uint32)
    sk, sv = orc.sort_pairs_stable(keys, vals)
    order = np.argsort(keys, kind="stable")
    np.testing.assert_array_equal(sk, keys[order])
    np.testing.assert_array_equal(sv, vals[order])
    uk, ss, nu = orc.unique_segments(sk)
    u, first = np.unique(sk, return_index=True)
    assert nu == u.size
    np.testing.assert_array_equal(uk, u)
    np.testing.assert_array_equal(ss[:-1], first)
    assert ss[-1] == 1000


def test_out_of_range_ids_are_skipped_and_reported(golden):
    g = golden("fwd")
    base, rows, col = g.table_meta()
    idx = g.indices.copy()
    idx[0] = rows[0]          # one past the end of table 0
    idx[2] = -1
    out, oob = orc.embedding_forward(g.weights, g.D, idx, g.offsets, None, base, rows, col, g.F, g.B, orc.POOL_SUM)
    assert oob
    keep = np.ones(idx.size, bool); keep[[0, 2]] = False
    # same as dropping those two ids from their bags
    psw = keep.astype(np.float32)
    idx2 = np.where(keep, idx, 0)
    want, _ = orc.embedding_forward(g.weights, g.D, idx2, g.offsets, psw, base, rows, col, g.F, g.B, orc.POOL_SUM)
    np.testing.assert_allclose(out, want, rtol=RTOL, atol=ATOL)


def test_weighted_backward_matches_torch_embedding_bag():
    """Pins the weighted-backward restatement on torch itself (live: torch is present wherever the
    tests run): nn.EmbeddingBag(mode="sum", per_sample_weights) + autograd + dense SGD step."""
    import torch
    from helpers import make_case
    c = make_case(seed=31, F=3, B=12, D=8, rows=[40, 9, 17], max_len=5, long_bag=30, zipf=True)
    rng = np.random.default_rng(3)
    psw = rng.uniform(-1.5, 2.0, size=c.L).astype(np.float32)
    grad = rng.normal(size=(c.B, c.F * c.D)).astype(np.float32)
    lr = 0.1
    want = c.weights.copy()
    for f in range(c.F):
        lo, hi = int(c.row_base[f]), int(c.row_base[f] + c.rows[f])
        bag = torch.nn.EmbeddingBag(int(c.rows[f]), c.D, mode="sum", include_last_offset=True)
        with torch.no_grad():
            bag.weight.copy_(torch.from_numpy(c.weights[lo:hi]))
        s, e = int(c.offsets[f * c.B]), int(c.offsets[(f + 1) * c.B])
        off = torch.from_numpy(c.offsets[f * c.B:(f + 1) * c.B + 1] - s)
        out = bag(torch.from_numpy(c.indices[s:e]), off, per_sample_weights=torch.from_numpy(psw[s:e]))
        out.backward(torch.from_numpy(grad[:, f * c.D:(f + 1) * c.D]))
        want[lo:hi] = (bag.weight - lr * bag.weight.grad).detach().numpy()
    w = c.weights.copy()
    res = orc.embedding_backward_update_weighted(w, c.D, None, None, grad, c.indices, c.offsets, psw, c.row_base,
                                                 c.rows, c.out_col, c.F, c.B, orc.POOL_SUM,
                                                 orc.OptArgs(orc.OPT_SGD, learning_rate=lr))
    assert not res["oob"]
    np.testing.assert_allclose(w, want, rtol=RTOL, atol=2e-6)
    # weights of 1 reproduce the unweighted path exactly
    w1, w2 = c.weights.copy(), c.weights.copy()
    orc.embedding_backward_update_weighted(w1, c.D, None, None, grad, c.indices, c.offsets, np.ones(c.L, np.float32),
                                           c.row_base, c.rows, c.out_col, c.F, c.B, orc.POOL_MEAN,
                                           orc.OptArgs(orc.OPT_SGD, learning_rate=lr))
    orc.embedding_backward_update(w2, c.D, None, None, grad, c.indices, c.offsets, c.row_base, c.rows, c.out_col,
                                  c.F, c.B, orc.POOL_MEAN, orc.OptArgs(orc.OPT_SGD, learning_rate=lr))
    np.testing.assert_allclose(w1, w2, rtol=RTOL, atol=1e-6)


def test_forward_matches_fbgemm_reference_spmdm():
    """FBGEMM's own CPU reference of the pooled lookup, `fbgemm::EmbeddingSpMDM_ref<float, int64, int64,
    float>` as exported by libtorch_cpu.so (the kernel family fbgemm_gpu's CPU TBE and nn.EmbeddingBag
    use): sum, mean (normalize_by_lengths) and per-sample-weighted sum, and the
    sequence lookup (no_bag)."""
    import ctypes
    import os
    import torch
    from helpers import make_case
    lib = ctypes.CDLL(os.path.join(os.path.dirname(torch.__file__), "lib", "libtorch_cpu.so"))
    try:
        fn = getattr(lib, "_ZN6fbgemm18EmbeddingSpMDM_refIfllfEEbllllPKT_PKT0_PKT1_PKfbPT2_bbllbbbb")
    except AttributeError:
        import pytest
        pytest.skip("this libtorch does not export fbgemm::EmbeddingSpMDM_ref<float, long, long, float>")
    i64, vp, cb = ctypes.c_int64, ctypes.c_void_p, ctypes.c_bool
    fn.restype = cb
    fn.argtypes = [i64, i64, i64, i64, vp, vp, vp, vp, cb, vp, cb, cb, i64, i64, cb, cb, cb, cb]
    c = make_case(seed=77, F=3, B=20, D=24, rows=[30, 5, 64], max_len=7)
    psw = np.random.default_rng(1).uniform(0.2, 2.0, size=c.L).astype(np.float32)
    ptr = lambda a: None if a is None else a.ctypes.data_as(vp)
    for mode, w in ((orc.POOL_SUM, None), (orc.POOL_MEAN, None), (orc.POOL_SUM, psw)):
        got, oob = orc.embedding_forward(c.weights, c.D, c.indices, c.offsets, w, c.row_base, c.rows, c.out_col,
                                         c.F, c.B, mode)
        assert not oob
        for f in range(c.F):
            s, e = int(c.offsets[f * c.B]), int(c.offsets[(f + 1) * c.B])
            table = np.ascontiguousarray(c.weights[c.row_base[f]:c.row_base[f] + c.rows[f]])
            idx = np.ascontiguousarray(c.indices[s:e])
            off = np.ascontiguousarray(c.offsets[f * c.B:(f + 1) * c.B + 1] - s)
            wf = None if w is None else np.ascontiguousarray(w[s:e])
            out = np.full((c.B, c.D), np.nan, dtype=np.float32)
            ok = fn(c.D, c.B, e - s, int(c.rows[f]), ptr(table), ptr(idx), ptr(off), ptr(wf), mode == orc.POOL_MEAN,
                    ptr(out), False, True, c.D, c.D, True, False, False, False)
            assert ok
            np.testing.assert_allclose(got[:, f * c.D:(f + 1) * c.D], out, rtol=RTOL, atol=ATOL, err_msg=f"mode {mode} f{f}")
    seq, _ = orc.embedding_forward(c.weights, c.D, c.indices, c.offsets, None, c.row_base, c.rows, None, c.F, c.B,
                                   orc.POOL_NONE)
    for f in range(c.F):
        s, e = int(c.offsets[f * c.B]), int(c.offsets[(f + 1) * c.B])
        table = np.ascontiguousarray(c.weights[c.row_base[f]:c.row_base[f] + c.rows[f]])
        idx = np.ascontiguousarray(c.indices[s:e])
        out = np.full((e - s, c.D), np.nan, dtype=np.float32)
        ok = fn(c.D, e - s, e - s, int(c.rows[f]), ptr(table), ptr(idx), None, None, False, ptr(out), False, True,
                c.D, c.D, True, True, False, False)
        assert ok
        np.testing.assert_array_equal(seq[s:e], out)


def _lazy_adam_case():
    """Three steps on a 12-row table with a different set of touched rows each step (rows 2 and 7 sit
    a step out and come back), some rows never touched, some touched twice in one step."""
    rng = np.random.default_rng(21)
    rows, D, B = 12, 8, 5
    steps = [np.array([0, 2, 2, 5, 7]), np.array([0, 1, 5, 9, 9]), np.array([2, 7, 7, 9, 11])]
    grads = [rng.normal(size=(B, D)).astype(np.float32) for _ in steps]
    w0 = rng.uniform(-1, 1, size=(rows, D)).astype(np.float32)
    return rows, D, B, steps, grads, w0


def _oracle_adam_steps(wd):
    rows, D, B, steps, grads, w0 = _lazy_adam_case()
    w, m, v = w0.copy(), np.zeros_like(w0), np.zeros_like(w0)
    off = np.arange(B + 1, dtype=np.int64)                     # one id per bag: grad row b belongs to ids[b]
    for t, (ids, g) in enumerate(zip(steps, grads), start=1):
        orc.embedding_backward_update(w, D, m, v, g, ids.astype(np.int64), off, [0], [rows], [0], 1, B, orc.POOL_SUM,
                                      orc.OptArgs(orc.OPT_ADAM, learning_rate=0.05, eps=1e-8, weight_decay=wd, step=t))
    return w, m, v


def test_lazy_adam_matches_torch_sparse_adam():
    """The fused ADAM is lazy: a row's moments move only in steps that touch the row, and the bias
    correction uses the global step count.  torch.optim.SparseAdam over nn.Embedding(sparse=True) is
    an implementation of exactly that contract that was not written here (it places eps after the
    bias correction of v instead of before, a 1e-8-relative difference at these gradient sizes)."""
    import torch
    rows, D, B, steps, grads, w0 = _lazy_adam_case()
    emb = torch.nn.Embedding(rows, D, sparse=True)
    with torch.no_grad():
        emb.weight.copy_(torch.from_numpy(w0))
    opt = torch.optim.SparseAdam(emb.parameters(), lr=0.05, betas=(0.9, 0.999), eps=1e-8)
    for ids, g in zip(steps, grads):
        opt.zero_grad()
        (emb(torch.from_numpy(ids)) * torch.from_numpy(g)).sum().backward()
        opt.step()
    w, m, v = _oracle_adam_steps(wd=0.0)
    np.testing.assert_allclose(w, emb.weight.detach().numpy(), rtol=RTOL, atol=ATOL)
    st = opt.state[emb.weight]
    np.testing.assert_allclose(m, st["exp_avg"].numpy(), rtol=RTOL, atol=ATOL)
    # beta2 crosses the ABI as a float (as it does in fbgemm's kernels [UPSTREAM]): 1 - 0.999f is 1.3e-5 below
    # 0.001, so v itself carries that factor; the step sees sqrt(v), half of it, and stays inside 1e-5
    np.testing.assert_allclose(v, st["exp_avg_sq"].numpy(), rtol=2e-5, atol=1e-9)
    untouched = [3, 4, 6, 8, 10]
    np.testing.assert_array_equal(w[untouched], w0[untouched])


def test_lazy_adam_weight_decay_against_masked_dense_restatement():
    """With weight decay: decoupled (AdamW-style) and applied to the touched rows only, next to stale
    moments for the rows that sat a step out.  Restated densely in float64 with a touched-row mask."""
    rows, D, B, steps, grads, w0 = _lazy_adam_case()
    lr, wd, b1, b2, eps = 0.05, 0.1, 0.9, 0.999, 1e-8
    w = w0.astype(np.float64)
    m, v = np.zeros_like(w), np.zeros_like(w)
    for t, (ids, g) in enumerate(zip(steps, grads), start=1):
        dense = np.zeros_like(w)
        np.add.at(dense, ids, g.astype(np.float64))
        mask = np.zeros(rows, bool)
        mask[ids] = True
        m[mask] = b1 * m[mask] + (1 - b1) * dense[mask]
        v[mask] = b2 * v[mask] + (1 - b2) * dense[mask] ** 2
        mh, vh = m / (1 - b1 ** t), v / (1 - b2 ** t)
        w[mask] -= lr * (mh[mask] / (np.sqrt(vh[mask]) + eps) + wd * w[mask])
    got_w, got_m, got_v = _oracle_adam_steps(wd=wd)
    np.testing.assert_allclose(got_w, w, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_m, m, rtol=RTOL, atol=ATOL)
    np.testing.assert_allclose(got_v, v, rtol=2e-5, atol=1e-9)     # 1 - 0.999f, see the test above
