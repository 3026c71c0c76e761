"""The torchrec-shaped surface on the GPU: EmbeddingBagCollection / EmbeddingCollection through
autograd with the fused optimizer, the reference's HistoryArch call pattern
(torchrec/models.py:163-178), and size-independent properties at BASELINE config-2 scale."""
import numpy as np
import pytest
import torch

import oracle as orc
from helpers import make_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def t():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import tdfo_b200
    return tdfo_b200


def test_ebc_forward_backward_matches_oracle(t):
    c = make_case(seed=4, F=3, B=32, D=64, rows=[100, 40, 7], max_len=8)
    tables = [t.EmbeddingBagConfig(num_embeddings=int(r), embedding_dim=c.D, name=f"t{i}", feature_names=[f"f{i}"],
                                   pooling=t.PoolingType.SUM) for i, r in enumerate(c.rows)]
    ebc = t.EmbeddingBagCollection(tables, device=torch.device("cuda"))
    ebc.load_state_dict({f"embedding_bags.t{i}.weight": torch.from_numpy(
        c.weights[c.row_base[i]:c.row_base[i] + c.rows[i]]) for i in range(c.F)})
    fused = ebc.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.1, eps=1e-8)
    assert not list(t.in_backward_optimizer_filter(ebc.named_parameters()))
    kjt = t.KeyedJaggedTensor.from_lengths_sync([f"f{i}" for i in range(c.F)], torch.from_numpy(c.indices),
                                                torch.from_numpy(c.lengths)).to("cuda")
    ebc.train()
    kt = ebc(kjt)
    assert kt.keys() == ["f0", "f1", "f2"] and kt.values().shape == (c.B, c.F * c.D)
    want, _ = orc.embedding_forward(c.weights, c.D, c.indices, c.offsets, None, c.row_base, c.rows, c.out_col,
                                    c.F, c.B, orc.POOL_SUM)
    np.testing.assert_allclose(kt.values().detach().cpu().numpy(), want, rtol=1e-5, atol=1e-5)
    np.testing.assert_array_equal(kt["f1"].detach().cpu().numpy(), kt.values()[:, c.D:2 * c.D].detach().cpu().numpy())

    proj = torch.randn(c.F * c.D, 1, device="cuda", requires_grad=True)
    loss = (kt.values() @ proj).sum()
    loss.backward()
    fused.step(); fused.zero_grad()
    grad_out = proj.detach().t().expand(c.B, -1).contiguous().cpu().numpy()
    w_o, h_o = c.weights.copy(), np.zeros(c.total_rows, np.float32)
    orc.embedding_backward_update(w_o, c.D, h_o, None, grad_out, c.indices, c.offsets, c.row_base, c.rows, c.out_col,
                                  c.F, c.B, orc.POOL_SUM, orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, 0.1, 1e-8))
    got = torch.cat([ebc.state_dict()[f"embedding_bags.t{i}.weight"] for i in range(c.F)]).cpu().numpy()
    np.testing.assert_allclose(got, w_o, rtol=1e-5, atol=2e-6)
    assert proj.grad is not None
    ebc.check_errors()


def test_weighted_ebc_trains_like_the_oracle(t):
    """EmbeddingBagCollection(is_weighted=True): KJT weights scale the pooled sum and, in the fused
    backward, each occurrence's gradient row (two tables of different D = two kernel groups)."""
    rows, dims, B = [60, 25], [32, 64], 24
    rng = np.random.default_rng(8)
    tables = [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=d, name=f"t{i}", feature_names=[f"f{i}"],
                                   pooling=t.PoolingType.SUM) for i, (r, d) in enumerate(zip(rows, dims))]
    ebc = t.EmbeddingBagCollection(tables, is_weighted=True, device=torch.device("cuda"))
    assert ebc.is_weighted()
    w0 = [rng.uniform(-1, 1, size=(r, d)).astype(np.float32) for r, d in zip(rows, dims)]
    ebc.load_state_dict({f"embedding_bags.t{i}.weight": torch.from_numpy(w0[i]) for i in range(2)})
    fused = ebc.set_fused_optimizer(t.EmbOptimType.SGD, learning_rate=0.2)
    ebc.train()
    w_o = [w.copy() for w in w0]
    for step in range(3):
        lens = [rng.integers(0, 6, size=B).astype(np.int32) for _ in rows]
        ids = [rng.integers(0, r, size=int(l.sum())) for r, l in zip(rows, lens)]
        psw = [rng.uniform(0.2, 1.8, size=i.size).astype(np.float32) for i in ids]
        kjt = t.KeyedJaggedTensor.from_lengths_sync(["f0", "f1"], torch.from_numpy(np.concatenate(ids)),
                                                    torch.from_numpy(np.concatenate(lens)),
                                                    weights=torch.from_numpy(np.concatenate(psw))).to("cuda")
        out = ebc(kjt).values()
        grad = rng.normal(size=(B, sum(dims))).astype(np.float32)
        out.backward(torch.from_numpy(grad).cuda())
        fused.step()
        col = 0
        for i, d in enumerate(dims):
            off = orc.complete_cumsum(lens[i])
            want, _ = orc.embedding_forward(w_o[i], d, ids[i], off, psw[i], [0], [rows[i]], [0], 1, B, orc.POOL_SUM)
            np.testing.assert_allclose(out[:, col:col + d].detach().cpu().numpy(), want, rtol=1e-5, atol=1e-5)
            orc.embedding_backward_update_weighted(w_o[i], d, None, None, np.ascontiguousarray(grad[:, col:col + d]),
                                                   ids[i], off, psw[i], [0], [rows[i]], [0], 1, B, orc.POOL_SUM,
                                                   orc.OptArgs(orc.OPT_SGD, 0.2))
            col += d
    sd = ebc.state_dict()
    for i in range(2):
        np.testing.assert_allclose(sd[f"embedding_bags.t{i}.weight"].cpu().numpy(), w_o[i], rtol=1e-5, atol=2e-6)
    ebc.check_errors()


def test_ebc_mixed_dims_and_key_order(t):
    rows, dims = [30, 20, 10], [16, 32, 16]
    tables = [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=d, name=f"t{i}", feature_names=[f"f{i}"],
                                   pooling=t.PoolingType.MEAN if i == 2 else t.PoolingType.SUM)
              for i, (r, d) in enumerate(zip(rows, dims))]
    ebc = t.EmbeddingBagCollection(tables, device=torch.device("cuda")).eval()
    rng = np.random.default_rng(0)
    B = 9
    # KJT arrives in a different key order than the module's
    order = ["f2", "f0", "f1"]
    lens = {k: rng.integers(0, 5, size=B).astype(np.int32) for k in order}
    ids = {k: rng.integers(0, rows[int(k[1])], size=int(lens[k].sum())) for k in order}
    kjt = t.KeyedJaggedTensor.from_lengths_sync(order, torch.from_numpy(np.concatenate([ids[k] for k in order])),
                                                torch.from_numpy(np.concatenate([lens[k] for k in order]))).to("cuda")
    out = ebc(kjt)
    sd = ebc.state_dict()
    for i, k in enumerate(["f0", "f1", "f2"]):
        w = sd[f"embedding_bags.t{i}.weight"].cpu().numpy()
        off = orc.complete_cumsum(lens[k])
        want, _ = orc.embedding_forward(np.ascontiguousarray(w), dims[i], ids[k], off, None, [0], [rows[i]], [0], 1, B,
                                        orc.POOL_MEAN if i == 2 else orc.POOL_SUM)
        np.testing.assert_allclose(out[k].cpu().numpy(), want, rtol=1e-5, atol=1e-6)


def test_history_arch_pattern(t):
    """EmbeddingCollection -> jagged_2d_to_dense -> loss -> backward, as torchrec/models.py:163-178
    and torchrec/train.py:98-105 do, against the oracle's sequence path."""
    V, D, B, T = 50, 16, 16, 20
    ec = t.EmbeddingCollection([t.EmbeddingConfig(name="item_embedding", embedding_dim=D, num_embeddings=V,
                                                  feature_names=["item"], weight_init_max=1.0, weight_init_min=-1.0)],
                               device=torch.device("cuda"))
    ec.set_fused_optimizer(t.EmbOptimType.SGD, learning_rate=3e-4, weight_decay=1e-4)
    w0 = ec.state_dict()["embeddings.item_embedding.weight"].cpu().numpy().copy()
    assert w0.min() >= -1 and w0.max() <= 1
    rng = np.random.default_rng(1)
    seqs = torch.from_numpy(rng.integers(0, V, size=(B, T)))
    seqs_list = list(seqs)                                    # _to_kjt, torchrec/train.py:33-41
    lengths = torch.IntTensor([v.size(0) for v in seqs_list])
    kjt = t.KeyedJaggedTensor.from_lengths_sync(keys=["item"], values=torch.cat(seqs_list, dim=0),
                                                lengths=lengths).to(torch.device("cuda"))
    dense_ids = kjt["item"].to_padded_dense(desired_length=T, padding_value=0)
    np.testing.assert_array_equal(dense_ids.cpu().numpy(), seqs.numpy())
    ec.train()
    jt = ec(kjt)
    padded = t.jagged_2d_to_dense(values=jt["item"].values(), offsets=jt["item"].offsets(),
                                  max_sequence_length=T).view(-1, T, D)
    assert padded.shape == (kjt.stride(), T, D)
    np.testing.assert_array_equal(padded.detach().cpu().numpy(), w0[seqs.numpy()])
    target = torch.randn(B, T, D, device="cuda")
    loss = ((padded - target) ** 2).sum()
    loss.backward()
    grad = (2 * (padded - target)).detach().view(-1, D).cpu().numpy()
    w_o = w0.copy()
    off = orc.complete_cumsum(lengths.numpy())
    orc.embedding_backward_update(w_o, D, None, None, grad, seqs.numpy().ravel(), off, [0], [V], None, 1, B,
                                  orc.POOL_NONE, orc.OptArgs(orc.OPT_SGD, 3e-4, weight_decay=1e-4))
    np.testing.assert_allclose(ec.state_dict()["embeddings.item_embedding.weight"].cpu().numpy(), w_o,
                               rtol=1e-5, atol=1e-6)


def test_full_size_values_config2_against_the_oracle(t):
    """BASELINE config 2 at its real size, values this time: random fp32 tables, Poisson-ish jagged
    lengths, one fused row-wise Adagrad step through the product path (forward emits the sort pairs,
    backward consumes them).  The oracle sees (a) 4096 sampled bags, fed with the very rows the GPU
    table holds, and (b) two whole tables (first and last): every row of them after the step, updated
    or not, must match the oracle's update of a host copy of that table."""
    free, _ = torch.cuda.mem_get_info()
    if free < 40 << 30:
        pytest.skip("needs 40 GB of free HBM")
    ops = t.ops
    F, rows, D, B, P = 26, 1_000_000, 128, 16384, 20
    dev = torch.device("cuda")
    g = torch.Generator(device=dev).manual_seed(7)
    w = torch.empty((F * rows, D), device=dev).uniform_(-0.05, 0.05, generator=g)
    h = torch.empty(F * rows, device=dev).uniform_(0.0, 0.01, generator=g)
    grp = ops.TableGroup(weights=w, feat_row_base=torch.arange(F, device=dev) * rows,
                         feat_rows=torch.full((F,), rows, device=dev, dtype=torch.int64),
                         feat_out_col=torch.arange(F, device=dev) * D, out_stride=F * D, state0=h)
    lengths = torch.poisson(torch.full((F * B,), float(P), device=dev), generator=g).to(torch.int32)
    lengths[::101] = 0
    offsets = ops.complete_cumsum(lengths)
    L = int(offsets[-1].item())
    idx = torch.randint(0, rows, (L,), device=dev, generator=g)
    grad = torch.randn((B, F * D), device=dev, generator=g)
    check_tables = (0, F - 1)
    before = {f: (w[f * rows:(f + 1) * rows].cpu().numpy(), h[f * rows:(f + 1) * rows].cpu().numpy()) for f in check_tables}

    lay = ops.backward_layout(L, F * rows)
    scratch = torch.empty(lay.total_bytes, dtype=torch.uint8, device=dev)
    out = ops.embedding_forward(grp, idx, offsets, B, orc.POOL_SUM, backward_scratch=scratch)

    # (a) sampled bags: gather their rows from the GPU table, pool them on the host in bag order
    rng = np.random.default_rng(3)
    bags = np.sort(rng.choice(F * B, size=4096, replace=False))
    off_h, idx_h = offsets.cpu().numpy(), idx.cpu().numpy()
    lens = (off_h[bags + 1] - off_h[bags]).astype(np.int32)
    pos = np.concatenate([np.arange(off_h[b_], off_h[b_ + 1]) for b_ in bags]).astype(np.int64)
    keys = (bags.repeat(lens) // B) * rows + idx_h[pos]
    mini = w[torch.from_numpy(keys).to(dev)].cpu().numpy()                       # one row per sampled occurrence
    want, _ = orc.embedding_forward(mini, D, np.arange(keys.size, dtype=np.int64), orc.complete_cumsum(lens), None,
                                    [0], [keys.size], [0], 1, bags.size, orc.POOL_SUM)
    got = out.view(B, F, D)[torch.from_numpy(bags % B).to(dev), torch.from_numpy(bags // B).to(dev)].cpu().numpy()
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6)

    ops.embedding_backward_update(grp, grad, idx, offsets, B, orc.POOL_SUM,
                                  ops.make_optimizer_args(orc.OPT_ROWWISE_ADAGRAD, 0.01, 1e-8), scratch=scratch,
                                  keys_in_scratch=True)
    torch.cuda.synchronize()
    assert int(grp.err_flag.item()) == 0
    # (b) whole tables against the oracle's update of their host copies
    grad_h = grad.cpu().numpy()
    for f in check_tables:
        w_o, h_o = before[f][0].copy(), before[f][1].copy()
        s, e = int(off_h[f * B]), int(off_h[(f + 1) * B])
        orc.embedding_backward_update(w_o, D, h_o, None, np.ascontiguousarray(grad_h[:, f * D:(f + 1) * D]), idx_h[s:e],
                                      off_h[f * B:(f + 1) * B + 1] - s, [0], [rows], [0], 1, B, orc.POOL_SUM,
                                      orc.OptArgs(orc.OPT_ROWWISE_ADAGRAD, 0.01, 1e-8))
        np.testing.assert_allclose(w[f * rows:(f + 1) * rows].cpu().numpy(), w_o, rtol=1e-5, atol=1e-7)
        np.testing.assert_allclose(h[f * rows:(f + 1) * rows].cpu().numpy(), h_o, rtol=1e-5, atol=1e-9)
        assert int((w_o != before[f][0]).any(axis=1).sum()) > 250_000           # the step did touch most ids' rows


def test_full_size_properties_config2(t):
    """BASELINE config 2 (26 tables x 1M rows x D128 fp32, B=16384, pooling 20): properties that
    do not need the oracle at this size.  Tables of ones => every pooled value is exactly the bag
    length; SGD with lr=1 and unit gradients subtracts each row's occurrence count, so the total
    decrement equals L*D exactly (small integers are exact in fp32)."""
    free, _ = torch.cuda.mem_get_info()
    if free < 40 << 30:
        pytest.skip("needs 40 GB of free HBM")
    ops = t.ops
    F, rows, D, B, P = 26, 1_000_000, 128, 16384, 20
    dev = torch.device("cuda")
    w = torch.ones((F * rows, D), device=dev)
    grp = ops.TableGroup(weights=w, feat_row_base=torch.arange(F, device=dev) * rows,
                         feat_rows=torch.full((F,), rows, device=dev, dtype=torch.int64),
                         feat_out_col=torch.arange(F, device=dev) * D, out_stride=F * D)
    g = torch.Generator(device=dev).manual_seed(42)
    lengths = torch.full((F * B,), P, dtype=torch.int32, device=dev)
    lengths[::97] = 0                                     # some empty bags
    lengths[5::1013] = 311                                # some long ones
    offsets = ops.complete_cumsum(lengths)
    assert torch.equal(offsets[1:], torch.cumsum(lengths.long(), 0))
    L = int(offsets[-1].item())
    idx = torch.randint(0, rows, (L,), device=dev, generator=g)
    out = ops.embedding_forward(grp, idx, offsets, B, orc.POOL_SUM)
    want = lengths.view(F, B).t().float()                 # [B, F]
    assert torch.equal(out.view(B, F, D), want.unsqueeze(-1).expand(B, F, D))
    grad = torch.ones((B, F * D), device=dev)
    scratch = ops.embedding_backward_update(grp, grad, idx, offsets, B, orc.POOL_SUM,
                                            ops.make_optimizer_args(orc.OPT_SGD, 1.0))
    lay = ops.backward_layout(L, F * rows)
    view = lambda off, n: scratch[off: off + 4 * n].view(torch.int32)
    sk = view(lay.sorted_keys, L)
    assert bool((sk[1:] >= sk[:-1]).all())                # sortedness (keys < 2^31 here)
    nu = int(view(lay.num_unique, 1).item())
    keys = (torch.arange(F, device=dev).repeat_interleave(lengths.view(F, B).sum(1).long()) * rows + idx)
    assert nu == int(torch.unique(keys).numel())
    ss = view(lay.seg_starts, nu + 1)
    assert int(ss[0]) == 0 and int(ss[-1]) == L and bool((ss[1:] > ss[:-1]).all())
    # the sorted payloads are a permutation of the bag ids
    sv = view(lay.sorted_vals, L)
    assert int(sv.long().sum()) == int(torch.arange(F * B, device=dev).repeat_interleave(lengths.long()).sum())
    dec = float(F * rows) * D - w.sum(dtype=torch.float64).item()
    assert dec == float(L) * D
    assert int(grp.err_flag.item()) == 0


def test_train_py_model_parallel_block(t):
    """The model-parallel block of the reference's main() (torchrec/train.py:235-254) and one step of
    its training loop (:97-106), with a small stand-in for Bert4Rec around the same HistoryArch calls."""
    import torch.nn as nn
    import torch.optim as optim
    V, D, T, B = 64, 16, 20, 8
    dev = torch.device("cuda")

    class Tiny(nn.Module):                                     # HistoryArch + output layer
        def __init__(self):
            super().__init__()
            self.embed_collection = t.EmbeddingCollection(
                tables=[t.EmbeddingConfig(name="item_embedding", embedding_dim=D, num_embeddings=V,
                                          feature_names=["item"], weight_init_max=1.0, weight_init_min=-1.0)],
                device=dev)
            self.out = nn.Linear(D, V, device=dev)

        def forward(self, kjt):
            jt = self.embed_collection(kjt)
            padded = [t.jagged_2d_to_dense(values=jt[e].values(), offsets=jt[e].offsets(),
                                           max_sequence_length=T).view(-1, T, D) for e in kjt.keys()]
            return self.out(torch.cat(padded, dim=1))

    fused_params = {"optimizer": t.EmbOptimType.ADAM, "learning_rate": 3e-4, "weight_decay": 1e-4}
    model = t.DistributedModelParallel(module=Tiny(), device=dev, sharders=[t.EmbeddingCollectionSharder(fused_params)])
    dense_optimizer = t.KeyedOptimizerWrapper(
        dict(t.in_backward_optimizer_filter(model.named_parameters())),
        lambda params: optim.Adam(params, lr=3e-4, weight_decay=1e-4))
    optimizer = t.CombinedOptimizer([model.fused_optimizer, dense_optimizer])
    assert sorted(dense_optimizer.params) == ["module.out.bias", "module.out.weight"]
    table0 = model.state_dict()["module.embed_collection.embeddings.item_embedding.weight"].clone()
    dense0 = model.module.out.weight.detach().clone()

    seqs = torch.randint(1, V, (B, T))
    lengths = torch.IntTensor([v.size(0) for v in list(seqs)])
    kjt = t.KeyedJaggedTensor.from_lengths_sync(keys=["item"], values=torch.cat(list(seqs), dim=0), lengths=lengths).to(dev)
    model.train()
    logits = model(kjt)
    loss = nn.CrossEntropyLoss(ignore_index=0, label_smoothing=0.1)(logits.view(-1, V), seqs.to(dev).view(-1))
    optimizer.zero_grad()
    loss.backward()
    optimizer.step()
    assert np.isfinite(loss.item())
    table1 = model.state_dict()["module.embed_collection.embeddings.item_embedding.weight"]
    touched = torch.zeros(V, dtype=torch.bool)
    touched[seqs.unique()] = True
    changed = (table1 != table0).any(dim=1).cpu()
    assert torch.equal(changed, touched)                       # exactly the looked-up rows moved (in backward)
    assert not torch.equal(model.module.out.weight, dense0)    # the dense optimizer stepped too


def test_bounds_check_modes(t):
    """Out-of-range ids never touch memory; who hears about them is selectable."""
    tables = [t.EmbeddingBagConfig(num_embeddings=10, embedding_dim=16, name="a", feature_names=["a"])]
    kjt = t.KeyedJaggedTensor.from_lengths_sync(["a"], torch.tensor([3, 10, 4]), torch.tensor([2, 1],
                                                                                               dtype=torch.int32)).to("cuda")
    ebc = t.EmbeddingBagCollection(tables, device=torch.device("cuda")).eval()
    w = ebc.state_dict()["embedding_bags.a.weight"]
    out = ebc(kjt).values()                                   # WARNING (default): deferred
    torch.testing.assert_close(out[0], w[3])                  # id 10 skipped
    torch.testing.assert_close(out[1], w[4])
    with pytest.raises(IndexError):
        ebc.check_errors()
    ebc.check_errors()                                        # the word was cleared by the raise
    ebc.bounds_check_mode = t.BoundsCheckMode.IGNORE
    ebc(kjt)
    ebc.check_errors()
    ebc.bounds_check_mode = t.BoundsCheckMode.FATAL
    with pytest.raises(IndexError):
        ebc(kjt)


def test_dense_gradient_mode_matches_torch_modules_under_a_dense_optimizer(t):
    """The reference's non-model-parallel branch (torchrec/train.py:255-260): the tables are
    plain parameters, backward yields dense gradients, `optim.Adam(model.parameters())` updates them.
    Checked against nn.Embedding / nn.EmbeddingBag + the same Adam: gradients to 1e-5, two steps."""
    rng = np.random.default_rng(12)
    B, T, V, D = 16, 7, 50, 16
    # ---- EmbeddingCollection vs nn.Embedding (the Bert4Rec item table pattern)
    ec = t.EmbeddingCollection([t.EmbeddingConfig(num_embeddings=V, embedding_dim=D, name="item",
                                                  feature_names=["item"])], device=torch.device("cuda"))
    ec.use_dense_gradients()
    assert [n for n, _ in t.in_backward_optimizer_filter(ec.named_parameters())] == [n for n, _ in ec.named_parameters()]
    ref = torch.nn.Embedding(V, D).cuda()
    with torch.no_grad():
        ref.weight.copy_(ec.state_dict()["embeddings.item.weight"])
    opt_a = torch.optim.Adam(ec.parameters(), lr=0.05, weight_decay=1e-4)
    opt_b = torch.optim.Adam(ref.parameters(), lr=0.05, weight_decay=1e-4)
    for step in range(2):
        seqs = torch.from_numpy(rng.integers(0, V, size=(B, T))).cuda()
        kjt = t.KeyedJaggedTensor.from_lengths_sync(["item"], seqs.reshape(-1),
                                                    torch.full((B,), T, dtype=torch.int32)).to("cuda")
        proj = torch.from_numpy(rng.normal(size=(B * T, D)).astype(np.float32)).cuda()
        opt_a.zero_grad(); opt_b.zero_grad()
        (ec(kjt)["item"].values() * proj).sum().backward()
        (ref(seqs.reshape(-1)) * proj).sum().backward()
        ga = next(ec.parameters()).grad
        torch.testing.assert_close(ga, ref.weight.grad, rtol=1e-5, atol=1e-6)
        opt_a.step(); opt_b.step()
        torch.testing.assert_close(ec.state_dict()["embeddings.item.weight"], ref.weight.detach(), rtol=1e-5, atol=1e-6)
    # ---- EmbeddingBagCollection (MEAN pooling, ragged bags) vs nn.EmbeddingBag
    ebc = t.EmbeddingBagCollection([t.EmbeddingBagConfig(num_embeddings=V, embedding_dim=D, name="g", feature_names=["g"],
                                                         pooling=t.PoolingType.MEAN)], device=torch.device("cuda"))
    ebc.use_dense_gradients()
    bag = torch.nn.EmbeddingBag(V, D, mode="mean", include_last_offset=True).cuda()
    with torch.no_grad():
        bag.weight.copy_(ebc.state_dict()["embedding_bags.g.weight"])
    lengths = torch.from_numpy(rng.integers(0, 6, size=B).astype(np.int32))
    ids = torch.from_numpy(rng.integers(0, V, size=int(lengths.sum())))
    kjt = t.KeyedJaggedTensor.from_lengths_sync(["g"], ids, lengths).to("cuda")
    proj = torch.from_numpy(rng.normal(size=(B, D)).astype(np.float32)).cuda()
    (ebc(kjt).values() * proj).sum().backward()
    (bag(ids.cuda(), kjt.offsets()) * proj).sum().backward()
    torch.testing.assert_close(next(ebc.parameters()).grad, bag.weight.grad, rtol=1e-5, atol=1e-6)
    ebc.check_errors()


def test_int32_ids_are_widened_on_the_device(t):
    """ids may cross PCIe as int32: a KeyedJaggedTensor built from int32 values gives the same lookup and the same
    fused update as the int64 one (tdfo_kjt_widen_ids runs once, on first use)."""
    dev = torch.device("cuda")
    rng = np.random.default_rng(8)
    rows, D, B = [300, 41], 32, 50
    tables = [t.EmbeddingBagConfig(num_embeddings=r, embedding_dim=D, name=f"t{i}", feature_names=[f"f{i}"])
              for i, r in enumerate(rows)]
    lengths = rng.integers(0, 7, size=2 * B).astype(np.int32)
    off = np.concatenate([[0], np.cumsum(lengths)])
    idx = np.concatenate([rng.integers(0, rows[f], size=int(off[(f + 1) * B] - off[f * B])) for f in range(2)])
    outs, tabs = [], []
    for dtype in (torch.int64, torch.int32):
        torch.manual_seed(3)
        ebc = t.EmbeddingBagCollection(tables, device=dev)
        ebc.set_fused_optimizer(t.EmbOptimType.EXACT_SGD, learning_rate=0.1)
        ebc.train()
        kjt = t.KeyedJaggedTensor.from_lengths_sync(["f0", "f1"], torch.from_numpy(idx).to(dtype),
                                                    torch.from_numpy(lengths)).to(dev)
        out = ebc(kjt).values()
        assert kjt.values().dtype == torch.int64 and kjt.values().numel() == idx.size
        out.sum().backward()
        outs.append(out.detach().clone())
        tabs.append(torch.cat([v for v in ebc.state_dict().values()]).clone())
    assert torch.equal(outs[0], outs[1]) and torch.equal(tabs[0], tabs[1])
