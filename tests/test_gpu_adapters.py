"""DLPackEmbeddingTables: the `lookup(ids) -> [B, D]` / sparse `apply_gradients` face for the
reference's jax-flax and tensorflow2 models (jax-flax/models.py:78-88, train_dp.py:63-65;
tensorflow2/models.py:48-58).  jax / tf are not installed here, so foreign arrays are stood in for by
an object that offers nothing but the DLPack protocol."""
import numpy as np
import pytest
import torch

import oracle as orc

pytestmark = pytest.mark.gpu

# the seven tables of the two-tower model (names from jax-flax/models.py:18-58), small cardinalities
TABLES = {"user_id": (300, 16), "item_id": (500, 16), "language": (12, 16), "is_ebook": (2, 16), "format": (9, 16),
          "publisher": (70, 32), "pub_decade": (15, 16)}


class DLPackOnly:
    """What a jax.Array looks like to a consumer: `__dlpack__` + `__dlpack_device__`, nothing else."""

    def __init__(self, t):
        self._t = t

    def __dlpack__(self, *a, **k):
        return self._t.__dlpack__(*a, **k)

    def __dlpack_device__(self):
        return self._t.__dlpack_device__()


@pytest.fixture(scope="module")
def t():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import tdfo_b200
    return tdfo_b200


def _ids(rng, B):
    return {n: rng.integers(0, r, size=B).astype(np.int32) for n, (r, _) in TABLES.items()}


def test_lookup_and_sparse_adamw_step(t):
    from tdfo_b200.adapters import DLPackEmbeddingTables
    rng = np.random.default_rng(2)
    B, lr, wd = 64, 0.01, 0.05
    tabs = DLPackEmbeddingTables(TABLES, "cuda", optimizer=t.EmbOptimType.ADAM, learning_rate=lr, weight_decay=wd)
    w0 = {n: v.cpu().numpy().copy() for n, v in tabs.state_dict().items()}
    w = {n: w0[f"embedding_bags.{n}.weight"].copy() for n in TABLES}
    m = {n: np.zeros_like(w[n]) for n in TABLES}
    v = {n: np.zeros_like(w[n]) for n in TABLES}
    for step in (1, 2, 3):
        ids = _ids(rng, B)
        rows = tabs.lookup({n: DLPackOnly(torch.from_numpy(a).cuda()) for n, a in ids.items()})
        assert list(rows) == list(TABLES)
        held = tabs.state_dict()
        for n, (r, d) in TABLES.items():
            assert rows[n].shape == (B, d)
            # a gather: bit-exact against the table the module holds, 1e-5 against the oracle's copy
            np.testing.assert_array_equal(rows[n].cpu().numpy(),
                                          held[f"embedding_bags.{n}.weight"].cpu().numpy()[ids[n]])
            np.testing.assert_allclose(rows[n].cpu().numpy(), w[n][ids[n]], rtol=1e-5, atol=2e-6)
            # the result can be handed on through DLPack without a copy
            assert torch.from_dlpack(rows[n]).data_ptr() == rows[n].data_ptr()
        grads = {n: rng.normal(size=(B, d)).astype(np.float32) for n, (_, d) in TABLES.items()}
        tabs.apply_gradients({n: torch.from_numpy(a).cuda() for n, a in ids.items()},
                             {n: DLPackOnly(torch.from_numpy(g).cuda()) for n, g in grads.items()})
        for n, (r, d) in TABLES.items():                  # oracle: one table = one sequence-free pooled feature
            off = np.arange(B + 1, dtype=np.int64)
            orc.embedding_backward_update(w[n], d, m[n], v[n], grads[n], ids[n].astype(np.int64), off, [0], [r], [0],
                                          1, B, orc.POOL_SUM,
                                          orc.OptArgs(orc.OPT_ADAM, lr, 1e-8, wd, 0.9, 0.999, step))
        sd = tabs.state_dict()
        for n in TABLES:
            np.testing.assert_allclose(sd[f"embedding_bags.{n}.weight"].cpu().numpy(), w[n], rtol=1e-5, atol=2e-6,
                                       err_msg=f"{n} step {step}")
    tabs.check_errors()


def test_first_step_equals_dense_torch_adamw_on_touched_rows(t):
    """Against the optimizer the reference actually runs (AdamW over the dense gradient): on the first
    step, rows that occur get exactly the dense update; the others differ only by the weight decay
    the lazy form skips."""
    from tdfo_b200.adapters import DLPackEmbeddingTables
    rng = np.random.default_rng(5)
    B, lr, wd = 40, 0.02, 0.1
    tabs = DLPackEmbeddingTables({"item_id": (200, 16)}, "cuda", optimizer=t.EmbOptimType.ADAM, learning_rate=lr,
                                 weight_decay=wd)
    w0 = tabs.state_dict()["embedding_bags.item_id.weight"].clone()
    emb = torch.nn.Embedding(200, 16).cuda()
    with torch.no_grad():
        emb.weight.copy_(w0)
    opt = torch.optim.AdamW(emb.parameters(), lr=lr, weight_decay=wd, eps=1e-8)
    ids = torch.from_numpy(rng.integers(0, 200, size=B)).cuda()
    g = torch.from_numpy(rng.normal(size=(B, 16)).astype(np.float32)).cuda()
    emb(ids).backward(g)
    opt.step()
    tabs.apply_gradients({"item_id": ids}, {"item_id": g})
    got = tabs.state_dict()["embedding_bags.item_id.weight"]
    touched = torch.zeros(200, dtype=torch.bool, device="cuda")
    touched[ids] = True
    torch.testing.assert_close(got[touched], emb.weight.detach()[touched], rtol=1e-5, atol=2e-6)
    torch.testing.assert_close(got[~touched], w0[~touched], rtol=0, atol=0)


def test_rejects_host_arrays_and_missing_tables(t):
    from tdfo_b200.adapters import DLPackEmbeddingTables
    tabs = DLPackEmbeddingTables({"a": (10, 16), "b": (10, 16)}, "cuda")
    with pytest.raises(RuntimeError, match="no host path"):
        tabs.lookup({"a": torch.zeros(4, dtype=torch.int64), "b": torch.zeros(4, dtype=torch.int64)})
    with pytest.raises(KeyError):
        tabs.lookup({"a": torch.zeros(4, dtype=torch.int64, device="cuda")})
    with pytest.raises(ValueError, match="same number"):
        tabs.lookup({"a": torch.zeros(4, dtype=torch.int64, device="cuda"),
                     "b": torch.zeros(5, dtype=torch.int64, device="cuda")})


def test_tf_side_capsules_and_indexed_slices(t):
    """The tensorflow2 face (tensorflow2/models.py:48-58, train_dp.py:71-72): ids arrive as raw DLPack capsules
    (what `tf.experimental.dlpack.to_dlpack` returns), gradients as IndexedSlices-shaped (indices, values) with
    duplicate indices and a different count per table, some tables absent.  One SGD step equals the dense
    scatter-add + SGD that keras applies, restated with torch index_add_."""
    from types import SimpleNamespace
    from tdfo_b200.adapters import DLPackEmbeddingTables
    dev = torch.device("cuda")
    rng = np.random.default_rng(4)
    lr = 0.1
    tabs = DLPackEmbeddingTables(TABLES, dev, optimizer=t.EmbOptimType.EXACT_SGD, learning_rate=lr)
    w0 = {n: tabs.state_dict()[f"embedding_bags.{n}.weight"].clone() for n in TABLES}
    # lookup through capsules
    ids = _ids(rng, 24)
    caps = {n: torch.utils.dlpack.to_dlpack(torch.from_numpy(v).to(dev)) for n, v in ids.items()}
    rows = tabs.lookup(caps)
    for n, v in ids.items():
        assert torch.equal(rows[n], w0[n][torch.from_numpy(v).long().to(dev)])
    # IndexedSlices: ragged counts, duplicates, two tables without a gradient this step
    slices, want = {}, {n: w.clone() for n, w in w0.items()}
    for k, (n, (r, d)) in enumerate(TABLES.items()):
        if n in ("is_ebook", "format"):
            continue
        cnt = 5 + 7 * k
        idx = torch.from_numpy(rng.integers(0, min(r, 6), size=cnt)).to(dev)            # few distinct rows: many duplicates
        val = torch.from_numpy(rng.normal(size=(cnt, d)).astype(np.float32)).to(dev)
        dense = torch.zeros_like(w0[n]).index_add_(0, idx, val)
        want[n] -= lr * dense
        payload = (torch.utils.dlpack.to_dlpack(idx.to(torch.int32)), torch.utils.dlpack.to_dlpack(val))
        slices[n] = SimpleNamespace(indices=payload[0], values=payload[1]) if k % 2 else payload
    tabs.apply_indexed_slices(slices)
    tabs.check_errors()
    for n in TABLES:
        np.testing.assert_allclose(tabs.state_dict()[f"embedding_bags.{n}.weight"].cpu().numpy(), want[n].cpu().numpy(),
                                   rtol=1e-5, atol=1e-6, err_msg=n)
