"""Host-side mirror of the torchrec surface the reference calls, on the CPU: the
KeyedJaggedTensor / JaggedTensor / KeyedTensor containers, the optimizer plumbing of
torchrec/train.py:248-254, and the sharding-plan helpers.  No kernel is involved."""
import numpy as np
import pytest
import torch
import torch.nn as nn

import tdfo_b200 as t
from tdfo_b200 import distributed as td


def to_kjt(seqs, device="cpu"):
    """The reference's `_to_kjt` (torchrec/train.py:33-41), word for word in behaviour: one key "item",
    values = concatenated sequences, lengths = their sizes."""
    seqs_list = list(seqs)
    lengths = torch.IntTensor([value.size(0) for value in seqs_list])
    values = torch.cat(seqs_list, dim=0)
    return t.KeyedJaggedTensor.from_lengths_sync(keys=["item"], values=values, lengths=lengths).to(device)


def test_to_kjt_pattern_of_the_reference():
    seqs = torch.arange(24).reshape(4, 6)
    kjt = to_kjt(seqs)
    assert kjt.keys() == ["item"] and kjt.stride() == 4
    assert kjt.lengths().dtype == torch.int32 and kjt.lengths().tolist() == [6] * 4
    assert kjt.offsets().dtype == torch.int64 and kjt.offsets().tolist() == [0, 6, 12, 18, 24]
    jt = kjt["item"]
    assert torch.equal(jt.values(), seqs.reshape(-1)) and jt.offsets().tolist() == [0, 6, 12, 18, 24]


def test_kjt_views_are_feature_major_with_rebased_offsets():
    lengths = torch.tensor([2, 0, 1, 3, 1, 0], dtype=torch.int32)           # keys a, b; B = 3
    values = torch.arange(7) + 10
    w = torch.arange(7, dtype=torch.float32) / 10
    kjt = t.KeyedJaggedTensor.from_lengths_sync(["a", "b"], values, lengths, weights=w)
    assert kjt.stride() == 3 and kjt.length_per_key() == [3, 4] and kjt.offset_per_key() == [0, 3, 7]
    b = kjt["b"]
    assert b.values().tolist() == [13, 14, 15, 16] and b.lengths().tolist() == [3, 1, 0]
    assert b.offsets().tolist() == [0, 3, 4, 4]                             # rebased to the slice
    assert torch.equal(b.weights(), w[3:])
    assert [x.tolist() for x in kjt["a"].to_dense()] == [[10, 11], [], [12]]
    assert list(kjt.to_dict()) == ["a", "b"]
    same = t.KeyedJaggedTensor.from_offsets_sync(["a", "b"], values, kjt.offsets())
    assert same.lengths().tolist() == lengths.tolist()
    first, second = kjt.split([1, 1])
    assert first.keys() == ["a"] and second.keys() == ["b"] and second.values().tolist() == [13, 14, 15, 16]
    assert second.offsets().tolist() == [0, 3, 4, 4] and torch.equal(second.weights(), w[3:])
    with pytest.raises(ValueError):
        t.KeyedJaggedTensor.from_lengths_sync(["a"], values, lengths[:3]).weights()
    with pytest.raises(RuntimeError, match="GPU"):
        kjt.permute([1, 0])                                                  # a kernel: no CPU path


def test_keyed_tensor_columns():
    v = torch.arange(2 * 7, dtype=torch.float32).reshape(2, 7)
    kt = t.KeyedTensor(["x", "y", "z"], [2, 4, 1], v)
    assert kt.keys() == ["x", "y", "z"] and kt.length_per_key() == [2, 4, 1] and kt.offset_per_key() == [0, 2, 6, 7]
    assert torch.equal(kt["y"], v[:, 2:6]) and torch.equal(kt.to_dict()["z"], v[:, 6:7]) and kt.values() is v


def test_optimizer_plumbing_of_train_py():
    """torchrec/train.py:248-254: fused parameters are filtered out of the dense optimizer, both sit
    behind one CombinedOptimizer."""
    ebc = t.EmbeddingBagCollection([t.EmbeddingBagConfig(num_embeddings=10, embedding_dim=8, name="a",
                                                         feature_names=["fa"])])
    fused = ebc.set_fused_optimizer(t.EmbOptimType.ADAM, learning_rate=3e-4, weight_decay=1e-4)
    model = nn.ModuleDict({"sparse": ebc, "dense": nn.Linear(8, 1)})
    dense_named = dict(t.in_backward_optimizer_filter(model.named_parameters()))
    assert sorted(dense_named) == ["dense.bias", "dense.weight"]
    assert [n for n, _ in t.in_backward_optimizer_filter(model.named_parameters(), include=True)] == \
        ["sparse." + ebc._groups[0].param_name]
    dense = t.KeyedOptimizerWrapper(dense_named, lambda params: torch.optim.Adam(params, lr=0.5))
    opt = t.CombinedOptimizer([fused, dense])
    assert len(opt.optimizers) == 2
    w0 = model["dense"].weight.detach().clone()
    model["dense"](torch.ones(3, 8)).sum().backward()
    opt.step()
    assert not torch.equal(model["dense"].weight, w0)                        # the dense half stepped
    opt.zero_grad()
    assert model["dense"].weight.grad is None or not model["dense"].weight.grad.any()
    sd = opt.state_dict()
    opt.load_state_dict(sd)                                                  # round trip
    assert fused.state_dict()["hyper"]["lr"] == pytest.approx(3e-4)
    assert fused.state_dict()["hyper"]["weight_decay"] == pytest.approx(1e-4)
    assert fused.args().optimizer == t.EmbOptimType.ADAM.code()


def test_sharding_plan_helpers():
    mk = lambda n, rows: t.EmbeddingBagConfig(num_embeddings=rows, embedding_dim=128, name=n, feature_names=[n])
    tables = [mk(f"t{i}", 1000) for i in range(26)]
    plan = td.balanced_plan(tables, 8)
    kinds = [plan[f"t{i}"].sharding_type for i in range(26)]
    assert kinds.count(td.ShardingType.TABLE_WISE) == 24 and kinds.count(td.ShardingType.COLUMN_WISE) == 2
    cols = [0] * 8                                                           # columns held per rank
    for i in range(26):
        ps = plan[f"t{i}"]
        if ps.sharding_type == td.ShardingType.TABLE_WISE:
            cols[ps.rank] += 128
        else:
            assert len(ps.ranks) == 4
            for r in ps.ranks:
                cols[r] += 128 // 4
    assert len(set(cols)) == 1                                               # perfectly balanced: 3 tables + 32 columns
    assert all(p.sharding_type == td.ShardingType.TABLE_WISE for p in td.balanced_plan(tables, 2).values())
    p4 = td.balanced_plan(tables, 4)
    assert sorted(len(p.ranks) for p in p4.values() if p.sharding_type == td.ShardingType.COLUMN_WISE) == [2, 2]
    # a shard row below 64 bytes is not worth it: falls back to table-wise
    small = [t.EmbeddingBagConfig(num_embeddings=10, embedding_dim=16, name=f"s{i}", feature_names=[f"s{i}"]) for i in range(3)]
    assert all(p.sharding_type == td.ShardingType.TABLE_WISE for p in td.balanced_plan(small, 2).values())
    # greedy size balance
    sized = [mk("big", 9000), mk("m1", 5000), mk("m2", 4000), mk("s", 100)]
    g = td.balanced_table_wise_plan(sized, 2)
    load = [sum(c.num_embeddings for c in sized if g[c.name].rank == r) for r in range(2)]
    assert sorted(load) == [9000, 9100]
    assert td.column_wise((3, 1)).ranks == [3, 1] and td.row_wise().sharding_type == td.ShardingType.ROW_WISE
