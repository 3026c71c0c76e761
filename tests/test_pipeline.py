"""DevicePrefetcher (the reference's prefetch_to_device(size=2), jax-flax/train_dp.py:211):
order, nesting and exhaustion on the CPU, stream safety on the GPU."""
import collections

import pytest
import torch

import tdfo_b200 as t

Batch = collections.namedtuple("Batch", ["kjt", "labels", "meta"])


def make(i, B=4):
    lengths = torch.tensor([(i + b) % 3 for b in range(2 * B)], dtype=torch.int32)
    values = torch.arange(int(lengths.sum())) + 100 * i
    kjt = t.KeyedJaggedTensor.from_lengths_sync(["a", "b"], values, lengths)
    return Batch(kjt, {"y": torch.full((B,), float(i))}, ("step", i))


def test_prefetcher_preserves_order_and_structure_on_cpu():
    seen = list(t.DevicePrefetcher((make(i) for i in range(7)), "cpu", depth=3))
    assert len(seen) == 7
    for i, b in enumerate(seen):
        assert isinstance(b, Batch) and b.meta == ("step", i)
        ref = make(i)
        assert torch.equal(b.kjt.values(), ref.kjt.values()) and torch.equal(b.kjt.lengths(), ref.kjt.lengths())
        assert b.kjt.keys() == ["a", "b"] and torch.equal(b.labels["y"], ref.labels["y"])
    assert list(t.DevicePrefetcher([], "cpu")) == []
    with pytest.raises(ValueError):
        t.DevicePrefetcher([], "cpu", depth=0)


@pytest.mark.gpu
def test_prefetcher_feeds_training_on_the_gpu():
    """Same losses with and without the prefetcher while the consumer keeps the GPU busy."""
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    tables = [t.EmbeddingBagConfig(num_embeddings=1000, embedding_dim=64, name=n, feature_names=[n]) for n in "ab"]

    def batches():
        g = torch.Generator().manual_seed(0)
        for i in range(12):
            lengths = torch.randint(0, 30, (2 * 512,), generator=g, dtype=torch.int32)
            values = torch.randint(0, 1000, (int(lengths.sum()),), generator=g)
            yield t.KeyedJaggedTensor.from_lengths_sync(["a", "b"], values, lengths), torch.randn(512, 128, generator=g)

    def run(prefetch):
        torch.manual_seed(1)
        ebc = t.EmbeddingBagCollection(tables, device=torch.device("cuda"))
        ebc.set_fused_optimizer(t.EmbOptimType.EXACT_ROWWISE_ADAGRAD, learning_rate=0.1)
        ebc.train()
        src = t.DevicePrefetcher(batches(), "cuda", depth=2) if prefetch else \
            ((k.to("cuda"), y.to("cuda")) for k, y in batches())
        losses = []
        for kjt, y in src:
            assert kjt.values().is_cuda and y.is_cuda
            loss = (ebc(kjt).values() * y).sum()
            loss.backward()
            losses.append(loss)
        ebc.check_errors()
        return torch.stack(losses).cpu()

    torch.testing.assert_close(run(True), run(False), rtol=0, atol=0)
