"""`tdfo_b200.compat.install()`: the reference's module paths resolve to this package, so its own source
imports unmodified.  Where /root/reference is present (the build
container, not the GPU box) the reference's `torchrec/models.py` itself is imported and its Bert4Rec built
on top of our EmbeddingCollection."""
import importlib.util
import sys
from pathlib import Path

import pytest
import torch

REF_MODELS = Path("/root/reference/torchrec/models.py")


@pytest.fixture(scope="module")
def compat():
    import tdfo_b200.compat as c
    c.install()
    c.install()                                   # idempotent
    return c


def test_reference_import_lines_resolve(compat):
    import tdfo_b200 as t
    from fbgemm_gpu.split_embedding_configs import EmbOptimType                       # torchrec/train.py:13
    from torchrec.distributed.embedding import EmbeddingCollectionSharder              # :17
    from torchrec.distributed.model_parallel import DistributedModelParallel as DMP    # :18
    from torchrec.distributed.types import ModuleSharder                               # :19
    from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizerWrapper          # :20
    from torchrec.optim.optimizers import in_backward_optimizer_filter                 # :21
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor                        # :22
    from torchrec.modules.embedding_configs import EmbeddingConfig                     # torchrec/models.py:6
    from torchrec.modules.embedding_modules import EmbeddingCollection                 # :7
    assert EmbOptimType is t.EmbOptimType and DMP is t.DistributedModelParallel
    assert EmbeddingCollectionSharder is t.EmbeddingCollectionSharder and KeyedJaggedTensor is t.KeyedJaggedTensor
    assert CombinedOptimizer is t.CombinedOptimizer and KeyedOptimizerWrapper is t.KeyedOptimizerWrapper
    assert in_backward_optimizer_filter is t.in_backward_optimizer_filter
    assert EmbeddingConfig is t.EmbeddingConfig and EmbeddingCollection is t.EmbeddingCollection
    from typing import cast
    import torch.nn as nn
    sharder = cast(ModuleSharder[nn.Module], EmbeddingCollectionSharder({"optimizer": EmbOptimType.ADAM}))   # :243-245
    assert isinstance(sharder, t.EmbeddingCollectionSharder)
    assert hasattr(torch.ops.fbgemm, "jagged_2d_to_dense")                             # torchrec/models.py:168
    with pytest.raises(RuntimeError, match="no CPU path"):                             # it is our op: CUDA only
        torch.ops.fbgemm.jagged_2d_to_dense(values=torch.zeros(3, 4), offsets=torch.tensor([0, 3]),
                                            max_sequence_length=5)


@pytest.mark.skipif(not REF_MODELS.exists(), reason="the reference checkout exists only in the build container")
def test_the_reference_models_py_builds_on_our_modules(compat):
    import tdfo_b200 as t
    spec = importlib.util.spec_from_file_location("tdfo_reference_models", REF_MODELS)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)                                                       # unmodified reference source
    model = ref.Bert4Rec(vocab_size=52, max_len=20, embed_dim=16, num_heads=2, num_layers=1)
    ec = model.history.embed_collection
    assert isinstance(ec, t.EmbeddingCollection)
    cfg = ec.embedding_configs()[0]
    assert (cfg.name, cfg.num_embeddings, cfg.embedding_dim, cfg.feature_names) == ("item_embedding", 52, 16, ["item"])
    w = ec.state_dict()["embeddings.item_embedding.weight"]
    assert w.shape == (52, 16) and float(w.min()) >= -1.0 and float(w.max()) <= 1.0     # weight_init_min / max honoured
    # the fused table is not handed to the dense optimizer (torchrec/train.py:248-249)
    dense = dict(t.in_backward_optimizer_filter(model.named_parameters()))
    assert not any("embed_collection" in k for k in dense) and "out.weight" in dense
    # and the model's forward reaches our module (which has no CPU path)
    seqs = torch.randint(1, 52, (4, 20))
    kjt = t.KeyedJaggedTensor.from_lengths_sync(keys=["item"], values=seqs.reshape(-1),
                                                lengths=torch.IntTensor([20] * 4))
    with pytest.raises(RuntimeError, match="no CPU path|CUDA"):
        model(kjt)


@pytest.mark.gpu
def test_fbgemm_op_alias_is_differentiable_on_the_gpu(compat):
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import numpy as np
    import oracle as orc
    offsets = torch.tensor([0, 3, 3, 7, 9], device="cuda")
    values = torch.randn(9, 8, device="cuda", requires_grad=True)
    dense = torch.ops.fbgemm.jagged_2d_to_dense(values=values, offsets=offsets, max_sequence_length=3)
    want = orc.jagged_2d_to_dense(values.detach().cpu().numpy(), offsets.cpu().numpy(), 3)
    np.testing.assert_array_equal(dense.detach().cpu().numpy(), want)
    g = torch.randn_like(dense)
    dense.backward(g)
    np.testing.assert_array_equal(values.grad.cpu().numpy(), orc.dense_to_jagged(g.cpu().numpy(), offsets.cpu().numpy()))
