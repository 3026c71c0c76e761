"""Import-level drop-in: `import tdfo_b200.compat; tdfo_b200.compat.install()` makes the reference's own
`torchrec/models.py` and `torchrec/train.py` import UNMODIFIED on a machine without torchrec /
fbgemm_gpu, by registering this package's mirrors under the module paths they import from
(torchrec/models.py:6-8, torchrec/train.py:13,17-22) and `torch.ops.fbgemm.jagged_2d_to_dense`
(torchrec/models.py:168-172).

    torchrec.modules.embedding_configs      EmbeddingConfig, EmbeddingBagConfig, PoolingType, DataType
    torchrec.modules.embedding_modules      EmbeddingCollection, EmbeddingBagCollection
    torchrec.sparse.jagged_tensor           KeyedJaggedTensor, JaggedTensor, KeyedTensor
    torchrec.distributed.embedding          EmbeddingCollectionSharder
    torchrec.distributed.embeddingbag       EmbeddingBagCollectionSharder
    torchrec.distributed.model_parallel     DistributedModelParallel
    torchrec.distributed.types              ModuleSharder, ShardingType, ParameterSharding
    torchrec.optim.keyed                    CombinedOptimizer, KeyedOptimizerWrapper
    torchrec.optim.optimizers               in_backward_optimizer_filter
    fbgemm_gpu.split_embedding_configs      EmbOptimType

A real torchrec / fbgemm_gpu installation is never shadowed: install() refuses if either imports.
Nothing here computes anything; it is a table of names.
"""
from __future__ import annotations

import importlib.util
import sys
import types
from typing import Generic, TypeVar

import torch

_M = TypeVar("_M")
_installed = False


class ModuleSharder(Generic[_M]):
    """Typing placeholder for `cast(ModuleSharder[nn.Module], ...)` at torchrec/train.py:243-245."""


def _module(name: str, package: bool = False) -> types.ModuleType:
    m = sys.modules.get(name)
    if m is None:
        m = types.ModuleType(name)
        m.__tdfo_b200_compat__ = True
        if package:
            m.__path__ = []                     # lets `import a.b.c` resolve through sys.modules
        sys.modules[name] = m
        parent, _, child = name.rpartition(".")
        if parent:
            setattr(_module(parent, package=True), child, m)
    return m


def _register_fbgemm_ops() -> None:
    from . import ops
    try:
        torch.ops.fbgemm.jagged_2d_to_dense      # a real fbgemm_gpu already defined it
        return
    except (AttributeError, RuntimeError):
        pass
    lib = torch.library.Library("fbgemm", "FRAGMENT")
    lib.define("jagged_2d_to_dense(Tensor values, Tensor offsets, int max_sequence_length) -> Tensor")
    # composite: the body is the package's own differentiable op, so autograd comes along; CPU tensors
    # are rejected by it (there is no CPU path)
    lib.impl("jagged_2d_to_dense", lambda values, offsets, max_sequence_length: ops.jagged_2d_to_dense(
        values, offsets, max_sequence_length), "CompositeImplicitAutograd")
    _register_fbgemm_ops._lib = lib              # keep the registration alive


def install() -> None:
    """Idempotent.  Raises if the real packages are present."""
    global _installed
    if _installed:
        return
    for real in ("torchrec", "fbgemm_gpu"):
        m = sys.modules.get(real)
        if (m is not None and not getattr(m, "__tdfo_b200_compat__", False)) or \
                (m is None and importlib.util.find_spec(real) is not None):
            raise RuntimeError(f"tdfo_b200.compat: a real `{real}` is installed; refusing to shadow it "
                               "(use the tdfo_b200 names directly, see INTEGRATION.md §4)")
    import tdfo_b200 as t
    from . import distributed as td
    table = {
        "torchrec.modules.embedding_configs": dict(EmbeddingConfig=t.EmbeddingConfig, EmbeddingBagConfig=t.EmbeddingBagConfig,
                                                   PoolingType=t.PoolingType, DataType=t.DataType),
        "torchrec.modules.embedding_modules": dict(EmbeddingCollection=t.EmbeddingCollection,
                                                   EmbeddingBagCollection=t.EmbeddingBagCollection),
        "torchrec.sparse.jagged_tensor": dict(KeyedJaggedTensor=t.KeyedJaggedTensor, JaggedTensor=t.JaggedTensor,
                                              KeyedTensor=t.KeyedTensor),
        "torchrec.distributed.embedding": dict(EmbeddingCollectionSharder=t.EmbeddingCollectionSharder),
        "torchrec.distributed.embeddingbag": dict(EmbeddingBagCollectionSharder=t.EmbeddingBagCollectionSharder),
        "torchrec.distributed.model_parallel": dict(DistributedModelParallel=t.DistributedModelParallel),
        "torchrec.distributed.types": dict(ModuleSharder=ModuleSharder, ShardingType=td.ShardingType,
                                           ParameterSharding=td.ParameterSharding),
        "torchrec.optim.keyed": dict(CombinedOptimizer=t.CombinedOptimizer, KeyedOptimizerWrapper=t.KeyedOptimizerWrapper),
        "torchrec.optim.optimizers": dict(in_backward_optimizer_filter=t.in_backward_optimizer_filter),
        "fbgemm_gpu.split_embedding_configs": dict(EmbOptimType=t.EmbOptimType),
    }
    for pkg in ("torchrec", "torchrec.modules", "torchrec.sparse", "torchrec.distributed", "torchrec.optim", "fbgemm_gpu"):
        _module(pkg, package=True)
    for name, attrs in table.items():
        m = _module(name)
        for k, v in attrs.items():
            setattr(m, k, v)
    top = sys.modules["torchrec"]
    for k in ("EmbeddingConfig", "EmbeddingBagConfig", "EmbeddingCollection", "EmbeddingBagCollection", "KeyedJaggedTensor",
              "JaggedTensor", "KeyedTensor", "PoolingType", "DataType"):
        setattr(top, k, getattr(t, k))
    _register_fbgemm_ops()
    _installed = True
