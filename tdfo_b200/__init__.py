"""tdfo_b200 — B200-native sparse-embedding hot path behind the torchrec-style surface that
massquantity/tdfo's ``torchrec/train.py`` and ``models.py`` call.

Everything on the path runs in hand-written sm_100a kernels exported through the C ABI in
``include/tdfo_b200.h`` (``libtdfo_b200.so``).  There is no CPU or PyTorch fallback.
"""
from . import ops
from .modules import (BoundsCheckMode, DataType, EmbeddingBagCollection, EmbeddingBagConfig, EmbeddingCollection,
                      EmbeddingConfig, PoolingType)
from .optim import (CombinedOptimizer, EmbOptimType, FusedOptimizer, KeyedOptimizerWrapper,
                    in_backward_optimizer_filter)
from .ops import jagged_2d_to_dense
from .sparse import JaggedTensor, KeyedJaggedTensor, KeyedTensor
from .adapters import DLPackEmbeddingTables
from .pipeline import DevicePrefetcher
from .distributed import (DistributedModelParallel, EmbeddingBagCollectionSharder, EmbeddingCollectionSharder,
                          ParameterSharding, ShardedEmbeddingBagCollection, ShardedEmbeddingCollection, ShardingType, balanced_plan,
                          balanced_table_wise_plan, column_wise, row_wise, table_wise)

__all__ = [
    "ops", "BoundsCheckMode", "DataType", "EmbeddingBagCollection", "EmbeddingBagConfig", "EmbeddingCollection", "EmbeddingConfig",
    "PoolingType", "CombinedOptimizer", "EmbOptimType", "FusedOptimizer", "KeyedOptimizerWrapper",
    "in_backward_optimizer_filter", "jagged_2d_to_dense", "DLPackEmbeddingTables", "DevicePrefetcher", "JaggedTensor", "KeyedJaggedTensor", "KeyedTensor",
    "DistributedModelParallel", "EmbeddingBagCollectionSharder", "EmbeddingCollectionSharder", "ParameterSharding",
    "ShardedEmbeddingBagCollection", "ShardedEmbeddingCollection", "ShardingType", "balanced_plan", "balanced_table_wise_plan", "column_wise", "row_wise", "table_wise",
]
