"""Optimizer plumbing with the names the reference imports (torchrec/train.py:13,20-21,248-254).

The embedding update itself is fused into the backward kernel (``tdfo_sparse_update``);
``FusedOptimizer`` is the handle ``model.fused_optimizer`` returns so the reference's
``CombinedOptimizer([model.fused_optimizer, dense_optimizer])`` idiom keeps working.
"""
from __future__ import annotations

import enum
from typing import Any, Callable, Dict, Iterable, Iterator, List, Tuple

import torch

from . import _lib


class EmbOptimType(enum.Enum):
    """fbgemm_gpu.split_embedding_configs.EmbOptimType members the path supports."""
    SGD = "sgd"
    EXACT_SGD = "exact_sgd"
    ADAM = "adam"
    EXACT_ROWWISE_ADAGRAD = "exact_row_wise_adagrad"
    ROWWISE_ADAGRAD = "row_wise_adagrad"

    def code(self) -> int:
        return {
            EmbOptimType.SGD: _lib.OPT_SGD, EmbOptimType.EXACT_SGD: _lib.OPT_SGD,
            EmbOptimType.ADAM: _lib.OPT_ADAM,
            EmbOptimType.EXACT_ROWWISE_ADAGRAD: _lib.OPT_ROWWISE_ADAGRAD,
            EmbOptimType.ROWWISE_ADAGRAD: _lib.OPT_ROWWISE_ADAGRAD,
        }[self]


class FusedOptimizer:
    """Handle on the optimizer that runs inside the embedding backward.

    ``step()`` / ``zero_grad()`` do nothing: rows were already updated by the time
    ``loss.backward()`` returns (torchrec/train.py:104-105).  Hyper-parameters live in
    ``param_groups`` (one group) and are read at every backward, so LR schedules work.
    """

    def __init__(self, params: Dict[str, torch.Tensor], optimizer: EmbOptimType = EmbOptimType.EXACT_SGD,
                 learning_rate: float = 0.01, eps: float = 1e-8, weight_decay: float = 0.0,
                 beta1: float = 0.9, beta2: float = 0.999):
        self.params = dict(params)
        self.optimizer = optimizer
        self.param_groups: List[Dict[str, Any]] = [{
            "params": list(self.params.values()), "lr": float(learning_rate), "eps": float(eps),
            "weight_decay": float(weight_decay), "beta1": float(beta1), "beta2": float(beta2),
        }]
        self.iteration = 0          # backward passes applied so far (ADAM bias correction)
        self.state: Dict[str, Any] = {}

    def args(self) -> "_lib.OptimizerArgs":
        g = self.param_groups[0]
        return _lib.OptimizerArgs(self.optimizer.code(), g["lr"], g["eps"], g["weight_decay"], g["beta1"],
                                  g["beta2"], max(self.iteration, 1))

    def zero_grad(self, set_to_none: bool = True) -> None:
        pass

    def step(self, closure: Any = None) -> None:
        pass

    def state_dict(self) -> Dict[str, Any]:
        g = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        return {"optimizer": self.optimizer.value, "iteration": self.iteration, "hyper": g,
                "state": {k: v for k, v in self.state.items()}}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        self.iteration = int(sd["iteration"])
        self.param_groups[0].update(sd["hyper"])
        dev = next(iter(self.params.values())).device if self.params else None
        for k, v in sd.get("state", {}).items():
            if not isinstance(v, torch.Tensor):
                continue
            if k in self.state:
                self.state[k].copy_(v)
            else:                       # state is created lazily by the first backward: adopt the loaded one
                self.state[k] = v.detach().clone().to(dev) if dev is not None else v.detach().clone()


class EmptyFusedOptimizer(FusedOptimizer):
    def __init__(self):
        super().__init__({})


class CombinedFusedOptimizer(FusedOptimizer):
    """Several fused handles (one per sharded/fused module) behind one handle."""

    def __init__(self, optims: List[FusedOptimizer]):
        self._optims = list(optims)
        params: Dict[str, torch.Tensor] = {}
        for i, o in enumerate(self._optims):
            for k, v in o.params.items():
                params[k if k not in params else f"{i}.{k}"] = v
        super().__init__(params)
        self.param_groups = [g for o in self._optims for g in o.param_groups]

    def state_dict(self) -> Dict[str, Any]:
        return {str(i): o.state_dict() for i, o in enumerate(self._optims)}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        for i, o in enumerate(self._optims):
            o.load_state_dict(sd[str(i)])


def in_backward_optimizer_filter(named_parameters: Iterable[Tuple[str, torch.nn.Parameter]],
                                 include: bool = False) -> Iterator[Tuple[str, torch.nn.Parameter]]:
    """Drop (or, with ``include=True``, keep only) parameters updated inside backward
    — torchrec/train.py:249."""
    for name, p in named_parameters:
        if hasattr(p, "_in_backward_optimizers") == include:
            yield name, p


class KeyedOptimizerWrapper:
    """``KeyedOptimizerWrapper(dict(named_params), lambda params: optim.Adam(params, ...))``
    — torchrec/train.py:248-253."""

    def __init__(self, params: Dict[str, torch.Tensor], optim_factory: Callable[[List[torch.Tensor]], torch.optim.Optimizer]):
        self.params = dict(params)
        self._optimizer = optim_factory(list(self.params.values())) if self.params else None

    @property
    def param_groups(self):
        return self._optimizer.param_groups if self._optimizer else []

    def zero_grad(self, set_to_none: bool = True) -> None:
        if self._optimizer:
            self._optimizer.zero_grad(set_to_none=set_to_none)

    def step(self, closure: Any = None) -> None:
        if self._optimizer:
            self._optimizer.step(closure)

    def state_dict(self) -> Dict[str, Any]:
        return self._optimizer.state_dict() if self._optimizer else {}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        if self._optimizer:
            self._optimizer.load_state_dict(sd)


class CombinedOptimizer:
    """``CombinedOptimizer([model.fused_optimizer, dense_optimizer])`` — torchrec/train.py:254."""

    def __init__(self, optims: List[Any]):
        self._optims = list(optims)

    @property
    def optimizers(self) -> List[Any]:
        return self._optims

    @property
    def param_groups(self) -> List[Dict[str, Any]]:
        return [g for o in self._optims for g in o.param_groups]

    def zero_grad(self, set_to_none: bool = True) -> None:
        for o in self._optims:
            o.zero_grad(set_to_none=set_to_none)

    def step(self, closure: Any = None) -> None:
        for o in self._optims:
            o.step(closure)

    def state_dict(self) -> Dict[str, Any]:
        return {str(i): o.state_dict() for i, o in enumerate(self._optims)}

    def load_state_dict(self, sd: Dict[str, Any]) -> None:
        for i, o in enumerate(self._optims):
            o.load_state_dict(sd[str(i)])
