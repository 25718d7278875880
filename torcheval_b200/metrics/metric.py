"""``Metric`` — the stateful base class, same lifecycle contract as the reference
(torcheval/metrics/metric.py:21-281): registered state variables, ``update / compute /
merge_state / reset / state_dict / load_state_dict / to / device``.

Instances stay picklable, deep-copyable and hashable (the reference's test harness checks all
three, utils/test_utils/metric_class_tester.py:148-154): no CUDA streams, library handles or
communicators are ever stored on the object — kernels resolve them per call.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from collections import defaultdict
from copy import deepcopy
from typing import Any, Dict, Generic, Iterable, List, Optional, TypeVar, Union

import torch

TSelf = TypeVar("TSelf", bound="Metric")
TComputeReturn = TypeVar("TComputeReturn")


def cat_all(tensors, dim: int = 0) -> torch.Tensor:
    """torch.cat without the copy when there is nothing to join: a sample-keeping metric that was
    updated once (or compacted by `_prepare_for_merge_state`) holds ONE tensor, and concatenating it
    with nothing used to duplicate the whole history (12.9 GB at 2^30 samples) on every compute().
    An empty list still raises what torch.cat raises, like the reference."""
    if len(tensors) == 1:
        return tensors[0]
    return torch.cat(tensors, dim)


def _copy_state(value: Any) -> Any:
    """An independent copy of a state value.  Tensors are cloned directly: copy.deepcopy of a
    tensor walks its storage through Python (~25 us each, six per binned metric constructed)."""
    if isinstance(value, torch.Tensor):
        return value.detach().clone()
    if isinstance(value, list) and all(isinstance(t, torch.Tensor) for t in value):
        return [t.detach().clone() for t in value]
    if isinstance(value, dict) and all(isinstance(t, torch.Tensor) for t in value.values()):
        out = value.__class__() if not isinstance(value, defaultdict) else defaultdict(value.default_factory)
        for k, t in value.items():
            out[k] = t.detach().clone()
        return out
    return deepcopy(value)
TState = Union[torch.Tensor, List[torch.Tensor], Dict[Any, torch.Tensor], int, float]


def _is_tensor_list(v: Any) -> bool:
    return isinstance(v, list) and all(isinstance(x, torch.Tensor) for x in v)


def _is_tensor_dict(v: Any) -> bool:
    return isinstance(v, dict) and all(isinstance(x, torch.Tensor) for x in v.values())


def _check_state_variable_type(name: str, value: Any) -> None:
    """A state is a Tensor, a list of Tensors, a dict with Tensor values, an int or a float
    (reference metric.py:260-281; same message)."""
    ok = isinstance(value, (torch.Tensor, int, float)) or _is_tensor_list(value) or _is_tensor_dict(value)
    if not ok:
        raise TypeError(
            "The value of state variable must be a ``torch.Tensor``, a list of ``torch.Tensor``, "
            f"a dictionary with ``torch.Tensor``, int, or float as values."
            f"Get {name}={value} instead."
        )


class Metric(Generic[TComputeReturn], ABC):
    """Base class of every metric.  Subclasses register their state with ``_add_state`` in
    ``__init__`` and implement ``update``, ``compute`` and ``merge_state``."""

    def __init__(self: TSelf, *, device: Optional[torch.device] = None) -> None:
        torch._C._log_api_usage_once(f"torcheval.metrics.{self.__class__.__name__}")
        self._state_name_to_default: Dict[str, TState] = {}
        self._device: torch.device = torch.device("cpu") if device is None else torch.device(device)

    def _add_state(self: TSelf, name: str, default: TState) -> None:
        """Register state ``name``; ``reset()`` restores a copy of ``default``."""
        _check_state_variable_type(name, default)
        setattr(self, name, _copy_state(default))
        self._state_name_to_default[name] = _copy_state(default)

    @abstractmethod
    @torch.inference_mode()
    def update(self: TSelf, *_: Any, **__: Any) -> TSelf:
        """Fold one batch into the state."""

    @abstractmethod
    @torch.inference_mode()
    def compute(self: TSelf) -> TComputeReturn:
        """Value of the metric for everything seen so far; does not consume the state."""

    @abstractmethod
    @torch.inference_mode()
    def merge_state(self: TSelf, metrics: Iterable[TSelf]) -> TSelf:
        """Fold the states of ``metrics`` into this one, leaving them unchanged."""

    @torch.inference_mode()
    def _prepare_for_merge_state(self: TSelf) -> None:
        """Hook called before a sync; list-state metrics compact their lists here."""

    def _default_on_device(self: TSelf, default: TState) -> TState:
        dev = self.device
        if isinstance(default, torch.Tensor):
            return default.clone().to(dev)
        if isinstance(default, list):
            return [t.clone().to(dev) for t in default]
        if isinstance(default, dict):
            return defaultdict(
                lambda: torch.tensor(0.0, device=dev),
                {k: t.clone().to(dev) for k, t in default.items()},
            )
        return default

    def reset(self: TSelf) -> TSelf:
        """Back to the registered defaults, on the device of the last ``to()`` call.  (As in the
        reference, int/float states keep their value: metric.py:126-146 only handles tensors,
        lists and dicts.)"""
        for name, default in self._state_name_to_default.items():
            if isinstance(default, (torch.Tensor, list, dict)):
                setattr(self, name, self._default_on_device(default))
        return self

    def state_dict(self: TSelf) -> Dict[str, TState]:
        out: Dict[str, TState] = {}
        for name in self._state_name_to_default:
            value = getattr(self, name)
            _check_state_variable_type(name, value)
            if isinstance(value, torch.Tensor):
                out[name] = value.detach().clone()
            elif isinstance(value, list):
                out[name] = [t.detach().clone() for t in value]
            elif isinstance(value, dict):
                out[name] = {k: t.detach().clone() for k, t in value.items()}
            else:
                out[name] = value
        return out

    def load_state_dict(self: TSelf, state_dict: Dict[str, Any], strict: bool = True) -> None:
        state_dict = deepcopy(state_dict)
        names = set(self._state_name_to_default.keys())
        for name in names:
            if name in state_dict:
                value = state_dict[name]
                _check_state_variable_type(name, value)
                setattr(self, name, value)
        if strict:
            keys = set(state_dict.keys())
            unexpected_keys = keys.difference(names)
            missing_keys = names.difference(keys)
            if missing_keys or unexpected_keys:
                raise RuntimeError(
                    f"Error(s) in loading state_dict for {self.__class__.__name__}. "
                    f"Encountered missing keys: {missing_keys} and unexpected "
                    f"keys: {unexpected_keys}."
                )

    def to(self: TSelf, device: Union[str, torch.device], *args: Any, **kwargs: Any) -> TSelf:
        device = torch.device(device) if isinstance(device, str) else device
        for name in self._state_name_to_default:
            value = getattr(self, name)
            _check_state_variable_type(name, value)
            if isinstance(value, torch.Tensor):
                setattr(self, name, value.to(device))
            elif isinstance(value, list):
                setattr(self, name, [t.to(device, *args, **kwargs) for t in value])
            elif isinstance(value, dict):
                setattr(
                    self,
                    name,
                    defaultdict(
                        lambda: torch.tensor(0.0, device=device),
                        {k: t.to(device, *args, **kwargs) for k, t in value.items()},
                    ),
                )
        self._device = device
        return self

    @property
    def device(self: TSelf) -> torch.device:
        """Device given at construction or to the last ``to()``; ``cpu`` by default."""
        return self._device
