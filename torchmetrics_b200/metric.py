"""`Metric`: the stateful boundary class of the hot path.

Observable behaviour follows the reference's `torchmetrics.metric.Metric` (src/torchmetrics/metric.py:50-1085):
`add_state` registry (:195-272), the two `forward` modes (:274-391), wrapped `update` / `compute` with result
caching (:459-481, :593-623), `reset` (:673-688), the `sync` / `unsync` / `sync_context` state machine with its
error strings (:490-591), device / dtype moves through `_apply` (:782-832), `state_dict` of persistent states
(:839-890), pickling (:694-713) and `_filter_kwargs` (:892-911).  Attribute names that other code relies on
(`_defaults`, `_reductions`, `_persistent`, `_update_count`, `_computed`, `_cache`, `_is_synced`, `_to_sync`,
`_should_unsync`, `_enable_grad`, `_forward_cache`) are kept.

What differs is the default cross-rank synchronisation: instead of barrier + 2 all_gathers + stack + reduce per
state (:427-457), every reducible tensor state goes through one packed all_reduce and every cat state through one
padded all_gather (`utilities.distributed.fused_state_sync`).  A user-supplied `dist_sync_fn` still gets the
reference's protocol: one call per state tensor, `fn(tensor, group=...) -> list of tensors`.
"""
from __future__ import annotations

import functools
import inspect
from abc import ABC, abstractmethod
from contextlib import contextmanager
from copy import deepcopy
from typing import Any, Callable, Dict, Generator, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from torch.nn import Module

from torchmetrics_b200.utilities.data import (
    _flatten,
    _squeeze_if_scalar,
    apply_to_collection,
    dim_zero_cat,
    dim_zero_max,
    dim_zero_mean,
    dim_zero_min,
    dim_zero_sum,
)
from torchmetrics_b200.utilities.distributed import (
    distributed_available,
    fused_cat_gather,
    fused_reduce,
    gather_all_tensors,
)
from torchmetrics_b200.utilities.exceptions import TorchMetricsUserError
from torchmetrics_b200.utilities.prints import rank_zero_warn

_NAMED_REDUCTIONS: Dict[str, Callable] = {
    "sum": dim_zero_sum,
    "mean": dim_zero_mean,
    "max": dim_zero_max,
    "min": dim_zero_min,
    "cat": dim_zero_cat,
}
_FUSABLE = {dim_zero_sum: "sum", dim_zero_mean: "mean", dim_zero_max: "max", dim_zero_min: "min"}
_CONSTANTS = ("higher_is_better", "is_differentiable", "full_state_update", "plot_lower_bound", "plot_upper_bound",
              "plot_legend_name")


def jit_distributed_available() -> bool:
    return distributed_available()


class Metric(Module, ABC):
    """Base class: registers states, accumulates them in `update`, merges them across ranks for `compute`."""

    __jit_ignored_attributes__ = ["device"]
    __jit_unused_properties__ = ["is_differentiable", "higher_is_better", "plot_lower_bound", "plot_upper_bound",
                                 "plot_legend_name", "metric_state", "_update_called", "update_called", "update_count",
                                 "dtype", "device"]
    is_differentiable: Optional[bool] = None
    higher_is_better: Optional[bool] = None
    full_state_update: Optional[bool] = None
    plot_lower_bound: Optional[float] = None
    plot_upper_bound: Optional[float] = None
    plot_legend_name: Optional[str] = None

    def __init__(self, **kwargs: Any) -> None:
        super().__init__()
        self._device = torch.device("cpu")
        self._dtype = torch.get_default_dtype()

        def pop_flag(name: str, default: bool, article: str = "a") -> bool:
            val = kwargs.pop(name, default)
            if not isinstance(val, bool):
                raise ValueError(f"Expected keyword argument `{name}` to be {article} `bool` but got {val}")
            return val

        self.compute_on_cpu = pop_flag("compute_on_cpu", False, "an")
        self.dist_sync_on_step = pop_flag("dist_sync_on_step", False, "an")
        self.process_group = kwargs.pop("process_group", None)
        self.dist_sync_fn = kwargs.pop("dist_sync_fn", None)
        if self.dist_sync_fn is not None and not callable(self.dist_sync_fn):
            raise ValueError(
                f"Expected keyword argument `dist_sync_fn` to be an callable function but got {self.dist_sync_fn}"
            )
        self.distributed_available_fn = kwargs.pop("distributed_available_fn", None) or jit_distributed_available
        self.sync_on_compute = pop_flag("sync_on_compute", True)
        self.compute_with_cache = pop_flag("compute_with_cache", True)
        if kwargs:
            raise ValueError("Unexpected keyword arguments: " + ", ".join(f"`{k}`" for k in sorted(kwargs)))

        self._install_wrappers()
        self._computed: Any = None
        self._forward_cache: Any = None
        self._update_count = 0
        self._to_sync = self.sync_on_compute
        self._should_unsync = True
        self._enable_grad = False
        self._dtype_convert = False

        self._defaults: Dict[str, Union[List, Tensor]] = {}
        self._persistent: Dict[str, bool] = {}
        self._reductions: Dict[str, Union[str, Callable, None]] = {}

        self._is_synced = False
        self._cache: Optional[Dict[str, Union[List[Tensor], Tensor]]] = None

    def _install_wrappers(self) -> None:
        self._update_signature = inspect.signature(self.update)
        self.update: Callable = self._wrap_update(self.update)  # type: ignore[method-assign]
        self.compute: Callable = self._wrap_compute(self.compute)  # type: ignore[method-assign]

    # ------------------------------------------------------------------ bookkeeping properties
    @property
    def _update_called(self) -> bool:
        rank_zero_warn("This property will be removed in 2.0.0. Use `Metric.updated_called` instead.",
                       DeprecationWarning, stacklevel=2)
        return self.update_called

    @property
    def update_called(self) -> bool:
        return self._update_count > 0

    @property
    def update_count(self) -> int:
        return self._update_count

    @property
    def metric_state(self) -> Dict[str, Union[List[Tensor], Tensor]]:
        return {name: getattr(self, name) for name in self._defaults}

    @property
    def device(self) -> "torch.device":
        return self._device

    @property
    def dtype(self) -> "torch.dtype":
        return self._dtype

    # ------------------------------------------------------------------ state registry
    def add_state(self, name: str, default: Union[list, Tensor], dist_reduce_fx: Optional[Union[str, Callable]] = None,
                  persistent: bool = False) -> None:
        """Register a state: a tensor, or an empty list that `update` appends tensors to (metric.py:195-272)."""
        if not isinstance(default, (Tensor, list)) or (isinstance(default, list) and default):
            raise ValueError("state variable must be a tensor or any empty list (where you can append tensors)")
        if isinstance(dist_reduce_fx, str) and dist_reduce_fx in _NAMED_REDUCTIONS:
            reduce_fn: Optional[Callable] = _NAMED_REDUCTIONS[dist_reduce_fx]
        elif dist_reduce_fx is None or callable(dist_reduce_fx):
            reduce_fn = dist_reduce_fx
        else:
            raise ValueError("`dist_reduce_fx` must be callable or one of ['mean', 'sum', 'cat', 'min', 'max', None]")
        if isinstance(default, Tensor):
            default = default.contiguous()
        setattr(self, name, default)
        self._defaults[name] = deepcopy(default)
        self._persistent[name] = persistent
        self._reductions[name] = reduce_fn

    # ------------------------------------------------------------------ forward
    @torch.jit.unused
    def forward(self, *args: Any, **kwargs: Any) -> Any:
        """Accumulate the batch into the global state and return the metric of the batch alone."""
        if self._is_synced:
            raise TorchMetricsUserError(
                "The Metric shouldn't be synced when performing ``forward``. HINT: Did you forget to call ``unsync`` ?."
            )
        if self.full_state_update or self.full_state_update is None or self.dist_sync_on_step:
            self._forward_cache = self._forward_full_state_update(*args, **kwargs)
        else:
            self._forward_cache = self._forward_reduce_state_update(*args, **kwargs)
        return self._forward_cache

    @contextmanager
    def _batch_scope(self) -> Generator:
        """Temporarily configure the metric for a batch-only compute inside forward."""
        on_cpu = self.compute_on_cpu
        self._to_sync = self.dist_sync_on_step
        self._should_unsync = False
        self.compute_on_cpu = False
        self._enable_grad = True
        try:
            yield
        finally:
            self._is_synced = False
            self._should_unsync = True
            self._to_sync = self.sync_on_compute
            self._computed = None
            self._enable_grad = False
            self.compute_on_cpu = on_cpu
            if on_cpu:
                self._move_list_states_to_cpu()

    def _forward_full_state_update(self, *args: Any, **kwargs: Any) -> Any:
        """Two `update` calls: one into the global state, one into a fresh state for the batch value."""
        self.update(*args, **kwargs)
        count = self._update_count
        with self._batch_scope():
            saved = {name: getattr(self, name) for name in self._defaults}
            self.reset()
            self.update(*args, **kwargs)
            batch_val = self.compute()
            for name, val in saved.items():
                setattr(self, name, val)
            self._update_count = count
        return batch_val

    def _forward_reduce_state_update(self, *args: Any, **kwargs: Any) -> Any:
        """One `update` into a fresh state, then merge that state into the saved global one."""
        global_state = {name: getattr(self, name) for name in self._defaults}
        count = self._update_count
        self.reset()
        with self._batch_scope():
            self.update(*args, **kwargs)
            batch_val = self.compute()
            self._update_count = count + 1
            with torch.no_grad():
                self._reduce_states(global_state)
        return batch_val

    def _reduce_states(self, incoming_state: Dict[str, Any]) -> None:
        """self.state <- merge(incoming_state, self.state) according to each state's reduction."""
        for name in self._defaults:
            local, incoming = getattr(self, name), incoming_state[name]
            fn = self._reductions[name]
            if fn is dim_zero_sum:
                merged = incoming + local
            elif fn is dim_zero_mean:
                merged = ((self._update_count - 1) * incoming + local).float() / self._update_count
            elif fn is dim_zero_max:
                merged = torch.max(incoming, local)
            elif fn is dim_zero_min:
                merged = torch.min(incoming, local)
            elif fn is dim_zero_cat:
                merged = torch.cat([incoming, local]) if isinstance(incoming, Tensor) else incoming + local
            elif fn is None and isinstance(incoming, Tensor):
                merged = torch.stack([incoming, local])
            elif fn is None and isinstance(incoming, list):
                merged = _flatten([incoming, local])
            elif callable(fn):
                merged = fn(torch.stack([incoming, local]))
            else:
                raise TypeError(f"Unsupported reduce_fn: {fn}")
            setattr(self, name, merged)

    # ------------------------------------------------------------------ update / compute wrappers
    def _wrap_update(self, update: Callable) -> Callable:
        @functools.wraps(update)
        def wrapped_update(*args: Any, **kwargs: Any) -> None:
            self._computed = None
            self._update_count += 1
            with torch.set_grad_enabled(self._enable_grad):
                try:
                    update(*args, **kwargs)
                except RuntimeError as err:
                    if "Expected all tensors to be on" in str(err):
                        cls = self.__class__.__name__
                        raise RuntimeError(
                            "Encountered different devices in metric calculation (see stacktrace for details)."
                            " This could be due to the metric class not being on the same device as input."
                            f" Instead of `metric={cls}(...)` try to do `metric={cls}(...).to(device)` where"
                            " device corresponds to the device of the input."
                        ) from err
                    raise
            if self.compute_on_cpu:
                self._move_list_states_to_cpu()

        return wrapped_update

    def _wrap_compute(self, compute: Callable) -> Callable:
        @functools.wraps(compute)
        def wrapped_compute(*args: Any, **kwargs: Any) -> Any:
            if not self.update_called:
                rank_zero_warn(
                    f"The ``compute`` method of metric {self.__class__.__name__} was called before the ``update`` "
                    "method which may lead to errors, as metric states have not yet been updated.",
                    UserWarning,
                )
            if self._computed is not None:
                return self._computed
            if self.__dict__.get("_pending_validation") is not None:
                # validate_args="deferred": the last update's verdict is still outstanding
                from torchmetrics_b200.functional.classification._checks import settle_deferred

                settle_deferred(self)
            with self.sync_context(dist_sync_fn=self.dist_sync_fn, should_sync=self._to_sync,
                                   should_unsync=self._should_unsync):
                value = _squeeze_if_scalar(compute(*args, **kwargs))
            if self.compute_with_cache:
                self._computed = value
            return value

        return wrapped_compute

    def _move_list_states_to_cpu(self) -> None:
        for name in self._defaults:
            val = getattr(self, name)
            if isinstance(val, Sequence):
                setattr(self, name, [v.to("cpu") for v in val])

    @abstractmethod
    def update(self, *_: Any, **__: Any) -> None:
        """Fold one batch into the state."""

    @abstractmethod
    def compute(self) -> Any:
        """Metric value from the (synchronised) state."""

    def plot(self, *_: Any, **__: Any) -> Any:
        raise NotImplementedError("plotting is outside the scope of torchmetrics_b200")

    # ------------------------------------------------------------------ cross-rank synchronisation
    def _sync_dist(self, dist_sync_fn: Callable = gather_all_tensors, process_group: Optional[Any] = None) -> None:
        """Reference protocol (metric.py:427-457): one `dist_sync_fn` call per state tensor, stack, reduce."""
        group = process_group or self.process_group
        gathered: Dict[str, Any] = {}
        for name, fn in self._reductions.items():
            val = getattr(self, name)
            if fn is dim_zero_cat and isinstance(val, list) and len(val) > 1:
                val = [dim_zero_cat(val)]  # fewer gathers
            gathered[name] = apply_to_collection(val, Tensor, dist_sync_fn, group=group)
        for name, fn in self._reductions.items():
            out = gathered[name]
            if isinstance(out, list) and len(out) == 0:
                setattr(self, name, [])
                continue
            if isinstance(out[0], Tensor):
                out = torch.stack(out)
            elif isinstance(out[0], list):
                out = _flatten(out)
            if not (callable(fn) or fn is None):
                raise TypeError("reduction_fn must be callable or None")
            setattr(self, name, fn(out) if fn is not None else out)

    def sync(self, dist_sync_fn: Optional[Callable] = None, process_group: Optional[Any] = None,
             should_sync: bool = True, distributed_available: Optional[Callable] = None) -> None:
        """Replace the local state by the state merged over all ranks (the local one is cached for `unsync`)."""
        if self._is_synced and should_sync:
            raise TorchMetricsUserError("The Metric has already been synced.")
        if distributed_available is None and self.distributed_available_fn is not None:
            distributed_available = self.distributed_available_fn
        is_distributed = distributed_available() if callable(distributed_available) else None
        if not should_sync or not is_distributed:
            return
        self._cache = {name: getattr(self, name) for name in self._defaults}
        if dist_sync_fn is None:
            fused_state_sync([self], process_group or self.process_group)
        else:
            self._sync_dist(dist_sync_fn, process_group=process_group)
        self._is_synced = True

    def unsync(self, should_unsync: bool = True) -> None:
        """Put the cached local state back so accumulation can continue."""
        if not should_unsync:
            return
        if not self._is_synced:
            raise TorchMetricsUserError("The Metric has already been un-synced.")
        if self._cache is None:
            raise TorchMetricsUserError("The internal cache should exist to unsync the Metric.")
        for name, val in self._cache.items():
            setattr(self, name, val)
        self._is_synced = False
        self._cache = None

    @contextmanager
    def sync_context(self, dist_sync_fn: Optional[Callable] = None, process_group: Optional[Any] = None,
                     should_sync: bool = True, should_unsync: bool = True,
                     distributed_available: Optional[Callable] = None) -> Generator:
        self.sync(dist_sync_fn=dist_sync_fn, process_group=process_group, should_sync=should_sync,
                  distributed_available=distributed_available)
        yield
        self.unsync(should_unsync=self._is_synced and should_unsync)

    # ------------------------------------------------------------------ reset / copy / pickle
    def reset(self) -> None:
        if self.__dict__.get("_pending_validation") is not None:
            from torchmetrics_b200.functional.classification._checks import settle_deferred

            settle_deferred(self)
        self._update_count = 0
        self._forward_cache = None
        self._computed = None
        for name, default in self._defaults.items():
            if isinstance(default, Tensor):
                setattr(self, name, default.detach().clone().to(getattr(self, name).device))
            else:
                setattr(self, name, [])
        self._cache = None
        self._is_synced = False

    def clone(self) -> "Metric":
        return deepcopy(self)

    def __getstate__(self) -> Dict[str, Any]:
        # the wrapped bound methods are closures over `self`; rebuild them on load
        skip = ("update", "compute", "_update_signature", "_pending_validation", "_validation_slots")
        return {k: v for k, v in self.__dict__.items() if k not in skip}

    def __setstate__(self, state: Dict[str, Any]) -> None:
        self.__dict__.update(state)
        self._install_wrappers()

    def __setattr__(self, name: str, value: Any) -> None:
        if name in _CONSTANTS:
            raise RuntimeError(f"Can't change const `{name}`.")
        # Metric states and bookkeeping flags are plain attributes that are rebound several times per update /
        # forward; nn.Module.__setattr__ walks its parameter / buffer / module tables on every one of them.  A name
        # that already lives directly in __dict__ is none of those, so a non-Parameter, non-Module value can go
        # straight in.
        if name in self.__dict__:
            # exact-type test first: `isinstance(x, Parameter)` goes through a Python-level __instancecheck__
            t = type(value)
            if t is Tensor or value is None or t in _PLAIN_TYPES or not isinstance(value, (torch.nn.Parameter,
                                                                                            torch.nn.Module)):
                object.__setattr__(self, name, value)
                return
        super().__setattr__(name, value)

    # ------------------------------------------------------------------ device / dtype
    def type(self, dst_type: Union[str, torch.dtype]) -> "Metric":  # noqa: A003
        return self  # dtype moves only through set_dtype

    def float(self) -> "Metric":
        return self

    def double(self) -> "Metric":
        return self

    def half(self) -> "Metric":
        return self

    def set_dtype(self, dst_type: Union[str, torch.dtype]) -> "Metric":
        self._dtype_convert = True
        out = super().type(dst_type)
        out._dtype_convert = False
        return out

    def _apply(self, fn: Callable, exclude_state: Sequence[str] = "") -> Module:
        """Moves states and their defaults together with parameters / buffers (metric.py:782-832)."""
        this = super()._apply(fn)
        is_cast = any(tag in str(fn) for tag in ("Module.type", "Module.half", "Module.float", "Module.double",
                                                 "Module.bfloat16"))
        if is_cast and not self._dtype_convert:
            return this
        for name, default in this._defaults.items():
            if name in exclude_state:
                continue
            if isinstance(default, Tensor):
                this._defaults[name] = fn(default)
            elif isinstance(default, Sequence):
                this._defaults[name] = [fn(v) for v in default]
            current = getattr(this, name)
            if isinstance(current, Tensor):
                setattr(this, name, fn(current))
            elif isinstance(current, Sequence):
                setattr(this, name, [fn(v) for v in current])
            else:
                raise TypeError(
                    f"Expected metric state to be either a Tensor or a list of Tensor, but encountered {current}"
                )
        probe = fn(torch.zeros(1, device=self.device))
        self._device = probe.device
        self._dtype = probe.dtype
        if this._computed is not None:
            this._computed = apply_to_collection(this._computed, Tensor, fn)
        if this._forward_cache is not None:
            this._forward_cache = apply_to_collection(this._forward_cache, Tensor, fn)
        return this

    # ------------------------------------------------------------------ checkpointing
    def persistent(self, mode: bool = False) -> None:
        for name in self._persistent:
            self._persistent[name] = mode

    def state_dict(self, destination: Optional[Dict[str, Any]] = None, prefix: str = "",  # type: ignore[override]
                   keep_vars: bool = False) -> Dict[str, Any]:
        destination = super().state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars)
        for name in self._defaults:
            if not self._persistent[name]:
                continue
            val = getattr(self, name)
            if not keep_vars:
                if isinstance(val, Tensor):
                    val = val.detach()
                elif isinstance(val, list):
                    val = [v.detach() if isinstance(v, Tensor) else v for v in val]
            destination[prefix + name] = deepcopy(val)
        return destination

    def _load_from_state_dict(self, state_dict: dict, prefix: str, local_metadata: dict, strict: bool,
                              missing_keys: List[str], unexpected_keys: List[str], error_msgs: List[str]) -> None:
        for name in self._defaults:
            key = prefix + name
            if key in state_dict:
                setattr(self, name, state_dict.pop(key))
        super()._load_from_state_dict(state_dict, prefix, local_metadata, True, missing_keys, unexpected_keys,
                                      error_msgs)

    def _filter_kwargs(self, **kwargs: Any) -> Dict[str, Any]:
        """Keep only the kwargs `update` can take (all of them if it has **kwargs)."""
        params = self._update_signature.parameters
        var_kinds = (inspect.Parameter.VAR_POSITIONAL, inspect.Parameter.VAR_KEYWORD)
        if any(p.kind == inspect.Parameter.VAR_KEYWORD for p in params.values()):
            return kwargs
        return {k: v for k, v in kwargs.items() if k in params and params[k].kind not in var_kinds}

    def __hash__(self) -> int:
        parts: List[Any] = [self.__class__.__name__, id(self)]
        for name in self._defaults:
            val = getattr(self, name)
            if isinstance(val, Sequence) and not isinstance(val, Tensor):
                parts.extend(id(v) for v in val)
            else:
                parts.append(id(val))
        return hash(tuple(parts))

    # `__getitem__` below is metric algebra, not a sequence protocol: without this, `for x in metric` would loop
    # forever over CompositionalMetrics.  None makes Python raise "'X' object is not iterable" (metric.py:1081).
    __iter__ = None

    # ------------------------------------------------------------------ operator algebra -> CompositionalMetric
    def __add__(self, other): return CompositionalMetric(torch.add, self, other)
    def __radd__(self, other): return CompositionalMetric(torch.add, other, self)
    def __sub__(self, other): return CompositionalMetric(torch.sub, self, other)
    def __rsub__(self, other): return CompositionalMetric(torch.sub, other, self)
    def __mul__(self, other): return CompositionalMetric(torch.mul, self, other)
    def __rmul__(self, other): return CompositionalMetric(torch.mul, other, self)
    def __truediv__(self, other): return CompositionalMetric(torch.true_divide, self, other)
    def __rtruediv__(self, other): return CompositionalMetric(torch.true_divide, other, self)
    def __floordiv__(self, other): return CompositionalMetric(torch.floor_divide, self, other)
    def __rfloordiv__(self, other): return CompositionalMetric(torch.floor_divide, other, self)
    def __mod__(self, other): return CompositionalMetric(torch.fmod, self, other)
    def __rmod__(self, other): return CompositionalMetric(torch.fmod, other, self)
    def __pow__(self, other): return CompositionalMetric(torch.pow, self, other)
    def __rpow__(self, other): return CompositionalMetric(torch.pow, other, self)
    def __matmul__(self, other): return CompositionalMetric(torch.matmul, self, other)
    def __rmatmul__(self, other): return CompositionalMetric(torch.matmul, other, self)
    def __and__(self, other): return CompositionalMetric(torch.bitwise_and, self, other)
    def __rand__(self, other): return CompositionalMetric(torch.bitwise_and, other, self)
    def __or__(self, other): return CompositionalMetric(torch.bitwise_or, self, other)
    def __ror__(self, other): return CompositionalMetric(torch.bitwise_or, other, self)
    def __xor__(self, other): return CompositionalMetric(torch.bitwise_xor, self, other)
    def __rxor__(self, other): return CompositionalMetric(torch.bitwise_xor, other, self)
    def __eq__(self, other): return CompositionalMetric(torch.eq, self, other)  # type: ignore[override]
    def __ne__(self, other): return CompositionalMetric(torch.ne, self, other)  # type: ignore[override]
    def __ge__(self, other): return CompositionalMetric(torch.ge, self, other)
    def __gt__(self, other): return CompositionalMetric(torch.gt, self, other)
    def __le__(self, other): return CompositionalMetric(torch.le, self, other)
    def __lt__(self, other): return CompositionalMetric(torch.lt, self, other)
    def __abs__(self): return CompositionalMetric(torch.abs, self, None)
    def __neg__(self): return CompositionalMetric(_neg, self, None)
    def __pos__(self): return CompositionalMetric(torch.abs, self, None)
    def __invert__(self): return CompositionalMetric(torch.bitwise_not, self, None)
    def __inv__(self): return CompositionalMetric(torch.bitwise_not, self, None)
    def __getitem__(self, idx): return CompositionalMetric(lambda x: x[idx], self, None)

    def __getnewargs__(self) -> Tuple:
        return tuple(Metric.__str__(self))


_PLAIN_TYPES = frozenset((bool, int, float, str, list, dict, tuple))


def _neg(x: Tensor) -> Tensor:
    return -torch.abs(x)


def fused_state_sync(metrics: Sequence[Metric], process_group: Optional[Any] = None) -> None:
    """Default sync of one or several metrics with a minimal number of collectives.

    All sum / mean / max / min tensor states of all `metrics` share one packed all_reduce per (op, dtype); all cat
    states share one descriptor exchange and take one padded all_gather each.  States with a custom callable or a
    `None` reduction keep the reference's gather + stack path.  Results are assigned to the state attributes; the
    caller is responsible for `_cache` / `_is_synced` (see `Metric.sync`, `MetricCollection.compute`).
    """
    reducible: List[Tuple[str, Tensor]] = []
    reducible_slots: List[Tuple[Metric, str]] = []
    cats: List[Optional[Tensor]] = []
    cat_slots: List[Tuple[Metric, str, bool]] = []
    legacy: List[Tuple[Metric, str]] = []
    for m in metrics:
        for name, fn in m._reductions.items():
            val = getattr(m, name)
            if fn in _FUSABLE and isinstance(val, Tensor) and not (fn is dim_zero_mean and not val.is_floating_point()):
                reducible.append((_FUSABLE[fn], val))
                reducible_slots.append((m, name))
            elif fn is dim_zero_cat and isinstance(val, list):
                cats.append(dim_zero_cat(val) if len(val) else None)
                cat_slots.append((m, name, True))
            else:
                # custom / None reductions, and the odd tensor state with "cat" (stays stacked (W, ...) as in the
                # reference): per-state gather
                legacy.append((m, name))
    if reducible:
        for (m, name), t in zip(reducible_slots, fused_reduce(reducible, process_group)):
            setattr(m, name, t)
    if cats:
        for (m, name, was_list), t in zip(cat_slots, fused_cat_gather(cats, process_group)):
            if t is None:
                setattr(m, name, [])
            else:
                setattr(m, name, t)
    for m, name in legacy:
        fn = m._reductions[name]
        val = getattr(m, name)
        out = apply_to_collection(val, Tensor, gather_all_tensors, group=process_group)
        if isinstance(out, list) and len(out) == 0:
            setattr(m, name, [])
            continue
        if isinstance(out[0], Tensor):
            out = torch.stack(out)
        elif isinstance(out[0], list):
            out = _flatten(out)
        if not (callable(fn) or fn is None):
            raise TypeError("reduction_fn must be callable or None")
        setattr(m, name, fn(out) if fn is not None else out)


class CompositionalMetric(Metric):
    """Lazy `op(metric_a, metric_b)`; update / reset fan out to the operands (metric.py:1088-1211)."""

    def __init__(self, operator: Callable, metric_a: Union[Metric, float, Tensor, None],
                 metric_b: Union[Metric, float, Tensor, None]) -> None:
        super().__init__()
        self.op = operator
        if isinstance(metric_a, Tensor):
            self.register_buffer("metric_a", metric_a, persistent=False)
        else:
            self.metric_a = metric_a
        if isinstance(metric_b, Tensor):
            self.register_buffer("metric_b", metric_b, persistent=False)
        else:
            self.metric_b = metric_b

    def _sync_dist(self, dist_sync_fn: Optional[Callable] = None, process_group: Optional[Any] = None) -> None:
        pass  # operands sync themselves

    def sync(self, *args: Any, **kwargs: Any) -> None:
        pass

    def unsync(self, should_unsync: bool = True) -> None:
        pass

    def update(self, *args: Any, **kwargs: Any) -> None:
        for m in (self.metric_a, self.metric_b):
            if isinstance(m, Metric):
                m.update(*args, **m._filter_kwargs(**kwargs))

    def compute(self) -> Any:
        a = self.metric_a.compute() if isinstance(self.metric_a, Metric) else self.metric_a
        b = self.metric_b.compute() if isinstance(self.metric_b, Metric) else self.metric_b
        if b is None:
            return self.op(a)
        return self.op(a, b)

    @torch.jit.unused
    def forward(self, *args: Any, **kwargs: Any) -> Any:
        a = (self.metric_a(*args, **self.metric_a._filter_kwargs(**kwargs))
             if isinstance(self.metric_a, Metric) else self.metric_a)
        b = (self.metric_b(*args, **self.metric_b._filter_kwargs(**kwargs))
             if isinstance(self.metric_b, Metric) else self.metric_b)
        if a is None:
            self._forward_cache = None
        elif b is None:
            self._forward_cache = None if isinstance(self.metric_b, Metric) else self.op(a)
        else:
            self._forward_cache = self.op(a, b)
        return self._forward_cache

    def reset(self) -> None:
        for m in (self.metric_a, self.metric_b):
            if isinstance(m, Metric):
                m.reset()

    def persistent(self, mode: bool = False) -> None:
        for m in (self.metric_a, self.metric_b):
            if isinstance(m, Metric):
                m.persistent(mode=mode)

    def __repr__(self) -> str:
        name = getattr(self.op, "__name__", repr(self.op))
        return f"{self.__class__.__name__}(\n  {name}(\n    {self.metric_a!r},\n    {self.metric_b!r}\n  )\n)"

    def _wrap_compute(self, compute: Callable) -> Callable:
        return compute
