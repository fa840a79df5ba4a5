"""`MetricCollection`: fan-out of update / forward / compute with compute groups.

Behaviour follows the reference (src/torchmetrics/collections.py:34-661): metrics whose states turn out equal after
the first `update` form a compute group (:228-287); members of a group share the state tensors of the first member
by reference (:289-308) so later `update`s touch one metric per group (:200-226); `forward` never uses groups;
accessing a member hands out a copy of the shared state (:515-550).

What differs is `compute()` under a distributed run: the reference lets every member run its own sync (16 gathers =
48 collectives for a 4-metric stat-scores group).  Here the states of ALL group leaders go through
one `fused_state_sync` (one packed all_reduce + one padded all_gather per cat state for the whole collection),
members compute from the synced tensors, then every leader is un-synced.  Metrics with a custom `dist_sync_fn`,
`sync_on_compute=False` or their own process group keep their individual path.
"""
from __future__ import annotations

from collections import OrderedDict
from copy import deepcopy
from typing import Any, Dict, Hashable, Iterable, Iterator, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from torch.nn import ModuleDict

from torchmetrics_b200.metric import Metric, fused_state_sync
from torchmetrics_b200.utilities.data import apply_to_collection  # noqa: F401
from torchmetrics_b200.utilities.prints import rank_zero_warn


def _flatten_dict(x: Dict) -> Tuple[Dict, bool]:
    flat: Dict = {}
    dup = False
    for key, value in x.items():
        if isinstance(value, dict):
            for k, v in value.items():
                dup = dup or k in flat
                flat[k] = v
        else:
            dup = dup or key in flat
            flat[key] = value
    return flat, dup


def _allclose_recursive(a: Any, b: Any) -> bool:
    if isinstance(a, Tensor) and isinstance(b, Tensor):
        return a.shape == b.shape and torch.allclose(a, b)
    if isinstance(a, (list, tuple)) and isinstance(b, (list, tuple)):
        return len(a) == len(b) and all(_allclose_recursive(x, y) for x, y in zip(a, b))
    return a == b


class MetricCollection(ModuleDict):
    """Dictionary of metrics that are updated / computed together."""

    _modules: Dict[str, Metric]  # type: ignore[assignment]
    _groups: Dict[int, List[str]]

    def __init__(self, metrics: Union[Metric, Sequence[Metric], Dict[str, Metric]], *additional_metrics: Metric,
                 prefix: Optional[str] = None, postfix: Optional[str] = None,
                 compute_groups: Union[bool, List[List[str]]] = True) -> None:
        super().__init__()
        self.prefix = self._check_arg(prefix, "prefix")
        self.postfix = self._check_arg(postfix, "postfix")
        self._enable_compute_groups = compute_groups
        self._groups_checked = False
        self._state_is_copy = False
        self.add_metrics(metrics, *additional_metrics)

    # ------------------------------------------------------------------ update / forward / compute
    @torch.jit.unused
    def forward(self, *args: Any, **kwargs: Any) -> Dict[str, Any]:
        return self._compute_and_reduce("forward", *args, **kwargs)

    def update(self, *args: Any, **kwargs: Any) -> None:
        if self._groups_checked:
            for members in self._groups.values():
                leader = getattr(self, members[0])
                leader.update(*args, **leader._filter_kwargs(**kwargs))
            if self._state_is_copy:
                self._compute_groups_create_state_ref()
                self._state_is_copy = False
        else:
            for m in self.values(copy_state=False):
                m.update(*args, **m._filter_kwargs(**kwargs))
            if self._enable_compute_groups:
                self._merge_compute_groups()
                self._compute_groups_create_state_ref()
                self._groups_checked = True

    def _merge_compute_groups(self) -> None:
        """Repeatedly merge two groups whose leaders hold equal states until nothing changes."""
        while True:
            merged = False
            keys = list(self._groups.keys())
            for i, k1 in enumerate(keys):
                for k2 in keys[i + 1:]:
                    a = getattr(self, self._groups[k1][0])
                    b = getattr(self, self._groups[k2][0])
                    if self._equal_metric_states(a, b):
                        self._groups[k1].extend(self._groups.pop(k2))
                        merged = True
                        break
                if merged:
                    break
            if not merged:
                break
        self._groups = {i: v for i, v in enumerate(deepcopy(list(self._groups.values())))}

    @staticmethod
    def _equal_metric_states(a: Metric, b: Metric) -> bool:
        if len(a._defaults) == 0 or len(b._defaults) == 0:
            return False
        if a._defaults.keys() != b._defaults.keys():
            return False
        for key in a._defaults:
            s1, s2 = getattr(a, key), getattr(b, key)
            if type(s1) != type(s2):  # noqa: E721
                return False
            if isinstance(s1, Tensor) and isinstance(s2, Tensor):
                if s1.shape != s2.shape or not torch.allclose(s1, s2):
                    return False
            elif isinstance(s1, list) and isinstance(s2, list):
                if not _allclose_recursive(s1, s2):
                    return False
        return True

    def _compute_groups_create_state_ref(self, copy: bool = False) -> None:
        """Members of a group get the leader's state objects (or deep copies of them when `copy`)."""
        if not self._state_is_copy:
            for members in self._groups.values():
                leader = getattr(self, members[0])
                # metrics whose state tensors are shared must not accumulate into them in place on their own
                # (classification/stat_scores.py `_forward_reduce_state_update`)
                leader.__dict__["_state_shared"] = not copy and len(members) > 1
                for name in members[1:]:
                    m = getattr(self, name)
                    for state in leader._defaults:
                        val = getattr(leader, state)
                        setattr(m, state, deepcopy(val) if copy else val)
                    m._update_count = leader._update_count
                    m._computed = deepcopy(leader._computed) if copy else leader._computed
                    m.__dict__["_state_shared"] = not copy
        self._state_is_copy = copy

    def compute(self) -> Dict[str, Any]:
        return self._compute_and_reduce("compute")

    def _fusable_leaders(self) -> List[Tuple[Metric, List[Metric]]]:
        """(leader, followers) pairs whose sync can be merged into one fused exchange."""
        out = []
        if not self._groups_checked or self._state_is_copy:
            groups: Iterable[List[str]] = [[k] for k in self.keys(keep_base=True)]
        else:
            groups = self._groups.values()
        for members in groups:
            leader = getattr(self, members[0])
            if not isinstance(leader, Metric) or isinstance(leader, _NO_FUSE):
                continue
            if (leader.dist_sync_fn is not None or not leader._to_sync or leader._is_synced
                    or leader.process_group is not None or leader._computed is not None
                    or not leader._defaults or not leader.distributed_available_fn()):
                continue
            out.append((leader, [getattr(self, n) for n in members[1:]]))
        return out

    def _compute_and_reduce(self, method_name: str, *args: Any, **kwargs: Any) -> Dict[str, Any]:
        if method_name not in ("compute", "forward"):
            raise ValueError(f"method_name should be either 'compute' or 'forward', but got {method_name}")
        result: Dict[str, Any] = {}
        fused: List[Tuple[Metric, List[Metric]]] = []
        if method_name == "compute":
            fused = self._fusable_leaders()
        if fused:
            leaders = [ld for ld, _ in fused]
            for ld in leaders:
                ld._cache = {name: getattr(ld, name) for name in ld._defaults}
            fused_state_sync(leaders)
            saved_flags = []
            for ld, followers in fused:
                ld._is_synced = True
                for m in [ld, *followers]:
                    saved_flags.append((m, m._to_sync, m._should_unsync))
                    m._to_sync = False       # states are already merged
                    m._should_unsync = False
                for m in followers:          # followers read the leader's synced tensors
                    for state in ld._defaults:
                        setattr(m, state, getattr(ld, state))
        # link the group members to their leader's state once; the loops below walk the module table directly
        self._compute_groups_create_state_ref(False)
        members = list(self._modules.items())
        if method_name == "forward":
            result.update(self._forward_groups(args, kwargs))
        try:
            for k, m in members:
                if k in result:
                    continue
                result[k] = m.compute() if method_name == "compute" else m(*args, **m._filter_kwargs(**kwargs))
        finally:
            if fused:
                for m, to_sync, should_unsync in saved_flags:
                    m._to_sync, m._should_unsync = to_sync, should_unsync
                for ld, followers in fused:
                    ld.unsync()
                    for m in followers:
                        for state in ld._defaults:
                            setattr(m, state, getattr(ld, state))

        if method_name == "forward" and self._enable_compute_groups and not self._groups_checked:
            # the first forward leaves every metric with the state of one batch, exactly what `update` inspects to form
            # the compute groups (collections.py:207-215 in the reference, which only does it there): from the second
            # forward on a group is served by one update (`_forward_groups`)
            self._merge_compute_groups()
            self._compute_groups_create_state_ref()
            self._groups_checked = True

        _, duplicates = _flatten_dict(result)
        flat: Dict[str, Any] = {}
        for k, m in members:
            res = result[k]
            if not isinstance(res, dict):
                flat[k] = res
                continue
            for key, v in res.items():
                if duplicates:
                    stripped = k.replace(getattr(m, "prefix", "") or "", "").replace(getattr(m, "postfix", "") or "", "")
                    key = f"{stripped}_{key}"
                if getattr(m, "_from_collection", None) and m.prefix is not None:
                    key = f"{m.prefix}{key}"
                if getattr(m, "_from_collection", None) and m.postfix is not None:
                    key = f"{key}{m.postfix}"
                flat[key] = v
        return {self._set_name(k): v for k, v in flat.items()}

    def _forward_groups(self, args: Any, kwargs: Any) -> Dict[str, Any]:
        """`forward` of whole compute groups with one update per group, where the leader knows how
        (`_forward_batch_values`: the stat-score family on packed states).  The reference forwards every member on its
        own (collections.py:194-231 via `_compute_and_reduce`); the values are the same, the members keep sharing the
        leader's state afterwards.  Returns {module key: batch value} for the members it served."""
        served: Dict[str, Any] = {}
        if not (self._enable_compute_groups and self._groups_checked) or self._state_is_copy:
            return served
        for names in self._groups.values():
            if len(names) < 2:
                continue
            group = [self._modules[n] for n in names]
            fast = getattr(group[0], "_forward_batch_values", None)
            if fast is None or any(getattr(m, "_forward_batch_values", None) is None for m in group):
                continue
            kw = group[0]._filter_kwargs(**kwargs)
            if any(m._filter_kwargs(**kwargs) != kw for m in group[1:]):
                continue
            values = fast(group, args, kw)
            if values is None:
                continue
            for n, m, v in zip(names, group, values):
                m._forward_cache = v
                served[n] = v
        return served

    def reset(self) -> None:
        for m in self.values(copy_state=False):
            m.reset()
        if self._enable_compute_groups and self._groups_checked:
            self._compute_groups_create_state_ref()

    def clone(self, prefix: Optional[str] = None, postfix: Optional[str] = None) -> "MetricCollection":
        mc = deepcopy(self)
        if prefix:
            mc.prefix = self._check_arg(prefix, "prefix")
        if postfix:
            mc.postfix = self._check_arg(postfix, "postfix")
        return mc

    def persistent(self, mode: bool = True) -> None:
        for m in self.values(copy_state=False):
            m.persistent(mode)

    # ------------------------------------------------------------------ construction
    def add_metrics(self, metrics: Union[Metric, Sequence[Metric], Dict[str, Metric]],
                    *additional_metrics: Metric) -> None:
        """Register metrics under their names (reference behaviour: collections.py:408-462).

        The three accepted shapes - one metric, a sequence (plus positional extras), a dict - are first flattened
        into one list of `(name, metric)` entries by `_named_entries`; nested collections contribute their members
        with the nested prefix / postfix carried along.  The entries are then inserted in one loop.
        """
        for name, metric in self._named_entries(metrics, additional_metrics):
            self[name] = metric
        self._groups_checked = False
        if self._enable_compute_groups:
            self._init_compute_groups()
        else:
            self._groups = {}

    def _named_entries(self, metrics: Any, extras: Sequence[Any]) -> List[Tuple[str, Metric]]:
        by_key = isinstance(metrics, dict)
        if by_key:
            if extras:
                raise ValueError(
                    f"You have passes extra arguments {extras} which are not compatible"
                    f" with first passed dictionary {metrics} so they will be ignored."
                )
            items = [(key, metrics[key]) for key in sorted(metrics)]
        elif isinstance(metrics, Metric) or (isinstance(metrics, Sequence) and not isinstance(metrics, str)):
            head = [metrics] if isinstance(metrics, Metric) else list(metrics)
            ignored = [m for m in extras if not isinstance(m, Metric)]
            if ignored:
                rank_zero_warn(
                    f"You have passes extra arguments {ignored} which are not `Metric` so they will be ignored."
                )
            items = [(None, m) for m in head + [m for m in extras if isinstance(m, Metric)]]
        else:
            raise ValueError(
                "Unknown input to MetricCollection. Expected, `Metric`, `MetricCollection` or `dict`/`sequence` of the"
                f" previous, but got {metrics}"
            )

        entries: List[Tuple[str, Metric]] = []
        taken = set(self.keys(keep_base=True))
        for key, obj in items:
            if isinstance(obj, MetricCollection):
                for inner, member in obj.items(keep_base=False):
                    member.prefix, member.postfix = obj.prefix, obj.postfix
                    member._from_collection = True
                    entries.append((inner if key is None else f"{key}_{inner}", member))
            elif isinstance(obj, Metric):
                name = key if key is not None else obj.__class__.__name__
                if key is None and (name in taken or any(name == n for n, _ in entries)):
                    raise ValueError(f"Encountered two metrics both named {name}")
                entries.append((name, obj))
            elif by_key:
                raise ValueError(
                    f"Value {obj} belonging to key {key} is not an instance of"
                    " `torchmetrics.Metric` or `torchmetrics.MetricCollection`"
                )
            else:
                raise ValueError(
                    f"Input {obj} to `MetricCollection` is not a instance of"
                    " `torchmetrics.Metric` or `torchmetrics.MetricCollection`"
                )
        return entries

    def _init_compute_groups(self) -> None:
        if isinstance(self._enable_compute_groups, list):
            self._groups = dict(enumerate(self._enable_compute_groups))
            for members in self._groups.values():
                for name in members:
                    if name not in self:
                        raise ValueError(
                            f"Input {name} in `compute_groups` argument does not match a metric in the collection."
                            f" Please make sure that {self._enable_compute_groups} matches {self.keys(keep_base=True)}"
                        )
            self._groups_checked = True
        else:
            self._groups = {i: [str(k)] for i, k in enumerate(self.keys(keep_base=True))}

    @property
    def compute_groups(self) -> Dict[int, List[str]]:
        return self._groups

    # ------------------------------------------------------------------ naming / access
    def _set_name(self, base: str) -> str:
        name = base if self.prefix is None else self.prefix + base
        return name if self.postfix is None else name + self.postfix

    def _to_renamed_ordered_dict(self) -> Dict[str, Metric]:
        # same mapping type as `_modules`, so keys()/items() views have one type whichever way they are asked for
        return type(self._modules)((self._set_name(k), v) for k, v in self._modules.items())

    def __iter__(self) -> Iterator[Hashable]:
        return iter(self.keys())

    def keys(self, keep_base: bool = False) -> Iterable[Hashable]:  # type: ignore[override]
        if keep_base:
            return self._modules.keys()
        return self._to_renamed_ordered_dict().keys()

    def items(self, keep_base: bool = False, copy_state: bool = True) -> Iterable[Tuple[str, Metric]]:  # type: ignore[override]
        self._compute_groups_create_state_ref(copy_state)
        if keep_base:
            return self._modules.items()
        return self._to_renamed_ordered_dict().items()

    def values(self, copy_state: bool = True) -> Iterable[Metric]:  # type: ignore[override]
        self._compute_groups_create_state_ref(copy_state)
        return self._modules.values()

    def __getitem__(self, key: str, copy_state: bool = True) -> Metric:  # type: ignore[override]
        self._compute_groups_create_state_ref(copy_state)
        if self.prefix:
            key = key.removeprefix(self.prefix)
        if self.postfix:
            key = key.removesuffix(self.postfix)
        return self._modules[key]

    @staticmethod
    def _check_arg(arg: Optional[str], name: str) -> Optional[str]:
        if arg is None or isinstance(arg, str):
            return arg
        raise ValueError(f"Expected input `{name}` to be a string, but got {type(arg)}")

    def __repr__(self) -> str:
        text = super().__repr__()[:-2]
        if self.prefix:
            text += f",\n  prefix={self.prefix}{',' if self.postfix else ''}"
        if self.postfix:
            text += f"{',' if not self.prefix else ''}\n  postfix={self.postfix}"
        return text + "\n)"

    def set_dtype(self, dst_type: Union[str, torch.dtype]) -> "MetricCollection":
        for m in self.values(copy_state=False):
            m.set_dtype(dst_type)
        return self


from torchmetrics_b200.metric import CompositionalMetric  # noqa: E402

_NO_FUSE = (CompositionalMetric,)
