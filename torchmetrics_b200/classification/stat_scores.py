"""Stateful tp/fp/tn/fn metrics (reference: classification/stat_scores.py:43-551)."""
from importlib import import_module
from typing import Any, Callable, List, Optional, Sequence, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
F = import_module("torchmetrics_b200.functional.classification.stat_scores")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.data import dim_zero_cat
from torchmetrics_b200.utilities.enums import ClassificationTask


class _AbstractStatScores(Metric):
    """States `tp, fp, tn, fn`: int64 tensors summed across ranks, or lists concatenated for samplewise averaging."""

    tp: Union[List[Tensor], Tensor]
    fp: Union[List[Tensor], Tensor]
    tn: Union[List[Tensor], Tensor]
    fn: Union[List[Tensor], Tensor]

    def _create_state(self, size: int, multidim_average: Literal["global", "samplewise"] = "global") -> None:
        default: Union[Callable[[], list], Callable[[], Tensor]]
        if multidim_average == "samplewise":
            default, reduce_fx = list, "cat"
        else:
            default, reduce_fx = (lambda: torch.zeros(size, dtype=torch.long)), "sum"
        for name in ("tp", "fp", "tn", "fn"):
            self.add_state(name, default(), dist_reduce_fx=reduce_fx)

    def _packed_states(self) -> Optional[Tensor]:
        """The four global count states as the rows of ONE (4, n) int64 tensor, so that the update kernel accumulates
        into them in place (the reference adds four freshly computed vectors to the four states, stat_scores.py:43-55:
        a zero-fill and four small additions per update).  States that are not consecutive in memory yet - after
        `reset`, `to`, a state sync - are re-pointed at a common buffer with `Tensor.set_`: the tensor OBJECTS stay,
        so compute-group members that alias them follow.  None when the states are not CUDA int64 tensors."""
        if self.multidim_average != "global":
            return None
        tp, fp, tn, fn = self.tp, self.fp, self.tn, self.fn
        if not (isinstance(tp, Tensor) and tp.is_cuda and tp.dtype == torch.int64 and tp.numel() > 0):
            return None
        n = tp.numel()
        for t in (tp, fp, tn, fn):
            if not (isinstance(t, Tensor) and t.dtype == torch.int64 and t.numel() == n and t.is_contiguous()
                    and t.device == tp.device):
                return None
        ptr, step = tp.data_ptr(), 8 * n
        if not (fp.data_ptr() == ptr + step and tn.data_ptr() == ptr + 2 * step and fn.data_ptr() == ptr + 3 * step
                and fn.untyped_storage().data_ptr() == tp.untyped_storage().data_ptr()):
            arena = torch.stack([tp.reshape(-1), fp.reshape(-1), tn.reshape(-1), fn.reshape(-1)])
            for i, t in enumerate((tp, fp, tn, fn)):
                t.set_(arena.untyped_storage(), i * n, t.shape, t.stride())
        return tp.as_strided((4, n), (n, 1))

    def _update_state(self, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, packed: Optional[Tensor] = None) -> None:
        if packed is not None and isinstance(tp, Tensor):
            if tp.data_ptr() == packed.data_ptr():
                return  # the kernel has accumulated into the states themselves
            n, ptr = packed.shape[1], tp.data_ptr()
            if (tp.dtype == torch.int64 and tp.numel() == n and fp.data_ptr() == ptr + 8 * n
                    and tn.data_ptr() == ptr + 16 * n and fn.data_ptr() == ptr + 24 * n
                    and fn.untyped_storage().data_ptr() == tp.untyped_storage().data_ptr()):
                packed.add_(tp.as_strided((4, n), (n, 1)))  # the rows of one (4, n) result: one addition, not four
                return
        if self.multidim_average == "samplewise":
            self.tp.append(tp)
            self.fp.append(fp)
            self.tn.append(tn)
            self.fn.append(fn)
        else:
            # in place, so compute-group members that alias these tensors see the update
            self.tp += tp
            self.fp += fp
            self.tn += tn
            self.fn += fn

    def _final_state(self) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
        return dim_zero_cat(self.tp), dim_zero_cat(self.fp), dim_zero_cat(self.tn), dim_zero_cat(self.fn)

    # ------------------------------------------------------------------ forward() on packed states
    def _forward_batch_values(self, group: Sequence["_AbstractStatScores"], args: Any, kwargs: Any) -> Optional[list]:
        """`forward` of every metric in `group` - this metric first, the others alias its state tensors - with ONE update.

        The generic route (metric.py `_forward_reduce_state_update`; reference metric.py:344-381) resets to fresh states
        (four clones), updates them, computes, and adds them to the saved running states (four additions) - per metric.
        Here the batch goes straight into the running states (they are one buffer the kernel accumulates into) and its
        own counts are the difference to a snapshot of that buffer: one clone and one subtraction, whatever the size of
        the group.  Returns None when the route does not apply (states not packed, a member synced on step, ...);
        nothing has been touched then."""
        for m in group:
            if (m.dist_sync_on_step or m.full_state_update is not False or m._is_synced or m.compute_on_cpu
                    or m.multidim_average != "global"):
                return None
        packed = self._packed_states()
        if packed is None:
            return None
        before = packed.clone()
        self.update(*args, **kwargs)  # raises before the states change when the batch fails validation
        batch = packed - before
        if len(group) > 1 and hasattr(self, "num_classes"):
            # the kernel's error word (label outside [0, C)) is read once for the group, not by every member's compute
            _native.raise_on_status(packed.device, self.num_classes)
        values = []
        for m in group:
            saved = (m.tp, m.fp, m.tn, m.fn)
            m._update_count = self._update_count
            with m._batch_scope():
                try:
                    m.__dict__["_status_settled"] = len(group) > 1
                    m.tp, m.fp, m.tn, m.fn = (batch[i].view(saved[i].shape) for i in range(4))
                    values.append(m.compute())
                finally:
                    m.tp, m.fp, m.tn, m.fn = saved
                    m.__dict__["_status_settled"] = False
        return values

    def _forward_reduce_state_update(self, *args: Any, **kwargs: Any) -> Any:
        if not self.__dict__.get("_state_shared", False):  # see MetricCollection._compute_groups_create_state_ref
            values = self._forward_batch_values([self], args, kwargs)
            if values is not None:
                return values[0]
        return super()._forward_reduce_state_update(*args, **kwargs)


class BinaryStatScores(_AbstractStatScores):
    is_differentiable = False
    higher_is_better = None
    full_state_update = False

    def __init__(self, threshold: float = 0.5, multidim_average: Literal["global", "samplewise"] = "global",
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        zero_division = kwargs.pop("zero_division", 0)  # noqa: F841  (accepted for signature compatibility)
        super().__init__(**kwargs)
        if validate_args:
            F._binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        self.threshold = threshold
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(size=1, multidim_average=multidim_average)

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, self.multidim_average == "global")
        if self.validate_args:
            with fv:
                F._binary_stat_scores_tensor_validation(preds, target, self.multidim_average, self.ignore_index)
        packed = self._packed_states()
        tp, fp, tn, fn = F._binary_stat_scores_update_fused(preds, target, self.threshold, self.multidim_average,
                                                            self.ignore_index, out=None if fv.enabled else packed)
        fv.finish(preds.device)  # before the state is touched, like the reference's up-front validation
        self._update_state(tp, fp, tn, fn, packed)

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return F._binary_stat_scores_compute(tp, fp, tn, fn, self.multidim_average)


class MulticlassStatScores(_AbstractStatScores):
    is_differentiable = False
    higher_is_better = None
    full_state_update = False

    def __init__(self, num_classes: int, top_k: int = 1,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        self.num_classes = num_classes
        self.top_k = top_k
        self.average = average
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(size=1 if (average == "micro" and top_k == 1) else num_classes,
                           multidim_average=multidim_average)

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, F._multiclass_stat_scores_reports_values(
            self.top_k, self.average, self.multidim_average))
        if self.validate_args:
            with fv:
                F._multiclass_stat_scores_tensor_validation(preds, target, self.num_classes, self.multidim_average,
                                                            self.ignore_index)
        packed = self._packed_states()
        tp, fp, tn, fn = F._multiclass_stat_scores_update_fused(preds, target, self.num_classes, self.top_k,
                                                                self.average, self.multidim_average,
                                                                self.ignore_index, out=None if fv.enabled else packed)
        fv.finish(preds.device, self.num_classes)
        self._update_state(tp, fp, tn, fn, packed)

    def _final_state(self) -> Tuple[Tensor, Tensor, Tensor, Tensor]:
        out = super()._final_state()
        if out[0].is_cuda and not self.__dict__.get("_status_settled", False):
            _native.raise_on_status(out[0].device, self.num_classes)
        return out

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return F._multiclass_stat_scores_compute(tp, fp, tn, fn, self.average, self.multidim_average)


class MultilabelStatScores(_AbstractStatScores):
    is_differentiable = False
    higher_is_better = None
    full_state_update = False

    def __init__(self, num_labels: int, threshold: float = 0.5,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        self.num_labels = num_labels
        self.threshold = threshold
        self.average = average
        self.multidim_average = multidim_average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._create_state(size=num_labels, multidim_average=multidim_average)

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, self.multidim_average == "global")
        if self.validate_args:
            with fv:
                F._multilabel_stat_scores_tensor_validation(preds, target, self.num_labels, self.multidim_average,
                                                            self.ignore_index)
        packed = self._packed_states()
        tp, fp, tn, fn = F._multilabel_stat_scores_update_fused(preds, target, self.num_labels, self.threshold,
                                                                self.multidim_average, self.ignore_index,
                                                                out=None if fv.enabled else packed)
        fv.finish(preds.device)
        self._update_state(tp, fp, tn, fn, packed)

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return F._multilabel_stat_scores_compute(tp, fp, tn, fn, self.average, self.multidim_average)


def _task_dispatch(task, binary_cls, multiclass_cls, multilabel_cls, threshold, num_classes, num_labels, average,
                   multidim_average, top_k, ignore_index, validate_args, extra_args=(), **kwargs):
    """Shared body of the `Metric(task=...)` wrappers for the stat-scores family."""
    task = ClassificationTask.from_str(task)
    assert multidim_average is not None
    kwargs.update({"multidim_average": multidim_average, "ignore_index": ignore_index, "validate_args": validate_args})
    if task == ClassificationTask.BINARY:
        return binary_cls(*extra_args, threshold, **kwargs)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        if not isinstance(top_k, int):
            raise ValueError(f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`")
        return multiclass_cls(*extra_args, num_classes, top_k, average, **kwargs)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_cls(*extra_args, num_labels, threshold, average, **kwargs)
    raise ValueError(f"Task {task} not supported!")


class StatScores(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _task_dispatch(task, BinaryStatScores, MulticlassStatScores, MultilabelStatScores, threshold,
                              num_classes, num_labels, average, multidim_average, top_k, ignore_index, validate_args,
                              **kwargs)
