"""Stateful precision-recall curves (reference: classification/precision_recall_curve.py:55-685).

`thresholds=None` keeps the raw (preds, target) as cat-reduced list states and computes the exact curve at the end
(K4); any other `thresholds` keeps a constant-size `(T[,C],2,2)` int64 `confmat` state that every `update` adds to in
place with one K3 pass.  `thresholds` itself is a non-persistent buffer, as in the reference (:157).
"""
from importlib import import_module
from typing import Any, List, Optional, Tuple, Union

import torch
from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200 import _native
from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
F = import_module("torchmetrics_b200.functional.classification.precision_recall_curve")
C = import_module("torchmetrics_b200.functional.classification._checks")
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.data import dim_zero_cat
from torchmetrics_b200.utilities.enums import ClassificationTask

_Thresholds = Optional[Union[int, List[float], Tensor]]


class _CurveBase(Metric):
    is_differentiable = False
    higher_is_better = None
    full_state_update = False
    preds: List[Tensor]
    target: List[Tensor]
    confmat: Tensor

    def _init_curve_state(self, thresholds: _Thresholds, state_shape: Tuple[int, ...]) -> None:
        """list states for the exact curve, or thresholds buffer + `confmat` state for the binned one."""
        thr = F._adjust_threshold_arg(thresholds)
        if thr is None:
            self.thresholds = thr
            self.add_state("preds", default=[], dist_reduce_fx="cat")
            self.add_state("target", default=[], dist_reduce_fx="cat")
            self._thresholds_ascending = True
            return
        self.register_buffer("thresholds", thr, persistent=False)
        # an int argument gives linspace (ascending); lists / tensors are checked once here so `update` never syncs
        self._thresholds_ascending = isinstance(thresholds, int) or bool((thr[1:] >= thr[:-1]).all())
        self.add_state("confmat", default=torch.zeros(len(thr), *state_shape, dtype=torch.long), dist_reduce_fx="sum")

    def _check_device_status(self) -> None:
        if self.thresholds is not None and self.confmat.is_cuda:
            _native.raise_on_status(self.confmat.device)


class BinaryPrecisionRecallCurve(_CurveBase):
    # AUROC / AveragePrecision can finish the exact curve without gathering every sample on every rank
    _supports_partition = False

    def __init__(self, thresholds: _Thresholds = None, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        # distributed_exact="partition" (extension): with thresholds=None, `compute` range-partitions
        # the samples by score across the ranks instead of all_gathering them; the states stay local
        mode = kwargs.pop("distributed_exact", "gather")
        if mode not in ("gather", "partition"):
            raise ValueError(f"Expected `distributed_exact` to be 'gather' or 'partition', got {mode}")
        if mode == "partition":
            if thresholds is not None or not self._supports_partition:
                raise ValueError("`distributed_exact='partition'` applies to BinaryAUROC / BinaryAveragePrecision with "
                                 "`thresholds=None` (binned states are a single all_reduce already)")
            kwargs["sync_on_compute"] = False
        super().__init__(**kwargs)
        self.distributed_exact = mode
        if validate_args:
            F._binary_precision_recall_curve_arg_validation(thresholds, ignore_index)
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, self.thresholds is not None)
        if self.validate_args:
            with fv:
                F._binary_precision_recall_curve_tensor_validation(preds, target, self.ignore_index)
        if self.thresholds is None:
            p, t, _ = F._binary_precision_recall_curve_format(preds, target, None, self.ignore_index)
            self.preds.append(p)
            # the exact curve only ever asks `target == 1` (precision_recall_curve.py:66): keep that answer as one
            # byte per sample - an eighth of the int64 labels in memory and on the wire of the cat-state gather
            self.target.append((t == 1).to(torch.uint8))
        else:
            def step(out: Tensor) -> None:
                F._binary_precision_recall_curve_update_fused(preds, target, self.thresholds, self.ignore_index,
                                                              out=out, assume_sorted=self._thresholds_ascending)

            step(self.confmat)
            fv.finish(preds.device, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def _state(self):
        if self.thresholds is None:
            return dim_zero_cat(self.preds), dim_zero_cat(self.target)
        self._check_device_status()
        return self.confmat

    def _partitioned_sums(self) -> Optional[Tuple[float, float, float, float]]:
        """(area2, ap_sum, positives, samples) over all ranks in partition mode, None otherwise."""
        import torch.distributed as dist

        if self.distributed_exact != "partition" or not (dist.is_available() and dist.is_initialized()):
            return None
        from torchmetrics_b200.functional.classification._exact import partitioned_exact_sums

        if len(self.preds):
            p, t = dim_zero_cat(self.preds), dim_zero_cat(self.target)
        else:
            p = torch.empty(0, device=self.device)
            t = torch.empty(0, dtype=torch.uint8, device=self.device)
        return partitioned_exact_sums(p, t, self.process_group)

    def compute(self) -> Tuple[Tensor, Tensor, Tensor]:
        return F._binary_precision_recall_curve_compute(self._state(), self.thresholds)


class MulticlassPrecisionRecallCurve(_CurveBase):
    def __init__(self, num_classes: int, thresholds: _Thresholds = None,
                 average: Optional[Literal["micro", "macro"]] = None, ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multiclass_precision_recall_curve_arg_validation(num_classes, thresholds, ignore_index, average)
        self.num_classes = num_classes
        self.average = average
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (2, 2) if average == "micro" else (num_classes, 2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, self.thresholds is not None)
        if self.validate_args:
            with fv:
                F._multiclass_precision_recall_curve_tensor_validation(preds, target, self.num_classes,
                                                                       self.ignore_index)
        if self.thresholds is None:
            p, t, _ = F._multiclass_precision_recall_curve_format(preds, target, self.num_classes, None,
                                                                  self.ignore_index, self.average)
            self.preds.append(p)
            self.target.append(t)
        else:
            def step(out: Tensor) -> None:
                F._multiclass_precision_recall_curve_update_fused(preds, target, self.num_classes, self.thresholds,
                                                                  self.ignore_index, self.average, out=out,
                                                                  assume_sorted=self._thresholds_ascending)

            step(self.confmat)
            fv.finish(preds.device, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def _state(self):
        if self.thresholds is None:
            return dim_zero_cat(self.preds), dim_zero_cat(self.target)
        return self.confmat

    def compute(self):
        return F._multiclass_precision_recall_curve_compute(self._state(), self.num_classes, self.thresholds,
                                                            self.average)


class MultilabelPrecisionRecallCurve(_CurveBase):
    def __init__(self, num_labels: int, thresholds: _Thresholds = None, ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(**kwargs)
        if validate_args:
            F._multilabel_precision_recall_curve_arg_validation(num_labels, thresholds, ignore_index)
        self.num_labels = num_labels
        self.ignore_index = ignore_index
        self.validate_args = validate_args
        self._init_curve_state(thresholds, (num_labels, 2, 2))

    def update(self, preds: Tensor, target: Tensor) -> None:
        fv = C.FusedValidation(self.validate_args, preds, target, self.thresholds is not None)
        if self.validate_args:
            with fv:
                F._multilabel_precision_recall_curve_tensor_validation(preds, target, self.num_labels,
                                                                       self.ignore_index)
        if self.thresholds is None:
            p, t, _ = F._multilabel_precision_recall_curve_format(preds, target, self.num_labels, None,
                                                                  self.ignore_index)
            self.preds.append(p)
            self.target.append(t)
        else:
            def step(out: Tensor) -> None:
                F._multilabel_precision_recall_curve_update_fused(preds, target, self.num_labels, self.thresholds,
                                                                  self.ignore_index, out=out,
                                                                  assume_sorted=self._thresholds_ascending)

            step(self.confmat)
            fv.finish(preds.device, undo=lambda: C.rollback(self.confmat, step), owner=self)

    def _state(self):
        if self.thresholds is None:
            return dim_zero_cat(self.preds), dim_zero_cat(self.target)
        self._check_device_status()
        return self.confmat

    def compute(self):
        return F._multilabel_precision_recall_curve_compute(self._state(), self.num_labels, self.thresholds,
                                                            self.ignore_index)


def _curve_task(task, binary_cls, multiclass_cls, multilabel_cls, thresholds, num_classes, num_labels, ignore_index,
                validate_args, kwargs, multiclass_extra=None, multilabel_extra=None, binary_extra=None):
    task = ClassificationTask.from_str(task)
    kwargs.update({"thresholds": thresholds, "ignore_index": ignore_index, "validate_args": validate_args})
    if task == ClassificationTask.BINARY:
        return binary_cls(**(binary_extra or {}), **kwargs)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        return multiclass_cls(num_classes, **(multiclass_extra or {}), **kwargs)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return multilabel_cls(num_labels, **(multilabel_extra or {}), **kwargs)
    raise ValueError(f"Task {task} not supported!")


class PrecisionRecallCurve(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], thresholds: _Thresholds = None,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _curve_task(task, BinaryPrecisionRecallCurve, MulticlassPrecisionRecallCurve,
                           MultilabelPrecisionRecallCurve, thresholds, num_classes, num_labels, ignore_index,
                           validate_args, kwargs)
