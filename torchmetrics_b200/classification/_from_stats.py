"""Factory for the metrics that only differ from `*StatScores` in their `compute` reducer
(reference: classification/accuracy.py, precision_recall.py, f_beta.py — ~2700 lines of near-identical classes)."""
from typing import Any, Callable, Optional

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.stat_scores import (
    BinaryStatScores,
    MulticlassStatScores,
    MultilabelStatScores,
    _task_dispatch,
)
from torchmetrics_b200.metric import Metric


def make_family(name: str, reducer: Callable[..., Tensor], doc: str):
    """Return (Binary<name>, Multiclass<name>, Multilabel<name>, <name>) with `reducer(tp, fp, tn, fn, average=...,
    multidim_average=..., multilabel=...)` as compute."""

    class _Binary(BinaryStatScores):
        is_differentiable = False
        higher_is_better = True
        full_state_update = False
        plot_lower_bound = 0.0
        plot_upper_bound = 1.0

        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(tp, fp, tn, fn, average="binary", multidim_average=self.multidim_average)

    class _Multiclass(MulticlassStatScores):
        is_differentiable = False
        higher_is_better = True
        full_state_update = False
        plot_lower_bound = 0.0
        plot_upper_bound = 1.0
        plot_legend_name = "Class"

        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(tp, fp, tn, fn, average=self.average, multidim_average=self.multidim_average)

    class _Multilabel(MultilabelStatScores):
        is_differentiable = False
        higher_is_better = True
        full_state_update = False
        plot_lower_bound = 0.0
        plot_upper_bound = 1.0
        plot_legend_name = "Label"

        def compute(self) -> Tensor:
            tp, fp, tn, fn = self._final_state()
            return reducer(tp, fp, tn, fn, average=self.average, multidim_average=self.multidim_average,
                           multilabel=True)

    class _Task(_ClassificationTaskWrapper):
        def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                    num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                    average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                    multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                    ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
            return _task_dispatch(task, _Binary, _Multiclass, _Multilabel, threshold, num_classes, num_labels, average,
                                  multidim_average, top_k, ignore_index, validate_args, **kwargs)

    for cls, prefix in ((_Binary, "Binary"), (_Multiclass, "Multiclass"), (_Multilabel, "Multilabel"), (_Task, "")):
        cls.__name__ = cls.__qualname__ = prefix + name
        cls.__doc__ = f"{prefix or 'Task-dispatching '}{name}: {doc}"
    return _Binary, _Multiclass, _Multilabel, _Task
