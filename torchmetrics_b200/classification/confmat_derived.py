"""Metrics whose state IS the confusion matrix and that only add a reducer in `compute`
(reference: classification/jaccard.py, cohen_kappa.py, matthews_corrcoef.py — subclasses of the *ConfusionMatrix
classes, so `update` is the K1 / K2 path of classification/confusion_matrix.py)."""
from importlib import import_module
from typing import Any, Optional

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.confusion_matrix import (
    BinaryConfusionMatrix,
    MulticlassConfusionMatrix,
    MultilabelConfusionMatrix,
)
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.enums import ClassificationTask, ClassificationTaskNoMultilabel

J = import_module("torchmetrics_b200.functional.classification.jaccard")
K = import_module("torchmetrics_b200.functional.classification.cohen_kappa")
M = import_module("torchmetrics_b200.functional.classification.matthews_corrcoef")



def _need_int(name: str, value) -> None:
    if not isinstance(value, int):
        raise ValueError(f"`{name}` is expected to be `int` but `{type(value)} was passed.`")


# ------------------------------------------------------------------------------------------- Jaccard index
class BinaryJaccardIndex(BinaryConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound = 0.0, 1.0

    def __init__(self, threshold: float = 0.5, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(threshold=threshold, ignore_index=ignore_index, normalize=None, validate_args=validate_args,
                         **kwargs)

    def compute(self) -> Tensor:
        self._check_device_status(2)
        return J._jaccard_index_reduce(self.confmat, average="binary")


class MulticlassJaccardIndex(MulticlassConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound, plot_legend_name = 0.0, 1.0, "Class"

    def __init__(self, num_classes: int, average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(num_classes=num_classes, ignore_index=ignore_index, normalize=None,
                         validate_args=validate_args, **kwargs)
        if validate_args:
            J._check_average(average)
        self.average = average

    def compute(self) -> Tensor:
        self._check_device_status(self.num_classes)
        return J._jaccard_index_reduce(self.confmat, average=self.average, ignore_index=self.ignore_index)


class MultilabelJaccardIndex(MultilabelConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound, plot_legend_name = 0.0, 1.0, "Label"

    def __init__(self, num_labels: int, threshold: float = 0.5,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(num_labels=num_labels, threshold=threshold, ignore_index=ignore_index, normalize=None,
                         validate_args=validate_args, **kwargs)
        if validate_args:
            J._check_average(average)
        self.average = average

    def compute(self) -> Tensor:
        return J._jaccard_index_reduce(self.confmat, average=self.average)


class JaccardIndex(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        task = ClassificationTask.from_str(task)
        kwargs.update({"ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTask.BINARY:
            return BinaryJaccardIndex(threshold, **kwargs)
        if task == ClassificationTask.MULTICLASS:
            _need_int("num_classes", num_classes)
            return MulticlassJaccardIndex(num_classes, average, **kwargs)
        if task == ClassificationTask.MULTILABEL:
            _need_int("num_labels", num_labels)
            return MultilabelJaccardIndex(num_labels, threshold, average, **kwargs)
        raise ValueError(f"Task {task} not supported!")


# ------------------------------------------------------------------------------------------- Cohen's kappa
class BinaryCohenKappa(BinaryConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound = 0.0, 1.0

    def __init__(self, threshold: float = 0.5, ignore_index: Optional[int] = None,
                 weights: Optional[Literal["linear", "quadratic", "none"]] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(threshold=threshold, ignore_index=ignore_index, normalize=None, validate_args=validate_args,
                         **kwargs)
        if validate_args:
            K._check_weights(weights)
        self.weights = weights

    def compute(self) -> Tensor:
        self._check_device_status(2)
        return K._cohen_kappa_reduce(self.confmat, self.weights)


class MulticlassCohenKappa(MulticlassConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound, plot_legend_name = 0.0, 1.0, "Class"

    def __init__(self, num_classes: int, ignore_index: Optional[int] = None,
                 weights: Optional[Literal["linear", "quadratic", "none"]] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(num_classes=num_classes, ignore_index=ignore_index, normalize=None,
                         validate_args=validate_args, **kwargs)
        if validate_args:
            K._check_weights(weights)
        self.weights = weights

    def compute(self) -> Tensor:
        self._check_device_status(self.num_classes)
        return K._cohen_kappa_reduce(self.confmat, self.weights)


class CohenKappa(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass"], threshold: float = 0.5, num_classes: Optional[int] = None,
                weights: Optional[Literal["linear", "quadratic", "none"]] = None, ignore_index: Optional[int] = None,
                validate_args: bool = True, **kwargs: Any) -> Metric:
        task = ClassificationTaskNoMultilabel.from_str(task)
        kwargs.update({"weights": weights, "ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTaskNoMultilabel.BINARY:
            return BinaryCohenKappa(threshold, **kwargs)
        if task == ClassificationTaskNoMultilabel.MULTICLASS:
            _need_int("num_classes", num_classes)
            return MulticlassCohenKappa(num_classes, **kwargs)
        raise ValueError(f"Task {task} not supported!")


# ------------------------------------------------------------------------------------------- Matthews corrcoef
class BinaryMatthewsCorrCoef(BinaryConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound = -1.0, 1.0

    def __init__(self, threshold: float = 0.5, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(threshold, ignore_index, normalize=None, validate_args=validate_args, **kwargs)

    def compute(self) -> Tensor:
        self._check_device_status(2)
        return M._matthews_corrcoef_reduce(self.confmat)


class MulticlassMatthewsCorrCoef(MulticlassConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound, plot_legend_name = -1.0, 1.0, "Class"

    def __init__(self, num_classes: int, ignore_index: Optional[int] = None, validate_args: bool = True,
                 **kwargs: Any) -> None:
        super().__init__(num_classes, ignore_index, normalize=None, validate_args=validate_args, **kwargs)

    def compute(self) -> Tensor:
        self._check_device_status(self.num_classes)
        return M._matthews_corrcoef_reduce(self.confmat)


class MultilabelMatthewsCorrCoef(MultilabelConfusionMatrix):
    is_differentiable, higher_is_better, full_state_update = False, True, False
    plot_lower_bound, plot_upper_bound, plot_legend_name = -1.0, 1.0, "Label"

    def __init__(self, num_labels: int, threshold: float = 0.5, ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(num_labels, threshold, ignore_index, normalize=None, validate_args=validate_args, **kwargs)

    def compute(self) -> Tensor:
        return M._matthews_corrcoef_reduce(self.confmat)


class MatthewsCorrCoef(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        task = ClassificationTask.from_str(task)
        kwargs.update({"ignore_index": ignore_index, "validate_args": validate_args})
        if task == ClassificationTask.BINARY:
            return BinaryMatthewsCorrCoef(threshold, **kwargs)
        if task == ClassificationTask.MULTICLASS:
            _need_int("num_classes", num_classes)
            return MulticlassMatthewsCorrCoef(num_classes, **kwargs)
        if task == ClassificationTask.MULTILABEL:
            _need_int("num_labels", num_labels)
            return MultilabelMatthewsCorrCoef(num_labels, threshold, **kwargs)
        raise ValueError(f"Not handled value: {task}")
