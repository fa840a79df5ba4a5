"""F-beta / F1 classes (reference: classification/f_beta.py)."""
from typing import Any, Optional

from torch import Tensor
from typing_extensions import Literal

from torchmetrics_b200.classification.base import _ClassificationTaskWrapper
from torchmetrics_b200.classification.stat_scores import BinaryStatScores, MulticlassStatScores, MultilabelStatScores
from torchmetrics_b200.functional.classification.f_beta import _check_beta, _fbeta_reduce
from torchmetrics_b200.metric import Metric
from torchmetrics_b200.utilities.enums import ClassificationTask


class BinaryFBetaScore(BinaryStatScores):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0

    def __init__(self, beta: float, threshold: float = 0.5, multidim_average: Literal["global", "samplewise"] = "global",
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(threshold=threshold, multidim_average=multidim_average, ignore_index=ignore_index,
                         validate_args=False, **kwargs)
        if validate_args:
            _check_beta(beta)
            from torchmetrics_b200.functional.classification.stat_scores import _binary_stat_scores_arg_validation

            _binary_stat_scores_arg_validation(threshold, multidim_average, ignore_index)
        self.validate_args = validate_args
        self.beta = beta

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _fbeta_reduce(tp, fp, tn, fn, self.beta, average="binary", multidim_average=self.multidim_average)


class MulticlassFBetaScore(MulticlassStatScores):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    plot_legend_name = "Class"

    def __init__(self, beta: float, num_classes: int, top_k: int = 1,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(num_classes=num_classes, top_k=top_k, average=average, multidim_average=multidim_average,
                         ignore_index=ignore_index, validate_args=False, **kwargs)
        if validate_args:
            _check_beta(beta)
            from torchmetrics_b200.functional.classification.stat_scores import _multiclass_stat_scores_arg_validation

            _multiclass_stat_scores_arg_validation(num_classes, top_k, average, multidim_average, ignore_index)
        self.validate_args = validate_args
        self.beta = beta

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _fbeta_reduce(tp, fp, tn, fn, self.beta, average=self.average, multidim_average=self.multidim_average)


class MultilabelFBetaScore(MultilabelStatScores):
    is_differentiable = False
    higher_is_better = True
    full_state_update = False
    plot_lower_bound = 0.0
    plot_upper_bound = 1.0
    plot_legend_name = "Label"

    def __init__(self, beta: float, num_labels: int, threshold: float = 0.5,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(num_labels=num_labels, threshold=threshold, average=average,
                         multidim_average=multidim_average, ignore_index=ignore_index, validate_args=False, **kwargs)
        if validate_args:
            _check_beta(beta)
            from torchmetrics_b200.functional.classification.stat_scores import _multilabel_stat_scores_arg_validation

            _multilabel_stat_scores_arg_validation(num_labels, threshold, average, multidim_average, ignore_index)
        self.validate_args = validate_args
        self.beta = beta

    def compute(self) -> Tensor:
        tp, fp, tn, fn = self._final_state()
        return _fbeta_reduce(tp, fp, tn, fn, self.beta, average=self.average, multidim_average=self.multidim_average,
                             multilabel=True)


class BinaryF1Score(BinaryFBetaScore):
    def __init__(self, threshold: float = 0.5, multidim_average: Literal["global", "samplewise"] = "global",
                 ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(beta=1.0, threshold=threshold, multidim_average=multidim_average, ignore_index=ignore_index,
                         validate_args=validate_args, **kwargs)


class MulticlassF1Score(MulticlassFBetaScore):
    def __init__(self, num_classes: int, top_k: int = 1,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(beta=1.0, num_classes=num_classes, top_k=top_k, average=average,
                         multidim_average=multidim_average, ignore_index=ignore_index, validate_args=validate_args,
                         **kwargs)


class MultilabelF1Score(MultilabelFBetaScore):
    def __init__(self, num_labels: int, threshold: float = 0.5,
                 average: Optional[Literal["micro", "macro", "weighted", "none"]] = "macro",
                 multidim_average: Literal["global", "samplewise"] = "global", ignore_index: Optional[int] = None,
                 validate_args: bool = True, **kwargs: Any) -> None:
        super().__init__(beta=1.0, num_labels=num_labels, threshold=threshold, average=average,
                         multidim_average=multidim_average, ignore_index=ignore_index, validate_args=validate_args,
                         **kwargs)


def _dispatch(task, beta, threshold, num_classes, num_labels, average, multidim_average, top_k, ignore_index,
              validate_args, kwargs, f1: bool) -> Metric:
    task = ClassificationTask.from_str(task)
    assert multidim_average is not None
    kwargs.update({"multidim_average": multidim_average, "ignore_index": ignore_index, "validate_args": validate_args})
    pre = () if f1 else (beta,)
    if task == ClassificationTask.BINARY:
        return (BinaryF1Score if f1 else BinaryFBetaScore)(*pre, threshold, **kwargs)
    if task == ClassificationTask.MULTICLASS:
        if not isinstance(num_classes, int):
            raise ValueError(f"`num_classes` is expected to be `int` but `{type(num_classes)} was passed.`")
        if not isinstance(top_k, int):
            raise ValueError(f"`top_k` is expected to be `int` but `{type(top_k)} was passed.`")
        return (MulticlassF1Score if f1 else MulticlassFBetaScore)(*pre, num_classes, top_k, average, **kwargs)
    if task == ClassificationTask.MULTILABEL:
        if not isinstance(num_labels, int):
            raise ValueError(f"`num_labels` is expected to be `int` but `{type(num_labels)} was passed.`")
        return (MultilabelF1Score if f1 else MultilabelFBetaScore)(*pre, num_labels, threshold, average, **kwargs)
    raise ValueError(f"Task {task} not supported!")


class FBetaScore(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], beta: float = 1.0, threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _dispatch(task, beta, threshold, num_classes, num_labels, average, multidim_average, top_k,
                         ignore_index, validate_args, kwargs, f1=False)


class F1Score(_ClassificationTaskWrapper):
    def __new__(cls, task: Literal["binary", "multiclass", "multilabel"], threshold: float = 0.5,
                num_classes: Optional[int] = None, num_labels: Optional[int] = None,
                average: Optional[Literal["micro", "macro", "weighted", "none"]] = "micro",
                multidim_average: Optional[Literal["global", "samplewise"]] = "global", top_k: Optional[int] = 1,
                ignore_index: Optional[int] = None, validate_args: bool = True, **kwargs: Any) -> Metric:
        return _dispatch(task, 1.0, threshold, num_classes, num_labels, average, multidim_average, top_k,
                         ignore_index, validate_args, kwargs, f1=True)
