"""Task-dispatch base (reference: classification/base.py:19-32)."""
from torchmetrics_b200.metric import Metric


class _ClassificationTaskWrapper(Metric):
    """`Wrapper(task=...)` returns the task-specific metric from `__new__`; it is never instantiated itself."""

    def update(self, *args, **kwargs) -> None:
        raise NotImplementedError(
            f"{self.__class__.__name__} metric does not have a global `update` method. Use the task specific metric."
        )

    def compute(self) -> None:
        raise NotImplementedError(
            f"{self.__class__.__name__} metric does not have a global `compute` method. Use the task specific metric."
        )
