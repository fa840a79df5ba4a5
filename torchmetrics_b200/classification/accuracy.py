"""Accuracy classes (reference: classification/accuracy.py)."""
from torchmetrics_b200.classification._from_stats import make_family
from torchmetrics_b200.functional.classification.accuracy import _accuracy_reduce

BinaryAccuracy, MulticlassAccuracy, MultilabelAccuracy, Accuracy = make_family(
    "Accuracy", _accuracy_reduce, "fraction of correct predictions")
for _c in (BinaryAccuracy, MulticlassAccuracy, MultilabelAccuracy, Accuracy):
    _c.__module__ = __name__
