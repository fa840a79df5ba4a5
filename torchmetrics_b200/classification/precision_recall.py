"""Precision and Recall classes (reference: classification/precision_recall.py)."""
from torchmetrics_b200.classification._from_stats import make_family
from torchmetrics_b200.functional.classification.precision_recall import _precision, _recall

BinaryPrecision, MulticlassPrecision, MultilabelPrecision, Precision = make_family(
    "Precision", _precision, "tp / (tp + fp)")
BinaryRecall, MulticlassRecall, MultilabelRecall, Recall = make_family("Recall", _recall, "tp / (tp + fn)")
for _c in (BinaryPrecision, MulticlassPrecision, MultilabelPrecision, Precision, BinaryRecall, MulticlassRecall,
           MultilabelRecall, Recall):
    _c.__module__ = __name__
