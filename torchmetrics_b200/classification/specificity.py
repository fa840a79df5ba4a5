"""Specificity classes (reference: classification/specificity.py)."""
from torchmetrics_b200.classification._from_stats import make_family
from torchmetrics_b200.functional.classification.specificity import _specificity_reduce

BinarySpecificity, MulticlassSpecificity, MultilabelSpecificity, Specificity = make_family(
    "Specificity", _specificity_reduce, "true negatives over all negatives")
for _c in (BinarySpecificity, MulticlassSpecificity, MultilabelSpecificity, Specificity):
    _c.__module__ = __name__
