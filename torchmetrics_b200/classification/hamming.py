"""Hamming distance classes (reference: classification/hamming.py)."""
from torchmetrics_b200.classification._from_stats import make_family
from torchmetrics_b200.functional.classification.hamming import _hamming_distance_reduce

BinaryHammingDistance, MulticlassHammingDistance, MultilabelHammingDistance, HammingDistance = make_family(
    "HammingDistance", _hamming_distance_reduce, "fraction of labels predicted wrongly")
for _c in (BinaryHammingDistance, MulticlassHammingDistance, MultilabelHammingDistance, HammingDistance):
    _c.__module__ = __name__
    _c.higher_is_better = False
