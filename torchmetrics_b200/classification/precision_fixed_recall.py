"""PrecisionAtFixedRecall classes (reference: classification/precision_fixed_recall.py); built by `_operating_point.make_classes`."""
from torchmetrics_b200.classification._operating_point import make_classes
from torchmetrics_b200.functional.classification._operating_point import PRECISION_AT_RECALL

BinaryPrecisionAtFixedRecall, MulticlassPrecisionAtFixedRecall, MultilabelPrecisionAtFixedRecall, PrecisionAtFixedRecall = make_classes(
    PRECISION_AT_RECALL, "PrecisionAtFixedRecall", "min_recall", __name__)
