"""RecallAtFixedPrecision classes (reference: classification/recall_fixed_precision.py); built by `_operating_point.make_classes`."""
from torchmetrics_b200.classification._operating_point import make_classes
from torchmetrics_b200.functional.classification._operating_point import RECALL_AT_PRECISION

BinaryRecallAtFixedPrecision, MulticlassRecallAtFixedPrecision, MultilabelRecallAtFixedPrecision, RecallAtFixedPrecision = make_classes(
    RECALL_AT_PRECISION, "RecallAtFixedPrecision", "min_precision", __name__)
