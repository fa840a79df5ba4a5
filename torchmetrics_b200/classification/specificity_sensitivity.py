"""SpecificityAtSensitivity classes (reference: classification/specificity_sensitivity.py); built by `_operating_point.make_classes`."""
from torchmetrics_b200.classification._operating_point import make_classes
from torchmetrics_b200.functional.classification._operating_point import SPECIFICITY_AT_SENSITIVITY

BinarySpecificityAtSensitivity, MulticlassSpecificityAtSensitivity, MultilabelSpecificityAtSensitivity, SpecificityAtSensitivity = make_classes(
    SPECIFICITY_AT_SENSITIVITY, "SpecificityAtSensitivity", "min_sensitivity", __name__)
