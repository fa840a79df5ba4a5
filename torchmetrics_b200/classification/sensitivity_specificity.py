"""SensitivityAtSpecificity classes (reference: classification/sensitivity_specificity.py); built by `_operating_point.make_classes`."""
from torchmetrics_b200.classification._operating_point import make_classes
from torchmetrics_b200.functional.classification._operating_point import SENSITIVITY_AT_SPECIFICITY

BinarySensitivityAtSpecificity, MulticlassSensitivityAtSpecificity, MultilabelSensitivityAtSpecificity, SensitivityAtSpecificity = make_classes(
    SENSITIVITY_AT_SPECIFICITY, "SensitivityAtSpecificity", "min_specificity", __name__)
