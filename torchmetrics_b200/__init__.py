"""B200-native metric-state hot path with the torchmetrics API.

    from torchmetrics_b200 import MetricCollection
    from torchmetrics_b200.classification import BinaryAUROC, MulticlassConfusionMatrix
    import torchmetrics_b200.functional as F

Count states are accumulated by hand-written sm_100a CUDA kernels (libtm_b200.so, C ABI in include/tm_b200.h);
there is no CPU fallback for them.
"""
__version__ = "0.1.0"

from torchmetrics_b200 import functional  # noqa: E402,F401
from torchmetrics_b200.classification import *  # noqa: E402,F401,F403
from torchmetrics_b200.classification import __all__ as _cls_all  # noqa: E402
from torchmetrics_b200.collections import MetricCollection  # noqa: E402
from torchmetrics_b200.nominal import CramersV, PearsonsContingencyCoefficient, TheilsU, TschuprowsT  # noqa: E402
from torchmetrics_b200.metric import CompositionalMetric, Metric  # noqa: E402

__all__ = ["Metric", "MetricCollection", "CompositionalMetric", "functional", "CramersV",
           "PearsonsContingencyCoefficient", "TheilsU", "TschuprowsT", *_cls_all]
