from torchmetrics_b200.functional.classification import *  # noqa: F401,F403
from torchmetrics_b200.functional.classification import __all__ as _cls_all

__all__ = list(_cls_all)
