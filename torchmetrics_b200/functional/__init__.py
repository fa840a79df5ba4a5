from torchmetrics_b200.functional.classification import *  # noqa: F401,F403
from torchmetrics_b200.functional.classification import __all__ as _cls_all
from torchmetrics_b200.functional.nominal import *  # noqa: F401,F403
from torchmetrics_b200.functional.nominal import __all__ as _nom_all

__all__ = list(_cls_all) + list(_nom_all)
