from torchmetrics_b200.utilities.data import dim_zero_cat, dim_zero_max, dim_zero_mean, dim_zero_min, dim_zero_sum
from torchmetrics_b200.utilities.distributed import gather_all_tensors
from torchmetrics_b200.utilities.prints import rank_zero_debug, rank_zero_info, rank_zero_warn

__all__ = [
    "dim_zero_cat", "dim_zero_max", "dim_zero_mean", "dim_zero_min", "dim_zero_sum",
    "gather_all_tensors", "rank_zero_debug", "rank_zero_info", "rank_zero_warn",
]
