"""State reducers and small tensor helpers (reference: utilities/data.py:28-55, :179-219)."""
from typing import Any, List, Optional, Sequence, Union

import torch
from torch import Tensor

from torchmetrics_b200.utilities.exceptions import TorchMetricsUserWarning  # noqa: F401


def dim_zero_cat(x: Union[Tensor, List[Tensor]]) -> Tensor:
    """Concatenate a list state along dim 0; 0-d entries become length-1 (data.py:28-35)."""
    if isinstance(x, Tensor):
        return x
    parts = [t.unsqueeze(0) if t.ndim == 0 else t for t in x]
    if not parts:
        raise ValueError("No samples to concatenate")
    return torch.cat(parts, dim=0)


def dim_zero_sum(x: Tensor) -> Tensor:
    return torch.sum(x, dim=0)


def dim_zero_mean(x: Tensor) -> Tensor:
    return torch.mean(x, dim=0)


def dim_zero_max(x: Tensor) -> Tensor:
    return torch.max(x, dim=0).values


def dim_zero_min(x: Tensor) -> Tensor:
    return torch.min(x, dim=0).values


def _flatten(x: Sequence) -> list:
    return [item for sub in x for item in sub]


def apply_to_collection(data: Any, dtype: type, function, *args: Any, **kwargs: Any) -> Any:
    """Map `function` over every `dtype` instance inside nested lists / tuples / dicts."""
    if isinstance(data, dtype):
        return function(data, *args, **kwargs)
    if isinstance(data, dict):
        return type(data)((k, apply_to_collection(v, dtype, function, *args, **kwargs)) for k, v in data.items())
    if isinstance(data, tuple) and hasattr(data, "_fields"):
        return type(data)(*(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data))
    if isinstance(data, (list, tuple)):
        return type(data)(apply_to_collection(v, dtype, function, *args, **kwargs) for v in data)
    return data


def _squeeze_scalar_element_tensor(x: Tensor) -> Tensor:
    return x.squeeze() if x.numel() == 1 else x


def _squeeze_if_scalar(data: Any) -> Any:
    return apply_to_collection(data, Tensor, _squeeze_scalar_element_tensor)


def _bincount(x: Tensor, minlength: Optional[int] = None) -> Tensor:
    """Occurrences of each value of `x` (data.py:179-207).

    CUDA tensors go through the K1 histogram kernel (integer atomics: deterministic by construction, so the
    reference's deterministic-mode detour through an N x minlength mesh is unnecessary). CPU tensors are only met
    in host-side bookkeeping and use torch.bincount.
    """
    if minlength is None:
        minlength = len(torch.unique(x))
    if x.is_cuda:
        from torchmetrics_b200 import _native

        out = torch.zeros(minlength, dtype=torch.int64, device=x.device)
        _native.bincount_update_(out, x)
        _native.raise_on_status(x.device)
        return out
    return torch.bincount(x, minlength=minlength)


def _cumsum(x: Tensor, dim: Optional[int] = 0, dtype: Optional[torch.dtype] = None) -> Tensor:
    return torch.cumsum(x, dim=dim, dtype=dtype)


def select_topk(prob_tensor: Tensor, topk: int = 1, dim: int = 1) -> Tensor:
    """0/1 int32 mask of the `topk` largest entries along `dim` (data.py:125-149)."""
    if topk == 1:
        idx = prob_tensor.argmax(dim=dim, keepdim=True)
    else:
        idx = prob_tensor.topk(k=topk, dim=dim).indices
    return torch.zeros_like(prob_tensor, dtype=torch.int32).scatter_(dim, idx, 1)


def to_onehot(label_tensor: Tensor, num_classes: Optional[int] = None) -> Tensor:
    """(N, d1, ...) labels -> (N, C, d1, ...) 0/1 tensor of the labels' dtype (data.py:80-111)."""
    if num_classes is None:
        num_classes = int(label_tensor.max().detach().item() + 1)
    out = torch.zeros(label_tensor.shape[0], num_classes, *label_tensor.shape[1:], dtype=label_tensor.dtype,
                      device=label_tensor.device)
    return out.scatter_(1, label_tensor.long().unsqueeze(1).expand_as(out), 1.0)
