"""Rank-zero logging helpers (reference: utilities/prints.py:22-57; the rank comes from LOCAL_RANK)."""
import logging
import os
import warnings
from functools import wraps
from typing import Any, Callable

log = logging.getLogger("torchmetrics_b200")


def rank_zero_only(fn: Callable) -> Callable:
    @wraps(fn)
    def only_on_rank_zero(*args: Any, **kwargs: Any) -> Any:
        return fn(*args, **kwargs) if rank_zero_only.rank == 0 else None

    return only_on_rank_zero


rank_zero_only.rank = getattr(rank_zero_only, "rank", int(os.environ.get("LOCAL_RANK", 0)))


@rank_zero_only
def rank_zero_warn(*args: Any, **kwargs: Any) -> None:
    warnings.warn(*args, **kwargs)  # noqa: B028


@rank_zero_only
def rank_zero_info(*args: Any, **kwargs: Any) -> None:
    log.info(*args, **kwargs)


@rank_zero_only
def rank_zero_debug(*args: Any, **kwargs: Any) -> None:
    log.debug(*args, **kwargs)
