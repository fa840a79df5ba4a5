"""Argument and tensor validation shared by the classification functionals.

Conditions and exception types follow the reference's `_*_arg_validation` / `_*_tensor_validation` functions
(functional/classification/confusion_matrix.py:62-116, :228-303; stat_scores.py:30-87, :216-322, :560-644;
precision_recall_curve.py:94-159, :360-420, :684-736).  Value checks avoid the reference's `torch.unique` sort on
the passing path: labels inside the legal set cannot have too many distinct values, so one min/max (or one masked
`any`) pass decides; only an input that is about to fail pays for the exact unique count used in the message.
"""
import threading
from typing import Callable, List, Optional, Sequence, Union

import torch
from torch import Tensor

from torchmetrics_b200 import _native
from torchmetrics_b200.utilities.checks import _check_same_shape


def check_ignore_index(ignore_index: Optional[int]) -> None:
    if ignore_index is not None and not isinstance(ignore_index, int):
        raise ValueError(f"Expected argument `ignore_index` to either be `None` or an integer, but got {ignore_index}")


def check_threshold(threshold: float) -> None:
    if not (isinstance(threshold, float) and (0 <= threshold <= 1)):
        raise ValueError(f"Expected argument `threshold` to be a float in the [0,1] range, but got {threshold}.")


def check_choice(name: str, value, allowed: Sequence) -> None:
    if value not in allowed:
        raise ValueError(f"Expected argument `{name}` to be one of {tuple(allowed)}, but got {value}")


def check_num(name: str, value, minimum: int = 2) -> None:
    if not isinstance(value, int) or value < minimum:
        raise ValueError(f"Expected argument `{name}` to be an integer larger than {minimum - 1}, but got {value}")


def check_thresholds_arg(thresholds: Optional[Union[int, List[float], Tensor]]) -> None:
    """precision_recall_curve.py:94-122."""
    if thresholds is not None and not isinstance(thresholds, (list, int, Tensor)):
        raise ValueError(
            "Expected argument `thresholds` to either be an integer, list of floats or"
            f" tensor of floats, but got {thresholds}"
        )
    if isinstance(thresholds, int) and thresholds < 2:
        raise ValueError(
            f"If argument `thresholds` is an integer, expected it to be larger than 1, but got {thresholds}"
        )
    if isinstance(thresholds, list) and not all(isinstance(t, float) and 0 <= t <= 1 for t in thresholds):
        raise ValueError(
            "If argument `thresholds` is a list, expected all elements to be floats in the [0,1] range,"
            f" but got {thresholds}"
        )
    if isinstance(thresholds, Tensor) and not thresholds.ndim == 1:
        raise ValueError("If argument `thresholds` is an tensor, expected the tensor to be 1d")


# ------------------------------------------------------------------------------------------------------------
# Fused value validation.  The reference validates label values in separate passes over the
# inputs before every update (torch.unique, e.g. stat_scores.py:67-87, confusion_matrix.py:98-115).  The update
# kernels look at every label anyway, so they note "target outside its value set" / "label preds outside theirs"
# in a device status word; the host then reads 8 bytes after the kernel instead of re-reading the inputs, and
# only when a bit is set runs the reference-style check below to produce the reference's verdict and message.
# ------------------------------------------------------------------------------------------------------------
_tls = threading.local()


class FusedValidation:
    """Collects the value checks of one `_tensor_validation` call and settles them after the update kernel.

        fv = FusedValidation(validate_args, preds, target)
        with fv:
            _xxx_tensor_validation(preds, target, ...)   # shape/dtype checks run now, value checks are recorded
        ... launch the fused update ...
        fv.finish(device, num_classes, undo)            # status word -> run the recorded checks only if flagged

    When not enabled (CPU tensors, or an update path whose kernel does not report values) the checks run
    immediately and `finish` does nothing.
    """

    def __init__(self, validate_args: bool, preds: Tensor, target: Tensor, fused_path: bool = True) -> None:
        self.enabled = bool(validate_args and fused_path and preds.is_cuda and target.is_cuda)
        self.pending: List[Callable[[], None]] = []

    def __enter__(self) -> "FusedValidation":
        if self.enabled:
            _tls.sink = self.pending
        return self

    def __exit__(self, *exc) -> None:
        _tls.sink = None

    def finish(self, dev: torch.device, num_classes: Optional[int] = None,
               undo: Optional[Callable[[], None]] = None, errors: bool = False, owner=None) -> None:
        """`errors`: also settle the kernel's error bits (label outside the matrix) when validation is off.
        `owner`: the metric; with `validate_args="deferred"` the verdict of this batch is collected at the owner's
        next update / compute / reset instead of now (see `DeferredVerdict`)."""
        if self.enabled and owner is not None and getattr(owner, "validate_args", True) == "deferred":
            slots = owner.__dict__.get("_validation_slots")
            if slots is None:  # two pinned status copies + events per metric, used alternately (creating them is slow)
                slots = [[torch.empty(2, dtype=torch.int32).pin_memory(), torch.cuda.Event()] for _ in range(2)] + [0]
                owner.__dict__["_validation_slots"] = slots
            slots[2] ^= 1
            verdict = DeferredVerdict(dev, self.pending, num_classes, undo, *slots[slots[2]])
            previous = owner.__dict__.get("_pending_validation")
            owner.__dict__["_pending_validation"] = verdict
            if previous is not None:
                previous.settle()
            return
        if not self.enabled:
            if errors and dev.type == "cuda":
                _native.raise_on_status(dev, num_classes)
            return
        err, val = _native.read_status_words(dev)
        if not (err or val):
            return
        try:
            if val:
                for chk in self.pending:
                    chk()
            _native.raise_for_bits(err, num_classes)
        except Exception:
            # the reference raises before it touches the state: take the batch out of the accumulator again
            if undo is not None:
                undo()
            raise


class DeferredVerdict:
    """`validate_args="deferred"`: the status words of one update, copied to pinned host memory behind the kernel.

    The synchronous mode reads the 8 status bytes right after every update, which drains the stream once per update
    (configs[1]: 0.114 ms instead of 0.069 ms per update).  Here the copy is only enqueued; the words are looked at
    when the owner's NEXT update has been launched (or at compute / reset), by which time the copy has long finished,
    so the GPU never waits for the host.  The price: a bad batch raises one call late.  Its counts are still taken
    out of the state first (`undo`), and the exception is the reference's.
    """

    def __init__(self, dev: torch.device, pending: List[Callable[[], None]], num_classes: Optional[int],
                 undo: Optional[Callable[[], None]], host: Tensor, event: "torch.cuda.Event") -> None:
        self.pending, self.num_classes, self.undo = pending, num_classes, undo
        words = _native.scratch(dev).status_words
        self.host = host
        self.host.copy_(words, non_blocking=True)
        words.zero_()  # stream-ordered behind the copy: the next update starts from clean words
        self.event = event
        self.event.record(torch.cuda.current_stream(dev))

    def settle(self) -> None:
        self.event.synchronize()
        err, val = self.host.tolist()
        if not (err or val):
            return
        try:
            if val:
                for chk in self.pending:
                    chk()
            _native.raise_for_bits(err, self.num_classes)
        except Exception:
            if self.undo is not None:
                self.undo()
            raise


def settle_deferred(owner) -> None:
    """Collect the outstanding deferred verdict of a metric, if any (called before compute / reset)."""
    verdict = owner.__dict__.get("_pending_validation")
    if verdict is not None:
        owner.__dict__["_pending_validation"] = None
        verdict.settle()


def _deferred(fn: Callable, t: Tensor, *args) -> bool:
    sink = getattr(_tls, "sink", None)
    if sink is None or not t.is_cuda:
        return False
    sink.append(lambda: fn(t, *args))
    return True


def rollback(state: Tensor, redo: Callable[[Tensor], object]) -> None:
    """Remove one batch from an int64 accumulator that a fused update already touched: its counts are a
    deterministic function of the inputs, so computing them again into zeros and subtracting is exact."""
    tmp = torch.zeros_like(state)
    redo(tmp)
    state -= tmp
    _native.read_status_words(state.device)


def check_binary_values(t: Tensor, name: str, ignore_index: Optional[int] = None, what: str = "") -> None:
    """Every value must be 0, 1 (or ignore_index); one fused elementwise pass + any()."""
    if _deferred(_check_binary_values_now, t, name, ignore_index, what):
        return
    _check_binary_values_now(t, name, ignore_index, what)


def _check_binary_values_now(t: Tensor, name: str, ignore_index: Optional[int] = None, what: str = "") -> None:
    bad = (t != 0) & (t != 1)
    if ignore_index is not None:
        bad &= t != ignore_index
    if bool(bad.any()):
        uniq = torch.unique(t)
        if name == "target":
            raise RuntimeError(
                f"Detected the following values in `target`: {uniq} but expected only"
                f" the following values {[0, 1] if ignore_index is None else [ignore_index]}."
            )
        raise RuntimeError(
            f"Detected the following values in `preds`: {uniq} but expected only"
            f" the following values [0,1] since {what or 'preds'} is a label tensor."
        )


def check_label_count(t: Tensor, name: str, num_classes: int, ignore_index: Optional[int] = None) -> None:
    """The reference rejects tensors with more distinct values than classes (+1 for ignore_index)."""
    if _deferred(_check_label_count_now, t, name, num_classes, ignore_index):
        return
    _check_label_count_now(t, name, num_classes, ignore_index)


def _check_label_count_now(t: Tensor, name: str, num_classes: int, ignore_index: Optional[int] = None) -> None:
    if t.numel() == 0:
        return
    allowed = num_classes + (1 if (ignore_index is not None and name == "target") else 0)
    lo, hi = torch.aminmax(t)
    lo, hi = int(lo), int(hi)
    span_ok = lo >= 0 and hi < num_classes
    if not span_ok and ignore_index is not None and name == "target":
        kept = t[t != ignore_index]
        if kept.numel() == 0:
            return
        lo, hi = (int(v) for v in torch.aminmax(kept))
        span_ok = lo >= 0 and hi < num_classes
    if span_ok:
        return
    n_unique = len(torch.unique(t))
    if n_unique > allowed:
        raise RuntimeError(
            f"Detected more unique values in `{name}` than `num_classes`. Expected only"
            f" {allowed} but found {n_unique} in `{name}`."
        )


def check_multiclass_shapes(preds: Tensor, target: Tensor, num_classes: int, multidim_average: str = "global") -> None:
    if preds.ndim == target.ndim + 1:
        if not preds.is_floating_point():
            raise ValueError("If `preds` have one dimension more than `target`, `preds` should be a float tensor.")
        if preds.shape[1] != num_classes:
            raise ValueError(
                "If `preds` have one dimension more than `target`, `preds.shape[1]` should be"
                " equal to number of classes."
            )
        if preds.shape[2:] != target.shape[1:]:
            raise ValueError(
                "If `preds` have one dimension more than `target`, the shape of `preds` should be"
                " (N, C, ...), and the shape of `target` should be (N, ...)."
            )
        if multidim_average != "global" and preds.ndim < 3:
            raise ValueError(
                "If `preds` have one dimension more than `target`, the shape of `preds` should "
                " at least 3D when multidim_average is set to `samplewise`"
            )
    elif preds.ndim == target.ndim:
        if preds.shape != target.shape:
            raise ValueError(
                "The `preds` and `target` should have the same shape,",
                f" got `preds` with shape={preds.shape} and `target` with shape={target.shape}.",
            )
        if multidim_average != "global" and preds.ndim < 2:
            raise ValueError(
                "When `preds` and `target` have the same shape, the shape of `preds` should "
                " at least 2D when multidim_average is set to `samplewise`"
            )
    else:
        raise ValueError(
            "Either `preds` and `target` both should have the (same) shape (N, ...), or `target` should be (N, ...)"
            " and `preds` should be (N, C, ...)."
        )


def check_multilabel_shapes(preds: Tensor, target: Tensor, num_labels: int) -> None:
    _check_same_shape(preds, target)
    if preds.shape[1] != num_labels:
        raise ValueError(
            "Expected both `target.shape[1]` and `preds.shape[1]` to be equal to the number of labels"
            f" but got {preds.shape[1]} and expected {num_labels}"
        )


def check_curve_target_dtype(target: Tensor) -> None:
    if target.is_floating_point():
        raise ValueError(
            "Expected argument `target` to be an int or long tensor with ground truth labels"
            f" but got tensor with dtype {target.dtype}"
        )


def check_float_preds(preds: Tensor) -> None:
    if not preds.is_floating_point():
        raise ValueError(
            "Expected argument `preds` to be an floating tensor with probability/logit scores,"
            f" but got tensor with dtype {preds.dtype}"
        )


def in_unit_interval(preds: Tensor) -> bool:
    """`torch.all((preds >= 0) * (preds <= 1))` of the reference's `_format` steps (e.g. precision_recall_curve.py:183)
    in one pass: min >= 0 and max <= 1.  NaN propagates through aminmax and fails both comparisons, like the
    elementwise form; an empty tensor passes, like `torch.all` of nothing."""
    if preds.numel() == 0:
        return True
    lo, hi = torch.aminmax(preds)
    return bool((lo >= 0) & (hi <= 1))
