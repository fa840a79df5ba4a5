"""ctypes binding of libtm_b200.so (C ABI declared in include/tm_b200.h).

This is the only door from the Python host layer to the CUDA kernels.  There is no CPU or eager-torch
fallback behind it: if the shared library is missing, or a tensor that reaches a kernel seam is not a CUDA
tensor, the call raises.
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

# TMB_B200_LIB: development hook, load an alternative build of the same library (kernel tuning experiments)
_LIB_PATH = os.environ.get("TMB_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "csrc",
                                                          "libtm_b200.so")

TMB_OK = 0
ERR_NAMES = {-1: "TMB_ERR_ARG", -2: "TMB_ERR_ALIGN", -3: "TMB_ERR_CUDA", -4: "TMB_ERR_TOO_LARGE"}

STATUS_LABEL_RANGE = 1
STATUS_THRESHOLD_ORDER = 2
STATUS_OUT_OF_UNIT_RANGE = 4
# fused-validation findings, reported in the second status word (include/tm_b200.h)
STATUS_TARGET_VALUES = 8
STATUS_PREDS_VALUES = 16
STATUS_PEER_TIMEOUT = 32

PROB_AS_IS, PROB_AUTO, PROB_SIGMOID, PROB_CHECK = 0, 1, 2, 3

_DTYPES = {
    torch.int64: 0,
    torch.int32: 1,
    torch.uint8: 4,
    torch.bool: 4,
    torch.float32: 8,
    torch.float16: 9,
    torch.bfloat16: 10,
    torch.float64: 11,
}

_c = ctypes
_SIGNATURES = {
    "tmb_abi_version": (_c.c_int, []),
    "tmb_last_cuda_error": (_c.c_char_p, []),
    "tmb_sm_count": (_c.c_int, []),
    "tmb_confmat_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_int,
                                      _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_int, _c.c_void_p]),
    "tmb_confmat_workspace_bytes": (_c.c_size_t, [_c.c_int64, _c.c_int64]),
    "tmb_confmat_update_ws": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_int,
                                         _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "tmb_bincount_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                       _c.c_void_p]),
    "tmb_match_count_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int,
                                          _c.c_int64, _c.c_void_p, _c.c_void_p]),
    "tmb_class_stats_workspace_bytes": (_c.c_size_t, [_c.c_int64]),
    "tmb_class_stats_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_int,
                                          _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p, _c.c_void_p]),
    "tmb_class_stats_topk_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                               _c.c_int64, _c.c_int, _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_size_t,
                                               _c.c_void_p, _c.c_void_p]),
    "tmb_class_stats_argmax_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                                 _c.c_int, _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_size_t,
                                                 _c.c_void_p, _c.c_void_p]),
    "tmb_rows_argmax": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p]),
    "tmb_binned_curve_workspace_bytes": (_c.c_size_t, [_c.c_int, _c.c_int64]),
    "tmb_binned_curve_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                           _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int64, _c.c_int,
                                           _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p, _c.c_void_p]),
    "tmb_binned_curve_update_static": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                                  _c.c_int, _c.c_void_p, _c.c_int, _c.c_int, _c.c_int64, _c.c_int,
                                                  _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "tmb_exact_curve_workspace_bytes": (_c.c_size_t, [_c.c_int64]),
    "tmb_exact_curve": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                   _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "tmb_exact_auroc_ap": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_void_p,
                                      _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "tmb_exact_auroc_ap_offset": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_int64,
                                             _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_size_t, _c.c_void_p]),
    "tmb_range_partition_workspace_bytes": (_c.c_size_t, [_c.c_int64, _c.c_int]),
    "tmb_range_partition": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64, _c.c_void_p,
                                       _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_size_t,
                                       _c.c_void_p]),
    "tmb_oneshot_buffer_bytes": (_c.c_size_t, []),
    "tmb_oneshot_allreduce": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int64,
                                         _c.c_int, _c.c_int, _c.c_uint32, _c.c_void_p, _c.c_void_p]),
    "tmb_oneshot_allreduce_packed": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_int, _c.c_void_p, _c.c_void_p, _c.c_int,
                                                _c.c_void_p, _c.c_int, _c.c_int, _c.c_uint32, _c.c_void_p,
                                                _c.c_void_p]),
    "tmb_stat_scores_workspace_bytes": (_c.c_size_t, [_c.c_int64]),
    "tmb_stat_scores_update": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                          _c.c_double, _c.c_int, _c.c_int, _c.c_int64, _c.c_int, _c.c_void_p,
                                          _c.c_void_p, _c.c_size_t, _c.c_void_p, _c.c_void_p]),
    "tmb_stat_scores_rows": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                        _c.c_double, _c.c_int, _c.c_int, _c.c_int64, _c.c_void_p, _c.c_void_p]),
    "tmb_class_stats_rows": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                        _c.c_int64, _c.c_int, _c.c_int64, _c.c_void_p, _c.c_void_p, _c.c_void_p]),
    "tmb_class_stats_rows_topk": (_c.c_int, [_c.c_void_p, _c.c_int, _c.c_void_p, _c.c_int, _c.c_int64, _c.c_int64,
                                             _c.c_int, _c.c_int64, _c.c_int, _c.c_int64, _c.c_void_p, _c.c_void_p,
                                             _c.c_void_p]),
    "tmb_stat_reduce": (_c.c_int, [_c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_void_p, _c.c_int64, _c.c_int64, _c.c_int,
                                   _c.c_int, _c.c_int, _c.c_double, _c.c_void_p, _c.c_void_p]),
}

_lib = None
_lock = threading.Lock()


class NativeLibraryError(RuntimeError):
    pass


def lib_path() -> str:
    return _LIB_PATH


def declared_symbols():
    return sorted(_SIGNATURES)


def load():
    """dlopen libtm_b200.so once; raises NativeLibraryError (never falls back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(_LIB_PATH):
            raise NativeLibraryError(
                f"{_LIB_PATH} not found. Build it with `python -c 'import __graft_entry__ as g; g.build()'` or "
                "`make -C torchmetrics_b200/csrc`. torchmetrics_b200 has no CPU or eager fallback for its kernels."
            )
        lib = ctypes.CDLL(_LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the .so is stale
            fn.restype = res
            fn.argtypes = args
        if lib.tmb_abi_version() != 1:
            raise NativeLibraryError(f"ABI version mismatch: library reports {lib.tmb_abi_version()}, binding expects 1")
        _lib = lib
    return _lib


def _check(rc: int, what: str) -> None:
    if rc == TMB_OK:
        return
    msg = ERR_NAMES.get(rc, str(rc))
    if rc == -3:
        msg += ": " + load().tmb_last_cuda_error().decode()
    raise RuntimeError(f"{what} failed with {msg}")


def require_cuda(*tensors: Tensor) -> torch.device:
    dev = None
    for t in tensors:
        if not t.is_cuda:
            raise RuntimeError(
                "torchmetrics_b200 kernels run on CUDA tensors only (there is no CPU fallback); got a tensor on "
                f"{t.device}. Move preds/target and the metric to a CUDA device."
            )
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"Expected all tensors on the same device, got {dev} and {t.device}")
    return dev


def dtype_code(t: Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype} at a kernel seam") from None


def as_kernel_labels(t: Tensor) -> Tensor:
    """Contiguous, 32-byte aligned label tensor in one of the dtypes the kernels read (int64/int32/uint8/bool)."""
    if t.dtype in (torch.int16, torch.int8):
        t = t.to(torch.int32)
    elif t.dtype not in (torch.int64, torch.int32, torch.uint8, torch.bool):
        t = t.to(torch.int64)
    return _aligned(t)


def as_kernel_scores(t: Tensor) -> Tensor:
    if t.dtype not in (torch.float32, torch.float16, torch.bfloat16, torch.float64):
        t = t.to(torch.float32)
    return _aligned(t)


def _aligned(t: Tensor) -> Tensor:
    t = t.contiguous()
    if t.data_ptr() % 32 != 0:
        t = t.clone(memory_format=torch.contiguous_format)
    return t


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream_ptr(dev: torch.device) -> int:
    """cudaStream_t of torch's current stream on `dev` (the raw getter skips building a Stream object: this is called
    twice per update)."""
    if _raw_stream is not None:
        return _raw_stream(dev.index if dev.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(dev).cuda_stream


# ---------------------------------------------------------------------------------------------
# per (device, stream) scratch: a status word and self-cleaning kernel workspaces
# ---------------------------------------------------------------------------------------------
class _Scratch:
    def __init__(self, dev: torch.device) -> None:
        self.status = torch.zeros(8, dtype=torch.int32, device=dev)
        self.status_words = self.status[:2]  # [errors, fused-validation findings]
        self.ws: Dict[str, Tensor] = {}

    def workspace(self, kind: str, nbytes: int, dev: torch.device) -> Tensor:
        w = self.ws.get(kind)
        if w is None or w.numel() < nbytes:
            # zero on first use; kernels leave it zero behind them
            w = torch.zeros(max(nbytes, 4096), dtype=torch.uint8, device=dev)
            self.ws[kind] = w
        return w


_scratch: Dict[Tuple[int, int], _Scratch] = {}


def scratch(dev: torch.device) -> _Scratch:
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), _stream_ptr(dev))
    s = _scratch.get(key)
    if s is None:
        s = _Scratch(dev)
        _scratch[key] = s
    return s


def read_status_words(dev: torch.device) -> Tuple[int, int]:
    """Synchronising read-and-clear of (error bits, fused-validation findings) for the current stream: one 8-byte
    device-to-host copy."""
    s = scratch(dev)
    err, val = s.status[:2].tolist()
    if err or val:
        s.status.zero_()
    return err, val


def read_status(dev: torch.device) -> int:
    """Synchronising read of the status words that returns and clears the error bits only: fused-validation
    findings (word 1) stay where they are for the `FusedValidation.finish` that follows."""
    s = scratch(dev)
    err = int(s.status[:2].tolist()[0])
    if err:
        s.status[0].zero_()
    return err


def raise_on_status(dev: torch.device, num_classes: Optional[int] = None) -> None:
    raise_for_bits(read_status(dev), num_classes)


def raise_for_bits(v: int, num_classes: Optional[int] = None) -> None:
    if v & STATUS_LABEL_RANGE:
        raise RuntimeError(
            "Detected label values outside the expected range in `preds`/`target`"
            + (f" (expected values in [0, {num_classes - 1}])" if num_classes else "")
            + "; the reference raises on the same input when it reshapes its bincount."
        )
    if v & STATUS_PEER_TIMEOUT:
        raise RuntimeError("one-shot all-reduce: a peer rank did not arrive within the time limit; the synced state "
                           "is invalid (set TM_B200_ONESHOT=0 to use NCCL for small states)")
    if v & STATUS_THRESHOLD_ORDER:
        raise RuntimeError("thresholds handed to the binned-curve kernel must be sorted ascending")


# ---------------------------------------------------------------------------------------------
# thin typed wrappers (tensor in, state tensor updated in place)
# ---------------------------------------------------------------------------------------------
def confmat_update_(confmat: Tensor, preds: Tensor, target: Tensor, num_classes: int,
                    ignore_index: Optional[int] = None, algo: int = 0) -> None:
    """confmat (C, C) int64 += confusion matrix of the label tensors (K1)."""
    dev = require_cuda(confmat, preds, target)
    assert confmat.dtype == torch.int64 and confmat.is_contiguous() and confmat.numel() == num_classes * num_classes
    preds, target = as_kernel_labels(preds.reshape(-1)), as_kernel_labels(target.reshape(-1))
    assert preds.numel() == target.numel()
    n = preds.numel()
    cc = num_classes * num_classes
    # large matrices and large batches: bucket partition + shared-memory histograms (algo 6); else direct kernels
    if algo == 6 or (algo == 0 and cc > 51200 and cc <= 4096 * 1024 and (1 << 23) <= n < (1 << 32)):
        lib = load()
        sc = scratch(dev)
        ws = sc.workspace("confmat", lib.tmb_confmat_workspace_bytes(n, num_classes), dev)
        with torch.cuda.device(dev):
            rc = lib.tmb_confmat_update_ws(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                           n, num_classes, ignore_index is not None,
                                           0 if ignore_index is None else int(ignore_index), confmat.data_ptr(),
                                           sc.status.data_ptr(), ws.data_ptr(), ws.numel(), _stream_ptr(dev))
        _check(rc, "tmb_confmat_update_ws")
        return
    with torch.cuda.device(dev):
        rc = load().tmb_confmat_update(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                       preds.numel(), num_classes, ignore_index is not None,
                                       0 if ignore_index is None else int(ignore_index), confmat.data_ptr(),
                                       scratch(dev).status.data_ptr(), algo, _stream_ptr(dev))
    _check(rc, "tmb_confmat_update")


def bincount_update_(out: Tensor, x: Tensor) -> None:
    """out (M,) int64 += histogram of the integer tensor x."""
    dev = require_cuda(out, x)
    assert out.dtype == torch.int64 and out.is_contiguous()
    x = as_kernel_labels(x.reshape(-1))
    with torch.cuda.device(dev):
        rc = load().tmb_bincount_update(x.data_ptr(), dtype_code(x), x.numel(), out.numel(), out.data_ptr(),
                                        scratch(dev).status.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_bincount_update")


def match_count_update_(out2: Tensor, preds: Tensor, target: Tensor, ignore_index: Optional[int] = None) -> None:
    dev = require_cuda(out2, preds, target)
    assert out2.dtype == torch.int64 and out2.numel() == 2 and out2.is_contiguous()
    preds, target = as_kernel_labels(preds.reshape(-1)), as_kernel_labels(target.reshape(-1))
    with torch.cuda.device(dev):
        rc = load().tmb_match_count_update(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                           preds.numel(), ignore_index is not None,
                                           0 if ignore_index is None else int(ignore_index), out2.data_ptr(),
                                           _stream_ptr(dev))
    _check(rc, "tmb_match_count_update")


def class_stats_update_(out: Tensor, preds: Tensor, target: Tensor, num_classes: int,
                        ignore_index: Optional[int] = None) -> None:
    """out (4, C) int64 [tp|fp|tn|fn] += per-class statistics of multiclass label tensors (K1c), no C x C matrix."""
    dev = require_cuda(out, preds, target)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == 4 * num_classes
    preds, target = as_kernel_labels(preds.reshape(-1)), as_kernel_labels(target.reshape(-1))
    assert preds.numel() == target.numel()
    lib = load()
    sc = scratch(dev)
    ws = sc.workspace("class_stats", lib.tmb_class_stats_workspace_bytes(num_classes), dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_class_stats_update(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                        preds.numel(), num_classes, ignore_index is not None,
                                        0 if ignore_index is None else int(ignore_index), out.data_ptr(),
                                        ws.data_ptr(), ws.numel(), sc.status.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_class_stats_update")


def class_stats_topk_update_(out: Tensor, topk: Tensor, target: Tensor, num_classes: int,
                             ignore_index: Optional[int] = None) -> None:
    """out (4, C) int64 [tp|fp|tn|fn] += per-class statistics with top-k predictions: topk (n, k) class ids."""
    dev = require_cuda(out, topk, target)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == 4 * num_classes
    assert topk.ndim == 2
    k = topk.shape[1]
    topk, target = as_kernel_labels(topk), as_kernel_labels(target.reshape(-1))
    assert topk.shape[0] == target.numel()
    lib = load()
    sc = scratch(dev)
    ws = sc.workspace("class_stats", lib.tmb_class_stats_workspace_bytes(num_classes), dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_class_stats_topk_update(topk.data_ptr(), dtype_code(topk), target.data_ptr(), dtype_code(target),
                                             target.numel(), num_classes, k, ignore_index is not None,
                                             0 if ignore_index is None else int(ignore_index), out.data_ptr(),
                                             ws.data_ptr(), ws.numel(), sc.status.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_class_stats_topk_update")


def class_stats_argmax_update_(out: Tensor, scores: Tensor, target: Tensor, num_classes: int,
                               ignore_index: Optional[int] = None) -> None:
    """out (4, C) int64 [tp|fp|tn|fn] += per-class statistics of (n, C) scores: argmax fused into the count (K1c)."""
    dev = require_cuda(out, scores, target)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == 4 * num_classes
    assert scores.ndim == 2 and scores.shape[1] == num_classes
    scores, target = as_kernel_scores(scores), as_kernel_labels(target.reshape(-1))
    assert scores.shape[0] == target.numel()
    lib = load()
    sc = scratch(dev)
    ws = sc.workspace("class_stats", lib.tmb_class_stats_workspace_bytes(num_classes), dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_class_stats_argmax_update(scores.data_ptr(), dtype_code(scores), target.data_ptr(),
                                               dtype_code(target), target.numel(), num_classes,
                                               ignore_index is not None,
                                               0 if ignore_index is None else int(ignore_index), out.data_ptr(),
                                               ws.data_ptr(), ws.numel(), sc.status.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_class_stats_argmax_update")


def rows_argmax(scores: Tensor) -> Tensor:
    """int32 class id of the first maximum of every row of (n, C) scores (torch.argmax(dim=1) semantics)."""
    dev = require_cuda(scores)
    assert scores.ndim == 2
    scores = as_kernel_scores(scores)
    n, C = scores.shape
    out = torch.empty(n, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = load().tmb_rows_argmax(scores.data_ptr(), dtype_code(scores), n, C, out.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_rows_argmax")
    return out


def binned_curve_update_(confmat: Tensor, preds: Tensor, target: Tensor, thresholds: Tensor, num_columns: int = 1,
                         target_is_class_index: bool = False, ignore_index: Optional[int] = None,
                         prob_mode: int = PROB_AUTO) -> None:
    """confmat (T[, C], 2, 2) int64 += multi-threshold confusion matrix (K3). thresholds must be ascending."""
    dev = require_cuda(confmat, preds, target, thresholds)
    T = thresholds.numel()
    assert confmat.dtype == torch.int64 and confmat.is_contiguous() and confmat.numel() == T * num_columns * 4
    assert thresholds.dtype == torch.float32 and thresholds.is_contiguous()
    preds = as_kernel_scores(preds.reshape(-1))
    target = as_kernel_labels(target.reshape(-1))
    n_rows = preds.numel() // num_columns
    lib = load()
    nbytes = lib.tmb_binned_curve_workspace_bytes(T, num_columns)
    sc = scratch(dev)
    ws = sc.workspace("binned", nbytes, dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_binned_curve_update(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                         n_rows, num_columns, bool(target_is_class_index), thresholds.data_ptr(), T,
                                         ignore_index is not None, 0 if ignore_index is None else int(ignore_index),
                                         prob_mode, confmat.data_ptr(), ws.data_ptr(), ws.numel(),
                                         sc.status.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_binned_curve_update")


def stat_scores_update_(out: Tensor, preds: Tensor, target: Tensor, num_labels: int, threshold: float = 0.5,
                        ignore_index: Optional[int] = None, prob_mode: int = PROB_AUTO,
                        range_check_ignored: bool = True, report_values: bool = True) -> None:
    """out (4, L) int64 [tp|fp|tn|fn] += per-label statistics of (n_rows, L) preds/target (K2).
    `report_values`: let the kernel note target/label-preds values outside {0, 1} in the validation status word."""
    dev = require_cuda(out, preds, target)
    assert out.dtype == torch.int64 and out.is_contiguous() and out.numel() == 4 * num_labels
    preds = as_kernel_scores(preds.reshape(-1)) if preds.is_floating_point() else as_kernel_labels(preds.reshape(-1))
    target = as_kernel_labels(target.reshape(-1))
    n_rows = preds.numel() // num_labels
    lib = load()
    nbytes = lib.tmb_stat_scores_workspace_bytes(num_labels)
    sc = scratch(dev)
    ws = sc.workspace("stat", nbytes, dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_stat_scores_update(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                        n_rows, num_labels, float(threshold), prob_mode, ignore_index is not None,
                                        0 if ignore_index is None else int(ignore_index), bool(range_check_ignored),
                                        out.data_ptr(), ws.data_ptr(), ws.numel(),
                                        sc.status.data_ptr() if report_values else None, _stream_ptr(dev))
    _check(rc, "tmb_stat_scores_update")


def stat_scores_rows(preds: Tensor, target: Tensor, width: int, threshold: float = 0.5,
                     ignore_index: Optional[int] = None, sigmoid: bool = False) -> Tensor:
    """(4, n_rows) int64 [tp|fp|tn|fn] of every row of (n_rows, width) preds / target (K2, row-wise form)."""
    dev = require_cuda(preds, target)
    preds = as_kernel_scores(preds.reshape(-1)) if preds.is_floating_point() else as_kernel_labels(preds.reshape(-1))
    target = as_kernel_labels(target.reshape(-1))
    assert preds.numel() == target.numel()
    n_rows = preds.numel() // width if width else 0
    out = torch.empty(4, n_rows, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        rc = load().tmb_stat_scores_rows(preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target),
                                         n_rows, width, float(threshold), PROB_SIGMOID if sigmoid else PROB_AS_IS,
                                         ignore_index is not None, 0 if ignore_index is None else int(ignore_index),
                                         out.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_stat_scores_rows")
    return out


def class_stats_rows(preds: Tensor, target: Tensor, num_classes: int, ignore_index: Optional[int] = None,
                     top_k: int = 1) -> Tensor:
    """(4, n_rows, C) int64 [tp|fp|tn|fn] per sample and class of (n_rows, width) label tensors (K1c, row-wise).
    With `top_k` > 1, `preds` is (n_rows, width, top_k): the class ids of the k highest scores of every element."""
    dev = require_cuda(preds, target)
    assert target.ndim == 2 and tuple(preds.shape) == (tuple(target.shape) + ((top_k,) if top_k > 1 else ()))
    n_rows, width = target.shape
    preds, target = as_kernel_labels(preds), as_kernel_labels(target)
    out = torch.empty(4, n_rows, num_classes, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        if top_k > 1:
            rc = load().tmb_class_stats_rows_topk(preds.data_ptr(), dtype_code(preds), target.data_ptr(),
                                                  dtype_code(target), n_rows, width, top_k, num_classes,
                                                  ignore_index is not None,
                                                  0 if ignore_index is None else int(ignore_index), out.data_ptr(),
                                                  scratch(dev).status.data_ptr(), _stream_ptr(dev))
        else:
            rc = load().tmb_class_stats_rows(preds.data_ptr(), dtype_code(preds), target.data_ptr(),
                                             dtype_code(target), n_rows, width, num_classes, ignore_index is not None,
                                             0 if ignore_index is None else int(ignore_index), out.data_ptr(),
                                             scratch(dev).status.data_ptr(), _stream_ptr(dev))
    if rc == -4:
        raise RuntimeError("tmb_class_stats_rows failed with TMB_ERR_TOO_LARGE: " + load().tmb_last_cuda_error().decode())
    _check(rc, "tmb_class_stats_rows")
    return out


STAT_PRECISION, STAT_RECALL, STAT_SPECIFICITY, STAT_ACCURACY, STAT_FBETA = range(5)
_AVERAGE_CODES = {None: 0, "none": 0, "binary": 0, "micro": 1, "macro": 2, "weighted": 3}


def stat_reduce_supported(tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str]) -> bool:
    """True when `stat_reduce` serves these states: int64 CUDA tensors of one shape, a known averaging mode."""
    return (tp.is_cuda and average in _AVERAGE_CODES and tp.dtype == torch.int64 and fp.dtype == torch.int64
            and tn.dtype == torch.int64 and fn.dtype == torch.int64 and tp.ndim <= 2
            and tp.shape == fp.shape == tn.shape == fn.shape and fp.device == tn.device == fn.device == tp.device)


def stat_reduce(kind: int, tp: Tensor, fp: Tensor, tn: Tensor, fn: Tensor, average: Optional[str],
                multilabel: bool = False, beta: float = 1.0) -> Tensor:
    """float32 score of the tp / fp / tn / fn states in one launch (`tmb_stat_reduce`).  The last dimension holds the
    classes / labels; "micro", "macro" and "weighted" reduce it, None / "none" / "binary" keep the shape."""
    dev = tp.device
    code = _AVERAGE_CODES[average]
    labels = tp.shape[-1] if tp.ndim else 1
    rows = tp.numel() // labels if labels else 0
    # micro reduces the labels inside the kernel; macro / weighted come back as terms that torch.sum adds up
    out = torch.empty(tp.shape[:-1] if code == 1 else tp.shape, dtype=torch.float32, device=dev)
    if out.numel() == 0:
        return out.sum(-1) if code >= 2 and tp.ndim else out
    if labels == 0:  # micro over no labels: 0 / 1
        return out.zero_()
    tp, fp, tn, fn = tp.contiguous(), fp.contiguous(), tn.contiguous(), fn.contiguous()
    with torch.cuda.device(dev):
        rc = load().tmb_stat_reduce(tp.data_ptr(), fp.data_ptr(), tn.data_ptr(), fn.data_ptr(), rows, labels, kind,
                                    code, bool(multilabel), float(beta), out.data_ptr(), _stream_ptr(dev))
    _check(rc, "tmb_stat_reduce")
    return out.sum(-1) if code >= 2 else out


def exact_curve(preds: Tensor, target: Tensor, pos_label: int = 1):
    """(fps int64, tps int64, thresholds float32) at every distinct score, descending (K4)."""
    dev = require_cuda(preds, target)
    preds = _aligned(preds.detach().reshape(-1).to(torch.float32))
    target = as_kernel_labels(target.reshape(-1))
    n = preds.numel()
    assert target.numel() == n
    lib = load()
    fps = torch.empty(n, dtype=torch.int64, device=dev)
    tps = torch.empty(n, dtype=torch.int64, device=dev)
    thr = torch.empty(n, dtype=torch.float32, device=dev)
    counts = torch.zeros(2, dtype=torch.int64, device=dev)
    ws = torch.empty(lib.tmb_exact_curve_workspace_bytes(n), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_exact_curve(preds.data_ptr(), target.data_ptr(), dtype_code(target), n, int(pos_label),
                                 fps.data_ptr(), tps.data_ptr(), thr.data_ptr(), counts.data_ptr(), ws.data_ptr(),
                                 ws.numel(), _stream_ptr(dev))
    _check(rc, "tmb_exact_curve")
    d = int(counts[1].item())  # the reference syncs at the same point (torch.where)
    return fps[:d], tps[:d], thr[:d]


def exact_auroc_ap(preds: Tensor, target: Tensor, pos_label: int = 1, tp_offset: int = 0,
                   fp_offset: int = 0, as_tensor: bool = False):
    """(sum_k dFP_k (TP_k + TP_{k-1}),  sum_k dTP_k TP_k / (TP_k + FP_k),  positives,  distinct scores) of the exact
    curve, reduced on the device (K4b): the curve itself is never written.  One 32-byte read back.
    `tp_offset` / `fp_offset`: counts ranked above this input when it is one shard of a score-range partition.
    `as_tensor`: return the float64[4] device tensor instead (no host sync)."""
    dev = require_cuda(preds, target)
    preds = _aligned(preds.detach().reshape(-1).to(torch.float32))
    target = as_kernel_labels(target.reshape(-1))
    n = preds.numel()
    assert target.numel() == n
    lib = load()
    out4 = torch.empty(4, dtype=torch.float64, device=dev)
    ws = torch.empty(lib.tmb_exact_curve_workspace_bytes(n), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_exact_auroc_ap_offset(preds.data_ptr(), target.data_ptr(), dtype_code(target), n, int(pos_label),
                                           int(tp_offset), int(fp_offset), out4.data_ptr(), ws.data_ptr(), ws.numel(),
                                           _stream_ptr(dev))
    _check(rc, "tmb_exact_auroc_ap_offset")
    if as_tensor:
        return out4
    a, b, c, d = out4.tolist()
    return a, b, c, d


def range_partition(preds: Tensor, target: Tensor, splitters: Tensor, pos_label: int = 1):
    """(scores grouped by score range, positives as uint8, int64 (2, R) [samples | positives] per range) - range 0
    holds the highest scores (tmb_range_partition)."""
    dev = require_cuda(preds, target, splitters)
    preds = preds.detach().reshape(-1).to(torch.float32).contiguous()
    target = as_kernel_labels(target.reshape(-1))
    splitters = splitters.to(torch.float32).contiguous()
    n, R = preds.numel(), splitters.numel() + 1
    lib = load()
    out_p = torch.empty(n, dtype=torch.float32, device=dev)
    out_t = torch.empty(n, dtype=torch.uint8, device=dev)
    totals = torch.empty(2, R, dtype=torch.int64, device=dev)
    ws = torch.empty(max(lib.tmb_range_partition_workspace_bytes(n, R), 256), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        rc = lib.tmb_range_partition(preds.data_ptr(), target.data_ptr(), dtype_code(target), n, int(pos_label),
                                     splitters.data_ptr(), R, out_p.data_ptr(), out_t.data_ptr(), totals.data_ptr(),
                                     ws.data_ptr(), ws.numel(), _stream_ptr(dev))
    _check(rc, "tmb_range_partition")
    return out_p, out_t, totals


def binned_curve_update_static_(confmat: Tensor, preds: Tensor, target: Tensor, thresholds: Tensor,
                                num_columns: int = 1, target_is_class_index: bool = False,
                                ignore_index: Optional[int] = None, prob_mode: int = PROB_AUTO) -> None:
    """`binned_curve_update_` for callers that must not allocate device memory: the kernel scratch is the fixed
    64 KiB array inside libtm_b200 instead of a torch-allocated workspace (tmb_binned_curve_update_static).

    The inputs must already be contiguous, aligned tensors of a kernel dtype, because this wrapper does not copy or
    cast (that would allocate); problems whose counters do not fit the array raise.  Calls on one device share the
    array, so do not let two of them overlap on the GPU.

    One allocation remains on the Python side: the 32-byte status tensor this binding keeps per (device, stream) is
    created by torch the first time ANY kernel wrapper runs on that stream.  Make one warm-up call (this function or
    any other update) on the stream before entering the region that must not allocate; a C caller passes its own
    `status` pointer and has no such step.
    """
    dev = require_cuda(confmat, preds, target, thresholds)
    T = thresholds.numel()
    assert confmat.dtype == torch.int64 and confmat.is_contiguous() and confmat.numel() == T * num_columns * 4
    assert thresholds.dtype == torch.float32 and thresholds.is_contiguous()
    for t in (preds, target):
        if not t.is_contiguous() or t.data_ptr() % 32 != 0:
            raise ValueError("the allocation-free entry point needs contiguous, 32-byte aligned inputs")
    n_rows = preds.numel() // num_columns
    with torch.cuda.device(dev):
        rc = load().tmb_binned_curve_update_static(
            preds.data_ptr(), dtype_code(preds), target.data_ptr(), dtype_code(target), n_rows, num_columns,
            bool(target_is_class_index), thresholds.data_ptr(), T, ignore_index is not None,
            0 if ignore_index is None else int(ignore_index), prob_mode, confmat.data_ptr(),
            scratch(dev).status.data_ptr(), _stream_ptr(dev))
    if rc == -4:
        raise RuntimeError("tmb_binned_curve_update_static failed with TMB_ERR_TOO_LARGE: "
                           + load().tmb_last_cuda_error().decode())
    _check(rc, "tmb_binned_curve_update_static")
