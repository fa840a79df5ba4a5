"""ctypes wrapper of oracle/liboracle.so (C/OpenMP restatement; test infrastructure only, see oracle.c)."""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def load():
    global _lib
    if _lib is None:
        path = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle` (or __graft_entry__.build())")
        _lib = ctypes.CDLL(path)
        _lib.orc_num_threads.restype = ctypes.c_int
        _lib.orc_confmat_update.restype = ctypes.c_int
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def num_threads() -> int:
    return load().orc_num_threads()


def set_num_threads(n: int) -> None:
    load().orc_set_num_threads(ctypes.c_int(int(n)))


def confmat_update(preds, target, num_classes, out=None):
    preds = np.ascontiguousarray(preds, dtype=np.int64)
    target = np.ascontiguousarray(target, dtype=np.int64)
    if out is None:
        out = np.zeros((num_classes, num_classes), dtype=np.int64)
    rc = load().orc_confmat_update(_p(preds), _p(target), ctypes.c_int64(preds.size), ctypes.c_int64(num_classes), _p(out))
    if rc == -1:
        raise RuntimeError("label key outside [0, C*C)")
    if rc != 0:
        raise MemoryError
    return out


def binned_update_loop(preds, target, thresholds, out=None):
    preds = np.ascontiguousarray(preds, dtype=np.float32)
    target = np.ascontiguousarray(target, dtype=np.int64)
    thr = np.ascontiguousarray(thresholds, dtype=np.float32)
    if out is None:
        out = np.zeros((thr.size, 2, 2), dtype=np.int64)
    load().orc_binned_update_loop(_p(preds), _p(target), ctypes.c_int64(preds.size), _p(thr), ctypes.c_int(thr.size),
                                  _p(out))
    return out


def multilabel_stats(preds, target, threshold=0.5, ignore_index=None, out=None):
    preds = np.ascontiguousarray(preds, dtype=np.float32)
    target = np.ascontiguousarray(target, dtype=np.int64)
    n, L = preds.shape
    if out is None:
        out = np.zeros((4, L), dtype=np.int64)
    load().orc_multilabel_stats(_p(preds), _p(target), ctypes.c_int64(n), ctypes.c_int64(L), ctypes.c_float(threshold),
                                ctypes.c_int(ignore_index is not None),
                                ctypes.c_int64(0 if ignore_index is None else ignore_index), _p(out))
    return out
