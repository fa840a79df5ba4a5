"""The C-ABI library loads on a CPU-only box and exports every symbol include/tm_b200.h declares."""
import ctypes
import os
import re

import pytest

from conftest import ROOT
from torchmetrics_b200 import _native


def _declared_in_header():
    text = open(os.path.join(ROOT, "include", "tm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(tmb_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert os.path.exists(_native.lib_path()), "run `python -c 'import __graft_entry__ as g; g.build()'` first"


def test_exports_every_declared_symbol():
    names = _declared_in_header()
    assert len(names) >= 8
    lib = ctypes.CDLL(_native.lib_path())
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/tm_b200.h but not exported"


def test_binding_covers_header():
    assert set(_native.declared_symbols()) == set(_declared_in_header())
    lib = _native.load()
    assert lib.tmb_abi_version() == 1


def test_cpu_tensor_at_kernel_seam_raises():
    import torch

    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        _native.confmat_update_(torch.zeros(4, dtype=torch.int64), torch.zeros(3, dtype=torch.int64),
                                torch.zeros(3, dtype=torch.int64), 2)
