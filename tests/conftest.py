import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


class Case(dict):
    """One golden case: arrays by name plus `.meta` (the parameters it was generated with)."""

    meta: dict


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    metas = json.loads(str(z["__meta__"]))
    cases = []
    for i, m in enumerate(metas):
        c = Case()
        c.meta = m
        pre = f"c{i}__"
        for k in z.files:
            if k.startswith(pre):
                c[k[len(pre):]] = z[k]
        cases.append(c)
    return cases


def case_id(c):
    return "-".join(f"{k}={v}" for k, v in c.meta.items())


@pytest.fixture(scope="session")
def golden():
    return load_golden


@pytest.fixture(autouse=True)
def _fresh_device_status(request):
    """GPU tests start from clean status words: findings left behind by an earlier test (an unvalidated call on bad
    labels, say) belong to that test, not to the next one's `== (0, 0)` assertion."""
    if request.node.get_closest_marker("gpu") is not None:
        import torch

        if torch.cuda.is_available():
            from torchmetrics_b200 import _native

            _native.read_status_words(torch.device("cuda:0"))
    yield
