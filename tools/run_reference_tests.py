"""Run the reference's own host-side test files against torchmetrics_b200.

The reference's `tests/unittests/bases/*` import `torchmetrics`; this script puts an alias package of that name
first on the path whose modules ARE the torchmetrics_b200 modules, plus the 4-symbol `lightning_utilities` stand-in
the golden-vector generator already uses, and runs pytest from the reference's tests directory (never written to).
Only usable where /root/reference exists (the build container); results are recorded in DESIGN.md section 4.

    python tools/run_reference_tests.py [pytest args / test files relative to tests/unittests]

Tests that build classification metrics on CPU tensors fail by design here (the kernels have no CPU path); the
host-logic tests (state machine, sync protocol, collections, composition, hashing, pickling) are the target.
"""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.environ.get("TM_REFERENCE", "/root/reference")
ALIAS = os.path.join(ROOT, "gpurun_out", "_alias")

ALIAS_INIT = '''
"""`torchmetrics` == torchmetrics_b200, module for module."""
import importlib
import pkgutil
import sys
import types

import torchmetrics_b200 as _pkg

_me = sys.modules[__name__]
for _m in pkgutil.walk_packages(_pkg.__path__, "torchmetrics_b200."):
    if "csrc" in _m.name:
        continue
    _mod = importlib.import_module(_m.name)
    sys.modules["torchmetrics" + _m.name[len("torchmetrics_b200"):]] = _mod
sys.modules["torchmetrics"] = _pkg

# out of scope (SURVEY.md section 2): the one regression metric test_metric.py imports at module level
import pytest
_reg = types.ModuleType("torchmetrics.regression")
class PearsonCorrCoef(_pkg.Metric):
    def __init__(self, *a, **k):
        pytest.skip("regression metrics are outside the hot path")
_reg.PearsonCorrCoef = PearsonCorrCoef
sys.modules["torchmetrics.regression"] = _reg

# helpers of the reference's test-suite that live in its utilities (test infrastructure, restated)
from collections.abc import Mapping, Sequence
import torch
def _allclose_recursive(res1, res2, atol=1e-6):
    if isinstance(res1, torch.Tensor):
        return torch.allclose(res1, res2, atol=atol)
    if isinstance(res1, str):
        return res1 == res2
    if isinstance(res1, Sequence):
        return all(_allclose_recursive(a, b) for a, b in zip(res1, res2))
    if isinstance(res1, Mapping):
        return all(_allclose_recursive(res1[k], res2[k]) for k in res1)
    return res1 == res2
sys.modules["torchmetrics.utilities.checks"]._allclose_recursive = _allclose_recursive
'''


def main() -> int:
    os.makedirs(os.path.join(ALIAS, "torchmetrics"), exist_ok=True)
    with open(os.path.join(ALIAS, "torchmetrics", "__init__.py"), "w") as f:
        f.write(textwrap.dedent(ALIAS_INIT))
    args = sys.argv[1:] or ["bases/test_metric.py", "bases/test_collections.py", "bases/test_ddp.py",
                            "bases/test_composition.py", "bases/test_hashing.py", "bases/test_saving_loading.py"]
    tests = os.path.join(REF, "tests")
    env = dict(os.environ)
    env["PYTHONPATH"] = os.pathsep.join([ALIAS, ROOT, os.path.join(ROOT, "tests", "golden", "_refshim"), tests])
    env.setdefault("USE_PYTEST_POOL", "1")
    cmd = [sys.executable, "-m", "pytest", "-c", "/dev/null", f"--rootdir={tests}", "-p", "no:cacheprovider", "-q"]
    cmd += [a if a.startswith("-") else os.path.join("unittests", a) for a in args]
    return subprocess.call(cmd, cwd=tests, env=env)


if __name__ == "__main__":
    sys.exit(main())
