"""Development micro-benchmark of the partition path only: python tools/kb_part.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchmetrics_b200 import _native as N  # noqa: E402
from tools.kbench import timeit, report, DEV  # noqa: E402

tag = ""
for C, n in [(1000, 1 << 24), (1000, 1 << 26), (1000, 1 << 23), (300, 1 << 24), (2000, 1 << 24), (1000, (1 << 24) + 12345)]:
    g = torch.Generator(device=DEV).manual_seed(1)
    p = torch.randint(0, C, (n,), device=DEV, generator=g)
    t = torch.randint(0, C, (n,), device=DEV, generator=g)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    ref = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(ref, p, t, C, algo=1)
    N.confmat_update_(cm, p, t, C, algo=6)
    assert torch.equal(cm, ref), "mismatch"
    med, best = timeit(lambda: N.confmat_update_(cm, p, t, C, algo=6))
    report(f"confmat C={C} n={n} algo=6 {tag}", n * 16, med, best, n)
    if C == 1000 and n == 1 << 24:
        p2 = torch.where(torch.rand(n, device=DEV) < 0.8, t, p)
        ref.zero_(); cm.zero_()
        N.confmat_update_(ref, p2, t, C, algo=1)
        N.confmat_update_(cm, p2, t, C, algo=6)
        assert torch.equal(cm, ref), "mismatch diag"
        med, best = timeit(lambda: N.confmat_update_(cm, p2, t, C, algo=6))
        report(f"confmat C=1000 n=2^24 diag-heavy algo=6 {tag}", n * 16, med, best, n)
        t4 = torch.where(torch.rand(n, device=DEV) < 0.5, torch.zeros_like(t), t)
        p4 = torch.where(torch.rand(n, device=DEV) < 0.9, t4, p)
        ref.zero_(); cm.zero_()
        N.confmat_update_(ref, p4, t4, C, algo=1)
        N.confmat_update_(cm, p4, t4, C, algo=6)
        assert torch.equal(cm, ref), "mismatch hot"
        med, best = timeit(lambda: N.confmat_update_(cm, p4, t4, C, algo=6))
        report(f"confmat C=1000 n=2^24 hot-cell algo=6 {tag}", n * 16, med, best, n)
        # one class holds 30 % of the targets
        t5 = torch.where(torch.rand(n, device=DEV) < 0.3, torch.full_like(t, 7), t)
        ref.zero_(); cm.zero_()
        N.confmat_update_(ref, p, t5, C, algo=1)
        N.confmat_update_(cm, p, t5, C, algo=6)
        assert torch.equal(cm, ref), "mismatch heavy row"
        med, best = timeit(lambda: N.confmat_update_(cm, p, t5, C, algo=6))
        report(f"confmat C=1000 n=2^24 heavy-row algo=6 {tag}", n * 16, med, best, n)
        p3, t3 = p.int(), t.int()
        med, best = timeit(lambda: N.confmat_update_(cm, p3, t3, C, algo=6))
        report(f"confmat C=1000 n=2^24 i32 algo=6 {tag}", n * 8, med, best, n)
