"""Run one kernel a few times (for ncu). Usage: python tools/prof_one.py binned|binned_multi|confmat|class_stats|stat|rows|partition|exact|exact_auroc|reduce [log2n]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchmetrics_b200 import _native as N  # noqa: E402

DEV = torch.device("cuda:0")
what = sys.argv[1]
n = 1 << int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 26
g = torch.Generator(device=DEV).manual_seed(1)
if what == "binned":
    thr = torch.linspace(0, 1, 200, device=DEV)
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    st = torch.zeros(200, 2, 2, dtype=torch.int64, device=DEV)
    for _ in range(4):
        N.binned_curve_update_(st, preds, target, thr, prob_mode=N.PROB_AS_IS)
elif what == "confmat":
    C = 1000
    p = torch.randint(0, C, (n,), device=DEV, generator=g)
    t = torch.randint(0, C, (n,), device=DEV, generator=g)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    for _ in range(4):
        N.confmat_update_(cm, p, t, C)
elif what == "stat":
    L = 100
    p = torch.rand(n // 128, L, device=DEV, generator=g)
    t = torch.randint(0, 2, (n // 128, L), device=DEV, generator=g)
    out = torch.zeros(4, L, dtype=torch.int64, device=DEV)
    for _ in range(4):
        N.stat_scores_update_(out, p, t, L, 0.5, None, N.PROB_AS_IS)
elif what == "binned_multi":
    L = 100
    rows = n // 64
    thr = torch.linspace(0, 1, 200, device=DEV)
    preds = torch.rand(rows, L, device=DEV, generator=g)
    tl = torch.randint(0, 2, (rows, L), device=DEV, generator=g)
    st = torch.zeros(200, L, 2, 2, dtype=torch.int64, device=DEV)
    for _ in range(4):
        N.binned_curve_update_(st, preds, tl, thr, num_columns=L, prob_mode=N.PROB_AS_IS)
elif what == "class_stats":
    C = 1000
    t = torch.randint(0, C, (n,), device=DEV, generator=g)
    p = torch.randint(0, C, (n,), device=DEV, generator=g)
    out = torch.zeros(4, C, dtype=torch.int64, device=DEV)
    for _ in range(4):
        N.class_stats_update_(out, p, t, C)
elif what == "reduce":
    # the fused reducer: (rows, 100) samplewise states, then the (100,) global ones of config 5
    rows = max(1, n // 100)
    st = [torch.randint(0, 1000, (rows, 100), device=DEV, generator=g) for _ in range(4)]
    for avg in ("macro", "micro", None):
        N.stat_reduce(N.STAT_FBETA, *st, avg, True, 1.0)
    one = [t[0].contiguous() for t in st]
    for kind in (N.STAT_PRECISION, N.STAT_RECALL, N.STAT_FBETA):
        N.stat_reduce(kind, *one, "macro", True, 1.0)
elif what == "exact_auroc":
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    for _ in range(2):
        N.exact_auroc_ap(preds, target, 1)
elif what == "rows":
    width = 1024
    p = torch.rand(n // width, width, device=DEV, generator=g)
    t = torch.randint(0, 2, (n // width, width), device=DEV, generator=g)
    for _ in range(4):
        N.stat_scores_rows(p, t, width, 0.5, None)
    C = 100
    pl = torch.randint(0, C, (n // width, width), device=DEV, generator=g)
    tl = torch.randint(0, C, (n // width, width), device=DEV, generator=g)
    for _ in range(4):
        N.class_stats_rows(pl, tl, C)
elif what == "partition":
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.4).to(torch.uint8)
    spl = torch.linspace(0, 1, 9, device=DEV)[1:-1].contiguous()
    for _ in range(4):
        N.range_partition(preds, target, spl, 1)
elif what == "exact":
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    for _ in range(2):
        N.exact_curve(preds, target, 1)
torch.cuda.synchronize()
print("done", what, n)
