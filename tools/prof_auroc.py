import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchmetrics_b200 import _native as N
DEV = torch.device("cuda:0")
n = 1 << int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
g = torch.Generator(device=DEV).manual_seed(1)
preds = torch.rand(n, device=DEV, generator=g)
target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
for _ in range(2):
    print(N.exact_auroc_ap(preds, target, 1))
torch.cuda.synchronize()
