import os, sys, time, cProfile, pstats
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchmetrics_b200 as tm
DEV = "cuda:0"
g = torch.Generator(device=DEV).manual_seed(0)
B, C = 256, 1000
logits = torch.randn(B, C, device=DEV, generator=g)
labels = torch.randint(0, C, (B,), device=DEV, generator=g)
m = tm.classification.MulticlassAccuracy(num_classes=C, validate_args=False).to(DEV)
def t(name, fn, reps=300):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); print(f"{name:40s} {(time.perf_counter()-t0)/reps*1e6:8.1f} us")
t("forward", lambda: m(logits, labels))
t("update", lambda: m.update(logits, labels))
def comp():
    m._computed = None
    return m.compute()
t("compute", comp)
t("reset", lambda: m.reset())
m.update(logits, labels)
pr = cProfile.Profile(); pr.enable()
for _ in range(200): m(logits, labels)
torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
