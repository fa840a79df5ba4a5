"""Per-call latency of typical small-batch updates (host wall clock, 300 calls, sync at the end)."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torchmetrics_b200 as tm  # noqa: E402

DEV = "cuda:0"
g = torch.Generator(device=DEV).manual_seed(0)


def bench(name, fn, reps=300):
    for _ in range(20):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    print(f"{name:70s} {(time.perf_counter() - t0) / reps * 1e6:8.1f} us/call", flush=True)


B, C = 256, 1000
logits = torch.randn(B, C, device=DEV, generator=g)
labels = torch.randint(0, C, (B,), device=DEV, generator=g)
pl = logits.argmax(1)
for va in (True, False):
    m = tm.classification.MulticlassAccuracy(num_classes=C, validate_args=va).to(DEV)
    bench(f"MulticlassAccuracy(C=1000).update(logits[256,1000]) validate={va}", lambda: m.update(logits, labels))
    m = tm.classification.MulticlassAccuracy(num_classes=C, validate_args=va).to(DEV)
    bench(f"MulticlassAccuracy(C=1000).update(labels[256]) validate={va}", lambda: m.update(pl, labels))
    m = tm.MulticlassConfusionMatrix(num_classes=100, validate_args=va).to(DEV)
    l100 = labels % 100
    bench(f"MulticlassConfusionMatrix(C=100).update(labels[256]) validate={va}", lambda: m.update(l100, l100))
    p = torch.rand(4096, device=DEV, generator=g)
    t = (torch.rand(4096, device=DEV, generator=g) < 0.5).long()
    m = tm.classification.BinaryAUROC(thresholds=200, validate_args=va).to(DEV)
    bench(f"BinaryAUROC(thresholds=200).update(p[4096]) validate={va}", lambda: m.update(p, t))
    m = tm.classification.BinaryF1Score(validate_args=va).to(DEV)
    bench(f"BinaryF1Score.update(p[4096]) validate={va}", lambda: m.update(p, t))
    coll = tm.MetricCollection([tm.classification.MulticlassAccuracy(num_classes=C, validate_args=va),
                                tm.classification.MulticlassPrecision(num_classes=C, validate_args=va),
                                tm.classification.MulticlassRecall(num_classes=C, validate_args=va),
                                tm.classification.MulticlassF1Score(num_classes=C, validate_args=va)]).to(DEV)
    bench(f"MetricCollection[Acc,P,R,F1](C=1000).update(logits[256,1000]) validate={va}", lambda: coll.update(logits, labels))
    bench(f"MetricCollection[Acc,P,R,F1](C=1000).forward(logits[256,1000]) validate={va}", lambda: coll(logits, labels))
for va in (True, False):
    m = tm.classification.MulticlassAccuracy(num_classes=C, top_k=5, validate_args=va).to(DEV)
    bench(f"MulticlassAccuracy(C=1000, top_k=5).update(logits[256,1000]) validate={va}", lambda: m.update(logits, labels))
big = torch.randn(8192, C, device=DEV, generator=g)
bigl = torch.randint(0, C, (8192,), device=DEV, generator=g)
m = tm.classification.MulticlassAccuracy(num_classes=C, top_k=5, validate_args=False).to(DEV)
bench("MulticlassAccuracy(C=1000, top_k=5).update(logits[8192,1000]) validate=False", lambda: m.update(big, bigl), reps=100)
from importlib import import_module
S = import_module("torchmetrics_b200.functional.classification.stat_scores")
p3, t2 = S._multiclass_stat_scores_format(big, bigl, 5)
bench("  reference's one-hot branch (torch ops on the device), same input", lambda: S._multiclass_stat_scores_onehot(p3, t2, C, 5, "global", None), reps=100)
bench("  torch.topk alone", lambda: big.topk(5, dim=1), reps=100)

# forward(): the call a training loop makes every step
for va in (True, False):
    m = tm.classification.MulticlassF1Score(num_classes=C, validate_args=va).to(DEV)
    bench(f"MulticlassF1Score(C=1000).forward(logits[256,1000]) validate={va}", lambda: m(logits, labels))
    coll2 = tm.MetricCollection([tm.classification.MulticlassAccuracy(num_classes=C, validate_args=va),
                                 tm.classification.MulticlassPrecision(num_classes=C, validate_args=va),
                                 tm.classification.MulticlassRecall(num_classes=C, validate_args=va),
                                 tm.classification.MulticlassF1Score(num_classes=C, validate_args=va)]).to(DEV)
    bench(f"MetricCollection[Acc,P,R,F1](C=1000).forward, never updated directly, validate={va}",
          lambda: coll2(logits, labels))

# the unmodified reference on the same GPU (torch ops on the device), when it travelled with the snapshot
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
try:
    from bench import _import_reference

    ref, where = _import_reference()
except Exception:
    ref = None
if ref is not None:
    R = ref.classification
    for va in (True, False):
        m = R.MulticlassF1Score(num_classes=C, validate_args=va).to(DEV)
        bench(f"reference MulticlassF1Score(C=1000).update(logits[256,1000]) validate={va}", lambda: m.update(logits, labels))
        bench(f"reference MulticlassF1Score(C=1000).forward(logits[256,1000]) validate={va}", lambda: m(logits, labels))
        rc = ref.MetricCollection([R.MulticlassAccuracy(num_classes=C, validate_args=va),
                                   R.MulticlassPrecision(num_classes=C, validate_args=va),
                                   R.MulticlassRecall(num_classes=C, validate_args=va),
                                   R.MulticlassF1Score(num_classes=C, validate_args=va)]).to(DEV)
        bench(f"reference MetricCollection[Acc,P,R,F1](C=1000).update validate={va}", lambda: rc.update(logits, labels))
        bench(f"reference MetricCollection[Acc,P,R,F1](C=1000).forward validate={va}", lambda: rc(logits, labels))
        m = R.BinaryAUROC(thresholds=200, validate_args=va).to(DEV)
        bench(f"reference BinaryAUROC(thresholds=200).update(p[4096]) validate={va}", lambda: m.update(p, t))
