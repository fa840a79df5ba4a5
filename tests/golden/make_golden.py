"""Generate golden input/output vectors by running the UNMODIFIED reference (torchmetrics 1.4.0dev).

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

It imports the reference from /root/reference/src with the 4-symbol `lightning_utilities` stand-in in
tests/golden/_refshim, runs the hot-path functions on seeded inputs on CPU and writes
tests/golden/*.npz. Each file holds a list of cases; every case stores its inputs, its parameters
(JSON) and the reference's outputs, so tests never need the reference at run time.
"""
import json
import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = os.environ.get("TM_REFERENCE_SRC", "/root/reference/src")
sys.path.insert(0, REF_SRC)
sys.path.insert(0, os.path.join(HERE, "_refshim"))

import torch  # noqa: E402
import torchmetrics  # noqa: E402
import torchmetrics.functional.classification as F  # noqa: E402
import importlib  # noqa: E402

# the package re-exports functions under the same names as its sub-modules, so fetch the modules explicitly
Fcm = importlib.import_module("torchmetrics.functional.classification.confusion_matrix")
Fprc = importlib.import_module("torchmetrics.functional.classification.precision_recall_curve")
Fss = importlib.import_module("torchmetrics.functional.classification.stat_scores")

assert torchmetrics.__version__.startswith("1.4.0"), torchmetrics.__version__
warnings.filterwarnings("ignore")


class Book:
    def __init__(self, name):
        self.name = name
        self.flat = {}
        self.metas = []

    def add(self, meta, **arrays):
        i = len(self.metas)
        self.metas.append(meta)
        for k, v in arrays.items():
            if isinstance(v, torch.Tensor):
                v = v.detach().cpu().numpy()
            self.flat[f"c{i}__{k}"] = np.asarray(v)

    def save(self):
        self.flat["__meta__"] = np.array(json.dumps(self.metas))
        path = os.path.join(HERE, f"{self.name}.npz")
        np.savez_compressed(path, **self.flat)
        print(f"{path}: {len(self.metas)} cases, {os.path.getsize(path) / 1024:.0f} KiB")


def rng(seed):
    g = torch.Generator()
    g.manual_seed(seed)
    return g


# ------------------------------------------------------------------ config 1: README example
def make_readme():
    b = Book("readme_accuracy")
    torch.manual_seed(0)
    micro = torchmetrics.Accuracy(task="multiclass", num_classes=5)
    macro = torchmetrics.classification.MulticlassAccuracy(num_classes=5)
    P, T, per_batch_micro, per_batch_macro = [], [], [], []
    for _ in range(10):
        preds = torch.randn(10, 5).softmax(dim=-1)
        target = torch.randint(5, (10,))
        per_batch_micro.append(micro(preds, target).item())
        per_batch_macro.append(macro(preds, target).item())
        P.append(preds)
        T.append(target)
    b.add({"num_classes": 5}, preds=torch.stack(P), target=torch.stack(T),
          per_batch_micro=np.array(per_batch_micro, np.float32), per_batch_macro=np.array(per_batch_macro, np.float32),
          final_micro=micro.compute(), final_macro=macro.compute())
    b.save()


# ------------------------------------------------------------------ confusion matrices
def make_confmat():
    b = Book("confusion_matrix")
    seed = 100
    for C, N in [(5, 257), (17, 4099), (128, 20000), (1000, 30011)]:
        for ignore_index in [None, -1, 0]:
            for dt in [torch.int64, torch.int32, torch.uint8]:
                if dt == torch.uint8 and C > 255:
                    continue
                seed += 1
                g = rng(seed)
                preds = torch.randint(0, C, (N,), generator=g).to(dt)
                target = torch.randint(0, C, (N,), generator=g)
                if ignore_index is not None:
                    m = torch.rand(N, generator=g) < 0.1
                    target = torch.where(m, torch.tensor(ignore_index), target)
                if dt == torch.uint8 and ignore_index == -1:
                    continue
                target = target.to(dt)
                p2, t2 = Fcm._multiclass_confusion_matrix_format(preds, target, ignore_index)
                upd = Fcm._multiclass_confusion_matrix_update(p2, t2, C)
                pub = F.multiclass_confusion_matrix(preds, target, C, ignore_index=ignore_index, validate_args=False)
                assert torch.equal(upd, pub)
                b.add({"kind": "multiclass_labels", "num_classes": C, "ignore_index": ignore_index},
                      preds=preds, target=target, confmat=upd)
    # float preds (N, C) and (N, C, X) -> argmax
    for C, shape in [(5, (64, 5)), (7, (33, 7, 3)), (100, (1000, 100))]:
        seed += 1
        g = rng(seed)
        preds = torch.randn(*shape, generator=g)
        tshape = (shape[0],) + tuple(shape[2:])
        target = torch.randint(0, C, tshape, generator=g)
        for ignore_index in [None, 1]:
            cm = F.multiclass_confusion_matrix(preds, target, C, ignore_index=ignore_index, validate_args=False)
            norm = F.multiclass_confusion_matrix(preds, target, C, ignore_index=ignore_index, normalize="true",
                                                 validate_args=False)
            b.add({"kind": "multiclass_scores", "num_classes": C, "ignore_index": ignore_index},
                  preds=preds, target=target, confmat=cm, confmat_norm_true=norm)
    # binary
    for kind in ["labels", "probs", "logits"]:
        for shape in [(301,), (40, 7)]:
            for ignore_index in [None, -1]:
                seed += 1
                g = rng(seed)
                if kind == "labels":
                    preds = torch.randint(0, 2, shape, generator=g)
                elif kind == "probs":
                    preds = torch.rand(*shape, generator=g)
                else:
                    preds = torch.randn(*shape, generator=g) * 3
                target = torch.randint(0, 2, shape, generator=g)
                if ignore_index is not None:
                    target = torch.where(torch.rand(*shape, generator=g) < 0.15, torch.tensor(ignore_index), target)
                cm = F.binary_confusion_matrix(preds, target, threshold=0.4, ignore_index=ignore_index,
                                               validate_args=False)
                b.add({"kind": "binary_" + kind, "threshold": 0.4, "ignore_index": ignore_index},
                      preds=preds, target=target, confmat=cm)
    # multilabel
    for kind in ["labels", "probs", "logits"]:
        for shape in [(203, 6), (50, 4, 5)]:
            for ignore_index in [None, -1]:
                seed += 1
                g = rng(seed)
                L = shape[1]
                if kind == "labels":
                    preds = torch.randint(0, 2, shape, generator=g)
                elif kind == "probs":
                    preds = torch.rand(*shape, generator=g)
                else:
                    preds = torch.randn(*shape, generator=g) * 3
                target = torch.randint(0, 2, shape, generator=g)
                if ignore_index is not None:
                    target = torch.where(torch.rand(*shape, generator=g) < 0.15, torch.tensor(ignore_index), target)
                cm = F.multilabel_confusion_matrix(preds, target, L, threshold=0.5, ignore_index=ignore_index,
                                                   validate_args=False)
                b.add({"kind": "multilabel_" + kind, "num_labels": L, "threshold": 0.5, "ignore_index": ignore_index},
                      preds=preds, target=target, confmat=cm)
    b.save()


# ------------------------------------------------------------------ stat scores + derived metrics
def make_stat_scores():
    b = Book("stat_scores")
    seed = 500
    # binary
    for kind in ["labels", "probs", "logits"]:
        for shape in [(311,), (30, 5, 2)]:
            for ignore_index in [None, -1, 0]:
                for mda in ["global", "samplewise"]:
                    if mda == "samplewise" and len(shape) == 1:
                        continue
                    seed += 1
                    g = rng(seed)
                    preds = {"labels": lambda: torch.randint(0, 2, shape, generator=g),
                             "probs": lambda: torch.rand(*shape, generator=g),
                             "logits": lambda: torch.randn(*shape, generator=g) * 2}[kind]()
                    target = torch.randint(0, 2, shape, generator=g)
                    if ignore_index is not None:
                        target = torch.where(torch.rand(*shape, generator=g) < 0.2, torch.tensor(ignore_index), target)
                    kw = dict(threshold=0.5, multidim_average=mda, ignore_index=ignore_index, validate_args=False)
                    ss = F.binary_stat_scores(preds, target, **kw)
                    out = {"stat_scores": ss}
                    for name in ["accuracy", "precision", "recall", "f1_score", "specificity"]:
                        out[name] = getattr(F, "binary_" + name)(preds, target, **kw)
                    out["fbeta2"] = F.binary_fbeta_score(preds, target, beta=2.0, **kw)
                    b.add({"kind": "binary_" + kind, "threshold": 0.5, "ignore_index": ignore_index,
                           "multidim_average": mda}, preds=preds, target=target, **out)
    # multilabel
    for kind in ["labels", "probs", "logits"]:
        for shape in [(257, 6), (40, 4, 3), (64, 100)]:
            for ignore_index in [None, -1]:
                for mda in ["global", "samplewise"]:
                    if mda == "samplewise" and len(shape) == 2:
                        continue
                    seed += 1
                    g = rng(seed)
                    L = shape[1]
                    preds = {"labels": lambda: torch.randint(0, 2, shape, generator=g),
                             "probs": lambda: torch.rand(*shape, generator=g),
                             "logits": lambda: torch.randn(*shape, generator=g) * 2}[kind]()
                    target = torch.randint(0, 2, shape, generator=g)
                    if ignore_index is not None:
                        target = torch.where(torch.rand(*shape, generator=g) < 0.2, torch.tensor(ignore_index), target)
                    out = {}
                    for avg in ["micro", "macro", "weighted", "none"]:
                        kw = dict(num_labels=L, threshold=0.5, average=avg, multidim_average=mda,
                                  ignore_index=ignore_index, validate_args=False)
                        out["stat_scores_" + avg] = F.multilabel_stat_scores(preds, target, **kw)
                        for name in ["accuracy", "precision", "recall", "f1_score"]:
                            out[f"{name}_{avg}"] = getattr(F, "multilabel_" + name)(preds, target, **kw)
                    b.add({"kind": "multilabel_" + kind, "num_labels": L, "threshold": 0.5,
                           "ignore_index": ignore_index, "multidim_average": mda}, preds=preds, target=target, **out)
    # multiclass
    for kind in ["labels", "scores"]:
        for C, shape in [(5, (203,)), (5, (31, 4)), (11, (500,))]:
            for ignore_index in [None, -1, 1]:
                for mda in ["global", "samplewise"]:
                    if mda == "samplewise" and len(shape) == 1:
                        continue
                    for top_k in [1, 2]:
                        if top_k != 1 and kind == "labels":
                            continue
                        seed += 1
                        g = rng(seed)
                        if kind == "labels":
                            preds = torch.randint(0, C, shape, generator=g)
                        else:
                            preds = torch.randn(shape[0], C, *shape[1:], generator=g)
                        target = torch.randint(0, C, shape, generator=g)
                        if ignore_index is not None:
                            target = torch.where(torch.rand(*shape, generator=g) < 0.2, torch.tensor(ignore_index),
                                                 target)
                        out = {}
                        for avg in ["micro", "macro", "weighted", "none"]:
                            kw = dict(num_classes=C, average=avg, top_k=top_k, multidim_average=mda,
                                      ignore_index=ignore_index, validate_args=False)
                            out["stat_scores_" + avg] = F.multiclass_stat_scores(preds, target, **kw)
                            for name in ["accuracy", "precision", "recall", "f1_score"]:
                                out[f"{name}_{avg}"] = getattr(F, "multiclass_" + name)(preds, target, **kw)
                        b.add({"kind": "multiclass_" + kind, "num_classes": C, "top_k": top_k,
                               "ignore_index": ignore_index, "multidim_average": mda},
                              preds=preds, target=target, **out)
    b.save()


# ------------------------------------------------------------------ binned curves
def make_binned():
    b = Book("binned_curve")
    seed = 900
    thr200 = torch.linspace(0, 1, 200)
    # binary: small (vectorised variant) and > 50 000 (loop variant, the one config 3 exercises)
    cases = [
        ("probs", 1000, 200, None), ("probs", 60_001, 200, None), ("probs", 60_001, 5, None),
        ("logits", 3000, 200, None), ("logits", 52_000, 50, -1), ("probs", 4000, "list_unsorted", -1),
        ("special", 0, 200, None), ("grid", 20_000, 200, None), ("probs", 2500, "tensor17", None),
    ]
    for kind, N, thr, ignore_index in cases:
        seed += 1
        g = rng(seed)
        if kind == "probs":
            preds = torch.rand(N, generator=g)
        elif kind == "logits":
            preds = torch.randn(N, generator=g) * 3
        elif kind == "grid":
            # every score is exactly one of the reference's own fp32 thresholds (or one ulp either side)
            idx = torch.randint(0, 200, (N,), generator=g)
            preds = thr200[idx].clone()
            bump = torch.randint(0, 3, (N,), generator=g)
            preds = torch.where(bump == 1, torch.nextafter(preds, torch.tensor(2.0)), preds)
            preds = torch.where(bump == 2, torch.nextafter(preds, torch.tensor(-1.0)), preds).clamp(0, 1)
        else:
            preds = torch.tensor([0.0, -0.0, 1.0, 0.5, 0.25, 0.75, 1e-45, 1 - 2**-24, 2**-24, 0.5000001, 0.4999999,
                                  float(thr200[57]), float(thr200[101]), float(thr200[199]), float(thr200[1])],
                                 dtype=torch.float32)
            N = preds.numel()
        target = (torch.rand(N, generator=g) < 0.3 + 0.4 * preds.clamp(0, 1)).long()
        if ignore_index is not None:
            target = torch.where(torch.rand(N, generator=g) < 0.1, torch.tensor(ignore_index), target)
        if thr == "list_unsorted":
            thr_arg = [0.7, 0.1, 0.5, 0.30000001192092896, 0.9, 0.0, 1.0, 0.5]
        elif thr == "tensor17":
            thr_arg = torch.sort(torch.rand(17, generator=g)).values
        else:
            thr_arg = thr
        p2, t2, thr_t = Fprc._binary_precision_recall_curve_format(preds, target, thr_arg, ignore_index)
        state = Fprc._binary_precision_recall_curve_update(p2, t2, thr_t)
        if 0 < p2.numel() <= 50_000:
            loop = Fprc._binary_precision_recall_curve_update_loop(p2, t2, thr_t)
            assert torch.equal(loop, state)
        kw = dict(thresholds=thr_arg, ignore_index=ignore_index, validate_args=False)
        prec, rec, _ = F.binary_precision_recall_curve(preds, target, **kw)
        fpr, tpr, _ = F.binary_roc(preds, target, **kw)
        b.add({"kind": "binary_" + kind, "ignore_index": ignore_index,
               "thresholds_arg": thr if isinstance(thr, (int, str)) else None},
              preds=preds, target=target, thresholds=thr_t, confmat=state,
              **({"formatted_preds": p2} if kind == "logits" else {}),
              auroc=F.binary_auroc(preds, target, **kw), ap=F.binary_average_precision(preds, target, **kw),
              auroc_maxfpr=F.binary_auroc(preds, target, max_fpr=0.3, **kw),
              precision=prec, recall=rec, fpr=fpr, tpr=tpr)
    # multiclass: vectorised (N*C*C <= 1e6) and loop
    for kind, N, C, T, ignore_index in [("probs", 500, 5, 50, None), ("logits", 700, 7, 20, -1),
                                        ("probs", 28_000, 6, 30, None), ("logits", 1200, 40, 10, 3)]:
        seed += 1
        g = rng(seed)
        preds = torch.randn(N, C, generator=g)
        if kind == "probs":
            preds = preds.softmax(1)
        target = torch.randint(0, C, (N,), generator=g)
        if ignore_index is not None:
            target = torch.where(torch.rand(N, generator=g) < 0.1, torch.tensor(ignore_index), target)
        out = {}
        for avg in [None, "micro"]:
            p2, t2, thr_t = Fprc._multiclass_precision_recall_curve_format(preds, target, C, T, ignore_index, avg)
            out["confmat_" + str(avg)] = Fprc._multiclass_precision_recall_curve_update(p2, t2, C, thr_t, avg)
            if kind == "logits" and avg is None:
                out["formatted_preds"] = p2
        kw = dict(num_classes=C, thresholds=T, ignore_index=ignore_index, validate_args=False)
        for avg in ["macro", "weighted", "none"]:
            out["auroc_" + avg] = F.multiclass_auroc(preds, target, average=avg, **kw)
            out["ap_" + avg] = F.multiclass_average_precision(preds, target, average=avg, **kw)
        b.add({"kind": "multiclass_" + kind, "num_classes": C, "ignore_index": ignore_index, "thresholds_arg": T},
              preds=preds, target=target, thresholds=thr_t, **out)
    # multilabel
    for kind, shape, T, ignore_index in [("probs", (400, 5), 50, None), ("logits", (300, 4, 3), 20, -1),
                                         ("probs", (2000, 100), 11, None)]:
        seed += 1
        g = rng(seed)
        L = shape[1]
        preds = torch.rand(*shape, generator=g) if kind == "probs" else torch.randn(*shape, generator=g) * 2
        target = torch.randint(0, 2, shape, generator=g)
        if ignore_index is not None:
            target = torch.where(torch.rand(*shape, generator=g) < 0.1, torch.tensor(ignore_index), target)
        p2, t2, thr_t = Fprc._multilabel_precision_recall_curve_format(preds, target, L, T, ignore_index)
        state = Fprc._multilabel_precision_recall_curve_update(p2, t2, L, thr_t)
        out = {}
        kw = dict(num_labels=L, thresholds=T, ignore_index=ignore_index, validate_args=False)
        for avg in ["micro", "macro", "weighted", "none"]:
            out["auroc_" + avg] = F.multilabel_auroc(preds, target, average=avg, **kw)
            out["ap_" + avg] = F.multilabel_average_precision(preds, target, average=avg, **kw)
        b.add({"kind": "multilabel_" + kind, "num_labels": L, "ignore_index": ignore_index, "thresholds_arg": T},
              preds=preds, target=target, thresholds=thr_t, confmat=state, **out)
    b.save()


# ------------------------------------------------------------------ exact curves
def make_exact():
    b = Book("exact_curve")
    seed = 1300
    for kind, N, ignore_index in [("probs", 1000, None), ("ties", 5000, None), ("logits", 3000, -1),
                                  ("probs", 70_000, None), ("allpos", 50, None), ("allneg", 50, None),
                                  ("probs", 1, None), ("ties", 40_000, -1)]:
        seed += 1
        g = rng(seed)
        if kind == "ties":
            preds = torch.randint(0, 64, (N,), generator=g).float() / 63
        elif kind == "logits":
            preds = torch.randn(N, generator=g) * 3
        else:
            preds = torch.rand(N, generator=g)
        target = (torch.rand(N, generator=g) < 0.3 + 0.4 * preds.clamp(0, 1)).long()
        if kind == "allpos":
            target = torch.ones(N, dtype=torch.long)
        if kind == "allneg":
            target = torch.zeros(N, dtype=torch.long)
        if ignore_index is not None:
            target = torch.where(torch.rand(N, generator=g) < 0.1, torch.tensor(ignore_index), target)
        p2, t2, _ = Fprc._binary_precision_recall_curve_format(preds, target, None, ignore_index)
        fps, tps, thr = Fprc._binary_clf_curve(p2, t2)
        kw = dict(thresholds=None, ignore_index=ignore_index, validate_args=False)
        prec, rec, pthr = F.binary_precision_recall_curve(preds, target, **kw)
        fpr, tpr, rthr = F.binary_roc(preds, target, **kw)
        b.add({"kind": "binary_" + kind, "ignore_index": ignore_index}, preds=preds, target=target,
              fps=fps, tps=tps, thr=thr,
              auroc=F.binary_auroc(preds, target, **kw), ap=F.binary_average_precision(preds, target, **kw),
              auroc_maxfpr=F.binary_auroc(preds, target, max_fpr=0.25, **kw),
              precision=prec, recall=rec, pr_thresholds=pthr, fpr=fpr, tpr=tpr, roc_thresholds=rthr)
    # multiclass / multilabel exact scalars
    for C, N, ignore_index in [(5, 600, None), (4, 900, -1)]:
        seed += 1
        g = rng(seed)
        preds = torch.randn(N, C, generator=g).softmax(1)
        target = torch.randint(0, C, (N,), generator=g)
        if ignore_index is not None:
            target = torch.where(torch.rand(N, generator=g) < 0.1, torch.tensor(ignore_index), target)
        out = {}
        kw = dict(num_classes=C, thresholds=None, ignore_index=ignore_index, validate_args=False)
        for avg in ["macro", "weighted", "none"]:
            out["auroc_" + avg] = F.multiclass_auroc(preds, target, average=avg, **kw)
            out["ap_" + avg] = F.multiclass_average_precision(preds, target, average=avg, **kw)
        b.add({"kind": "multiclass", "num_classes": C, "ignore_index": ignore_index}, preds=preds, target=target,
              **out)
    for L, N, ignore_index in [(5, 600, None), (3, 900, -1)]:
        seed += 1
        g = rng(seed)
        preds = torch.rand(N, L, generator=g)
        target = torch.randint(0, 2, (N, L), generator=g)
        if ignore_index is not None:
            target = torch.where(torch.rand(N, L, generator=g) < 0.1, torch.tensor(ignore_index), target)
        out = {}
        kw = dict(num_labels=L, thresholds=None, ignore_index=ignore_index, validate_args=False)
        for avg in ["micro", "macro", "weighted", "none"]:
            out["auroc_" + avg] = F.multilabel_auroc(preds, target, average=avg, **kw)
            out["ap_" + avg] = F.multilabel_average_precision(preds, target, average=avg, **kw)
        b.add({"kind": "multilabel", "num_labels": L, "ignore_index": ignore_index}, preds=preds, target=target,
              **out)
    b.save()


# ------------------------------------------------------------------ doctest known answers (reference docstrings)
def make_doctests():
    b = Book("doctest_known_answers")
    # functional/classification/confusion_matrix.py:388-408
    t = torch.tensor([2, 1, 0, 0]); p = torch.tensor([2, 1, 0, 1])
    b.add({"fn": "multiclass_confusion_matrix", "num_classes": 3}, preds=p, target=t,
          out=F.multiclass_confusion_matrix(p, t, num_classes=3))
    # functional/classification/confusion_matrix.py:204-219
    t = torch.tensor([1, 1, 0, 0]); p = torch.tensor([0.35, 0.85, 0.48, 0.01])
    b.add({"fn": "binary_confusion_matrix"}, preds=p, target=t, out=F.binary_confusion_matrix(p, t))
    # functional/classification/confusion_matrix.py:579-596
    t = torch.tensor([[0, 1, 0], [1, 0, 1]]); p = torch.tensor([[0.11, 0.22, 0.84], [0.73, 0.33, 0.92]])
    b.add({"fn": "multilabel_confusion_matrix", "num_labels": 3}, preds=p, target=t,
          out=F.multilabel_confusion_matrix(p, t, num_labels=3))
    # functional/classification/precision_recall_curve.py:340-351
    p = torch.tensor([0, 0.5, 0.7, 0.8]); t = torch.tensor([0, 1, 1, 0])
    for thr in [None, 5]:
        pr, rc, th = F.binary_precision_recall_curve(p, t, thresholds=thr)
        b.add({"fn": "binary_precision_recall_curve", "thresholds": thr}, preds=p, target=t, precision=pr, recall=rc,
              thresholds=th)
        b.add({"fn": "binary_auroc", "thresholds": thr}, preds=p, target=t, out=F.binary_auroc(p, t, thresholds=thr))
    # functional/classification/stat_scores.py:772-809
    t = torch.tensor([[0, 1, 0], [1, 0, 1]]); p = torch.tensor([[0, 0, 1], [1, 0, 1]])
    b.add({"fn": "multilabel_stat_scores", "num_labels": 3}, preds=p, target=t,
          out=F.multilabel_stat_scores(p, t, num_labels=3),
          out_none=F.multilabel_stat_scores(p, t, num_labels=3, average=None))
    b.save()


# ------------------------------------------------------------------ consumers of the same update kernels
def make_derived():
    """specificity / hamming distance (stat-scores state) and jaccard / cohen kappa / matthews corrcoef (confusion-matrix
    state): reference outputs of the public functionals on seeded inputs, every task, averaging and ignore_index."""
    b = Book("derived_metrics")
    g = rng(77)
    n, C, L = 257, 6, 5
    # binary: probabilities, logits, labels
    bt = torch.randint(0, 2, (n,), generator=g)
    binary_inputs = {"probs": torch.rand(n, generator=g), "logits": torch.randn(n, generator=g) * 3,
                     "labels": torch.randint(0, 2, (n,), generator=g)}
    for kind, bp in binary_inputs.items():
        for ii in (None, -1):
            t = bt.clone()
            if ii is not None:
                t[::9] = ii
            out = {
                "specificity": F.binary_specificity(bp, t, ignore_index=ii),
                "hamming": F.binary_hamming_distance(bp, t, ignore_index=ii),
                "jaccard": F.binary_jaccard_index(bp, t, ignore_index=ii),
                "kappa_none": F.binary_cohen_kappa(bp, t, ignore_index=ii),
                "kappa_linear": F.binary_cohen_kappa(bp, t, weights="linear", ignore_index=ii),
                "kappa_quadratic": F.binary_cohen_kappa(bp, t, weights="quadratic", ignore_index=ii),
                "mcc": F.binary_matthews_corrcoef(bp, t, ignore_index=ii),
            }
            b.add({"task": "binary", "kind": kind, "ignore_index": ii}, preds=bp, target=t, **out)
    # multiclass: labels and (N, C) scores; ignore_index outside and inside the class range
    mt = torch.randint(0, C, (n,), generator=g)
    multiclass_inputs = {"labels": torch.where(torch.rand(n, generator=g) < 0.6, mt, torch.randint(0, C, (n,), generator=g)),
                         "scores": torch.randn(n, C, generator=g)}
    for kind, mp in multiclass_inputs.items():
        for ii in (None, -1, 2):
            t = mt.clone()
            if ii is not None:
                t[::7] = ii
            out = {}
            for avg in ("micro", "macro", "weighted", "none"):
                out[f"specificity_{avg}"] = F.multiclass_specificity(mp, t, C, average=avg, ignore_index=ii)
                out[f"hamming_{avg}"] = F.multiclass_hamming_distance(mp, t, C, average=avg, ignore_index=ii)
                out[f"jaccard_{avg}"] = F.multiclass_jaccard_index(mp, t, C, average=avg, ignore_index=ii)
            for w in ("none", "linear", "quadratic"):
                out[f"kappa_{w}"] = F.multiclass_cohen_kappa(mp, t, C, weights=w, ignore_index=ii)
            out["mcc"] = F.multiclass_matthews_corrcoef(mp, t, C, ignore_index=ii)
            b.add({"task": "multiclass", "kind": kind, "num_classes": C, "ignore_index": ii}, preds=mp, target=t, **out)
    # multilabel
    lt = torch.randint(0, 2, (n, L), generator=g)
    multilabel_inputs = {"probs": torch.rand(n, L, generator=g), "logits": torch.randn(n, L, generator=g) * 2}
    for kind, lp in multilabel_inputs.items():
        for ii in (None, -1):
            t = lt.clone()
            if ii is not None:
                t[::5, 1] = ii
            out = {}
            for avg in ("micro", "macro", "weighted", "none"):
                out[f"specificity_{avg}"] = F.multilabel_specificity(lp, t, L, average=avg, ignore_index=ii)
                out[f"hamming_{avg}"] = F.multilabel_hamming_distance(lp, t, L, average=avg, ignore_index=ii)
                out[f"jaccard_{avg}"] = F.multilabel_jaccard_index(lp, t, L, average=avg, ignore_index=ii)
            out["mcc"] = F.multilabel_matthews_corrcoef(lp, t, L, ignore_index=ii)
            b.add({"task": "multilabel", "kind": kind, "num_labels": L, "ignore_index": ii}, preds=lp, target=t, **out)
    # degenerate tables the reducers special-case: perfect, all wrong, single class
    ones, zeros = torch.ones(40, dtype=torch.long), torch.zeros(40, dtype=torch.long)
    half = torch.cat([ones[:20], zeros[:20]])
    for name, (p, t) in {"perfect": (half, half), "all_wrong": (half, 1 - half), "only_pos": (ones, ones),
                         "pred_const": (ones, half)}.items():
        b.add({"task": "binary", "kind": "degenerate_" + name, "ignore_index": None}, preds=p, target=t,
              specificity=F.binary_specificity(p, t), hamming=F.binary_hamming_distance(p, t),
              jaccard=F.binary_jaccard_index(p, t), kappa_none=F.binary_cohen_kappa(p, t),
              mcc=F.binary_matthews_corrcoef(p, t))
    b.save()


def make_consumers2():
    """Second batch of consumers of the same update kernels: operating points on the binned / exact
    curves, group fairness on the binary stat scores, nominal association measures on the confusion matrix."""
    import torchmetrics.functional.nominal as FN

    b = Book("consumers2")
    g = rng(91)
    n, C, L = 311, 4, 3
    minima = (0.0, 0.25, 0.5, 0.9, 1.0)
    fams = {"recall_at_fixed_precision": "min_precision", "precision_at_fixed_recall": "min_recall",
            "sensitivity_at_specificity": "min_specificity", "specificity_at_sensitivity": "min_sensitivity"}
    bt = torch.randint(0, 2, (n,), generator=g)
    bp = (torch.rand(n, generator=g) * 0.6 + 0.4 * bt * torch.rand(n, generator=g)).clamp(0, 1)
    mt = torch.randint(0, C, (n,), generator=g)
    mp = (torch.randn(n, C, generator=g) + 1.5 * torch.nn.functional.one_hot(mt, C)).softmax(1)
    lt = torch.randint(0, 2, (n, L), generator=g)
    lp = (torch.rand(n, L, generator=g) * 0.7 + 0.3 * lt * torch.rand(n, L, generator=g)).clamp(0, 1)
    for thresholds in (None, 17):
        for ii in (None, -1):
            tb, tm_, tl = bt.clone(), mt.clone(), lt.clone()
            if ii is not None:
                tb[::9] = ii
                tm_[::7] = ii
                tl[::5, 1] = ii
            for task, (p, t, extra) in {"binary": (bp, tb, {}), "multiclass": (mp, tm_, {"num_classes": C}),
                                        "multilabel": (lp, tl, {"num_labels": L})}.items():
                out = {}
                for fam, arg in fams.items():
                    fn = getattr(F, f"{task}_{fam}")
                    for m in minima:
                        val, thr = fn(p, t, **extra, **{arg: m}, thresholds=thresholds, ignore_index=ii)
                        out[f"{fam}@{m}"] = torch.stack([val.float().reshape(-1), thr.float().reshape(-1)])
                b.add({"what": "operating_point", "task": task, "thresholds": thresholds, "ignore_index": ii, **extra},
                      preds=p, target=t, **out)
    # group fairness: three groups (one of them rare), probabilities and logits, with and without ignore_index
    groups = torch.randint(0, 3, (n,), generator=g)
    groups[groups == 2] = torch.where(torch.rand(int((groups == 2).sum()), generator=g) < 0.8, 0, 2)
    for kind, p in {"probs": bp, "logits": torch.randn(n, generator=g) * 2}.items():
        for ii in (None, -1):
            t = bt.clone()
            if ii is not None:
                t[::9] = ii
            rates = F.binary_groups_stat_rates(p, t, groups, 3, ignore_index=ii)
            fair = F.binary_fairness(p, t, groups, "all", ignore_index=ii)
            dp = F.demographic_parity(p, groups, ignore_index=ii)
            b.add({"what": "group_fairness", "kind": kind, "ignore_index": ii,
                   "fair_keys": sorted(fair.keys()), "dp_keys": sorted(dp.keys())}, preds=p, target=t, groups=groups,
                  rates=torch.stack([rates[f"group_{i}"] for i in range(len(rates))]),
                  fair=torch.stack([fair[k] for k in sorted(fair.keys())]),
                  dp=torch.stack([dp[k] for k in sorted(dp.keys())]))
    # nominal association: label tensors (int and float with NaN), a 2 x 2 table (Yates), a constant variable
    cases = {}
    x = torch.randint(0, 5, (n,), generator=g)
    y = torch.where(torch.rand(n, generator=g) < 0.5, x, torch.randint(0, 5, (n,), generator=g))
    cases["labels5"] = (x, y)
    xf = x.float()
    xf[::11] = float("nan")
    cases["nan"] = (xf, y.float())
    cases["two_by_two"] = (torch.randint(0, 2, (n,), generator=g), torch.randint(0, 2, (n,), generator=g))
    cases["constant"] = (torch.zeros(n, dtype=torch.long), torch.randint(0, 3, (n,), generator=g))
    for name, (p, t) in cases.items():
        for strat in ("replace", "drop"):
            def stateful(cls, **kw):
                # the functional form of the reference fails on a 2 x 2 table with bias correction (in-place float add
                # into its int64 matrix, utils.py:56); its stateful form keeps a float table and works
                m = cls(num_classes=len(torch.cat([p, t]).unique()), nan_strategy=strat, **kw)
                m.update(p, t)
                return m.compute()

            import torchmetrics.nominal as N_

            if name == "two_by_two":
                cb, tb_ = stateful(N_.CramersV, bias_correction=True), stateful(N_.TschuprowsT, bias_correction=True)
            else:
                cb, tb_ = FN.cramers_v(p, t, True, strat), FN.tschuprows_t(p, t, True, strat)
            out = {"cramers_bias": cb, "cramers": FN.cramers_v(p, t, False, strat),
                   "tschuprows_bias": tb_, "tschuprows": FN.tschuprows_t(p, t, False, strat),
                   "pearson": FN.pearsons_contingency_coefficient(p, t, strat), "theils_u": FN.theils_u(p, t, strat),
                   "theils_u_rev": FN.theils_u(t, p, strat)}
            b.add({"what": "nominal", "kind": name, "nan_strategy": strat}, preds=p, target=t,
                  **{k: v.float() for k, v in out.items()})
    mat = torch.randint(0, 4, (200, 4), generator=g)
    b.add({"what": "nominal_matrix"}, matrix=mat, cramers=FN.cramers_v_matrix(mat), theils_u=FN.theils_u_matrix(mat),
          tschuprows=FN.tschuprows_t_matrix(mat), pearson=FN.pearsons_contingency_coefficient_matrix(mat))
    b.save()


def make_dice():
    """Legacy `_stat_scores_update` + Dice: every input case the reference's test_dice.py / legacy
    inputs cover, with the raw statistics and the functional's value for each valid parameter combination."""
    from torchmetrics.functional.classification.dice import dice as ref_dice

    b = Book("dice_legacy")
    g = rng(123)
    n, C, X = 64, 5, 7
    inputs = {
        "binary_probs": (torch.rand(n, generator=g), torch.randint(0, 2, (n,), generator=g), {}),
        "multiclass_labels": (torch.randint(0, C, (n,), generator=g), torch.randint(0, C, (n,), generator=g),
                              {"num_classes": C}),
        "multiclass_probs": (torch.randn(n, C, generator=g).softmax(1), torch.randint(0, C, (n,), generator=g),
                             {"num_classes": C}),
        "multilabel_probs": (torch.rand(n, C, generator=g), torch.randint(0, 2, (n, C), generator=g),
                             {"num_classes": C}),
        "mdmc_labels": (torch.randint(0, C, (n, X), generator=g), torch.randint(0, C, (n, X), generator=g),
                        {"num_classes": C}),
        "mdmc_probs": (torch.randn(n, C, X, generator=g).softmax(1), torch.randint(0, C, (n, X), generator=g),
                       {"num_classes": C}),
    }
    for kind, (p, t, base) in inputs.items():
        for reduce in ("micro", "macro", "samples"):
            for mdmc in ((None,) if "mdmc" not in kind else ("global", "samplewise")):
                for ii in (None, 0):
                    for top_k in ((None, 2) if kind in ("multiclass_probs", "mdmc_probs") else (None,)):
                        kw = dict(base, reduce=reduce, mdmc_reduce=mdmc, ignore_index=ii, top_k=top_k)
                        try:
                            tp, fp, tn, fn = Fss._stat_scores_update(p, t, **kw)
                        except Exception:
                            continue
                        out = {"tp": tp, "fp": fp, "tn": tn, "fn": fn}
                        for avg in ("micro", "macro", "weighted", "none", "samples"):
                            if ("macro" if avg in ("weighted", "none") else avg) != reduce:
                                continue
                            try:
                                out[f"dice_{avg}"] = ref_dice(p, t, average=avg, mdmc_average=mdmc, ignore_index=ii,
                                                              top_k=top_k, **base).float()
                            except Exception:
                                pass
                        b.add({"kind": kind, "reduce": reduce, "mdmc_reduce": mdmc, "ignore_index": ii, "top_k": top_k,
                               **base}, preds=p, target=t, **out)
    # the reference's doctest (dice.py:196-199)
    p, t = torch.tensor([2, 0, 2, 1]), torch.tensor([1, 1, 2, 0])
    b.add({"kind": "doctest", "reduce": "micro", "mdmc_reduce": None, "ignore_index": None, "top_k": None},
          preds=p, target=t, dice_micro=ref_dice(p, t, average="micro").float())
    b.save()


if __name__ == "__main__":
    make_readme()
    make_confmat()
    make_stat_scores()
    make_binned()
    make_exact()
    make_doctests()
    make_derived()
    make_consumers2()
    make_dice()
