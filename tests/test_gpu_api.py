"""The reference-facing API (functional + Metric classes + MetricCollection) on the GPU, checked against golden
vectors produced by the unmodified reference and against scikit-learn on seeded inputs."""
import numpy as np
import pytest
import torch

from conftest import case_id, load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _tm():
    import torchmetrics_b200 as tm

    return tm


def _t(a):
    return torch.from_numpy(np.asarray(a)).to(DEV)


def _close(got, want, atol=1e-6):
    got = got.detach().cpu().numpy() if isinstance(got, torch.Tensor) else np.asarray(got)
    want = np.asarray(want)
    assert got.shape == want.shape, (got.shape, want.shape)
    if want.dtype.kind in "iub":
        assert np.array_equal(got, want)
    else:
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=atol, equal_nan=True)  # rtol: counts averaged in fp32


# ------------------------------------------------------------------ config 1: README example
def test_readme_accuracy_config1():
    tm = _tm()
    c = load_golden("readme_accuracy")[0]
    micro = tm.Accuracy(task="multiclass", num_classes=5).to(DEV)
    macro = tm.classification.MulticlassAccuracy(num_classes=5).to(DEV)
    for i in range(10):
        p, t = _t(c["preds"][i]), _t(c["target"][i])
        assert abs(float(micro(p, t)) - float(c["per_batch_micro"][i])) < 1e-7
        assert abs(float(macro(p, t)) - float(c["per_batch_macro"][i])) < 1e-7
    assert float(micro.compute()) == float(c["final_micro"])
    assert float(macro.compute()) == float(c["final_macro"])
    assert abs(float(micro.compute()) - 0.20000000298023224) < 1e-9
    assert abs(float(macro.compute()) - 0.19858768582344055) < 1e-9


# ------------------------------------------------------------------ confusion matrix
@pytest.mark.parametrize("c", load_golden("confusion_matrix"), ids=case_id)
def test_confusion_matrix_functional_and_class(c):
    tm = _tm()
    F = tm.functional
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    p, t = _t(c["preds"]), _t(c["target"])
    if kind.startswith("multiclass"):
        C = c.meta["num_classes"]
        got = F.multiclass_confusion_matrix(p, t, C, ignore_index=ii, validate_args=False)
        cls = tm.MulticlassConfusionMatrix(C, ignore_index=ii, validate_args=False).to(DEV)
        if "confmat_norm_true" in c:
            _close(F.multiclass_confusion_matrix(p, t, C, ignore_index=ii, normalize="true", validate_args=False),
                   c["confmat_norm_true"])
    elif kind.startswith("binary"):
        got = F.binary_confusion_matrix(p, t, c.meta["threshold"], ignore_index=ii, validate_args=False)
        cls = tm.BinaryConfusionMatrix(c.meta["threshold"], ignore_index=ii, validate_args=False).to(DEV)
    else:
        L = c.meta["num_labels"]
        got = F.multilabel_confusion_matrix(p, t, L, c.meta["threshold"], ignore_index=ii, validate_args=False)
        cls = tm.MultilabelConfusionMatrix(L, c.meta["threshold"], ignore_index=ii, validate_args=False).to(DEV)
    slack = 1 if kind.endswith("logits") else 0  # CPU-vs-CUDA sigmoid can move a score across the threshold
    assert np.abs(got.cpu().numpy() - c["confmat"]).max() <= slack
    half = p.shape[0] // 2
    cls.update(p[:half], t[:half])
    cls.update(p[half:], t[half:])
    assert torch.equal(cls.compute(), got)


def test_confusion_matrix_validation_errors():
    tm = _tm()
    F = tm.functional
    p = torch.randint(0, 5, (100,), device=DEV)
    with pytest.raises(ValueError, match="Expected argument `num_classes` to be an integer larger than 1"):
        F.multiclass_confusion_matrix(p, p, 1)
    with pytest.raises(RuntimeError, match="Detected more unique values in `target` than `num_classes`"):
        F.multiclass_confusion_matrix(p % 3, torch.arange(100, device=DEV), 3)
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        F.binary_confusion_matrix(torch.rand(10, device=DEV), torch.arange(10, device=DEV))
    with pytest.raises(ValueError, match="Expected argument `threshold` to be a float"):
        F.binary_confusion_matrix(torch.rand(10, device=DEV), torch.zeros(10, device=DEV).long(), threshold=2)
    # validate_args=False: out-of-range labels surface as RuntimeError (the reference's reshape error)
    with pytest.raises(RuntimeError):
        F.multiclass_confusion_matrix(p, p + 3, 5, validate_args=False)
    m = tm.MulticlassConfusionMatrix(5, validate_args=False).to(DEV)
    m.update(p, p + 3)
    with pytest.raises(RuntimeError):
        m.compute()
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        F.multiclass_confusion_matrix(p.cpu(), p.cpu(), 5)


def test_confusion_matrix_nan_warning():
    tm = _tm()
    t = torch.tensor([0, 0, 1, 1], device=DEV)
    with pytest.warns(UserWarning, match="NaN values found in confusion matrix have been replaced with zeros"):
        out = tm.functional.multiclass_confusion_matrix(t, t, 3, normalize="true")
    assert float(out[2].sum()) == 0


# ------------------------------------------------------------------ stat scores and derived metrics
@pytest.mark.parametrize("c", load_golden("stat_scores"), ids=case_id)
def test_stat_scores_family(c):
    tm = _tm()
    F = tm.functional
    kind, ii, mda = c.meta["kind"], c.meta["ignore_index"], c.meta["multidim_average"]
    p, t = _t(c["preds"]), _t(c["target"])
    tol = 1e-6
    if kind.startswith("binary"):
        kw = dict(threshold=c.meta["threshold"], multidim_average=mda, ignore_index=ii, validate_args=False)
        _close(F.binary_stat_scores(p, t, **kw), c["stat_scores"])
        for name in ["accuracy", "precision", "recall", "f1_score"]:
            _close(getattr(F, "binary_" + name)(p, t, **kw), c[name], tol)
        _close(F.binary_fbeta_score(p, t, beta=2.0, **kw), c["fbeta2"], tol)
        m = tm.BinaryF1Score(**kw).to(DEV)
        m.update(p, t)
        _close(m.compute(), c["f1_score"], tol)
    elif kind.startswith("multilabel"):
        L = c.meta["num_labels"]
        for avg in ["micro", "macro", "weighted", "none"]:
            kw = dict(num_labels=L, threshold=c.meta["threshold"], average=avg, multidim_average=mda, ignore_index=ii,
                      validate_args=False)
            _close(F.multilabel_stat_scores(p, t, **kw), c["stat_scores_" + avg], tol)
            for name in ["accuracy", "precision", "recall", "f1_score"]:
                _close(getattr(F, "multilabel_" + name)(p, t, **kw), c[f"{name}_{avg}"], tol)
        m = tm.MultilabelRecall(L, c.meta["threshold"], "macro", multidim_average=mda, ignore_index=ii).to(DEV)
        m.update(p, t)
        _close(m.compute(), c["recall_macro"], tol)
    else:
        C, top_k = c.meta["num_classes"], c.meta["top_k"]
        for avg in ["micro", "macro", "weighted", "none"]:
            kw = dict(num_classes=C, average=avg, top_k=top_k, multidim_average=mda, ignore_index=ii,
                      validate_args=False)
            _close(F.multiclass_stat_scores(p, t, **kw), c["stat_scores_" + avg], tol)
            for name in ["accuracy", "precision", "recall", "f1_score"]:
                _close(getattr(F, "multiclass_" + name)(p, t, **kw), c[f"{name}_{avg}"], tol)
        m = tm.MulticlassPrecision(C, top_k, "weighted", multidim_average=mda, ignore_index=ii).to(DEV)
        m.update(p, t)
        _close(m.compute(), c["precision_weighted"], tol)


# ------------------------------------------------------------------ binned curves
@pytest.mark.parametrize("c", load_golden("binned_curve"), ids=case_id)
def test_binned_curve_family(c):
    tm = _tm()
    F = tm.functional
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    arg = c.meta["thresholds_arg"]
    thr = arg if isinstance(arg, int) else _t(c["thresholds"])
    p, t = _t(c["preds"]), _t(c["target"])
    tol = 2e-4 if kind.endswith("logits") else 1e-6
    if p.numel() == 0:
        pytest.skip("empty")
    if kind.startswith("binary"):
        kw = dict(thresholds=thr, ignore_index=ii, validate_args=False)
        _close(F.binary_auroc(p, t, **kw), c["auroc"], tol)
        _close(F.binary_average_precision(p, t, **kw), c["ap"], tol)
        _close(F.binary_auroc(p, t, max_fpr=0.3, **kw), c["auroc_maxfpr"], 10 * tol)
        pr, rc, th = F.binary_precision_recall_curve(p, t, **kw)
        _close(pr, c["precision"], tol)
        _close(rc, c["recall"], tol)
        _close(th, c["thresholds"], 0)
        fpr, tpr, _ = F.binary_roc(p, t, **kw)
        _close(fpr, c["fpr"], tol)
        _close(tpr, c["tpr"], tol)
        # class API over two batches, sharing one state between AUROC and AP in a collection
        mc = tm.MetricCollection([tm.BinaryAUROC(thresholds=thr, ignore_index=ii, validate_args=False),
                                  tm.BinaryAveragePrecision(thresholds=thr, ignore_index=ii, validate_args=False)]).to(DEV)
        h = p.shape[0] // 2
        if kind.endswith("logits") or h == 0:
            mc.update(p, t)  # the sigmoid rule is per batch, keep the batch as the reference saw it
        else:
            mc.update(p[:h], t[:h])
            mc.update(p[h:], t[h:])
        out = mc.compute()
        assert len(mc.compute_groups) == 1
        _close(out["BinaryAUROC"], c["auroc"], tol)
        _close(out["BinaryAveragePrecision"], c["ap"], tol)
        if not kind.endswith("logits"):
            st = mc.__getitem__("BinaryAUROC", copy_state=False).confmat
            assert np.array_equal(st.cpu().numpy(), c["confmat"])
    elif kind.startswith("multiclass"):
        C = c.meta["num_classes"]
        kw = dict(num_classes=C, thresholds=thr, ignore_index=ii, validate_args=False)
        for avg in ["macro", "weighted", "none"]:
            _close(F.multiclass_auroc(p, t, average=avg, **kw), c["auroc_" + avg], tol)
            _close(F.multiclass_average_precision(p, t, average=avg, **kw), c["ap_" + avg], tol)
        m = tm.MulticlassAUROC(C, "macro", thresholds=thr, ignore_index=ii, validate_args=False).to(DEV)
        m.update(p, t)
        _close(m.compute(), c["auroc_macro"], tol)
        if not kind.endswith("logits"):
            assert np.array_equal(m.confmat.cpu().numpy(), c["confmat_None"])
            mm = tm.MulticlassPrecisionRecallCurve(C, thresholds=thr, average="micro", ignore_index=ii,
                                                   validate_args=False).to(DEV)
            mm.update(p, t)
            assert np.array_equal(mm.confmat.cpu().numpy(), c["confmat_micro"])
    else:
        L = c.meta["num_labels"]
        kw = dict(num_labels=L, thresholds=thr, ignore_index=ii, validate_args=False)
        for avg in ["micro", "macro", "weighted", "none"]:
            _close(F.multilabel_auroc(p, t, average=avg, **kw), c["auroc_" + avg], tol)
            _close(F.multilabel_average_precision(p, t, average=avg, **kw), c["ap_" + avg], tol)
        m = tm.MultilabelAveragePrecision(L, "macro", thresholds=thr, ignore_index=ii, validate_args=False).to(DEV)
        m.update(p, t)
        _close(m.compute(), c["ap_macro"], tol)


# ------------------------------------------------------------------ exact curves
@pytest.mark.parametrize("c", load_golden("exact_curve"), ids=case_id)
def test_exact_curve_family(c):
    tm = _tm()
    F = tm.functional
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    p, t = _t(c["preds"]), _t(c["target"])
    tol = 1e-5 if kind.endswith("logits") else 1e-6
    if kind.startswith("binary"):
        kw = dict(thresholds=None, ignore_index=ii, validate_args=False)
        _close(F.binary_auroc(p, t, **kw), c["auroc"], tol)
        _close(F.binary_average_precision(p, t, **kw), c["ap"], tol)
        _close(F.binary_auroc(p, t, max_fpr=0.25, **kw), c["auroc_maxfpr"], 10 * tol)
        fpr, tpr, rthr = F.binary_roc(p, t, **kw)
        _close(fpr, c["fpr"], tol)
        _close(tpr, c["tpr"], tol)
        _close(rthr, c["roc_thresholds"], 1e-6)
        pr, rc, pthr = F.binary_precision_recall_curve(p, t, **kw)
        _close(pr, c["precision"], tol)
        _close(rc, c["recall"], tol)
        _close(pthr, c["pr_thresholds"], 1e-6)
        if not kind.endswith("logits"):
            from torchmetrics_b200.functional.classification.precision_recall_curve import (
                _binary_clf_curve,
                _binary_precision_recall_curve_format,
            )

            fp_, ft_, _ = _binary_precision_recall_curve_format(p, t, None, ii)
            fps, tps, thr = _binary_clf_curve(fp_, ft_)
            _close(fps, c["fps"], 0)
            _close(tps, c["tps"], 0)
            _close(thr, c["thr"], 0)
        m = tm.BinaryAUROC(thresholds=None, ignore_index=ii, validate_args=False).to(DEV)
        h = max(1, p.shape[0] // 3)
        if kind.endswith("logits"):
            m.update(p, t)
        else:
            m.update(p[:h], t[:h])
            if p.shape[0] > h:
                m.update(p[h:], t[h:])
        _close(m.compute(), c["auroc"], tol)
    elif kind == "multiclass":
        C = c.meta["num_classes"]
        kw = dict(num_classes=C, thresholds=None, ignore_index=ii, validate_args=False)
        for avg in ["macro", "weighted", "none"]:
            _close(F.multiclass_auroc(p, t, average=avg, **kw), c["auroc_" + avg], tol)
            _close(F.multiclass_average_precision(p, t, average=avg, **kw), c["ap_" + avg], tol)
    else:
        L = c.meta["num_labels"]
        kw = dict(num_labels=L, thresholds=None, ignore_index=ii, validate_args=False)
        for avg in ["micro", "macro", "weighted", "none"]:
            _close(F.multilabel_auroc(p, t, average=avg, **kw), c["auroc_" + avg], tol)
            _close(F.multilabel_average_precision(p, t, average=avg, **kw), c["ap_" + avg], tol)


def test_exact_sort_large_random_vs_torch():
    """K4 against torch.sort-based restatement on the device at a size the oracle would take minutes for."""
    from torchmetrics_b200 import _native as N

    g = torch.Generator(device=DEV).manual_seed(3)
    for n in [1, 2, 4095, 4096, 4097, 1 << 20, (1 << 22) + 77]:
        preds = torch.rand(n, device=DEV, generator=g)
        if n > 10000:
            preds = (preds * 50000).floor() / 50000  # plenty of ties
        target = (torch.rand(n, device=DEV, generator=g) < 0.4).long()
        fps, tps, thr = N.exact_curve(preds, target, 1)
        sp, order = torch.sort(preds, descending=True, stable=True)
        st = target[order]
        ends = torch.cat([torch.nonzero(sp[1:] != sp[:-1]).flatten(), torch.tensor([n - 1], device=DEV)])
        want_tps = torch.cumsum(st, 0)[ends]
        assert torch.equal(thr, sp[ends])
        assert torch.equal(tps, want_tps)
        assert torch.equal(fps, ends + 1 - want_tps)
    # negative scores, infinities and both zeros sort correctly too
    preds = torch.tensor([0.0, -0.0, -1.5, 3.0, float("inf"), -float("inf"), 1e-40, -1e-40, 3.0], device=DEV)
    target = torch.tensor([1, 0, 1, 0, 1, 0, 1, 0, 1], device=DEV)
    fps, tps, thr = N.exact_curve(preds, target, 1)
    assert thr.tolist() == [float("inf"), 3.0, 1e-40, 0.0, -1e-40, -1.5, -float("inf")] or \
        [abs(x) for x in thr.tolist()][3] == 0.0
    assert tps.tolist() == [1, 2, 3, 4, 4, 5, 5] and fps.tolist() == [0, 1, 1, 2, 3, 3, 4]


def test_binned_with_all_scores_as_thresholds_equals_exact():
    """The reference's own cross-path identity (test_precision_recall_curve.py:132-143, test_auroc.py:129-138)."""
    tm = _tm()
    F = tm.functional
    g = torch.Generator(device=DEV).manual_seed(8)
    p = torch.rand(500, device=DEV, generator=g)
    t = torch.randint(0, 2, (500,), device=DEV, generator=g)
    thr = torch.unique(p)
    assert abs(float(F.binary_auroc(p, t, thresholds=None)) - float(F.binary_auroc(p, t, thresholds=thr))) < 1e-6
    a = F.binary_average_precision(p, t, thresholds=None)
    b = F.binary_average_precision(p, t, thresholds=thr)
    assert abs(float(a) - float(b)) < 1e-6


# ------------------------------------------------------------------ scikit-learn cross-checks on fresh inputs
def test_against_sklearn():
    sk = pytest.importorskip("sklearn.metrics")
    tm = _tm()
    F = tm.functional
    rng = np.random.default_rng(0)
    n = 20_000
    p = rng.random(n).astype(np.float32)
    t = (rng.random(n) < 0.3 + 0.4 * p).astype(np.int64)
    assert abs(float(F.binary_auroc(_t(p), _t(t))) - sk.roc_auc_score(t, p)) < 1e-5
    assert abs(float(F.binary_average_precision(_t(p), _t(t))) - sk.average_precision_score(t, p)) < 1e-5
    assert abs(float(F.binary_auroc(_t(p), _t(t), thresholds=2000)) - sk.roc_auc_score(t, p)) < 2e-3
    C = 37
    yp, yt = rng.integers(0, C, n), rng.integers(0, C, n)
    got = F.multiclass_confusion_matrix(_t(yp), _t(yt), C).cpu().numpy()
    assert np.array_equal(got, sk.confusion_matrix(yt, yp, labels=list(range(C))))
    assert abs(float(F.multiclass_f1_score(_t(yp), _t(yt), C, average="macro"))
               - sk.f1_score(yt, yp, average="macro")) < 1e-6
    assert abs(float(F.multiclass_accuracy(_t(yp), _t(yt), C, average="micro")) - sk.accuracy_score(yt, yp)) < 1e-6
    L = 9
    mp, mt = rng.random((n, L)).astype(np.float32), rng.integers(0, 2, (n, L))
    assert abs(float(F.multilabel_precision(_t(mp), _t(mt), L, average="macro"))
               - sk.precision_score(mt, mp > 0.5, average="macro")) < 1e-6
    assert abs(float(F.multilabel_auroc(_t(mp), _t(mt), L, average="macro"))
               - sk.roc_auc_score(mt, mp, average="macro")) < 1e-5


# ------------------------------------------------------------------ config 5: collection of multilabel metrics
def test_multilabel_collection_config5_shape():
    tm = _tm()
    L, n = 100, 4096
    g = torch.Generator(device=DEV).manual_seed(5)
    mc = tm.MetricCollection([tm.MultilabelStatScores(num_labels=L), tm.MultilabelPrecision(num_labels=L),
                              tm.MultilabelRecall(num_labels=L), tm.MultilabelF1Score(num_labels=L)]).to(DEV)
    P, T = [], []
    for _ in range(3):
        p = torch.rand(n, L, device=DEV, generator=g)
        t = torch.randint(0, 2, (n, L), device=DEV, generator=g)
        mc.update(p, t)
        P.append(p)
        T.append(t)
    assert len(mc.compute_groups) == 1
    out = mc.compute()
    p, t = torch.cat(P), torch.cat(T)
    F = tm.functional
    _close(out["MultilabelStatScores"], F.multilabel_stat_scores(p, t, L).cpu().numpy(), 1e-6)
    _close(out["MultilabelPrecision"], F.multilabel_precision(p, t, L).cpu().numpy(), 1e-6)
    _close(out["MultilabelRecall"], F.multilabel_recall(p, t, L).cpu().numpy(), 1e-6)
    _close(out["MultilabelF1Score"], F.multilabel_f1_score(p, t, L).cpu().numpy(), 1e-6)
    pb = (p > 0.5).long()
    tp = ((pb == 1) & (t == 1)).sum(0)
    st = mc.__getitem__("MultilabelStatScores", copy_state=False)
    assert torch.equal(st.tp, tp)
    # reset keeps the group, state goes back to zero on the same device
    mc.reset()
    assert int(st.tp.sum()) == 0 and st.tp.is_cuda


def test_metric_device_moves_and_pickle_on_gpu():
    import pickle

    tm = _tm()
    m = tm.BinaryAUROC(thresholds=50)
    assert m.confmat.device.type == "cpu"
    m = m.to(DEV)
    assert m.confmat.is_cuda and m.thresholds.is_cuda
    p = torch.rand(1000, device=DEV)
    t = torch.randint(0, 2, (1000,), device=DEV)
    m.update(p, t)
    v = float(m.compute())
    m2 = pickle.loads(pickle.dumps(m))
    assert abs(float(m2.compute()) - v) < 1e-7
    m3 = m.clone()
    m3.update(p, t)
    assert abs(float(m3.compute()) - v) < 1e-6  # same data twice: same curve
    assert m.state_dict() == {}
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        tm.BinaryAUROC(thresholds=50).update(p.cpu(), t.cpu())


# ------------------------------------------------------------------ fused value validation (validate_args=True)
def _count_slow_checks(monkeypatch):
    """Instrument the reference-style value checks so a test can tell whether the slow path ran."""
    from importlib import import_module

    C = import_module("torchmetrics_b200.functional.classification._checks")
    calls = {"n": 0}
    for name in ("_check_binary_values_now", "_check_label_count_now"):
        orig = getattr(C, name)

        def wrapped(*a, _orig=orig, **k):
            calls["n"] += 1
            return _orig(*a, **k)

        monkeypatch.setattr(C, name, wrapped)
    return calls


def test_fused_validation_skips_input_passes_on_good_data(monkeypatch):
    tm = _tm()
    calls = _count_slow_checks(monkeypatch)
    g = torch.Generator().manual_seed(3)
    n = 4096
    t5 = torch.randint(0, 5, (n,), generator=g).to(DEV)
    p5 = torch.randint(0, 5, (n,), generator=g).to(DEV)
    tb = torch.randint(0, 2, (n,), generator=g).to(DEV)
    ps = torch.rand(n, generator=g).to(DEV)
    tl = torch.randint(0, 2, (n, 7), generator=g).to(DEV)
    pl = torch.rand(n, 7, generator=g).to(DEV)
    pc = torch.rand(n, 5, generator=g).to(DEV)
    F = tm.functional
    F.multiclass_confusion_matrix(p5, t5, 5)
    F.binary_confusion_matrix(ps, tb)
    F.binary_confusion_matrix((ps > 0.5).long(), tb)
    F.multilabel_confusion_matrix(pl, tl, 7)
    F.binary_stat_scores(ps, tb)
    F.multiclass_stat_scores(p5, t5, 5, average="macro")
    F.multilabel_stat_scores(pl, tl, 7)
    F.multiclass_accuracy(p5, t5, 5, average="macro")
    F.binary_auroc(ps, tb, thresholds=100)
    F.multiclass_auroc(pc, t5, 5, thresholds=50)
    F.multilabel_average_precision(pl, tl, 7, thresholds=50)
    for m in (tm.MulticlassConfusionMatrix(5), tm.classification.MulticlassF1Score(5),
              tm.classification.BinaryAUROC(thresholds=64)):
        m = m.to(DEV)
        if isinstance(m, tm.classification.BinaryAUROC):
            m.update(ps, tb)
        else:
            m.update(p5, t5)
        m.compute()
    assert calls["n"] == 0, "value checks re-read the inputs although the kernels reported nothing"
    # the paths whose kernels do not report values still validate up front
    F.multiclass_stat_scores(p5, t5, 5, average="micro")
    F.binary_auroc(ps, tb, thresholds=None)
    assert calls["n"] > 0


def test_fused_validation_raises_like_the_reference_and_restores_state(monkeypatch):
    tm = _tm()
    F = tm.functional
    calls = _count_slow_checks(monkeypatch)
    g = torch.Generator().manual_seed(4)
    n = 2048
    tb = torch.randint(0, 2, (n,), generator=g).to(DEV)
    ps = torch.rand(n, generator=g).to(DEV)
    bad_tb = tb.clone()
    bad_tb[17] = 3
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        F.binary_stat_scores(ps, bad_tb)
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        F.binary_confusion_matrix(ps, bad_tb)
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        F.binary_auroc(ps, bad_tb, thresholds=20)
    with pytest.raises(RuntimeError, match="Detected the following values in `preds`"):
        F.binary_stat_scores(torch.full((n,), 2, device=DEV), tb)
    assert calls["n"] >= 4
    # with ignore_index the same value is legal
    assert torch.equal(F.binary_stat_scores(ps, bad_tb, ignore_index=3),
                       F.binary_stat_scores(ps, bad_tb, ignore_index=3, validate_args=False))

    # multiclass: too many distinct values
    t = torch.arange(n, device=DEV) % 9
    p = torch.arange(n, device=DEV) % 5
    with pytest.raises(RuntimeError, match="Detected more unique values in `target` than `num_classes`"):
        F.multiclass_confusion_matrix(p, t, 5)
    with pytest.raises(RuntimeError, match="Detected more unique values in `preds` than `num_classes`"):
        F.multiclass_stat_scores(t, p, 5, average="macro")

    # stateful metrics: a rejected batch leaves the state as it was
    good_t = torch.randint(0, 5, (n,), generator=g).to(DEV)
    good_p = torch.randint(0, 5, (n,), generator=g).to(DEV)
    cm = tm.MulticlassConfusionMatrix(5).to(DEV)
    cm.update(good_p, good_t)
    before = cm.confmat.clone()
    with pytest.raises(RuntimeError, match="Detected more unique values"):
        cm.update(p, t)
    assert torch.equal(cm.confmat, before)
    cm.update(good_p, good_t)
    assert torch.equal(cm.compute(), 2 * before)

    au = tm.classification.BinaryAUROC(thresholds=32).to(DEV)
    au.update(ps, tb)
    before = au.confmat.clone()
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        au.update(ps, bad_tb)
    assert torch.equal(au.confmat, before)

    f1 = tm.classification.BinaryF1Score().to(DEV)
    f1.update(ps, tb)
    before = (f1.tp.clone(), f1.fn.clone())
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        f1.update(ps, bad_tb)
    assert torch.equal(f1.tp, before[0]) and torch.equal(f1.fn, before[1])

    # a stale finding left by an unvalidated call costs one slow check but raises nothing
    F.binary_stat_scores(ps, bad_tb, validate_args=False)
    F.binary_stat_scores(ps, tb)


def test_deferred_validation_raises_one_call_late_and_restores_state():
    """validate_args="deferred": no host sync per update; a bad batch is reported (with the reference's message) at the
    next update / compute / reset, after its counts have been taken out of the state again."""
    tm = _tm()
    g = torch.Generator().manual_seed(4)
    n = 2048
    good_t = torch.randint(0, 5, (n,), generator=g).to(DEV)
    good_p = torch.randint(0, 5, (n,), generator=g).to(DEV)
    bad_t = torch.arange(n, device=DEV) % 9
    cm = tm.MulticlassConfusionMatrix(5, validate_args="deferred").to(DEV)
    cm.update(good_p, good_t)
    one = cm.confmat.clone()
    cm.update(good_p, bad_t)          # nothing raised yet
    with pytest.raises(RuntimeError, match="Detected more unique values in `target` than `num_classes`"):
        cm.update(good_p, good_t)     # the verdict of the previous batch; this batch itself is applied
    assert torch.equal(cm.confmat, 2 * one)
    cm.update(good_p, bad_t)
    with pytest.raises(RuntimeError, match="Detected more unique values"):
        cm.compute()
    assert torch.equal(cm.compute(), 2 * one)
    cm.update(good_p, bad_t)
    with pytest.raises(RuntimeError, match="Detected more unique values"):
        cm.reset()
    # curve classes take the same route
    tb = torch.randint(0, 2, (n,), generator=g).to(DEV)
    ps = torch.rand(n, generator=g).to(DEV)
    bad_tb = tb.clone()
    bad_tb[17] = 3
    au = tm.classification.BinaryAUROC(thresholds=32, validate_args="deferred").to(DEV)
    au.update(ps, tb)
    one = au.confmat.clone()
    au.update(ps, bad_tb)
    with pytest.raises(RuntimeError, match="Detected the following values in `target`"):
        au.compute()
    assert torch.equal(au.confmat, one)
    import pickle

    au.update(ps, tb)
    assert torch.equal(pickle.loads(pickle.dumps(au)).confmat, 2 * one)  # an outstanding verdict is not pickled


def test_multiclass_binned_curve_rejects_targets_outside_the_classes():
    """The multiclass binned family reads the error word in the middle of its update (softmax decision); the
    fused-validation finding next to it must survive until `finish` (reference: the `torch.unique` check in
    precision_recall_curve.py:382-420 raises before any state is touched)."""
    tm = _tm()
    F = tm.functional
    g = torch.Generator().manual_seed(8)
    n, C = 4096, 5
    p = torch.rand(n, C, generator=g).softmax(1).to(DEV)
    logits = (torch.randn(n, C, generator=g) * 3).to(DEV)
    t = torch.randint(0, C, (n,), generator=g).to(DEV)
    bad = t.clone()
    bad[::7] = 7
    msg = "Detected more unique values in `target` than `num_classes`"
    for scores in (p, logits):
        with pytest.raises(RuntimeError, match=msg):
            F.multiclass_auroc(scores, bad, C, thresholds=50)
        with pytest.raises(RuntimeError, match=msg):
            F.multiclass_precision_recall_curve(scores, bad, C, thresholds=50)
        with pytest.raises(RuntimeError, match=msg):
            F.multiclass_average_precision(scores, bad, C, thresholds=50)
        for cls in (tm.classification.MulticlassPrecisionRecallCurve, tm.classification.MulticlassAUROC,
                    tm.classification.MulticlassROC, tm.classification.MulticlassAveragePrecision):
            m = cls(num_classes=C, thresholds=50).to(DEV)
            m.update(scores, t)
            before = m.confmat.clone()
            with pytest.raises(RuntimeError, match=msg):
                m.update(scores, bad)
            assert torch.equal(m.confmat, before)
            m.update(scores, t)
            assert torch.equal(m.confmat, 2 * before)
    # the same value is legal as ignore_index
    a = F.multiclass_auroc(p, bad, C, thresholds=50, ignore_index=7)
    b = F.multiclass_auroc(p, bad, C, thresholds=50, ignore_index=7, validate_args=False)
    assert torch.equal(a, b)


# ------------------------------------------------------------------ N2: consumers of the same update kernels
_AVGS = ("micro", "macro", "weighted", "none")


@pytest.mark.parametrize("c", load_golden("derived_metrics"), ids=case_id)
def test_derived_metrics_match_reference(c):
    """specificity / hamming (stat-scores state), jaccard / cohen kappa / matthews (confusion-matrix state): functional
    and class results against the unmodified reference's outputs, plus two-batch accumulation in the classes."""
    tm = _tm()
    F = tm.functional
    task, ii = c.meta["task"], c.meta["ignore_index"]
    p, t = _t(c["preds"]), _t(c["target"])
    h = p.shape[0] // 2

    def cls_value(metric):
        metric = metric.to(DEV)
        metric.update(p[:h], t[:h])
        metric.update(p[h:], t[h:])
        return metric.compute()

    def check(name, fn_value, metric):
        want = c[name]
        _close(fn_value, want, atol=2e-6)
        _close(cls_value(metric), want, atol=2e-6)

    if task == "binary":
        check("specificity", F.binary_specificity(p, t, ignore_index=ii), tm.BinarySpecificity(ignore_index=ii))
        check("hamming", F.binary_hamming_distance(p, t, ignore_index=ii), tm.BinaryHammingDistance(ignore_index=ii))
        check("jaccard", F.binary_jaccard_index(p, t, ignore_index=ii), tm.BinaryJaccardIndex(ignore_index=ii))
        check("mcc", F.binary_matthews_corrcoef(p, t, ignore_index=ii), tm.BinaryMatthewsCorrCoef(ignore_index=ii))
        for w in ("none", "linear", "quadratic"):
            if f"kappa_{w}" in c:
                check(f"kappa_{w}", F.binary_cohen_kappa(p, t, weights=w, ignore_index=ii),
                      tm.BinaryCohenKappa(weights=w, ignore_index=ii))
        # task wrappers reach the same classes
        _close(F.jaccard_index(p, t, task="binary", ignore_index=ii), c["jaccard"], atol=2e-6)
        assert isinstance(tm.MatthewsCorrCoef(task="binary"), tm.BinaryMatthewsCorrCoef)
    elif task == "multiclass":
        C = c.meta["num_classes"]
        for a in _AVGS:
            check(f"specificity_{a}", F.multiclass_specificity(p, t, C, average=a, ignore_index=ii),
                  tm.MulticlassSpecificity(C, average=a, ignore_index=ii))
            check(f"hamming_{a}", F.multiclass_hamming_distance(p, t, C, average=a, ignore_index=ii),
                  tm.MulticlassHammingDistance(C, average=a, ignore_index=ii))
            check(f"jaccard_{a}", F.multiclass_jaccard_index(p, t, C, average=a, ignore_index=ii),
                  tm.MulticlassJaccardIndex(C, average=a, ignore_index=ii))
        for w in ("none", "linear", "quadratic"):
            check(f"kappa_{w}", F.multiclass_cohen_kappa(p, t, C, weights=w, ignore_index=ii),
                  tm.MulticlassCohenKappa(C, weights=w, ignore_index=ii))
        check("mcc", F.multiclass_matthews_corrcoef(p, t, C, ignore_index=ii),
              tm.MulticlassMatthewsCorrCoef(C, ignore_index=ii))
        _close(F.cohen_kappa(p, t, task="multiclass", num_classes=C, ignore_index=ii), c["kappa_none"], atol=2e-6)
        _close(F.specificity(p, t, task="multiclass", num_classes=C, average="macro", ignore_index=ii),
               c["specificity_macro"], atol=2e-6)
    else:
        L = c.meta["num_labels"]
        for a in _AVGS:
            check(f"specificity_{a}", F.multilabel_specificity(p, t, L, average=a, ignore_index=ii),
                  tm.MultilabelSpecificity(L, average=a, ignore_index=ii))
            check(f"hamming_{a}", F.multilabel_hamming_distance(p, t, L, average=a, ignore_index=ii),
                  tm.MultilabelHammingDistance(L, average=a, ignore_index=ii))
            check(f"jaccard_{a}", F.multilabel_jaccard_index(p, t, L, average=a, ignore_index=ii),
                  tm.MultilabelJaccardIndex(L, average=a, ignore_index=ii))
        check("mcc", F.multilabel_matthews_corrcoef(p, t, L, ignore_index=ii),
              tm.MultilabelMatthewsCorrCoef(L, ignore_index=ii))
        _close(F.hamming_distance(p, t, task="multilabel", num_labels=L, average="macro", ignore_index=ii),
               c["hamming_macro"], atol=2e-6)


def test_derived_metrics_share_compute_groups_and_validate():
    tm = _tm()
    g = torch.Generator().manual_seed(8)
    C = 7
    t = torch.randint(0, C, (500,), generator=g).to(DEV)
    p = torch.randint(0, C, (500,), generator=g).to(DEV)
    coll = tm.MetricCollection([tm.MulticlassConfusionMatrix(C), tm.MulticlassJaccardIndex(C),
                                tm.MulticlassCohenKappa(C), tm.MulticlassMatthewsCorrCoef(C),
                                tm.MulticlassSpecificity(C), tm.MulticlassHammingDistance(C),
                                tm.MulticlassAccuracy(C)]).to(DEV)
    coll.update(p, t)
    coll.update(p, t)
    groups = sorted(len(v) for v in coll.compute_groups.values())
    assert groups == [3, 4]  # one confusion-matrix group, one stat-scores group: two update kernels per batch
    out = coll.compute()
    F = tm.functional
    _close(out["MulticlassJaccardIndex"], F.multiclass_jaccard_index(p, t, C).cpu().numpy())
    _close(out["MulticlassSpecificity"], F.multiclass_specificity(p, t, C).cpu().numpy())
    with pytest.raises(ValueError, match="Expected argument `weight` to be one of"):
        tm.MulticlassCohenKappa(C, weights="cubic")
    with pytest.raises(ValueError, match="Expected argument `average` to be one of"):
        tm.MulticlassJaccardIndex(C, average="samples")
    with pytest.raises(RuntimeError, match="Detected more unique values in `target`"):
        F.multiclass_matthews_corrcoef(p, t + C * (torch.arange(500, device=DEV) % 2), C)


# ------------------------------------------------------------------ packed count states (in-place accumulation)
@pytest.mark.gpu
@pytest.mark.parametrize("validate", [False, True])
def test_stat_scores_states_accumulate_in_place(validate):
    """The four count states are re-pointed at one buffer the kernel accumulates into; everything that replaces or
    copies the state tensors (reset, forward, deepcopy, pickle, collections) must keep giving the reference's counts."""
    import copy
    import pickle

    tm = _tm()
    F = tm.functional
    L = 64  # 512-byte states: separately allocated results can sit next to each other in the allocator
    g = torch.Generator().manual_seed(11)
    batches = [(torch.rand(96, L, generator=g).to(DEV), torch.randint(0, 2, (96, L), generator=g).to(DEV))
               for _ in range(4)]

    def want(idx):
        p = torch.cat([batches[i][0] for i in idx])
        t = torch.cat([batches[i][1] for i in idx])
        return F.multilabel_stat_scores(p, t, L, average=None, validate_args=False)

    m = tm.MultilabelStatScores(num_labels=L, average=None, validate_args=validate).to(DEV)
    m.update(*batches[0])
    m.update(*batches[1])
    assert torch.equal(m.compute(), want([0, 1]))
    if not validate:  # consecutive rows of one buffer
        assert m.fp.data_ptr() == m.tp.data_ptr() + 8 * L and m.fn.data_ptr() == m.tp.data_ptr() + 24 * L
    batch_value = m(*batches[2])  # forward: batch value, and the batch joins the running state
    assert torch.equal(batch_value, want([2]))
    assert torch.equal(m.compute(), want([0, 1, 2]))
    for clone in (copy.deepcopy(m), pickle.loads(pickle.dumps(m))):
        clone.update(*batches[3])
        assert torch.equal(clone.compute(), want([0, 1, 2, 3]))
    assert torch.equal(m.compute(), want([0, 1, 2]))  # the copies did not write into the original
    with torch.inference_mode():  # states created here are inference tensors: keep updating them in here
        m.reset()
        m.update(*batches[3])
        m.update(*batches[0])
        m.update(*batches[1])
        assert torch.equal(m.compute(), want([3, 0, 1]))
    m.reset()
    with torch.inference_mode():  # ordinary states, updated under inference mode and after it
        m.update(*batches[3])
    m.update(*batches[0])
    assert torch.equal(m.compute(), want([3, 0]))

    mc = tm.MetricCollection([tm.MultilabelStatScores(num_labels=L, average=None, validate_args=validate),
                              tm.MultilabelF1Score(num_labels=L, average=None, validate_args=validate),
                              tm.MultilabelRecall(num_labels=L, average=None, validate_args=validate)]).to(DEV)
    for rounds in range(2):
        mc.update(*batches[0])
        mc.update(*batches[1])
        out = mc.compute()
        assert torch.equal(out["MultilabelStatScores"], want([0, 1]))
        p = torch.cat([batches[0][0], batches[1][0]])
        t = torch.cat([batches[0][1], batches[1][1]])
        assert torch.equal(out["MultilabelF1Score"], F.multilabel_f1_score(p, t, L, average=None, validate_args=False))
        assert torch.equal(out["MultilabelRecall"], F.multilabel_recall(p, t, L, average=None, validate_args=False))
        mc.reset()

    # binary (one number per state) and multiclass (labels, scores, micro) take the same route
    b = tm.BinaryStatScores(validate_args=validate).to(DEV)
    mcl = tm.MulticlassStatScores(num_classes=L, average=None, validate_args=validate).to(DEV)
    mcs = tm.MulticlassStatScores(num_classes=L, average=None, validate_args=validate).to(DEV)
    mic = tm.MulticlassStatScores(num_classes=L, average="micro", validate_args=validate).to(DEV)
    lab = [(torch.randint(0, L, (500,), generator=g).to(DEV), torch.randint(0, L, (500,), generator=g).to(DEV))
           for _ in range(2)]
    sc = [(torch.randn(500, L, generator=g).to(DEV), lab[i][1]) for i in range(2)]
    for i in range(2):
        b.update(batches[i][0][:, 0], batches[i][1][:, 0])
        mcl.update(*lab[i])
        mcs.update(*sc[i])
        mic.update(*lab[i])
    bp, bt = torch.cat([batches[0][0][:, 0], batches[1][0][:, 0]]), torch.cat([batches[0][1][:, 0], batches[1][1][:, 0]])
    assert torch.equal(b.compute(), F.binary_stat_scores(bp, bt, validate_args=False))
    lp, lt = torch.cat([lab[0][0], lab[1][0]]), torch.cat([lab[0][1], lab[1][1]])
    assert torch.equal(mcl.compute(), F.multiclass_stat_scores(lp, lt, L, average=None, validate_args=False))
    assert torch.equal(mic.compute(), F.multiclass_stat_scores(lp, lt, L, average="micro", validate_args=False))
    assert torch.equal(mcs.compute(), F.multiclass_stat_scores(torch.cat([sc[0][0], sc[1][0]]), lt, L, average=None,
                                                               validate_args=False))


@pytest.mark.gpu
def test_stat_scores_bad_batch_leaves_state_untouched():
    """validate_args=True: a batch that fails validation is reported before the states change (the kernel then counts
    into a scratch tensor, not into the states)."""
    tm = _tm()
    L = 8
    m = tm.MultilabelStatScores(num_labels=L, average=None).to(DEV)
    good_p, good_t = torch.rand(32, L, device=DEV), torch.randint(0, 2, (32, L), device=DEV)
    m.update(good_p, good_t)
    before = m.compute().clone()
    bad_t = good_t.clone()
    bad_t[3, 2] = 7
    with pytest.raises(RuntimeError):
        m.update(good_p, bad_t)
    assert torch.equal(m.compute(), before)


@pytest.mark.gpu
@pytest.mark.parametrize("validate", [False, True])
@pytest.mark.parametrize("groups_first", [False, True])
def test_collection_forward_batch_values_and_running_state(groups_first, validate):
    """forward() of the stat-score family (alone, and as a compute group served by one update) returns the batch's own
    value and leaves the running state equal to plain accumulation."""
    tm = _tm()
    F = tm.functional
    C = 40
    g = torch.Generator().manual_seed(5)
    batches = [(torch.randn(200, C, generator=g).to(DEV), torch.randint(0, C, (200,), generator=g).to(DEV))
               for _ in range(4)]
    mc = tm.MetricCollection([tm.MulticlassAccuracy(C, validate_args=validate),
                              tm.MulticlassPrecision(C, validate_args=validate),
                              tm.MulticlassRecall(C, average="micro", validate_args=validate),
                              tm.MulticlassF1Score(C, average=None, validate_args=validate)]).to(DEV)
    fns = {"MulticlassAccuracy": lambda p, t: F.multiclass_accuracy(p, t, C, validate_args=False),
           "MulticlassPrecision": lambda p, t: F.multiclass_precision(p, t, C, validate_args=False),
           "MulticlassRecall": lambda p, t: F.multiclass_recall(p, t, C, average="micro", validate_args=False),
           "MulticlassF1Score": lambda p, t: F.multiclass_f1_score(p, t, C, average=None, validate_args=False)}
    seen = []
    if groups_first:  # compute groups are formed by the first update: Accuracy / Precision / F1 share their states
        mc.update(*batches[0])
        seen.append(0)
        assert any(len(v) > 1 for v in mc.compute_groups.values())
    for i in (1, 2, 3):
        out = mc(*batches[i])
        seen.append(i)
        for k, fn in fns.items():
            assert torch.equal(out[k], fn(*batches[i])), (k, i)
    total = mc.compute()
    p = torch.cat([batches[i][0] for i in seen])
    t = torch.cat([batches[i][1] for i in seen])
    for k, fn in fns.items():
        assert torch.equal(total[k], fn(p, t)), k
    # a single metric outside any collection
    m = tm.MulticlassF1Score(C, validate_args=validate).to(DEV)
    for i in range(3):
        assert torch.equal(m(*batches[i]), F.multiclass_f1_score(*batches[i], C, validate_args=False))
    assert torch.equal(m.compute(), F.multiclass_f1_score(torch.cat([b[0] for b in batches[:3]]),
                                                         torch.cat([b[1] for b in batches[:3]]), C, validate_args=False))
    assert m._update_count == 3
