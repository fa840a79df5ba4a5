"""Pin the numpy oracle against golden vectors produced by the unmodified reference
(tests/golden/make_golden.py). CPU only."""
import numpy as np
import pytest

from conftest import case_id, load_golden
from oracle import oracle_np as O


def _eq(a, b):
    a, b = np.asarray(a), np.asarray(b)
    assert a.shape == b.shape, (a.shape, b.shape)
    assert np.array_equal(a, b)


# ---------------------------------------------------------------- confusion matrix
@pytest.mark.parametrize("c", load_golden("confusion_matrix"), ids=case_id)
def test_confusion_matrix(c):
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    if kind.startswith("multiclass"):
        p, t = O.multiclass_confusion_matrix_format(c["preds"], c["target"], ii)
        _eq(O.multiclass_confusion_matrix_update(p, t, c.meta["num_classes"]), c["confmat"])
    elif kind.startswith("binary"):
        p, t = O.binary_confusion_matrix_format(c["preds"], c["target"], c.meta["threshold"], ii)
        _eq(O.binary_confusion_matrix_update(p, t), c["confmat"])
    else:
        L = c.meta["num_labels"]
        p, t = O.multilabel_confusion_matrix_format(c["preds"], c["target"], L, c.meta["threshold"], ii)
        _eq(O.multilabel_confusion_matrix_update(p, t, L), c["confmat"])


def test_confusion_matrix_out_of_range_is_an_error():
    with pytest.raises(RuntimeError):
        O.multiclass_confusion_matrix_update(np.array([0, 1]), np.array([1, 5]), 5)
    # like the reference, a too-large *pred* whose key still lands inside C*C is counted at the aliased cell
    cm = O.multiclass_confusion_matrix_update(np.array([0, 5]), np.array([1, 1]), 5)
    assert cm[1, 0] == 1 and cm[2, 0] == 1
    with pytest.raises(RuntimeError):
        O.multiclass_confusion_matrix_update(np.array([0, -1]), np.array([0, 0]), 5)


# ---------------------------------------------------------------- stat scores
def _ref_stack(tp, fp, tn, fn):
    return np.stack([tp, fp, tn, fn, tp + fn], axis=-1)


@pytest.mark.parametrize("c", load_golden("stat_scores"), ids=case_id)
def test_stat_scores(c):
    kind, ii, mda = c.meta["kind"], c.meta["ignore_index"], c.meta["multidim_average"]
    if kind.startswith("binary"):
        p, t = O.binary_stat_scores_format(c["preds"], c["target"], c.meta["threshold"], ii)
        got = _ref_stack(*O.binary_stat_scores_update(p, t, mda))
        _eq(got, c["stat_scores"])
    elif kind.startswith("multilabel"):
        p, t = O.multilabel_stat_scores_format(c["preds"], c["target"], c.meta["num_labels"], c.meta["threshold"], ii)
        got = _ref_stack(*O.multilabel_stat_scores_update(p, t, mda))
        _eq(got, c["stat_scores_none"])
        _eq(got.sum(0 if mda == "global" else 1), c["stat_scores_micro"])
    else:
        C, top_k = c.meta["num_classes"], c.meta["top_k"]
        if mda == "global" and top_k == 1:
            p, t = O.multiclass_stat_scores_format(c["preds"], c["target"])
            got = _ref_stack(*O.multiclass_stat_scores_update(p, t, C, "macro", ii))
            _eq(got, c["stat_scores_none"])
            micro = _ref_stack(*O.multiclass_stat_scores_update(p, t, C, "micro", ii))
            _eq(micro, c["stat_scores_micro"])
        else:
            preds, target = c["preds"], c["target"]
            if top_k == 1:
                preds, target = O.multiclass_stat_scores_format(preds, target)
            else:
                preds = preds.reshape(preds.shape[0], preds.shape[1], -1)
                target = target.reshape(target.shape[0], -1)
            got = _ref_stack(*O.multiclass_stat_scores_update_samplewise(preds, target, C, top_k, mda, ii))
            _eq(got, c["stat_scores_none"])


# ---------------------------------------------------------------- binned curves
@pytest.mark.parametrize("c", load_golden("binned_curve"), ids=case_id)
def test_binned_curve(c):
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    thr = c["thresholds"]
    if kind.startswith("binary"):
        p, t = O.binary_precision_recall_curve_format(c["preds"], c["target"], ii)
        if "formatted_preds" in c:
            # numpy's float32 exp is not guaranteed bit-identical to ATen's: compare the sigmoid within a few ulp
            # and then use the reference's own formatted scores so the integer part is tested exactly
            assert np.allclose(p, c["formatted_preds"], rtol=2e-6, atol=1e-9)
            p = c["formatted_preds"]
        loop = O.binary_precision_recall_curve_update_loop(p, t, thr)
        _eq(loop, c["confmat"])
        _eq(O.binary_precision_recall_curve_update_hist(p, t, thr), c["confmat"])
        if p.size <= 5000:
            _eq(O.binary_precision_recall_curve_update_vectorized(p, t, thr), c["confmat"])
        np.testing.assert_allclose(O.binary_auroc_binned(loop, thr), c["auroc"], rtol=0, atol=1e-6)
        np.testing.assert_allclose(O.binary_average_precision_binned(loop, thr), c["ap"], rtol=0, atol=1e-6)
        fpr, tpr, _ = O.binary_roc_compute_binned(loop, thr)
        np.testing.assert_array_equal(fpr, c["fpr"])
        np.testing.assert_array_equal(tpr, c["tpr"])
        pr, rc, _ = O.binary_precision_recall_binned(loop, thr)
        np.testing.assert_array_equal(pr, c["precision"])
        np.testing.assert_array_equal(rc, c["recall"])
    elif kind.startswith("multiclass"):
        C = c.meta["num_classes"]
        p, t = O.multiclass_precision_recall_curve_format(c["preds"], c["target"], C, ii, None)
        if "formatted_preds" in c:
            assert np.allclose(p, c["formatted_preds"], rtol=1e-6, atol=1e-9)
            p = c["formatted_preds"]
        _eq(O.multiclass_precision_recall_curve_update_loop(p, t, C, thr), c["confmat_None"])
        pm = p.reshape(-1)
        tm = np.eye(C, dtype=np.int64)[t.astype(np.int64)].reshape(-1)
        _eq(O.binary_precision_recall_curve_update_hist(pm, tm, thr), c["confmat_micro"])
    else:
        L = c.meta["num_labels"]
        p, t, ign = O.multilabel_precision_recall_curve_format(c["preds"], c["target"], L, ii)
        if kind.endswith("logits"):
            import torch  # reference-grade sigmoid for the float step; integer logic stays numpy

            p = torch.from_numpy(np.swapaxes(c["preds"], 0, 1).reshape(L, -1).T.copy()).sigmoid().numpy()
        _eq(O.multilabel_precision_recall_curve_update(p, t, ign, L, thr), c["confmat"])


# ---------------------------------------------------------------- exact curves
@pytest.mark.parametrize("c", [c for c in load_golden("exact_curve") if c.meta["kind"].startswith("binary")],
                         ids=case_id)
def test_exact_curve(c):
    import torch

    ii = c.meta["ignore_index"]
    p, t = O.binary_precision_recall_curve_format(c["preds"], c["target"], ii)
    if c.meta["kind"].endswith("logits"):
        keep = c["target"] != ii if ii is not None else slice(None)
        p = torch.from_numpy(c["preds"][keep]).sigmoid().numpy()
    fps, tps, thr = O.binary_clf_curve(p, t)
    np.testing.assert_array_equal(thr, c["thr"])
    np.testing.assert_array_equal(tps, c["tps"])
    np.testing.assert_array_equal(fps, c["fps"])
    # exact integer counts agree with the fp32 ones at these sizes (< 2**24)
    fe, te, _ = O.binary_clf_curve_exact(p, t)
    np.testing.assert_array_equal(te.astype(np.float32), c["tps"])
    np.testing.assert_array_equal(fe.astype(np.float32), c["fps"])
    np.testing.assert_allclose(O.binary_auroc_exact(p, t), c["auroc"], rtol=0, atol=1e-6)
    np.testing.assert_allclose(O.binary_auroc_exact(p, t, np.float64), c["auroc"], rtol=0, atol=1e-6)
    if not np.isnan(c["ap"]):
        np.testing.assert_allclose(O.binary_average_precision_exact(p, t), c["ap"], rtol=0, atol=1e-6)
        np.testing.assert_allclose(O.binary_average_precision_exact(p, t, np.float64), c["ap"], rtol=0, atol=1e-6)
    fpr, tpr, rthr = O.binary_roc_compute_exact(p, t)
    np.testing.assert_allclose(fpr, c["fpr"], rtol=0, atol=1e-7)
    np.testing.assert_allclose(tpr, c["tpr"], rtol=0, atol=1e-7)
    np.testing.assert_array_equal(rthr, c["roc_thresholds"])


# ---------------------------------------------------------------- doctest known answers
def test_doctest_known_answers():
    cases = {(c.meta["fn"], c.meta.get("thresholds")): c for c in load_golden("doctest_known_answers")}
    c = cases[("multiclass_confusion_matrix", None)]
    _eq(c["out"], [[1, 1, 0], [0, 1, 0], [0, 0, 1]])  # confusion_matrix.py:388-408
    _eq(O.multiclass_confusion_matrix_update(c["preds"], c["target"], 3), c["out"])
    c = cases[("binary_confusion_matrix", None)]
    _eq(c["out"], [[2, 0], [1, 1]])  # confusion_matrix.py:204-219
    p, t = O.binary_confusion_matrix_format(c["preds"], c["target"])
    _eq(O.binary_confusion_matrix_update(p, t), c["out"])
    c = cases[("multilabel_confusion_matrix", None)]
    p, t = O.multilabel_confusion_matrix_format(c["preds"], c["target"], 3)
    _eq(O.multilabel_confusion_matrix_update(p, t, 3), c["out"])
    c = cases[("binary_precision_recall_curve", 5)]
    thr = c["thresholds"]
    cm = O.binary_precision_recall_curve_update_loop(c["preds"], c["target"], thr)
    pr, rc, _ = O.binary_precision_recall_binned(cm, thr)
    np.testing.assert_allclose(pr, [0.5, 0.6667, 0.6667, 0.0, 0.0, 1.0], atol=1e-4)  # precision_recall_curve.py:340-351
    np.testing.assert_array_equal(pr, c["precision"])
    np.testing.assert_array_equal(rc, c["recall"])
    c = cases[("multilabel_stat_scores", None)]
    p, t = O.multilabel_stat_scores_format(c["preds"], c["target"], 3)
    _eq(_ref_stack(*O.multilabel_stat_scores_update(p, t)), c["out_none"])


# ---------------------------------------------------------------- sync semantics
def test_sync_state_restatement():
    a = [np.arange(6).reshape(2, 3), np.arange(6).reshape(2, 3) * 10]
    _eq(O.sync_state(a, "sum"), a[0] + a[1])
    b = [np.ones((2,)), np.ones((3,)) * 2]
    _eq(O.sync_state(b, "cat"), [1, 1, 2, 2, 2])


# ---------------------------------------------------------------- C/OpenMP restatement (timed CPU baseline)
def test_c_oracle_against_golden_and_numpy():
    from oracle import oracle_c as OC

    assert OC.num_threads() >= 1
    for c in load_golden("confusion_matrix"):
        if c.meta["kind"] == "multiclass_labels" and c.meta["ignore_index"] is None:
            _eq(OC.confmat_update(c["preds"], c["target"], c.meta["num_classes"]), c["confmat"])
    with pytest.raises(RuntimeError):
        OC.confmat_update(np.array([0, 1]), np.array([1, 5]), 5)
    for c in load_golden("binned_curve"):
        if c.meta["kind"] in ("binary_probs", "binary_grid", "binary_special") and c.meta["ignore_index"] is None:
            _eq(OC.binned_update_loop(c["preds"], c["target"], c["thresholds"]), c["confmat"])
    rng = np.random.default_rng(0)
    p = rng.random((5000, 12)).astype(np.float32)
    t = rng.integers(-1, 2, (5000, 12))
    fp_, ft_ = O.multilabel_stat_scores_format(p[:, :, None], t[:, :, None], 12, 0.37, -1)
    tp, fp, tn, fn = O.multilabel_stat_scores_update(fp_, ft_)
    _eq(OC.multilabel_stats(p, t, 0.37, -1), np.stack([tp, fp, tn, fn]))
