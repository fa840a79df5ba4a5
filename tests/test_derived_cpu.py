"""CPU checks of the O(C) reducers added for the consumers of the update kernels: the state is built
by the numpy oracle (the CUDA kernels are checked against the same oracle in the -m gpu suite), the reducer is ours, the
expected value is the unmodified reference's (tests/golden/derived_metrics.npz)."""
from importlib import import_module

import numpy as np
import pytest
import torch

from conftest import case_id, load_golden
from oracle import oracle_np as O

J = import_module("torchmetrics_b200.functional.classification.jaccard")
K = import_module("torchmetrics_b200.functional.classification.cohen_kappa")
M = import_module("torchmetrics_b200.functional.classification.matthews_corrcoef")
S = import_module("torchmetrics_b200.functional.classification.specificity")
H = import_module("torchmetrics_b200.functional.classification.hamming")

AVGS = ("micro", "macro", "weighted", "none")


def _close(got, want):
    np.testing.assert_allclose(got.numpy(), np.asarray(want), rtol=2e-6, atol=2e-6, equal_nan=True)


def _states(c):
    """(confusion matrix, (tp, fp, tn, fn)) of the case, as int64 tensors, from the numpy oracle."""
    task, ii = c.meta["task"], c.meta["ignore_index"]
    p, t = c["preds"], c["target"]
    if task == "binary":
        cm = O.binary_confusion_matrix_update(*O.binary_confusion_matrix_format(p, t, 0.5, ii))
        stats = (cm[1, 1], cm[0, 1], cm[0, 0], cm[1, 0])
    elif task == "multiclass":
        C = c.meta["num_classes"]
        cm = O.multiclass_confusion_matrix_update(*O.multiclass_confusion_matrix_format(p, t, ii), C)
        tp = np.diag(cm)
        fp, fn = cm.sum(0) - tp, cm.sum(1) - tp
        stats = (tp, fp, cm.sum() - tp - fp - fn, fn)
    else:
        L = c.meta["num_labels"]
        cm = O.multilabel_confusion_matrix_update(*O.multilabel_confusion_matrix_format(p, t, L, 0.5, ii), L)
        stats = (cm[:, 1, 1], cm[:, 0, 1], cm[:, 0, 0], cm[:, 1, 0])
    return torch.from_numpy(np.ascontiguousarray(cm)), tuple(torch.as_tensor(np.array(s)) for s in stats)


@pytest.mark.parametrize("c", load_golden("derived_metrics"), ids=case_id)
def test_reducers_on_oracle_states(c):
    task, ii = c.meta["task"], c.meta["ignore_index"]
    cm, (tp, fp, tn, fn) = _states(c)
    if task == "binary":
        _close(S._specificity_reduce(tp, fp, tn, fn, average="binary"), c["specificity"])
        _close(H._hamming_distance_reduce(tp, fp, tn, fn, average="binary"), c["hamming"])
        _close(J._jaccard_index_reduce(cm, average="binary"), c["jaccard"])
        _close(M._matthews_corrcoef_reduce(cm), c["mcc"])
        for w in ("none", "linear", "quadratic"):
            if f"kappa_{w}" in c:
                _close(K._cohen_kappa_reduce(cm, w), c[f"kappa_{w}"])
        return
    multilabel = task == "multilabel"
    for a in AVGS:
        _close(S._specificity_reduce(tp, fp, tn, fn, average=a, multilabel=multilabel), c[f"specificity_{a}"])
        _close(H._hamming_distance_reduce(tp, fp, tn, fn, average=a, multilabel=multilabel), c[f"hamming_{a}"])
        _close(J._jaccard_index_reduce(cm, average=a, ignore_index=None if multilabel else ii), c[f"jaccard_{a}"])
    _close(M._matthews_corrcoef_reduce(cm), c["mcc"])
    if not multilabel:
        for w in ("none", "linear", "quadratic"):
            _close(K._cohen_kappa_reduce(cm, w), c[f"kappa_{w}"])


def test_argument_errors_match_the_reference_messages():
    with pytest.raises(ValueError, match="The `average` has to be one of"):
        J._jaccard_index_reduce(torch.zeros(3, 3), average="samples")
    with pytest.raises(ValueError, match="should be either None, 'linear' or 'quadratic'"):
        K._cohen_kappa_reduce(torch.ones(3, 3), weights="cubic")
    import torchmetrics_b200 as tm

    with pytest.raises(ValueError, match="`num_classes` is expected to be `int`"):
        tm.JaccardIndex(task="multiclass")
    with pytest.raises(ValueError, match="Invalid Classification"):
        tm.CohenKappa(task="multilabel", num_classes=3)
    assert tm.classification.BinaryHammingDistance.higher_is_better is False
    assert tm.classification.MulticlassSpecificity.__name__ == "MulticlassSpecificity"
