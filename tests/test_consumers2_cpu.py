"""CPU checks of the reducers of the second batch of consumers: states come from the numpy oracle
(the CUDA kernels are checked against the same oracle in the -m gpu suite), the O(T) / O(C) reducers are ours, the
expected values are the unmodified reference's (tests/golden/consumers2.npz)."""
from importlib import import_module

import numpy as np
import pytest
import torch

from conftest import case_id, load_golden
from oracle import oracle_np as O

OP = import_module("torchmetrics_b200.functional.classification._operating_point")
GF = import_module("torchmetrics_b200.functional.classification.group_fairness")
CR = import_module("torchmetrics_b200.functional.nominal.cramers")
TS = import_module("torchmetrics_b200.functional.nominal.tschuprows")
PE = import_module("torchmetrics_b200.functional.nominal.pearson")
TH = import_module("torchmetrics_b200.functional.nominal.theils_u")

CASES = load_golden("consumers2")
FAMS = {"recall_at_fixed_precision": OP.RECALL_AT_PRECISION, "precision_at_fixed_recall": OP.PRECISION_AT_RECALL,
        "sensitivity_at_specificity": OP.SENSITIVITY_AT_SPECIFICITY,
        "specificity_at_sensitivity": OP.SPECIFICITY_AT_SENSITIVITY}


def _binned_state(c):
    """(T[, C], 2, 2) int64 state of an operating-point case from the oracle, and the threshold tensor."""
    T, ii, task = c.meta["thresholds"], c.meta["ignore_index"], c.meta["task"]
    thr = torch.linspace(0, 1, T).numpy()  # the tensor the package itself would build
    p, t = c["preds"], c["target"]
    if task == "binary":
        p, t = O.binary_precision_recall_curve_format(p, t, ii)
        st = O.binary_precision_recall_curve_update_hist(p, t, thr)
    elif task == "multiclass":
        p, t = O.multiclass_precision_recall_curve_format(p, t, c.meta["num_classes"], ii)
        st = O.multiclass_precision_recall_curve_update_loop(p, t, c.meta["num_classes"], thr)
    else:
        p, t, ign = O.multilabel_precision_recall_curve_format(p, t, c.meta["num_labels"], ii)
        st = O.multilabel_precision_recall_curve_update(p, t, ign, c.meta["num_labels"], thr)
    return torch.from_numpy(np.ascontiguousarray(st)), torch.from_numpy(thr)


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "operating_point" and c.meta["thresholds"]],
                         ids=case_id)
def test_operating_points_on_binned_oracle_state(c):
    state, thr = _binned_state(c)
    task = c.meta["task"]
    for key, want in c.items():
        if "@" not in key:
            continue
        fam, m = key.split("@")
        m = float(m)
        F = FAMS[fam]
        if task == "binary":
            val, at = F.binary_compute(state, thr, m)
        elif task == "multiclass":
            val, at = F.multiclass_compute(state, c.meta["num_classes"], thr, m)
        else:
            val, at = F.multilabel_compute(state, c.meta["num_labels"], thr, c.meta["ignore_index"], m)
        got = torch.stack([val.float().reshape(-1), at.float().reshape(-1)]).numpy()
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6, err_msg=key)


def test_selection_rules_edge_cases():
    p = torch.tensor([0.5, 0.5, 1.0, 1.0])
    r = torch.tensor([1.0, 0.5, 0.5, 0.0])
    t = torch.tensor([0.1, 0.4, 0.7])
    # ties on recall are broken by precision, then by threshold (lexicographic maximum)
    val, at = OP._recall_at_precision(p, r, t, 0.5)
    assert float(val) == 1.0 and abs(float(at) - 0.1) < 1e-7
    val, at = OP._precision_at_recall(p, r, t, 0.5)
    assert float(val) == 1.0 and abs(float(at) - 0.7) < 1e-7
    # nothing qualifies -> (0, 1e6)
    val, at = OP._recall_at_precision(p * 0.1, r, t, 0.9)
    assert float(val) == 0.0 and float(at) == 1e6
    val, at = OP._sensitivity_at_specificity(r[:3], p[:3] * 0.1, t, 0.9)
    assert float(val) == 0.0 and float(at) == 1e6


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "group_fairness"], ids=case_id)
def test_group_fairness_reducers(c):
    """Per-group counts from the oracle's binary stat scores, then our rate / parity reducers."""
    ii = c.meta["ignore_index"]
    stats = []
    for gid in np.unique(c["groups"]):
        sel = c["groups"] == gid
        p, t = O.binary_stat_scores_format(c["preds"][sel], c["target"][sel], 0.5, ii)
        stats.append(tuple(torch.as_tensor(int(v)) for v in O.binary_stat_scores_update(p, t)))
    rates = GF._groups_reduce(stats)
    np.testing.assert_allclose(torch.stack([rates[f"group_{i}"] for i in range(len(stats))]).numpy(), c["rates"],
                               rtol=1e-6)
    tr = GF._groups_stat_transform(stats)
    fair = {**GF._compute_binary_demographic_parity(**tr), **GF._compute_binary_equal_opportunity(**tr)}
    assert sorted(fair) == c.meta["fair_keys"]
    np.testing.assert_allclose(torch.stack([fair[k] for k in sorted(fair)]).numpy(), c["fair"], rtol=1e-6)


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "nominal"], ids=case_id)
def test_nominal_reducers(c):
    p, t = torch.from_numpy(c["preds"]), torch.from_numpy(c["target"])
    C = len(torch.cat([p, t]).unique())
    if c.meta["nan_strategy"] == "replace":
        p, t = p.nan_to_num(0.0), t.nan_to_num(0.0)
    elif p.is_floating_point():
        keep = ~(p.isnan() | t.isnan())
        p, t = p[keep], t[keep]
    cm = torch.from_numpy(O.multiclass_confusion_matrix_update(p.long().numpy(), t.long().numpy(), C))

    def close(got, key):
        np.testing.assert_allclose(np.asarray(float(got)), c[key], rtol=2e-6, atol=2e-6, equal_nan=True, err_msg=key)

    close(CR._cramers_v_compute(cm, True), "cramers_bias")
    close(CR._cramers_v_compute(cm, False), "cramers")
    close(TS._tschuprows_t_compute(cm, True), "tschuprows_bias")
    close(TS._tschuprows_t_compute(cm, False), "tschuprows")
    close(PE._pearsons_contingency_coefficient_compute(cm), "pearson")
    close(TH._theils_u_compute(cm), "theils_u")
    close(TH._theils_u_compute(cm.T), "theils_u_rev")


def test_cpu_states_keep_the_torch_op_reducers_and_the_generic_forward():
    """The fused reducer kernel and the packed-state routes only exist for CUDA states: on CPU tensors the reducers
    are the reference's torch ops and the stat-score modules report no packed view (so `forward` and `_update_state`
    take the generic route)."""
    import torch

    import torchmetrics_b200 as tm
    from torchmetrics_b200.functional.classification.f_beta import _fbeta_reduce
    from torchmetrics_b200.functional.classification.precision_recall import _precision_recall_reduce
    from torchmetrics_b200.utilities.compute import _fused_stat_reduce

    tp, fp = torch.tensor([3, 0, 5]), torch.tensor([1, 0, 5])
    tn, fn = torch.tensor([9, 9, 2]), torch.tensor([0, 4, 1])
    assert _fused_stat_reduce("precision", tp, fp, tn, fn, "macro") is None
    # compute.py:46-69 by hand: classes 0 and 2 score 3/4 and 5/10, class 1 never predicted (0), all three occur
    assert torch.allclose(_precision_recall_reduce("precision", tp, fp, tn, fn, average="macro"),
                          torch.tensor((0.75 + 0.0 + 0.5) / 3))
    assert torch.allclose(_precision_recall_reduce("recall", tp, fp, tn, fn, average="micro"), torch.tensor(8 / 13))
    assert torch.allclose(_fbeta_reduce(tp, fp, tn, fn, 1.0, average="none"),
                          torch.tensor([6 / 7, 0.0, 10 / 16]))
    m = tm.MultilabelF1Score(num_labels=3, validate_args=False)
    assert m._packed_states() is None
    m.tp, m.fp, m.tn, m.fn = tp, fp, tn, fn
    m._update_count = 1
    assert torch.allclose(m.compute(), torch.tensor((6 / 7 + 0.0 + 10 / 16) / 3))
