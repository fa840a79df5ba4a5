"""End to end on the GPU: operating-point metrics, group fairness and nominal association measures through the public
functional and class APIs against the unmodified reference's outputs (tests/golden/consumers2.npz)."""
import numpy as np
import pytest
import torch

from conftest import case_id, load_golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
CASES = load_golden("consumers2")
FAMS = {"recall_at_fixed_precision": ("min_precision", "RecallAtFixedPrecision"),
        "precision_at_fixed_recall": ("min_recall", "PrecisionAtFixedRecall"),
        "sensitivity_at_specificity": ("min_specificity", "SensitivityAtSpecificity"),
        "specificity_at_sensitivity": ("min_sensitivity", "SpecificityAtSensitivity")}


def _tm():
    import torchmetrics_b200 as tm

    return tm


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "operating_point"], ids=case_id)
def test_operating_points(c):
    tm = _tm()
    F = tm.functional
    task, thresholds, ii = c.meta["task"], c.meta["thresholds"], c.meta["ignore_index"]
    extra = {k: c.meta[k] for k in ("num_classes", "num_labels") if k in c.meta}
    p, t = torch.from_numpy(c["preds"]).to(DEV), torch.from_numpy(c["target"]).to(DEV)
    for key, want in c.items():
        if "@" not in key:
            continue
        fam, m = key.split("@")
        arg, cls_stem = FAMS[fam]
        kw = dict(extra, thresholds=thresholds, ignore_index=ii, **{arg: float(m)})
        val, at = getattr(F, f"{task}_{fam}")(p, t, **kw)
        got = torch.stack([val.float().reshape(-1), at.float().reshape(-1)]).cpu().numpy()
        np.testing.assert_allclose(got, want, rtol=1e-6, atol=1e-6, err_msg=key)
        # dispatcher and class (two updates on halves of the data)
        val2, at2 = getattr(F, fam)(p, t, task=task, **kw)
        assert torch.equal(val2, val) and torch.equal(at2, at)
        metric = getattr(tm, task.capitalize() + cls_stem)(**kw).to(DEV)
        h = p.shape[0] // 2
        metric.update(p[:h], t[:h])
        metric.update(p[h:], t[h:])
        val3, at3 = metric.compute()
        np.testing.assert_allclose(torch.stack([val3.float().reshape(-1), at3.float().reshape(-1)]).cpu().numpy(),
                                   want, rtol=1e-6, atol=1e-6, err_msg=key + " (class)")


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "group_fairness"], ids=case_id)
def test_group_fairness(c):
    tm = _tm()
    F = tm.functional
    ii = c.meta["ignore_index"]
    p, t, g = (torch.from_numpy(c[k]).to(DEV) for k in ("preds", "target", "groups"))
    rates = F.binary_groups_stat_rates(p, t, g, 3, ignore_index=ii)
    np.testing.assert_allclose(torch.stack([rates[f"group_{i}"] for i in range(len(rates))]).cpu().numpy(),
                               c["rates"], rtol=1e-6)
    fair = F.binary_fairness(p, t, g, "all", ignore_index=ii)
    assert sorted(fair) == c.meta["fair_keys"]
    np.testing.assert_allclose(torch.stack([fair[k] for k in sorted(fair)]).cpu().numpy(), c["fair"], rtol=1e-6)
    dp = F.demographic_parity(p, g, ignore_index=ii)
    assert sorted(dp) == c.meta["dp_keys"]
    np.testing.assert_allclose(torch.stack([dp[k] for k in sorted(dp)]).cpu().numpy(), c["dp"], rtol=1e-6)
    # stateful forms, two batches
    h = p.shape[0] // 2
    m = tm.BinaryGroupStatRates(3, ignore_index=ii).to(DEV)
    m.update(p[:h], t[:h], g[:h])
    m.update(p[h:], t[h:], g[h:])
    out = m.compute()
    np.testing.assert_allclose(torch.stack([out[f"group_{i}"] for i in range(3)]).cpu().numpy(), c["rates"], rtol=1e-6)
    bf = tm.BinaryFairness(3, ignore_index=ii).to(DEV)
    bf.update(p[:h], t[:h], g[:h])
    bf.update(p[h:], t[h:], g[h:])
    out = bf.compute()
    np.testing.assert_allclose(torch.stack([out[k] for k in sorted(out)]).cpu().numpy(), c["fair"], rtol=1e-6)


@pytest.mark.parametrize("c", [c for c in CASES if c.meta.get("what") == "nominal"], ids=case_id)
def test_nominal(c):
    tm = _tm()
    F = tm.functional
    strat = c.meta["nan_strategy"]
    p, t = torch.from_numpy(c["preds"]).to(DEV), torch.from_numpy(c["target"]).to(DEV)

    def close(got, key):
        np.testing.assert_allclose(float(got), c[key], rtol=2e-6, atol=2e-6, equal_nan=True, err_msg=key)

    close(F.cramers_v(p, t, True, strat), "cramers_bias")
    close(F.cramers_v(p, t, False, strat), "cramers")
    close(F.tschuprows_t(p, t, True, strat), "tschuprows_bias")
    close(F.tschuprows_t(p, t, False, strat), "tschuprows")
    close(F.pearsons_contingency_coefficient(p, t, strat), "pearson")
    close(F.theils_u(p, t, strat), "theils_u")
    close(F.theils_u(t, p, strat), "theils_u_rev")
    C = len(torch.cat([p, t]).unique())
    h = p.shape[0] // 2
    for cls, key, kw in [(tm.CramersV, "cramers_bias", {"bias_correction": True}),
                         (tm.TschuprowsT, "tschuprows", {"bias_correction": False}),
                         (tm.PearsonsContingencyCoefficient, "pearson", {}), (tm.TheilsU, "theils_u", {})]:
        m = cls(num_classes=C, nan_strategy=strat, **kw).to(DEV)
        m.update(p[:h], t[:h])
        m.update(p[h:], t[h:])
        assert m.confmat.dtype == torch.int64 and int(m.confmat.sum()) > 0
        close(m.compute(), key)


def test_nominal_matrix():
    F = _tm().functional
    c = [c for c in CASES if c.meta.get("what") == "nominal_matrix"][0]
    mat = torch.from_numpy(c["matrix"]).to(DEV)
    for fn, key in [(F.cramers_v_matrix, "cramers"), (F.theils_u_matrix, "theils_u"),
                    (F.tschuprows_t_matrix, "tschuprows"), (F.pearsons_contingency_coefficient_matrix, "pearson")]:
        np.testing.assert_allclose(fn(mat).cpu().numpy(), c[key], rtol=2e-6, atol=2e-6, err_msg=key)
