"""Parity of the CUDA kernels (called through the C ABI via torchmetrics_b200._native) with the oracle and
with golden vectors from the reference. Integer states must be bit-exact."""
import numpy as np
import pytest
import torch

from conftest import case_id, load_golden
from oracle import oracle_np as O

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _native():
    from torchmetrics_b200 import _native

    return _native


def _np(t):
    return t.detach().cpu().numpy()


# =========================================================================== K1 confusion matrix
@pytest.mark.parametrize("c", [c for c in load_golden("confusion_matrix") if c.meta["kind"] == "multiclass_labels"],
                         ids=case_id)
@pytest.mark.parametrize("algo", [0, 1, 2, 6])
def test_confmat_golden(c, algo):
    N = _native()
    C = c.meta["num_classes"]
    if algo == 2 and C * C * 4 > 200 * 1024:
        pytest.skip("does not fit shared memory")
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, torch.from_numpy(c["preds"]).to(DEV), torch.from_numpy(c["target"]).to(DEV), C,
                      c.meta["ignore_index"], algo=algo)
    assert np.array_equal(_np(cm), c["confmat"])
    assert N.read_status(torch.device(DEV)) == 0


@pytest.mark.parametrize("C,n", [(2, 1), (2, 3), (3, 4), (5, 1023), (10, 65537), (64, 1 << 20), (1000, 1 << 22),
                                 (1000, (1 << 22) + 3), (3000, 1 << 20)])
@pytest.mark.parametrize("dt", [torch.int64, torch.int32, torch.uint8])
def test_confmat_seeded_vs_oracle(C, n, dt):
    if dt == torch.uint8 and C > 255:
        pytest.skip("labels do not fit")
    N = _native()
    g = torch.Generator().manual_seed(C * 31 + n)
    p = torch.randint(0, C, (n,), generator=g).to(dt)
    t = torch.randint(0, C, (n,), generator=g).to(dt)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, p.to(DEV), t.to(DEV), C)
    N.confmat_update_(cm, p.to(DEV), t.to(DEV), C)  # accumulates
    want = O.multiclass_confusion_matrix_update(p.numpy(), t.numpy(), C) * 2
    assert np.array_equal(_np(cm), want)
    assert int(cm.sum()) == 2 * n


def test_confmat_diagonal_heavy_and_single_cell():
    N = _native()
    C, n = 1000, 1 << 21
    g = torch.Generator().manual_seed(5)
    t = torch.randint(0, C, (n,), generator=g)
    p = torch.where(torch.rand(n, generator=g) < 0.8, t, torch.randint(0, C, (n,), generator=g))
    for pp, tt in [(p, t), (torch.full((n,), 7), torch.full((n,), 7))]:
        for algo, CC in [(1, C), (2, 16)]:
            pp2, tt2 = pp % CC, tt % CC
            cm = torch.zeros(CC, CC, dtype=torch.int64, device=DEV)
            N.confmat_update_(cm, pp2.to(DEV), tt2.to(DEV), CC, algo=algo)
            assert np.array_equal(_np(cm), O.multiclass_confusion_matrix_update(pp2.numpy(), tt2.numpy(), CC))


@pytest.mark.parametrize("C", [2, 64, 65, 1000, 2048])
@pytest.mark.parametrize("n", [1, 8191, 8192, 8193, (1 << 21) + 3])
def test_confmat_partition_path(C, n):
    """algo 6 (bucket partition + shared-memory histograms): uniform, diagonal-heavy, single-cell (queue overflow)
    and ignore_index inputs must all give the oracle's matrix."""
    N = _native()
    g = torch.Generator().manual_seed(C * 13 + n)
    t = torch.randint(0, C, (n,), generator=g)
    uniform = torch.randint(0, C, (n,), generator=g)
    diag = torch.where(torch.rand(n, generator=g) < 0.8, t, uniform)
    hot_t = torch.where(torch.rand(n, generator=g) < 0.9, torch.full((n,), C - 1), t)
    for p, tt, ii in [(uniform, t, None), (diag, t, None), (torch.full((n,), C - 1), hot_t, None), (uniform, t, 0)]:
        cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
        N.confmat_update_(cm, p.to(DEV), tt.to(DEV), C, ii, algo=6)
        N.confmat_update_(cm, p.to(DEV), tt.to(DEV), C, ii, algo=6)  # accumulates, workspace reused
        pp, tf = O.multiclass_confusion_matrix_format(p.numpy(), tt.numpy(), ii)
        assert np.array_equal(_np(cm), 2 * O.multiclass_confusion_matrix_update(pp, tf, C))
    assert N.read_status(torch.device(DEV)) == 0
    # out-of-range labels are flagged, in-range ones still counted
    if n > 4:
        bad_t = t.clone()
        bad_t[3] = C
        cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
        N.confmat_update_(cm, uniform.to(DEV), bad_t.to(DEV), C, algo=6)
        assert int(cm.sum()) == n - 1
        assert N.read_status(torch.device(DEV)) == N.STATUS_LABEL_RANGE


def test_confmat_out_of_range_sets_status():
    N = _native()
    dev = torch.device(DEV)
    C = 5
    p = torch.tensor([0, 1, 2, 3, 4, 1, 1, 1, 2], device=DEV)
    t = torch.tensor([0, 1, 2, 3, 5, 1, -1, 1, 2], device=DEV)  # 5 and -1 are out of range
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, p, t, C)
    with pytest.raises(RuntimeError, match="outside the expected range"):
        N.raise_on_status(dev, C)
    assert N.read_status(dev) == 0  # cleared
    # like the reference, a too-large pred whose key still lands inside C*C is counted at the aliased cell
    cm.zero_()
    N.confmat_update_(cm, torch.tensor([0, 5], device=DEV), torch.tensor([1, 1], device=DEV), C)
    assert np.array_equal(_np(cm), O.multiclass_confusion_matrix_update(np.array([0, 5]), np.array([1, 1]), C))
    assert N.read_status(dev) == 0


def test_confmat_empty_and_unaligned():
    N = _native()
    cm = torch.zeros(3, 3, dtype=torch.int64, device=DEV)
    e = torch.zeros(0, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, e, e, 3)
    assert int(cm.sum()) == 0
    base = torch.randint(0, 3, (1001,), device=DEV)
    p, t = base[1:], base[:-1]  # p is an 8-byte-offset view: the binding must realign it
    N.confmat_update_(cm, p, t, 3)
    assert np.array_equal(_np(cm), O.multiclass_confusion_matrix_update(_np(p), _np(t), 3))


def test_match_count():
    N = _native()
    g = torch.Generator().manual_seed(3)
    for n in [1, 5, 1000, (1 << 20) + 1]:
        p = torch.randint(0, 7, (n,), generator=g)
        t = torch.randint(-1, 7, (n,), generator=g)
        out = torch.zeros(2, dtype=torch.int64, device=DEV)
        N.match_count_update_(out, p.to(DEV), t.to(DEV), ignore_index=-1)
        valid = t != -1
        assert out.tolist() == [int(valid.sum()), int(((p == t) & valid).sum())]


# =========================================================================== K3 binned curve
def _binned(N, preds, target, thr, ignore_index=None, prob_mode=None, num_columns=1, class_index=False):
    T = thr.numel()
    shape = (T, 2, 2) if num_columns == 1 else (T, num_columns, 2, 2)
    state = torch.zeros(shape, dtype=torch.int64, device=DEV)
    N.binned_curve_update_(state, preds.to(DEV), target.to(DEV), thr.to(DEV), num_columns=num_columns,
                           target_is_class_index=class_index, ignore_index=ignore_index,
                           prob_mode=N.PROB_AUTO if prob_mode is None else prob_mode)
    return state


@pytest.mark.parametrize("c", [c for c in load_golden("binned_curve") if c.meta["kind"].startswith("binary")],
                         ids=case_id)
def test_binned_binary_golden(c):
    N = _native()
    thr = torch.from_numpy(c["thresholds"])
    sthr, order = torch.sort(thr)
    preds, target = torch.from_numpy(c["preds"]), torch.from_numpy(c["target"])
    state = _binned(N, preds, target, sthr, c.meta["ignore_index"])
    got = torch.empty_like(state)
    got[order.to(DEV)] = state
    if "formatted_preds" in c:
        # logits: the golden state used ATen's CPU sigmoid; ours must equal ATen's CUDA sigmoid bit for bit,
        # so rebuild the expectation from torch.sigmoid on the device and the (pinned) oracle
        keep = c["target"] != c.meta["ignore_index"] if c.meta["ignore_index"] is not None else slice(None)
        sp = _np(torch.from_numpy(c["preds"][keep]).to(DEV).sigmoid())
        want = O.binary_precision_recall_curve_update_loop(sp, c["target"][keep], c["thresholds"])
        assert np.array_equal(_np(got), want)
        # and it can differ from the CPU-sigmoid golden only where a score sits within an ulp of a threshold
        assert np.abs(_np(got) - c["confmat"]).max() <= 2
    else:
        assert np.array_equal(_np(got), c["confmat"])
    assert N.read_status(torch.device(DEV)) == 0


@pytest.mark.parametrize("n", [0, 1, 3, 4, 5, 2047, 2048, 2049, 100_000, (1 << 22) + 1])
@pytest.mark.parametrize("T", [1, 2, 5, 200, 427, 428, 1000, 3000])
def test_binned_binary_seeded(n, T):
    N = _native()
    g = torch.Generator().manual_seed(n * 7 + T)
    preds = torch.rand(n, generator=g)
    target = (torch.rand(n, generator=g) < 0.3 + 0.4 * preds).long()
    thr = torch.linspace(0, 1, T) if T > 1 else torch.tensor([0.5])
    state = _binned(N, preds, target, thr)
    want = O.binary_precision_recall_curve_update_hist(preds.numpy(), target.numpy(), thr.numpy())
    assert np.array_equal(_np(state), want)


@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16, torch.float64])
@pytest.mark.parametrize("tdt", [torch.int64, torch.int32, torch.uint8, torch.bool])
def test_binned_binary_dtypes(dt, tdt):
    N = _native()
    g = torch.Generator().manual_seed(11)
    n = 50_001
    preds = torch.rand(n, generator=g, dtype=torch.float64).to(dt)
    target = (torch.rand(n, generator=g) < 0.5).to(tdt)
    thr = torch.linspace(0, 1, 200)
    state = _binned(N, preds, target, thr)
    # type promotion: half/bfloat16 compare in float32, float64 in float64
    pn = preds.double().numpy() if dt == torch.float64 else preds.float().numpy()
    want = O.binary_precision_recall_curve_update_hist(pn, target.long().numpy(), thr.numpy())
    assert np.array_equal(_np(state), want)


def test_binned_special_values_and_sigmoid_rule():
    N = _native()
    thr = torch.linspace(0, 1, 200)
    nan, inf = float("nan"), float("inf")
    target = torch.tensor([1, 0, 1, 0, 1, 0, 1, 0])
    # in range: -0.0, 0, 1, denormal
    p = torch.tensor([-0.0, 0.0, 1.0, 1e-45, 0.5, thr[100].item(), thr[199].item(), thr[1].item()])
    want = O.binary_precision_recall_curve_update_loop(p.numpy(), target.numpy(), thr.numpy())
    assert np.array_equal(_np(_binned(N, p, target, thr)), want)
    # one NaN / inf / value > 1 anywhere flips the WHOLE batch to sigmoid (precision_recall_curve.py:183-184)
    for bad in [nan, inf, -inf, 1.0000001, -1e-30]:
        p2 = p.clone()
        p2[3] = bad
        sp = _np(p2.to(DEV).sigmoid())
        want = O.binary_precision_recall_curve_update_loop(sp, target.numpy(), thr.numpy())
        assert np.array_equal(_np(_binned(N, p2, target, thr)), want), bad
    # ... unless that sample is ignored (binary format drops ignored samples before the range check)
    p2 = p.clone()
    p2[3] = 7.0
    t2 = target.clone()
    t2[3] = -1
    want = O.binary_precision_recall_curve_update_loop(np.delete(p2.numpy(), 3), np.delete(t2.numpy(), 3), thr.numpy())
    assert np.array_equal(_np(_binned(N, p2, t2, thr, ignore_index=-1)), want)
    # AS_IS never applies it; NaN compares false with every threshold
    p3 = torch.tensor([nan, 2.0, -1.0, 0.5])
    t3 = torch.tensor([1, 1, 0, 0])
    want = O.binary_precision_recall_curve_update_loop(p3.numpy(), t3.numpy(), thr.numpy())
    assert np.array_equal(_np(_binned(N, p3, t3, thr, prob_mode=N.PROB_AS_IS)), want)


def test_binned_logits_large_matches_device_sigmoid():
    N = _native()
    g = torch.Generator().manual_seed(2)
    n = 1 << 20
    logits = torch.randn(n, generator=g) * 4
    target = (torch.rand(n, generator=g) < 0.5).long()
    thr = torch.linspace(0, 1, 200)
    sp = _np(logits.to(DEV).sigmoid())
    want = O.binary_precision_recall_curve_update_hist(sp, target.numpy(), thr.numpy())
    assert np.array_equal(_np(_binned(N, logits, target, thr)), want)
    # the workspace is left clean: a probability batch right after gives the plain result
    probs = torch.rand(n, generator=g)
    want = O.binary_precision_recall_curve_update_hist(probs.numpy(), target.numpy(), thr.numpy())
    assert np.array_equal(_np(_binned(N, probs, target, thr)), want)


def test_binned_unsorted_thresholds_flagged_and_arbitrary_sorted_ok():
    N = _native()
    dev = torch.device(DEV)
    p = torch.rand(1000)
    t = torch.randint(0, 2, (1000,))
    _binned(N, p, t, torch.tensor([0.5, 0.1, 0.9]))
    with pytest.raises(RuntimeError, match="sorted ascending"):
        N.raise_on_status(dev)
    thr = torch.tensor([-3.0, 0.1, 0.1, 0.10000001, 0.5, 0.5, 7.0])  # duplicates, outside [0,1]
    want = O.binary_precision_recall_curve_update_loop(p.numpy(), t.numpy(), thr.numpy())
    assert np.array_equal(_np(_binned(N, p, t, thr)), want)
    assert N.read_status(dev) == 0


def test_binned_accumulates_in_place():
    N = _native()
    g = torch.Generator().manual_seed(9)
    thr = torch.linspace(0, 1, 200).to(DEV)
    state = torch.zeros(200, 2, 2, dtype=torch.int64, device=DEV)
    total = np.zeros((200, 2, 2), dtype=np.int64)
    for n in [1000, 77, 123_456]:
        p = torch.rand(n, generator=g)
        t = torch.randint(0, 2, (n,), generator=g)
        N.binned_curve_update_(state, p.to(DEV), t.to(DEV), thr)
        total += O.binary_precision_recall_curve_update_hist(p.numpy(), t.numpy(), _np(thr))
    assert np.array_equal(_np(state), total)


@pytest.mark.parametrize("c", [c for c in load_golden("binned_curve") if c.meta["kind"].startswith("multilabel")],
                         ids=case_id)
def test_binned_multilabel_golden(c):
    N = _native()
    L = c.meta["num_labels"]
    preds = torch.from_numpy(c["preds"])
    target = torch.from_numpy(c["target"])
    # host-side layout step of the reference's _format (precision_recall_curve.py:754-755)
    p2 = preds.transpose(0, 1).reshape(L, -1).T.contiguous()
    t2 = target.transpose(0, 1).reshape(L, -1).T.contiguous()
    state = _binned(N, p2, t2, torch.from_numpy(c["thresholds"]), c.meta["ignore_index"], num_columns=L)
    if c.meta["kind"].endswith("logits"):
        sp = _np(p2.to(DEV).sigmoid())
        ign = _np(t2) == c.meta["ignore_index"] if c.meta["ignore_index"] is not None else np.zeros(t2.shape, bool)
        want = O.multilabel_precision_recall_curve_update(sp, _np(t2), ign, L, c["thresholds"])
        assert np.array_equal(_np(state), want)
        assert np.abs(_np(state) - c["confmat"]).max() <= 2
    else:
        assert np.array_equal(_np(state), c["confmat"])


@pytest.mark.parametrize("c", [c for c in load_golden("binned_curve") if c.meta["kind"] == "multiclass_probs"],
                         ids=case_id)
def test_binned_multiclass_golden(c):
    N = _native()
    C = c.meta["num_classes"]
    state = _binned(N, torch.from_numpy(c["preds"]), torch.from_numpy(c["target"]), torch.from_numpy(c["thresholds"]),
                    c.meta["ignore_index"], prob_mode=N.PROB_CHECK, num_columns=C, class_index=True)
    assert np.array_equal(_np(state), c["confmat_None"])
    assert N.read_status(torch.device(DEV)) == 0


def test_binned_multiclass_check_mode_reports_logits():
    N = _native()
    dev = torch.device(DEV)
    p = torch.randn(100, 4)
    t = torch.randint(0, 4, (100,))
    state = _binned(N, p, t, torch.linspace(0, 1, 10), prob_mode=N.PROB_CHECK, num_columns=4, class_index=True)
    assert int(state.abs().sum()) == 0  # discarded
    assert N.read_status(dev) == N.STATUS_OUT_OF_UNIT_RANGE


def test_binned_large_columns_global_bins():
    N = _native()
    g = torch.Generator().manual_seed(4)
    L, T, n = 300, 200, 2000  # 300 * 2 * 201 bins do not fit shared memory -> global scratch path
    p = torch.rand(n, L, generator=g)
    t = torch.randint(0, 2, (n, L), generator=g)
    thr = torch.linspace(0, 1, T)
    state = _binned(N, p, t, thr, num_columns=L)
    want = O.multilabel_precision_recall_curve_update(p.numpy(), t.numpy(), np.zeros((n, L), bool), L, thr.numpy())
    assert np.array_equal(_np(state), want)


@pytest.mark.parametrize("C,n,T", [(3, 1001, 7), (7, 4099, 200), (10, 50000, 200), (130, 3000, 50), (100, 20011, 200)])
@pytest.mark.parametrize("uniform", [True, False])
def test_binned_multicolumn_seeded(C, n, T, uniform):
    """The shared-memory multi-column kernel + warp-per-column finalize: rows that wrap inside a 4-element vector,
    ignore_index, evenly spaced and arbitrary sorted thresholds, both target layouts, two updates into one state."""
    N = _native()
    g = torch.Generator().manual_seed(C * 1000 + T)
    thr = torch.linspace(0, 1, T) if uniform else torch.sort(torch.rand(T, generator=g)).values
    # multiclass: class-index target, scores already in [0, 1]
    p = torch.rand(n, C, generator=g)
    p[::17, 0] = thr[min(3, T - 1)]  # scores exactly on a threshold
    t = torch.randint(0, C, (n,), generator=g)
    t[::13] = -1
    state = torch.zeros(T, C, 2, 2, dtype=torch.int64, device=DEV)
    for _ in range(2):
        N.binned_curve_update_(state, p.to(DEV), t.to(DEV), thr.to(DEV), num_columns=C, target_is_class_index=True,
                               ignore_index=-1, prob_mode=N.PROB_CHECK)
    keep = t != -1
    want = O.multiclass_precision_recall_curve_update_loop(p.numpy()[keep.numpy()], t.numpy()[keep.numpy()], C,
                                                           thr.numpy())
    assert np.array_equal(_np(state), 2 * want)
    assert N.read_status_words(torch.device(DEV)) == (0, 0)
    # multilabel: per-element targets, logits -> the all-or-nothing sigmoid pass
    pl = torch.randn(n, C, generator=g) * 3
    tl = torch.randint(0, 2, (n, C), generator=g)
    tl[::11, C // 2] = -1
    state = torch.zeros(T, C, 2, 2, dtype=torch.int64, device=DEV)
    N.binned_curve_update_(state, pl.to(DEV), tl.to(DEV), thr.to(DEV), num_columns=C, ignore_index=-1)
    sp = _np(pl.to(DEV).sigmoid())
    want = O.multilabel_precision_recall_curve_update(sp, tl.numpy(), tl.numpy() == -1, C, thr.numpy())
    assert np.array_equal(_np(state), want)
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


# =========================================================================== K2 stat scores
def _stat(N, preds, target, L, threshold=0.5, ignore_index=None, prob_mode=None):
    out = torch.zeros(4, L, dtype=torch.int64, device=DEV)
    N.stat_scores_update_(out, preds.to(DEV), target.to(DEV), L, threshold, ignore_index,
                          N.PROB_AUTO if prob_mode is None else prob_mode)
    return _np(out)


@pytest.mark.parametrize("c", [c for c in load_golden("stat_scores")
                               if c.meta["multidim_average"] == "global" and not c.meta["kind"].startswith("multiclass")],
                         ids=case_id)
def test_stat_scores_golden(c):
    N = _native()
    kind, ii = c.meta["kind"], c.meta["ignore_index"]
    preds, target = torch.from_numpy(c["preds"]), torch.from_numpy(c["target"])
    if kind.startswith("binary"):
        got = _stat(N, preds.reshape(-1), target.reshape(-1), 1, c.meta["threshold"], ii)
        tp, fp, tn, fn = got[:, 0]
        want = c["stat_scores"]
        ours = np.array([tp, fp, tn, fn, tp + fn])
    else:
        L = c.meta["num_labels"]
        # (N, L, X) -> (N*X, L): host-side layout step
        p2 = preds.reshape(preds.shape[0], L, -1).movedim(1, -1).reshape(-1, L).contiguous()
        t2 = target.reshape(target.shape[0], L, -1).movedim(1, -1).reshape(-1, L).contiguous()
        got = _stat(N, p2, t2, L, c.meta["threshold"], ii)
        ours = np.stack([got[0], got[1], got[2], got[3], got[0] + got[3]], axis=-1)
        want = c["stat_scores_none"]
    if kind.endswith("logits"):
        # rebuild the expectation with the device sigmoid (bit-exact target), then compare with the golden loosely
        if kind.startswith("binary"):
            sp = preds.to(DEV).sigmoid().cpu().numpy()
            p, t = O.binary_stat_scores_format(sp, c["target"], c.meta["threshold"], ii)
            tp, fp, tn, fn = O.binary_stat_scores_update(p, t)
            exp = np.array([tp, fp, tn, fn, tp + fn])
        else:
            sp = preds.to(DEV).sigmoid().cpu().numpy()
            p, t = O.multilabel_stat_scores_format(sp, c["target"], c.meta["num_labels"], c.meta["threshold"], ii)
            tp, fp, tn, fn = O.multilabel_stat_scores_update(p, t)
            exp = np.stack([tp, fp, tn, fn, tp + fn], axis=-1)
        assert np.array_equal(ours, exp)
        assert np.abs(ours - want).max() <= 1
    else:
        assert np.array_equal(ours, want)


@pytest.mark.parametrize("L", [1, 2, 3, 4, 5, 7, 8, 100, 513, 2048, 2052])
@pytest.mark.parametrize("n", [1, 3, 1000, 40_003])
def test_stat_scores_seeded(L, n):
    N = _native()
    g = torch.Generator().manual_seed(L * 1000 + n)
    p = torch.rand(n, L, generator=g)
    t = torch.randint(-1, 2, (n, L), generator=g)
    got = _stat(N, p, t, L, 0.37, ignore_index=-1)
    fp_, ft_ = O.multilabel_stat_scores_format(p.numpy()[:, :, None], t.numpy()[:, :, None], L, 0.37, -1)
    tp, fp, tn, fn = O.multilabel_stat_scores_update(fp_, ft_)
    assert np.array_equal(got, np.stack([tp, fp, tn, fn]))


@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16, torch.float64, torch.int64, torch.int32,
                                torch.uint8])
def test_stat_scores_dtypes(dt):
    N = _native()
    g = torch.Generator().manual_seed(21)
    n, L = 5000, 12
    if dt.is_floating_point:
        p = torch.rand(n, L, generator=g, dtype=torch.float64).to(dt)
    else:
        p = torch.randint(0, 2, (n, L), generator=g).to(dt)
    t = torch.randint(0, 2, (n, L), generator=g)
    got = _stat(N, p, t, L, 0.3)
    if dt.is_floating_point:
        pb = (p > 0.3).long()  # torch's own scalar-comparison semantics for this dtype
    else:
        pb = p.long()
    tp, fp, tn, fn = O.multilabel_stat_scores_update(pb.numpy()[:, :, None], t.numpy()[:, :, None])
    assert np.array_equal(got, np.stack([tp, fp, tn, fn]))


def test_stat_scores_sigmoid_rule_counts_ignored_positions():
    N = _native()
    # stat-scores format range-checks the whole tensor, ignored positions included (stat_scores.py:104)
    p = torch.tensor([[0.2, 0.9], [5.0, 0.4], [0.6, 0.1]])
    t = torch.tensor([[1, 0], [-1, 1], [0, 1]])
    got = _stat(N, p, t, 2, 0.5, ignore_index=-1)
    sp = p.to(DEV).sigmoid().cpu().numpy()
    fp_, ft_ = O.multilabel_stat_scores_format(sp[:, :, None], t.numpy()[:, :, None], 2, 0.5, -1)
    tp, fp, tn, fn = O.multilabel_stat_scores_update(fp_, ft_)
    assert np.array_equal(got, np.stack([tp, fp, tn, fn]))
    # followed by a clean batch: flag and scratch were reset
    p2 = torch.rand(1000, 2)
    t2 = torch.randint(0, 2, (1000, 2))
    got = _stat(N, p2, t2, 2)
    tp, fp, tn, fn = O.multilabel_stat_scores_update((p2 > 0.5).long().numpy()[:, :, None], t2.numpy()[:, :, None])
    assert np.array_equal(got, np.stack([tp, fp, tn, fn]))


# =========================================================================== full-size properties
def test_full_size_properties():
    """BASELINE.json sizes (scaled to fit the test budget): totals and linearity instead of an oracle pass."""
    N = _native()
    n = 1 << 24
    g = torch.Generator(device=DEV).manual_seed(1)
    C = 1000
    p = torch.randint(0, C, (n,), device=DEV, generator=g)
    t = torch.randint(0, C, (n,), device=DEV, generator=g)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, p, t, C)
    assert int(cm.sum()) == n
    assert torch.equal(cm.sum(1), torch.bincount(t, minlength=C))
    assert torch.equal(cm.sum(0), torch.bincount(p, minlength=C))
    assert torch.equal(cm.diag(), torch.bincount(t[p == t], minlength=C))
    # linearity: two halves add up to the whole
    cm2 = torch.zeros_like(cm)
    N.confmat_update_(cm2, p[: n // 2], t[: n // 2], C)
    N.confmat_update_(cm2, p[n // 2:], t[n // 2:], C)
    assert torch.equal(cm, cm2)

    n = 1 << 26
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    thr = torch.linspace(0, 1, 200, device=DEV)
    st = torch.zeros(200, 2, 2, dtype=torch.int64, device=DEV)
    N.binned_curve_update_(st, preds, target, thr)
    assert torch.all(st.sum((1, 2)) == n)
    npos = int(target.sum())
    assert torch.all(st[:, 1, :].sum(1) == npos)
    assert torch.all(st[1:, 1, 1] <= st[:-1, 1, 1]) and torch.all(st[1:, 0, 1] <= st[:-1, 0, 1])
    for i in [0, 1, 57, 100, 199]:
        above = preds >= thr[i]
        assert int(st[i, 1, 1]) == int((above & (target == 1)).sum())
        assert int(st[i, 0, 1]) == int((above & (target == 0)).sum())


# --------------------------------------------------------------------------------- fused validation words
def test_validation_findings_go_to_the_second_status_word():
    N = _native()
    dev = torch.device(DEV)
    N.read_status_words(dev)
    n = 1 << 12
    g = torch.Generator().manual_seed(9)
    t = torch.randint(0, 7, (n,), generator=g).to(DEV)
    p = torch.randint(0, 7, (n,), generator=g).to(DEV)
    cm = torch.zeros(7, 7, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, p, t, 7)
    assert N.read_status_words(dev) == (0, 0)
    t2 = t.clone()
    t2[5] = 7  # aliases into the next row's cell: counted like the reference, reported as a finding only
    p2 = p.clone()
    p2[5] = 0
    N.confmat_update_(cm, p2, t2, 7)
    assert N.read_status_words(dev) == (N.STATUS_LABEL_RANGE, N.STATUS_TARGET_VALUES)  # key 49 is outside 7x7
    p3 = p.clone()
    p3[9] = 8
    t3 = t.clone()
    t3[9] = 0
    N.confmat_update_(cm, p3, t3, 7)
    assert N.read_status_words(dev) == (0, N.STATUS_PREDS_VALUES)
    # ignored samples are never findings
    t4 = t.clone()
    t4[3] = -100
    N.confmat_update_(cm, p, t4, 7, ignore_index=-100)
    assert N.read_status_words(dev) == (0, 0)
    # partition path (algo 6)
    C = 300
    n = 1 << 16
    tt = torch.randint(0, C, (n,), generator=g).to(DEV)
    pp = torch.randint(0, C, (n,), generator=g).to(DEV)
    big = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(big, pp, tt, C, algo=6)
    assert N.read_status_words(dev) == (0, 0)
    pp[77] = C + 3
    tt[77] = 1
    N.confmat_update_(big, pp, tt, C, algo=6)
    assert N.read_status_words(dev) == (0, N.STATUS_PREDS_VALUES)

    # K2
    L = 5
    tl = torch.randint(0, 2, (1000, L), generator=g).to(DEV)
    sl = torch.rand(1000, L, generator=g).to(DEV)
    out = torch.zeros(4, L, dtype=torch.int64, device=DEV)
    N.stat_scores_update_(out, sl, tl, L)
    assert N.read_status_words(dev) == (0, 0)
    tl2 = tl.clone()
    tl2[3, 2] = 2
    N.stat_scores_update_(out, sl, tl2, L)
    assert N.read_status_words(dev) == (0, N.STATUS_TARGET_VALUES)
    N.stat_scores_update_(out, sl, tl2, L, ignore_index=2)
    assert N.read_status_words(dev) == (0, 0)
    N.stat_scores_update_(out, sl, tl2, L, report_values=False)
    assert N.read_status_words(dev) == (0, 0)
    N.stat_scores_update_(out, (sl > 0.5).long() * 3, tl, L)
    assert N.read_status_words(dev) == (0, N.STATUS_PREDS_VALUES)

    # K3, both kernels
    thr = torch.linspace(0, 1, 50, device=DEV)
    st = torch.zeros(50, 2, 2, dtype=torch.int64, device=DEV)
    tb = tl[:, 0].contiguous()
    N.binned_curve_update_(st, sl[:, 0].contiguous(), tb, thr)
    assert N.read_status_words(dev) == (0, 0)
    tb2 = tb.clone()
    tb2[11] = 5
    N.binned_curve_update_(st, sl[:, 0].contiguous(), tb2, thr)
    assert N.read_status_words(dev) == (0, N.STATUS_TARGET_VALUES)
    stc = torch.zeros(50, L, 2, 2, dtype=torch.int64, device=DEV)
    cls = torch.randint(0, L, (1000,), generator=g).to(DEV)
    N.binned_curve_update_(stc, sl, cls, thr, num_columns=L, target_is_class_index=True, prob_mode=N.PROB_AS_IS)
    assert N.read_status_words(dev) == (0, 0)
    cls[4] = L
    N.binned_curve_update_(stc, sl, cls, thr, num_columns=L, target_is_class_index=True, prob_mode=N.PROB_AS_IS)
    assert N.read_status_words(dev) == (0, N.STATUS_TARGET_VALUES)
    N.binned_curve_update_(stc, sl, tl2, thr, num_columns=L)
    assert N.read_status_words(dev) == (0, N.STATUS_TARGET_VALUES)


# =========================================================================== K1c per-class stats without the matrix
@pytest.mark.parametrize("C,n", [(2, 1), (5, 1000), (10, 4097), (1000, 100_003), (3000, 50_000), (20_000, 70_001)])
@pytest.mark.parametrize("ignore_index", [None, -1, 1])
@pytest.mark.parametrize("dt", [torch.int64, torch.int32])
def test_class_stats_matches_confusion_matrix_reduction(C, n, ignore_index, dt):
    """tp/fp/tn/fn from three length-C histograms == the reference's reduction of the C x C matrix
    (stat_scores.py:395-419), including ignore_index inside and outside [0, C) and two accumulating updates."""
    N = _native()
    g = torch.Generator().manual_seed(C + n)
    t = torch.randint(0, C, (n,), generator=g)
    p = torch.where(torch.rand(n, generator=g) < 0.6, t, torch.randint(0, C, (n,), generator=g))
    if ignore_index is not None:
        t[::7] = ignore_index
    out = torch.zeros(4, C, dtype=torch.int64, device=DEV)
    for _ in range(2):
        N.class_stats_update_(out, p.to(dt).to(DEV), t.to(dt).to(DEV), C, ignore_index)
    keep = (t != ignore_index) if ignore_index is not None else torch.ones(n, dtype=torch.bool)
    tp, fp, tn, fn = O.multiclass_stat_scores_update(p[keep].numpy(), t[keep].numpy(), C, average="macro")
    want = 2 * np.stack([tp, fp, tn, fn])
    assert np.array_equal(_np(out), want)
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


def test_class_stats_out_of_range_labels_follow_the_key_arithmetic():
    N = _native()
    dev = torch.device(DEV)
    C = 7
    t = torch.tensor([0, 1, 2, 6, 3, 3])
    p = torch.tensor([1, 1, 9, 0, -2, 3])  # (2, 9) aliases to cell (3, 2); (3, -2) to cell (2, 6)
    out = torch.zeros(4, C, dtype=torch.int64, device=DEV)
    N.class_stats_update_(out, p.to(DEV), t.to(DEV), C)
    cm = np.bincount(t.numpy() * C + p.numpy(), minlength=C * C).reshape(C, C)
    tp = np.diag(cm)
    fp, fn = cm.sum(0) - tp, cm.sum(1) - tp
    want = np.stack([tp, fp, cm.sum() - tp - fp - fn, fn])
    assert np.array_equal(_np(out), want)
    assert N.read_status_words(dev) == (0, N.STATUS_PREDS_VALUES)
    # a key outside the matrix is what makes the reference's reshape fail: flagged, not counted
    N.class_stats_update_(out, torch.tensor([0, 50], device=DEV), torch.tensor([0, 6], device=DEV), C)
    assert N.read_status_words(dev) == (N.STATUS_LABEL_RANGE, N.STATUS_PREDS_VALUES)
    assert int(out.sum()) == int(want.sum()) + C  # one more counted sample adds C to tp+fp+tn+fn


@pytest.mark.parametrize("C,k,shape", [(5, 2, (100,)), (10, 3, (64, 7)), (1000, 5, (2048,)), (3000, 10, (300,)),
                                       (20_000, 5, (513,))])
@pytest.mark.parametrize("ignore_index", [None, -1, 1])
def test_class_stats_topk_matches_the_one_hot_branch(C, k, shape, ignore_index):
    """top-k form of K1c == the reference's select_topk / one_hot / four masked sums (stat_scores.py:363-393),
    restated with torch ops on the device (`_multiclass_stat_scores_onehot`), for (N, C) and (N, C, X) scores."""
    from importlib import import_module

    S = import_module("torchmetrics_b200.functional.classification.stat_scores")
    N = _native()
    g = torch.Generator().manual_seed(C * 31 + k)
    n = shape[0]
    preds = torch.randn(n, C, *shape[1:], generator=g).to(DEV)
    preds[: n // 4] = preds[: n // 4].round()  # ties among the scores
    target = torch.randint(0, C, shape, generator=g).to(DEV)
    if ignore_index is not None:
        target[::5] = ignore_index
    p3, t2 = S._multiclass_stat_scores_format(preds, target, k)
    want = S._multiclass_stat_scores_onehot(p3, t2, C, k, "global", ignore_index)
    got = S._multiclass_stat_scores_update(p3, t2, C, k, "macro", "global", ignore_index)
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    assert N.read_status_words(torch.device(DEV)) == (0, 0)
    # a target outside [0, C) is what the reference's one_hot rejects
    bad = target.clone()
    bad.view(-1)[3] = C + 2
    p3, t2 = S._multiclass_stat_scores_format(preds, bad, k)
    S._multiclass_stat_scores_update(p3, t2, C, k, "macro", "global", ignore_index)
    assert N.read_status_words(torch.device(DEV))[0] == N.STATUS_LABEL_RANGE


@pytest.mark.parametrize("C,n", [(2, 5), (5, 1000), (10, 4099), (1000, 3001), (1001, 777), (20_000, 130)])
@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16, torch.float64])
@pytest.mark.parametrize("ignore_index", [None, -1])
def test_class_stats_argmax_matches_argmax_then_count(C, n, dt, ignore_index):
    """argmax form of K1c == torch.argmax followed by the label kernel: ties (first maximum wins), NaN (counts as the
    maximum), +-inf, vector (C % 4 == 0) and scalar rows, ignore_index."""
    N = _native()
    g = torch.Generator().manual_seed(C + n)
    s = torch.randn(n, C, generator=g)
    s[::3] = s[::3].round()  # many exact ties, more so in the 16-bit dtypes
    s[1::17, C // 2] = float("nan")
    s[2::19, 0] = float("inf")
    s[4::23] = float("-inf")  # whole rows of -inf: the first index wins
    s = s.to(dt).to(DEV)
    t = torch.randint(0, C, (n,), generator=g).to(DEV)
    if ignore_index is not None:
        t[::6] = ignore_index
    got = torch.zeros(4, C, dtype=torch.int64, device=DEV)
    N.class_stats_argmax_update_(got, s, t, C, ignore_index)
    N.class_stats_argmax_update_(got, s, t, C, ignore_index)
    want = torch.zeros(4, C, dtype=torch.int64, device=DEV)
    N.class_stats_update_(want, s.argmax(dim=1), t, C, ignore_index)
    assert torch.equal(got, 2 * want)
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


@pytest.mark.parametrize("C,n", [(1, 9), (3, 1000), (10, 4099), (64, 513), (1000, 3001), (1003, 64)])
@pytest.mark.parametrize("dt", [torch.float32, torch.float16, torch.bfloat16, torch.float64])
def test_rows_argmax_is_torch_argmax(C, n, dt):
    N = _native()
    g = torch.Generator().manual_seed(C * 7 + n)
    s = torch.randn(n, C, generator=g)
    s[::2] = s[::2].round()
    if C > 1:
        s[1::17, C // 2] = float("nan")
        s[3::17, 0] = float("nan")
    s[2::19, C - 1] = float("inf")
    s[4::23] = float("-inf")
    s[5::29] = 0.0
    s[5::29, ::2] = -0.0  # +0.0 and -0.0 compare equal: the first index wins
    s = s.to(dt).to(DEV)
    got = N.rows_argmax(s)
    assert got.dtype == torch.int32
    assert torch.equal(got.long(), s.argmax(dim=1))


# =========================================================================== K4b fused exact AUROC / AP
def _auroc_ap_from_curve(N, p, t):
    """The same two scalars from the materialised curve, in float64 (what the fused kernel must reproduce)."""
    fps, tps, _ = N.exact_curve(p, t, 1)
    fps, tps = fps.double().cpu().numpy(), tps.double().cpu().numpy()
    fp0, tp0 = np.concatenate([[0.0], fps[:-1]]), np.concatenate([[0.0], tps[:-1]])
    area2 = float(((fps - fp0) * (tps + tp0)).sum())
    apsum = float(((tps - tp0) * tps / (tps + fps)).sum())
    return area2, apsum, float(tps[-1]), float(len(tps))


@pytest.mark.parametrize("case", ["random", "ties", "constant", "long_runs", "tiny", "ragged", "nan_inf", "two_values"])
def test_exact_auroc_ap_fused_matches_curve(case):
    N = _native()
    g = torch.Generator().manual_seed(11)
    if case == "random":
        n = 200_003
        p = torch.rand(n, generator=g)
    elif case == "ties":
        n = 100_000
        p = torch.randint(0, 50, (n,), generator=g).float() / 50
    elif case == "constant":
        n = 30_000
        p = torch.full((n,), 0.25)
    elif case == "long_runs":  # runs of equal scores far longer than a 4096-key tile, plus singletons
        n = 150_000
        p = torch.cat([torch.full((60_000,), 0.9), torch.rand(10_000, generator=g) * 0.5,
                       torch.full((70_000,), 0.1), torch.rand(10_000, generator=g) * 0.05])
    elif case == "tiny":
        n = 3
        p = torch.tensor([0.2, 0.9, 0.2])
    elif case == "ragged":
        n = 4096 * 3 + 1
        p = torch.rand(n, generator=g).round(decimals=3)
    elif case == "nan_inf":
        n = 10_000
        p = torch.randn(n, generator=g)
        p[::97] = float("nan")
        p[5::101] = float("inf")
        p[7::103] = float("-inf")
    else:
        n = 50_000
        p = (torch.rand(n, generator=g) < 0.5).float()
    t = (torch.rand(n, generator=g) < 0.4).long()
    if case == "tiny":
        t = torch.tensor([0, 1, 1])
    p, t = p.to(DEV), t.to(DEV)
    got = N.exact_auroc_ap(p, t, 1)
    want = _auroc_ap_from_curve(N, p, t)
    assert got[2] == want[2] and got[3] == want[3]
    assert got[0] == want[0], (got[0], want[0])  # integer-valued terms below 2^53: exact in float64
    assert abs(got[1] - want[1]) <= 1e-9 * max(1.0, abs(want[1]))
    again = N.exact_auroc_ap(p, t, 1)
    assert again == got  # fixed summation order: bit-reproducible


def test_exact_auroc_ap_functional_agrees_with_curve_path_and_sklearn():
    import torchmetrics_b200.functional as F
    from sklearn.metrics import average_precision_score, roc_auc_score

    g = torch.Generator().manual_seed(5)
    n = 300_000
    p = torch.rand(n, generator=g).round(decimals=4)
    t = (torch.rand(n, generator=g) < 0.2 + 0.5 * p).long()
    au = float(F.binary_auroc(p.to(DEV), t.to(DEV), thresholds=None))
    ap = float(F.binary_average_precision(p.to(DEV), t.to(DEV), thresholds=None))
    assert abs(au - roc_auc_score(t.numpy(), p.numpy())) < 1e-6
    assert abs(ap - average_precision_score(t.numpy(), p.numpy())) < 1e-6
    # the curve path (what runs for max_fpr / degenerate targets) gives the same numbers
    fpr, tpr, _ = F.binary_roc(p.to(DEV), t.to(DEV), thresholds=None)
    assert abs(au - float(torch.trapz(tpr.double(), fpr.double()))) < 1e-6
    # degenerate targets fall back to the reference behaviour
    with pytest.warns(UserWarning):
        assert float(F.binary_auroc(p.to(DEV), torch.zeros(n, dtype=torch.long, device=DEV), thresholds=None)) == 0.0


# =========================================================================== row-wise (samplewise) forms
@pytest.mark.parametrize("width", [1, 2, 3, 4, 7, 8, 31, 32, 33, 100, 128, 1000, 4096 + 4])
@pytest.mark.parametrize("pdt", [torch.float32, torch.float16, torch.int64, torch.uint8])
def test_stat_scores_rows_vs_masked_sums(width, pdt):
    """K2 row-wise form against the reference's four masked sums over dim 1 (stat_scores.py:120-131)."""
    N = _native()
    n_rows = 257 if width < 1000 else 19
    g = torch.Generator().manual_seed(width * 7 + 1)
    if pdt.is_floating_point:
        preds = torch.rand(n_rows, width, generator=g).to(pdt)
        hard = preds > torch.tensor(0.5, dtype=pdt)  # the comparison happens in the preds dtype
    else:
        preds = torch.randint(0, 2, (n_rows, width), generator=g).to(pdt)
        hard = preds.bool()
    target = torch.randint(0, 2, (n_rows, width), generator=g)
    for ignore_index in (None, -1):
        t = target.clone()
        if ignore_index is not None:
            t[torch.rand(n_rows, width, generator=g) < 0.2] = ignore_index
        got = N.stat_scores_rows(preds.to(DEV), t.to(DEV), width, 0.5, ignore_index).cpu()
        valid = t != ignore_index if ignore_index is not None else torch.ones_like(t, dtype=torch.bool)
        want = torch.stack([(hard & (t == 1) & valid).sum(1), (hard & (t == 0) & valid).sum(1),
                            (~hard & (t == 0) & valid).sum(1), (~hard & (t == 1) & valid).sum(1)])
        assert torch.equal(got, want)
    if pdt.is_floating_point:  # logits through the kernel's sigmoid
        logits = (torch.randn(n_rows, width, generator=g) * 3).to(pdt)
        got = N.stat_scores_rows(logits.to(DEV), target.to(DEV), width, 0.5, None, sigmoid=True).cpu()
        hard = logits.to(DEV).sigmoid().cpu() > torch.tensor(0.5, dtype=pdt)
        want = torch.stack([(hard & (target == 1)).sum(1), (hard & (target == 0)).sum(1),
                            (~hard & (target == 0)).sum(1), (~hard & (target == 1)).sum(1)])
        assert torch.equal(got, want)


@pytest.mark.parametrize("C", [2, 5, 100, 1000, 5000])
@pytest.mark.parametrize("width", [1, 17, 64, 1000])
@pytest.mark.parametrize("ignore_index", [None, -1, 1])
def test_class_stats_rows_vs_onehot(C, width, ignore_index):
    """K1c row-wise form against the one-hot formulation (stat_scores.py:363-393) evaluated with torch ops."""
    N = _native()
    n_rows = 64 if C <= 1000 else 5
    g = torch.Generator().manual_seed(C + width)
    p = torch.randint(0, C, (n_rows, width), generator=g)
    t = torch.randint(0, C, (n_rows, width), generator=g)
    p = torch.where(torch.rand(n_rows, width, generator=g) < 0.5, t, p)
    if ignore_index == -1:
        t[torch.rand(n_rows, width, generator=g) < 0.2] = -1
    got = N.class_stats_rows(p.to(DEV), t.to(DEV), C, ignore_index).cpu()
    valid = t != ignore_index if ignore_index is not None else torch.ones_like(t, dtype=torch.bool)
    poh = torch.nn.functional.one_hot(p, C).bool()
    toh = torch.nn.functional.one_hot(t.clamp(min=0), C).bool()
    v = valid[..., None]
    want = torch.stack([(poh & toh & v).sum(1), (poh & ~toh & v).sum(1), (~poh & ~toh & v).sum(1),
                        (~poh & toh & v).sum(1)])
    assert torch.equal(got, want)
    assert N.read_status(torch.device(DEV)) == 0


def test_samplewise_paths_are_native(monkeypatch):
    """The samplewise branches must go through the row-wise kernels, not through eager one-hot sums."""
    import torchmetrics_b200 as tm
    from importlib import import_module

    S = import_module("torchmetrics_b200.functional.classification.stat_scores")

    def boom(*a, **k):
        raise AssertionError("eager masked sums used")

    monkeypatch.setattr(S, "_four_masked_sums", boom)
    g = torch.Generator().manual_seed(2)
    F = tm.functional
    pb = torch.rand(6, 50, generator=g).to(DEV)
    tb = torch.randint(0, 2, (6, 50), generator=g).to(DEV)
    assert F.binary_stat_scores(pb, tb, multidim_average="samplewise").shape == (6, 5)
    pl = torch.rand(6, 4, 50, generator=g).to(DEV)
    tl = torch.randint(0, 2, (6, 4, 50), generator=g).to(DEV)
    assert F.multilabel_stat_scores(pl, tl, 4, multidim_average="samplewise", average=None).shape == (6, 4, 5)
    pc = torch.randn(6, 7, 50, generator=g).to(DEV)
    tc = torch.randint(0, 7, (6, 50), generator=g).to(DEV)
    assert F.multiclass_stat_scores(pc, tc, 7, multidim_average="samplewise", average=None).shape == (6, 7, 5)
    with pytest.raises(RuntimeError, match="CUDA tensors only"):
        F.binary_stat_scores(pb.cpu(), tb.cpu(), multidim_average="samplewise")


# =========================================================================== score-range partition (N4)
@pytest.mark.parametrize("n", [0, 1, 33, 1025, 100_003, (1 << 21) + 5])
@pytest.mark.parametrize("R", [1, 2, 8, 64])
def test_range_partition_vs_bucketize(n, R):
    """tmb_range_partition against torch.bucketize + a stable sort by range: same grouping, same order, NaN on top,
    ties on a splitter stay together."""
    N = _native()
    g = torch.Generator().manual_seed(n * 7 + R)
    p = torch.rand(n, generator=g)
    p = (p * 50).floor() / 50  # many ties, some exactly on splitters
    if n > 40:
        p[::37] = float("nan")
        p[5] = float("inf")
        p[6] = -float("inf")
    t = torch.randint(0, 3, (n,), generator=g)  # positives are the samples with target == 1
    splitters = torch.sort(torch.rand(R - 1, generator=g)).values
    if R > 2:
        splitters[1] = 0.5  # a splitter equal to many scores
        splitters = torch.sort(splitters).values
    out_p, out_t, totals = N.range_partition(p.to(DEV), t.to(DEV), splitters.to(DEV), 1)
    dest = (R - 1) - torch.bucketize(p, splitters)
    dest[p.isnan()] = 0
    order = torch.argsort(dest, stable=True)
    want_p, want_t = p[order], (t[order] == 1).to(torch.uint8)
    got_p = out_p.cpu()
    assert torch.equal(got_p.isnan(), want_p.isnan())
    assert torch.equal(got_p.nan_to_num(7.0), want_p.nan_to_num(7.0))
    assert torch.equal(out_t.cpu(), want_t)
    assert totals[0].cpu().tolist() == torch.bincount(dest, minlength=R).tolist()
    assert totals[1].cpu().tolist() == torch.bincount(dest[t == 1], minlength=R).tolist()


def test_exact_auroc_ap_offsets_add_up():
    """K4b per score range with rank offsets == K4b of the whole input (what the distributed mode relies on)."""
    N = _native()
    g = torch.Generator(device=DEV).manual_seed(5)
    n = 300_000
    p = (torch.rand(n, device=DEV, generator=g) * 4096).floor() / 4096
    t = (torch.rand(n, device=DEV, generator=g) < 0.2 + 0.5 * p).to(torch.uint8)
    whole = N.exact_auroc_ap(p, t, 1)
    splitters = torch.tensor([0.2, 0.55, 0.8], device=DEV)
    sp, st, totals = N.range_partition(p, t, splitters, 1)
    counts, pos = totals[0].tolist(), totals[1].tolist()
    a = b = 0.0
    start = tp_off = fp_off = 0
    for r in range(4):
        part = N.exact_auroc_ap(sp[start:start + counts[r]].clone(), st[start:start + counts[r]].clone(), 1, tp_off,
                                fp_off)
        a, b = a + part[0], b + part[1]
        start, tp_off, fp_off = start + counts[r], tp_off + pos[r], fp_off + counts[r] - pos[r]
    assert a == whole[0]  # integer-valued terms below 2^53: exact
    assert abs(b - whole[1]) <= 1e-9 * whole[1]
    assert sum(pos) == whole[2]


# ---------------------------------------------------------------------------------------------
# fused reducers (tmb_stat_reduce): the CUDA path against the torch-op path of the same functions on CPU copies
# ---------------------------------------------------------------------------------------------
def _reduce_states(shape, big, seed):
    g = torch.Generator().manual_seed(seed)
    hi = (1 << 40) if big else 50
    st = [torch.randint(0, hi, shape, generator=g, dtype=torch.int64) for _ in range(4)]
    if len(shape):  # classes that never occur, zero denominators
        for t in st:
            t[..., 0] = 0
        if shape[-1] > 2:
            st[0][..., 2] = 0
            st[1][..., 2] = 0
    return st


@pytest.mark.gpu
@pytest.mark.parametrize("big", [False, True])
@pytest.mark.parametrize("shape,mda", [((), "global"), ((1,), "global"), ((5,), "global"), ((100,), "global"),
                                       ((1000,), "global"), ((37, 7), "samplewise"), ((3, 100), "samplewise"),
                                       ((37,), "samplewise")])
def test_stat_reduce_matches_torch_ops(shape, mda, big):
    from torchmetrics_b200.functional.classification.accuracy import _accuracy_reduce
    from torchmetrics_b200.functional.classification.f_beta import _fbeta_reduce
    from torchmetrics_b200.functional.classification.precision_recall import _precision_recall_reduce
    from torchmetrics_b200.functional.classification.specificity import _specificity_reduce

    tp, fp, tn, fn = _reduce_states(shape, big, 7 + len(shape))
    dev = [t.to(DEV) for t in (tp, fp, tn, fn)]
    reducers = [
        ("precision", lambda *a, **k: _precision_recall_reduce("precision", *a, **k)),
        ("recall", lambda *a, **k: _precision_recall_reduce("recall", *a, **k)),
        ("specificity", _specificity_reduce),
        ("accuracy", _accuracy_reduce),
        ("f1", lambda tp, fp, tn, fn, **k: _fbeta_reduce(tp, fp, tn, fn, 1.0, **k)),
        ("f0.5", lambda tp, fp, tn, fn, **k: _fbeta_reduce(tp, fp, tn, fn, 0.5, **k)),
        ("f2", lambda tp, fp, tn, fn, **k: _fbeta_reduce(tp, fp, tn, fn, 2.0, **k)),
    ]
    if len(shape) == 0 or (mda == "samplewise" and len(shape) == 1):
        averages = ["binary"]
    else:
        averages = ["micro", "macro", "weighted", "none", None]
    from torchmetrics_b200 import _native

    supported = _native.stat_reduce_supported
    for name, fn_ in reducers:
        for average in averages:
            for multilabel in (False, True):
                kw = dict(average=average, multidim_average=mda, multilabel=multilabel)
                want = fn_(tp, fp, tn, fn, **kw)
                got = fn_(*dev, **kw)
                _native.stat_reduce_supported = lambda *a: False  # the same torch ops on the GPU
                try:
                    torch_ops = fn_(*dev, **kw)
                finally:
                    _native.stat_reduce_supported = supported
                assert got.is_cuda and got.dtype == want.dtype and got.shape == want.shape, (name, kw)
                assert torch.equal(got, torch_ops), (name, kw, got, torch_ops)  # bit for bit
                if average in ("binary", "none", None, "micro"):
                    assert torch.equal(got.cpu(), want), (name, kw)
                else:  # torch adds the terms in a different order on the CPU
                    assert torch.allclose(got.cpu(), want, rtol=2e-6, atol=1e-7), (name, kw, got.cpu(), want)


@pytest.mark.gpu
def test_stat_reduce_is_one_launch_per_metric():
    """compute() of the precision / recall / F1 modules goes through the fused kernel (same values as the functional
    form evaluated from CPU states)."""
    import torchmetrics_b200 as tm
    from torchmetrics_b200 import _native

    L = 100
    g = torch.Generator().manual_seed(3)
    p = torch.rand(4096, L, generator=g)
    t = torch.randint(0, 2, (4096, L), generator=g)
    calls = []
    orig = _native.stat_reduce
    _native.stat_reduce = lambda *a, **k: (calls.append(a[0]), orig(*a, **k))[1]
    try:
        for cls in (tm.MultilabelPrecision, tm.MultilabelRecall, tm.MultilabelF1Score, tm.MultilabelAccuracy,
                    tm.MultilabelSpecificity):
            m = cls(num_labels=L).to(DEV)
            m.update(p.to(DEV), t.to(DEV))
            got = m.compute()
            tp, fp, tn, fn = (getattr(m, k).cpu() for k in ("tp", "fp", "tn", "fn"))
            ref = cls(num_labels=L)
            ref.tp, ref.fp, ref.tn, ref.fn = tp, fp, tn, fn
            ref._update_count = 1
            assert torch.allclose(got.cpu(), ref.compute(), rtol=2e-6, atol=1e-7), cls.__name__
    finally:
        _native.stat_reduce = orig
    assert len(calls) == 5  # one launch per compute() (plus torch.sum over the labels for the macro average)


@pytest.mark.gpu
@pytest.mark.parametrize("ignore_index", [None, 3, -1])
@pytest.mark.parametrize("top_k", [2, 5])
def test_class_stats_rows_topk_matches_onehot_branch(top_k, ignore_index):
    """samplewise multiclass statistics with top_k > 1: the row-wise kernel on the top-k class ids against the
    reference's one-hot formulation (stat_scores.py:363-393) evaluated with torch ops on the same inputs."""
    from importlib import import_module

    S = import_module("torchmetrics_b200.functional.classification.stat_scores")
    C, N, X = 11, 37, 53
    g = torch.Generator().manual_seed(17 + top_k)
    preds = torch.randn(N, C, X, generator=g)
    preds[:, :, ::7] = preds[:, :1, ::7]  # ties between all classes at some positions
    target = torch.randint(0, C, (N, X), generator=g)
    if ignore_index is not None:
        target[torch.rand(N, X, generator=g) < 0.2] = ignore_index
    preds, target = preds.to(DEV), target.to(DEV)
    want = S._multiclass_stat_scores_onehot(preds, target, C, top_k, "samplewise", ignore_index)
    got = S._multiclass_stat_scores_update(preds, target, C, top_k, "macro", "samplewise", ignore_index)
    for w, g_, name in zip(want, got, ("tp", "fp", "tn", "fn")):
        assert torch.equal(g_, w), name
    # through the public function and the module (cat states)
    import torchmetrics_b200 as tm

    f = tm.functional.multiclass_stat_scores(preds, target, C, average=None, top_k=top_k, multidim_average="samplewise",
                                             ignore_index=ignore_index, validate_args=False)
    assert torch.equal(f, torch.stack([*want, want[0] + want[3]], dim=-1))
    m = tm.MulticlassStatScores(C, top_k=top_k, average=None, multidim_average="samplewise", ignore_index=ignore_index,
                                validate_args=False).to(DEV)
    m.update(preds[:20], target[:20])
    m.update(preds[20:], target[20:])
    assert torch.equal(m.compute(), f)
