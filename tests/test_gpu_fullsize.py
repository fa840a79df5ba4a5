"""Parity at BASELINE.json's full sizes, where the numpy oracle would take minutes: the CUDA path (through the C ABI)
is compared with device restatements built from torch ops that follow the same reference lines as the oracle
(`torch.bincount(t*C+p)` = confusion_matrix.py:333-337; `preds >= thr[i]` counts = precision_recall_curve.py:228-250;
`argsort` + `cumsum` = precision_recall_curve.py:28-80).  Every comparison is `torch.equal` on the complete state."""
import numpy as np
import pytest
import torch

from oracle import oracle_np as O

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def _native():
    from torchmetrics_b200 import _native

    return _native


def _bincount_matrix(p, t, C, ignore_index=None):
    if ignore_index is not None:
        keep = t != ignore_index
        p, t = p[keep], t[keep]
    return torch.bincount(t.long() * C + p.long(), minlength=C * C).view(C, C)


def _confmat_inputs(kind, n, C, g):
    t = torch.randint(0, C, (n,), device=DEV, generator=g)
    if kind == "uniform":
        p = torch.randint(0, C, (n,), device=DEV, generator=g)
    elif kind == "diagonal":  # 80 % correct predictions
        p = torch.where(torch.rand(n, device=DEV, generator=g) < 0.8, t,
                        torch.randint(0, C, (n,), device=DEV, generator=g))
    else:  # hot cell: half of the batch lands in one cell
        p = torch.randint(0, C, (n,), device=DEV, generator=g)
        hot = torch.rand(n, device=DEV, generator=g) < 0.5
        p = torch.where(hot, torch.full_like(p, 3), p)
        t = torch.where(hot, torch.full_like(t, 7), t)
    return p, t


@pytest.mark.parametrize("kind", ["uniform", "diagonal", "hot_cell"])
@pytest.mark.parametrize("dt", [torch.int64, torch.int32])
def test_confmat_c1000_2p24_full_matrix(kind, dt):
    """configs[1]: the whole 1000 x 1000 matrix of the default path for 2^24 samples."""
    N = _native()
    n, C = 1 << 24, 1000
    g = torch.Generator(device=DEV).manual_seed(11)
    p, t = _confmat_inputs(kind, n, C, g)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.read_status_words(torch.device(DEV))  # findings left behind by earlier tests are not this call's
    N.confmat_update_(cm, p.to(dt), t.to(dt), C)
    assert torch.equal(cm, _bincount_matrix(p, t, C))
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


@pytest.mark.parametrize("ignore_index", [-1, 0, 999])
def test_confmat_c1000_2p24_ignore_index(ignore_index):
    N = _native()
    n, C = 1 << 24, 1000
    g = torch.Generator(device=DEV).manual_seed(12)
    p, t = _confmat_inputs("uniform", n, C, g)
    if ignore_index < 0:
        t = torch.where(torch.rand(n, device=DEV, generator=g) < 0.1, torch.full_like(t, ignore_index), t)
    cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    N.confmat_update_(cm, p, t, C, ignore_index=ignore_index)
    assert torch.equal(cm, _bincount_matrix(p, t, C, ignore_index))
    N.read_status_words(torch.device(DEV))


def test_confmat_metric_class_2p24():
    """The same through `MulticlassConfusionMatrix.update` (what bench.py times), two batches."""
    import torchmetrics_b200 as tm

    n, C = 1 << 24, 1000
    g = torch.Generator(device=DEV).manual_seed(13)
    m = tm.MulticlassConfusionMatrix(C, validate_args=False).to(DEV)
    want = torch.zeros(C, C, dtype=torch.int64, device=DEV)
    for _ in range(2):
        p, t = _confmat_inputs("uniform", n, C, g)
        m.update(p, t)
        want += _bincount_matrix(p, t, C)
    assert torch.equal(m.compute(), want)


def _binned_restatement(preds, target, thr):
    """(T,2,2) int64 from a histogram of `#thresholds <= score` and a suffix sum: tp[i] = #(positive, b > i)."""
    T = thr.numel()
    hist = torch.zeros(2, T + 1, dtype=torch.int64, device=preds.device)
    step = 1 << 26
    for s in range(0, preds.numel(), step):
        b = torch.bucketize(preds[s:s + step], thr, right=True)
        pos = target[s:s + step] == 1
        hist[1] += torch.bincount(b[pos], minlength=T + 1)
        hist[0] += torch.bincount(b[~pos], minlength=T + 1)
    above = hist.flip(1).cumsum(1).flip(1)[:, 1:]  # [:, i] = #(b >= i + 1)
    tot = hist.sum(1)
    st = torch.zeros(T, 2, 2, dtype=torch.int64, device=preds.device)
    st[:, 1, 1] = above[1]
    st[:, 0, 1] = above[0]
    st[:, 1, 0] = tot[1] - above[1]
    st[:, 0, 0] = tot[0] - above[0]
    return st


def test_binned_t200_2p28_all_thresholds():
    """configs[2] on one GPU: every one of the 200 x 2 x 2 counters for 2^28 scores."""
    N = _native()
    n = 1 << 28
    g = torch.Generator(device=DEV).manual_seed(21)
    preds = torch.rand(n, device=DEV, generator=g)
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    thr = torch.linspace(0, 1, 200, device=DEV)
    st = torch.zeros(200, 2, 2, dtype=torch.int64, device=DEV)
    N.read_status_words(torch.device(DEV))
    N.binned_curve_update_(st, preds, target, thr)
    assert torch.equal(st, _binned_restatement(preds, target, thr))
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


def test_binned_collection_2p26_logits_all_thresholds():
    """Scores outside [0,1] (the all-or-nothing sigmoid branch) through MetricCollection.update, 2^26 samples."""
    import torchmetrics_b200 as tm

    n = 1 << 26
    g = torch.Generator(device=DEV).manual_seed(22)
    logits = torch.randn(n, device=DEV, generator=g) * 2
    target = (torch.rand(n, device=DEV, generator=g) < torch.sigmoid(logits)).long()
    mc = tm.MetricCollection([tm.BinaryAUROC(thresholds=200), tm.BinaryAveragePrecision(thresholds=200)]).to(DEV)
    mc.update(logits, target)
    state = mc.__getitem__("BinaryAUROC", copy_state=False).confmat
    thr = torch.linspace(0, 1, 200, device=DEV)
    assert torch.equal(state, _binned_restatement(torch.sigmoid(logits), target, thr))


def _sorted_curve_restatement(preds, target):
    n = preds.numel()
    sp, order = torch.sort(preds, descending=True, stable=True)
    st = target[order]
    # run boundaries as the reference finds them: `torch.where(preds[1:] - preds[:-1])` (inf - inf is NaN, i.e.
    # non-zero, so equal infinities do NOT merge; -0.0 and 0.0 do)
    ends = torch.cat([torch.nonzero((sp[1:] - sp[:-1]) != 0).flatten(), torch.tensor([n - 1], device=preds.device)])
    tps = torch.cumsum(st, 0)[ends]
    return ends + 1 - tps, tps, sp[ends]


@pytest.mark.parametrize("ties", ["rand", "heavy"])
def test_exact_curve_2p26_vs_sort(ties):
    """configs[3] after the gather: 2^26 scores; `torch.rand` has 2^24 distinct values, so ties are everywhere."""
    N = _native()
    n = 1 << 26
    g = torch.Generator(device=DEV).manual_seed(31)
    preds = torch.rand(n, device=DEV, generator=g)
    if ties == "heavy":
        preds = (preds * 1000).floor() / 1000
    target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
    fps, tps, thr = N.exact_curve(preds, target, 1)
    wf, wt, wthr = _sorted_curve_restatement(preds, target)
    assert torch.equal(thr, wthr)
    assert torch.equal(tps, wt)
    assert torch.equal(fps, wf)
    # K4b: the fused reductions against the same sums taken from the curve in float64
    a, b, npos, nd = N.exact_auroc_ap(preds, target, 1)
    assert int(npos) == int(wt[-1]) and int(nd) == wthr.numel()
    tpd, fpd = wt.double(), wf.double()
    tp0 = torch.cat([torch.zeros(1, device=DEV, dtype=torch.float64), tpd[:-1]])
    fp0 = torch.cat([torch.zeros(1, device=DEV, dtype=torch.float64), fpd[:-1]])
    want_a = float(((fpd - fp0) * (tpd + tp0)).sum())
    want_b = float(((tpd - tp0) * (tpd / (tpd + fpd))).sum())
    assert abs(a - want_a) <= 1e-12 * want_a
    assert abs(b - want_b) <= 1e-9 * want_b


@pytest.mark.parametrize("n", [0, 1, 31, 4096, 8191, 8192, 8193, 65536 + 5, (1 << 20) + 1])
@pytest.mark.parametrize("dist", ["rand", "const", "two_values", "normal", "sorted", "reversed", "specials"])
def test_exact_curve_shapes_and_distributions(n, dist):
    """Tile boundaries, degenerate digit histograms (one bin holds everything), pre-sorted input, NaN/inf/-0."""
    N = _native()
    g = torch.Generator(device=DEV).manual_seed(1000 + n)
    if dist == "rand":
        preds = torch.rand(n, device=DEV, generator=g)
    elif dist == "const":
        preds = torch.full((n,), 0.25, device=DEV)
    elif dist == "two_values":
        preds = (torch.rand(n, device=DEV, generator=g) < 0.5).float()
    elif dist == "normal":
        preds = torch.randn(n, device=DEV, generator=g) * 1e3
    elif dist == "sorted":
        preds = torch.sort(torch.rand(n, device=DEV, generator=g)).values
    elif dist == "reversed":
        preds = torch.sort(torch.rand(n, device=DEV, generator=g), descending=True).values
    else:
        pool = torch.tensor([0.0, -0.0, float("inf"), -float("inf"), 1e-45, -1e-45, 3.4e38, -3.4e38, 1.0, 0.5],
                            device=DEV)
        preds = pool[torch.randint(0, pool.numel(), (n,), device=DEV, generator=g)]
    target = (torch.rand(n, device=DEV, generator=g) < 0.4).long()
    if n == 0:
        fps, tps, thr = N.exact_curve(preds, target, 1)
        assert fps.numel() == 0 and tps.numel() == 0 and thr.numel() == 0
        return
    fps, tps, thr = N.exact_curve(preds, target, 1)
    wf, wt, wthr = _sorted_curve_restatement(preds, target)
    # -0.0 and 0.0 form one run; which of the two zeros labels it is unspecified (the reference reports whichever
    # argsort left last), so thresholds are compared by value
    assert thr.shape == wthr.shape and bool((thr == wthr).all())
    assert torch.equal(tps, wt) and torch.equal(fps, wf)


@pytest.mark.parametrize("dt", [torch.int64, torch.int32, torch.uint8])
@pytest.mark.parametrize("n,m", [(0, 4), (1, 1), (1000, 7), (65537, 1000), ((1 << 20) + 3, 100_000)])
def test_bincount_update_vs_oracle(dt, n, m):
    """Row A1: `_bincount` (utilities/data.py:179-207) called directly through the C ABI."""
    if dt == torch.uint8 and m > 256:
        pytest.skip("values do not fit")
    N = _native()
    g = torch.Generator().manual_seed(n + m)
    x = torch.randint(0, m, (n,), generator=g).to(dt)
    out = torch.zeros(m, dtype=torch.int64, device=DEV)
    N.read_status_words(torch.device(DEV))
    N.bincount_update_(out, x.to(DEV))
    N.bincount_update_(out, x.to(DEV))  # accumulates
    assert np.array_equal(out.cpu().numpy(), 2 * O.bincount(x.numpy().astype(np.int64), m))
    assert N.read_status_words(torch.device(DEV)) == (0, 0)


def test_bincount_seam_matches_torch():
    from torchmetrics_b200.utilities.data import _bincount

    g = torch.Generator(device=DEV).manual_seed(4)
    x = torch.randint(0, 5000, (1 << 22,), device=DEV, generator=g)
    assert torch.equal(_bincount(x, minlength=5000), torch.bincount(x, minlength=5000))
