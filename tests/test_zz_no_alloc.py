"""The allocation-free curve update (tmb_binned_curve_update_static) must give exactly what the workspace-based
entry point gives.  It is for callers that cannot allocate device memory at update time (no-allocation-after-warm-up
loops, processes whose memory belongs to another pool, pointer-only foreign bindings), so its scratch is a fixed
array inside the library; this test drives it through the C ABI twice in a row, because the second call starts from
whatever the first one left in that shared array.  Collected last on purpose (file name), and written so that it
passes both alone in a fresh process and after every other GPU test has used the library."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _pair(N, preds, target, thr, **kw):
    """(static-scratch result, workspace result) for the same inputs, both through the C ABI."""
    T = thr.numel()
    cols = kw.get("num_columns", 1)
    shape = (T, 2, 2) if cols == 1 else (T, cols, 2, 2)
    got = torch.zeros(shape, dtype=torch.int64, device=DEV)
    want = torch.zeros(shape, dtype=torch.int64, device=DEV)
    N.binned_curve_update_static_(got, preds, target, thr, **kw)
    N.binned_curve_update_(want, preds, target, thr, **kw)
    return got, want


def test_static_scratch_matches_workspace_entry_on_consecutive_calls():
    from torchmetrics_b200 import _native as N

    dev = torch.device(DEV)
    g = torch.Generator(device=DEV).manual_seed(2024)
    thr = torch.linspace(0, 1, 200, device=DEV)
    n = (1 << 20) + 5
    # call 1: probabilities, with some ignored samples
    p1 = torch.rand(n, device=DEV, generator=g)
    t1 = torch.randint(-1, 2, (n,), device=DEV, generator=g)
    got, want = _pair(N, p1, t1, thr, ignore_index=-1)
    assert torch.equal(got, want) and int(got.sum()) == 200 * int((t1 != -1).sum())
    # call 2, straight after: logits, so the kernel abandons its first pass half-way and redoes the batch with
    # sigmoid — the path that leaves the most behind in the scratch
    p2 = torch.randn(n, device=DEV, generator=g) * 3
    t2 = torch.randint(0, 2, (n,), device=DEV, generator=g)
    got, want = _pair(N, p2, t2, thr)
    assert torch.equal(got, want)
    # and once more with several columns and arbitrary (non-uniform) thresholds, accumulating into a used state
    thr3 = torch.sort(torch.rand(50, device=DEV, generator=g)).values
    p3 = torch.rand(4096, 8, device=DEV, generator=g)
    t3 = torch.randint(0, 2, (4096, 8), device=DEV, generator=g)
    got, want = _pair(N, p3, t3, thr3, num_columns=8)
    assert torch.equal(got, want)
    N.binned_curve_update_static_(got, p3, t3, thr3, num_columns=8)
    assert torch.equal(got, 2 * want)
    assert N.read_status(dev) == 0


def test_static_scratch_rejects_problems_beyond_its_size():
    from torchmetrics_b200 import _native as N

    thr = torch.linspace(0, 1, 5000, device=DEV)  # 2 * 5001 counters * 8 B > 64 KiB
    p = torch.rand(1024, device=DEV)
    t = torch.randint(0, 2, (1024,), device=DEV)
    out = torch.zeros(5000, 2, 2, dtype=torch.int64, device=DEV)
    with pytest.raises(RuntimeError, match="TMB_ERR_TOO_LARGE.*built-in array holds 65536"):
        N.binned_curve_update_static_(out, p, t, thr)
    assert int(out.sum()) == 0
    # the workspace-based entry point handles the same problem
    N.binned_curve_update_(out, p, t, thr)
    assert int(out.sum()) == 5000 * 1024
