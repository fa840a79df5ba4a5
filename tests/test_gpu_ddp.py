"""Cross-GPU synchronisation over NCCL (one process per GPU), at every world size in {2, 4, 8} the box has GPUs
for; skipped on a single-GPU box.  Shards are uneven, and from 4 ranks on one rank holds nothing at all.
Run with: gpurun --gpus 2 -- python -m pytest tests/test_gpu_ddp.py -m gpu -q   (logs: profiles/r2_nccl_parity_*.log)"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

pytestmark = pytest.mark.gpu
WORLDS = [2, 4, 8]


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, body):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        globals()[body](rank, world)
        torch.cuda.synchronize()
    finally:
        dist.destroy_process_group()


def _data(rank_or_all, n_per_rank, world):
    """Deterministic per-rank shards; rank None -> everything, in rank order."""
    out = []
    for r in range(world):
        g = torch.Generator().manual_seed(100 + r)
        n = n_per_rank + 17 * r  # uneven shards
        if world > 2 and r == 1:
            n = 0  # a rank that saw no data still takes part in every collective
        p = torch.rand(n, generator=g)
        t = (torch.rand(n, generator=g) < 0.3 + 0.4 * p).long()
        lp = torch.randint(0, 1000, (n,), generator=g)
        lt = torch.randint(0, 1000, (n,), generator=g)
        out.append((p, t, lp, lt))
    if rank_or_all is None:
        return tuple(torch.cat([o[i] for o in out]) for i in range(4))
    return out[rank_or_all]


def _body_sum_and_cat_states(rank, world):
    import numpy as np

    import torchmetrics_b200 as tm
    from oracle import oracle_np as O

    dev = torch.device("cuda", rank)
    p, t, lp, lt = (x.to(dev) for x in _data(rank, 50_000, world))
    ap, at, alp, alt = _data(None, 50_000, world)

    cm = tm.MulticlassConfusionMatrix(1000, validate_args=False).to(dev)
    cm.update(lp, lt)
    got = cm.compute()  # one packed all_reduce over NVLink
    assert np.array_equal(got.cpu().numpy(), O.multiclass_confusion_matrix_update(alp.numpy(), alt.numpy(), 1000))
    # local state restored after compute
    assert int(cm.confmat.sum()) == lp.numel()

    coll = tm.MetricCollection([tm.BinaryAUROC(thresholds=200), tm.BinaryAveragePrecision(thresholds=200)]).to(dev)
    coll.update(p, t)
    out = coll.compute()
    thr = torch.linspace(0, 1, 200)
    want = O.binary_precision_recall_curve_update_hist(ap.numpy(), at.numpy(), thr.numpy())
    assert abs(float(out["BinaryAUROC"]) - float(O.binary_auroc_binned(want, thr.numpy()))) < 1e-6
    assert abs(float(out["BinaryAveragePrecision"]) - float(O.binary_average_precision_binned(want, thr.numpy()))) < 1e-6

    exact = tm.BinaryAUROC(thresholds=None).to(dev)
    exact.update(p, t)  # uneven cat states: padded all_gather
    val = float(exact.compute())
    assert abs(val - float(O.binary_auroc_exact(ap.numpy(), at.numpy(), np.float64))) < 1e-6
    assert isinstance(exact.preds, list) and exact.preds[0].numel() == p.numel()  # un-synced again

    # the exact curve itself after the gather: counts bit-exact against the oracle on the concatenated shards
    prc = tm.BinaryPrecisionRecallCurve(thresholds=None).to(dev)
    prc.update(p, t)
    with prc.sync_context():
        from torchmetrics_b200.functional.classification.precision_recall_curve import _binary_clf_curve
        from torchmetrics_b200.utilities.data import dim_zero_cat

        fps, tps, thr = _binary_clf_curve(dim_zero_cat(prc.preds), dim_zero_cat(prc.target))
    wf, wt, wthr = O.binary_clf_curve_exact(ap.numpy(), at.numpy())
    assert np.array_equal(fps.cpu().numpy(), wf) and np.array_equal(tps.cpu().numpy(), wt)
    assert np.array_equal(thr.cpu().numpy(), wthr)


def _body_partitioned_exact(rank, world):
    """distributed_exact="partition": range partition + local sort, against the gathered result and the oracle."""
    import numpy as np

    import torchmetrics_b200 as tm
    from oracle import oracle_np as O

    dev = torch.device("cuda", rank)
    for quantise in (False, True):  # heavy ties must stay on one rank
        p, t, _, _ = (x.to(dev) for x in _data(rank, 200_000, world))
        ap_, at_, _, _ = _data(None, 200_000, world)
        if quantise:
            p, ap_ = (p * 64).floor() / 64, (ap_ * 64).floor() / 64
        part_auc = tm.BinaryAUROC(thresholds=None, distributed_exact="partition").to(dev)
        part_ap = tm.BinaryAveragePrecision(thresholds=None, distributed_exact="partition").to(dev)
        gath_auc = tm.BinaryAUROC(thresholds=None).to(dev)
        gath_ap = tm.BinaryAveragePrecision(thresholds=None).to(dev)
        for m in (part_auc, part_ap, gath_auc, gath_ap):
            m.update(p[: p.numel() // 2], t[: p.numel() // 2])
            m.update(p[p.numel() // 2:], t[p.numel() // 2:])
        a, b = float(part_auc.compute()), float(part_ap.compute())
        assert abs(a - float(gath_auc.compute())) < 1e-6 and abs(b - float(gath_ap.compute())) < 1e-6
        assert abs(a - float(O.binary_auroc_exact(ap_.numpy(), at_.numpy(), np.float64))) < 1e-6
        assert abs(b - float(O.binary_average_precision_exact(ap_.numpy(), at_.numpy(), np.float64))) < 1e-6
        # the states stayed local: nothing was gathered
        assert sum(x.numel() for x in part_auc.preds) == p.numel()


def _body_oneshot_allreduce(rank, world):
    """K5 against NCCL: every dtype / op, sizes from 1 element to the 64 KB limit, many calls in a row (the two-slot
    epoch scheme), bit-identical results on every rank."""
    from torchmetrics_b200 import _native
    from torchmetrics_b200.utilities import distributed as D

    dev = torch.device("cuda", rank)
    g = torch.Generator(device=dev).manual_seed(1000 + rank)
    used = 0
    for it in range(40):
        for dtype, n in ((torch.int64, 1), (torch.int64, 800), (torch.int64, 8192), (torch.float32, 777),
                         (torch.float64, 100), (torch.int32, 16384)):
            for op in ("sum", "max", "min"):
                if dtype.is_floating_point:
                    x = torch.randn(n, device=dev, generator=g, dtype=dtype)
                else:
                    x = torch.randint(-1000, 1000, (n,), device=dev, generator=g, dtype=dtype)
                want = x.clone()
                dist.all_reduce(want, op={"sum": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}[op])
                got = D.fused_reduce([(op, x)])[0]
                wire_bytes = n * D._wire_dtype(op, dtype).itemsize  # narrow integer sums travel as int64
                if wire_bytes <= 65536 and wire_bytes * world <= 131072:
                    used += 1
                if dtype.is_floating_point and op == "sum":
                    assert torch.allclose(got, want, rtol=1e-5, atol=1e-5)
                    # rank-ordered summation: the bits agree on every rank
                    ref = got.clone()
                    dist.broadcast(ref, 0)
                    assert torch.equal(got, ref)
                else:
                    assert torch.equal(got, want), (it, dtype, n, op)
    assert used and D._ONESHOT and all(v is not False for v in D._ONESHOT.values()), "one-shot path was not taken"
    assert next(iter(D._ONESHOT.values())).epoch == used
    assert _native.read_status(dev) == 0
    # above the slot size the packed reduce goes through NCCL as before
    big = torch.ones(1 << 16, dtype=torch.int64, device=dev)
    assert int(D.fused_reduce([("sum", big)])[0][0]) == world


def _body_multilabel_collection(rank, world):
    import torchmetrics_b200 as tm

    dev = torch.device("cuda", rank)
    L = 100
    g = torch.Generator().manual_seed(5)
    allp = torch.rand(world * 4096, L, generator=g)
    allt = torch.randint(0, 2, (world * 4096, L), generator=g)
    mine = slice(rank * 4096, (rank + 1) * 4096)
    mc = tm.MetricCollection([tm.MultilabelStatScores(num_labels=L), tm.MultilabelPrecision(num_labels=L),
                              tm.MultilabelRecall(num_labels=L), tm.MultilabelF1Score(num_labels=L)]).to(dev)
    mc.update(allp[mine].to(dev), allt[mine].to(dev))
    mc.update(allp[mine].to(dev), allt[mine].to(dev))
    out = mc.compute()
    F = tm.functional
    both_p, both_t = torch.cat([allp, allp]).to(dev), torch.cat([allt, allt]).to(dev)
    assert torch.equal(out["MultilabelStatScores"], F.multilabel_stat_scores(both_p, both_t, L))
    assert torch.allclose(out["MultilabelF1Score"], F.multilabel_f1_score(both_p, both_t, L), atol=1e-6)


@pytest.mark.parametrize("world", WORLDS)
@pytest.mark.parametrize("body", ["_body_sum_and_cat_states", "_body_multilabel_collection", "_body_partitioned_exact", "_body_oneshot_allreduce"])
def test_nccl_sync(body, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    mp.spawn(_worker, args=(world, _free_port(), body), nprocs=world, join=True)
