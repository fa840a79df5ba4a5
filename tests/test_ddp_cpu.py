"""world_size-2 gloo tests of the sync layer on CPU (what the reference pins in tests/unittests/bases/test_ddp.py)."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from helpers import DummyCounts, DummyList, DummySum  # noqa: E402
from torchmetrics_b200.collections import MetricCollection  # noqa: E402
from torchmetrics_b200.metric import Metric  # noqa: E402
from torchmetrics_b200.utilities.distributed import gather_all_tensors  # noqa: E402

WORLD = 2


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, port, fn_name):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=WORLD)
    try:
        globals()[fn_name](rank)
    finally:
        dist.destroy_process_group()


def _run(fn_name):
    mp.spawn(_worker, args=(_free_port(), fn_name), nprocs=WORLD, join=True)


# ---------------------------------------------------------------- bodies (run inside the workers)
def _body_gather_uneven(rank):
    t = torch.ones(rank + 1, 2 - rank) * (rank + 1)
    out = gather_all_tensors(t)
    assert len(out) == WORLD
    for r in range(WORLD):
        assert out[r].shape == (r + 1, 2 - r) and bool((out[r] == r + 1).all())
    s = gather_all_tensors(torch.tensor(float(rank)))
    assert [float(x) for x in s] == [0.0, 1.0]


def _body_fused_reductions(rank):
    m = DummyCounts(3)
    m.update(torch.tensor([1, 2, 3]) * (rank + 1))
    local_tp = m.tp.clone()
    m.sync()
    assert m._is_synced
    assert m.tp.tolist() == [3, 6, 9] and m.fp.tolist() == [6, 12, 18]  # sum over ranks, exact int64
    assert m.tp.dtype == torch.int64
    assert m.hi.tolist() == [2.0, 4.0, 6.0] and m.lo.tolist() == [1.0, 2.0, 3.0]
    assert torch.allclose(m.avg, torch.tensor([1.5, 3.0, 4.5]))
    m.unsync()
    assert torch.equal(m.tp, local_tp) and not m._is_synced
    assert int(m.compute()) == 18 + 36  # compute syncs and restores
    assert torch.equal(m.tp, local_tp)


class _Counting:
    """Count collectives issued through torch.distributed while active."""

    def __enter__(self):
        self.calls = {"all_reduce": 0, "all_gather": 0, "all_gather_into_tensor": 0, "barrier": 0}
        self._orig = {k: getattr(dist, k) for k in self.calls}
        for k in self.calls:
            def make(k):
                def wrapped(*a, **kw):
                    self.calls[k] += 1
                    return self._orig[k](*a, **kw)
                return wrapped
            setattr(dist, k, make(k))
        return self

    def __exit__(self, *exc):
        for k, v in self._orig.items():
            setattr(dist, k, v)


def _body_one_allreduce_per_metric(rank):
    m = DummyCounts(3)
    m.update(torch.tensor([1, 2, 3]))
    with _Counting() as c:
        m.sync()
    # two int64 sum states -> 1 all_reduce; max / min / mean-as-sum float states -> one each (different ops)
    assert c.calls == {"all_reduce": 4, "all_gather": 0, "all_gather_into_tensor": 0, "barrier": 0}, c.calls


def _body_cat_states(rank):
    m = DummyList()
    for i in range(rank + 1):  # uneven number of rows per rank
        m.update(torch.full((2, 3), float(10 * rank + i)))
    m.sync()
    assert isinstance(m.x, torch.Tensor) and m.x.shape == (2 * 3, 3)
    assert m.x[:, 0].tolist() == [0.0, 0.0, 10.0, 10.0, 11.0, 11.0]
    m.unsync()
    assert isinstance(m.x, list) and len(m.x) == rank + 1
    # all ranks empty -> stays an empty list; one rank empty -> that rank still takes part
    e = DummyList()
    e.update()
    e.sync()
    assert e.x == []
    e.unsync()
    h = DummyList()
    if rank == 1:
        h.update(torch.tensor([1, 2, 3]))
    else:
        h.update()
    h.sync()
    assert h.x.tolist() == [1, 2, 3]


def _body_state_dict_when_synced(rank):
    m = DummySum()
    m.persistent(True)
    m.update(rank + 1)
    assert m.state_dict()["x"] == rank + 1
    m.sync()
    assert m.state_dict()["x"] == 3
    m.unsync()
    with m.sync_context():
        assert m.state_dict()["x"] == 3
    assert m.state_dict()["x"] == rank + 1


class _NoneRed(Metric):
    full_state_update = False

    def __init__(self):
        super().__init__()
        self.add_state("a", torch.zeros(2), dist_reduce_fx=None)
        self.add_state("b", torch.zeros(2), dist_reduce_fx=lambda x: x.sum(0) * 10)

    def update(self, x):
        self.a = self.a + x
        self.b = self.b + x

    def compute(self):
        return self.a, self.b


def _body_none_and_custom_reduction(rank):
    m = _NoneRed()
    m.update(torch.tensor([1.0, 2.0]) * (rank + 1))
    m.sync()
    assert m.a.shape == (WORLD, 2) and m.a.tolist() == [[1.0, 2.0], [2.0, 4.0]]
    assert m.b.tolist() == [30.0, 60.0]


def _body_custom_sync_fn_protocol(rank):
    calls = []

    def my_sync(t, group=None):
        calls.append(t.shape)
        return gather_all_tensors(t, group)

    m = DummyCounts(3, dist_sync_fn=my_sync)
    m.update(torch.tensor([1, 2, 3]))
    assert int(m.compute()) == 2 * 18
    assert len(calls) == 5  # once per state tensor, as the reference does


class _Narrow(Metric):
    full_state_update = False

    def __init__(self, **kwargs):
        super().__init__(**kwargs)
        self.add_state("flag", torch.zeros(3, dtype=torch.bool), dist_reduce_fx="sum")
        self.add_state("small", torch.zeros(2, dtype=torch.int32), dist_reduce_fx="sum")
        self.add_state("byte", torch.zeros(1, dtype=torch.uint8), dist_reduce_fx="sum")
        self.add_state("half", torch.zeros(2, dtype=torch.float16), dist_reduce_fx="mean")
        self.add_state("seen", torch.zeros(2, dtype=torch.bool), dist_reduce_fx="max")

    def update(self, rank):
        self.flag = torch.tensor([True, rank == 1, False])
        self.small = torch.tensor([2**31 - 1, rank], dtype=torch.int32)
        self.byte = torch.tensor([200], dtype=torch.uint8)
        self.half = torch.tensor([60000.0, 1.0 + rank], dtype=torch.float16)
        self.seen = torch.tensor([rank == 0, False])

    def compute(self):
        return self.flag


def _body_narrow_dtypes_follow_the_reference(rank):
    """Sums of bool / narrow integer states are int64 counts (torch.sum of the stacked gather, metric.py:449-456),
    half-precision means do not overflow on the way; the per-state protocol gives the same dtypes and values."""
    m = _Narrow()
    m.update(rank)
    m.sync()
    ref = _Narrow(dist_sync_fn=gather_all_tensors)
    ref.update(rank)
    ref.sync()
    for name in ("flag", "small", "byte", "half", "seen"):
        a, b = getattr(m, name), getattr(ref, name)
        assert a.dtype == b.dtype, (name, a.dtype, b.dtype)
        assert torch.equal(a, b), (name, a, b)
    assert m.flag.tolist() == [2, 1, 0] and m.small.tolist() == [2 * (2**31 - 1), 1] and m.byte.tolist() == [400]
    assert m.half.tolist() == [60000.0, 1.5] and m.seen.tolist() == [True, False]


class _SS(DummyCounts):
    def compute(self):
        return self.tp


class _Prec(DummyCounts):
    def compute(self):
        return self.tp.float() / (self.tp + self.fp).float()


def _body_collection_single_exchange(rank):
    mc = MetricCollection([_SS(3), _Prec(3), DummySum()])
    mc.update(torch.tensor([1, 2, 3]) * (rank + 1))
    mc.update(torch.tensor([1, 1, 1]))
    assert mc.compute_groups == {0: ["DummySum"], 1: ["_SS", "_Prec"]} or len(mc.compute_groups) == 2
    with _Counting() as c:
        out = mc.compute()
    assert out["_SS"].tolist() == [5, 8, 11]
    assert torch.allclose(out["_Prec"], torch.full((3,), 1 / 3))
    # one all_reduce per (op, dtype) bucket for the WHOLE collection: int64-sum, f32-sum(mean + DummySum), max, min
    assert c.calls["all_reduce"] == 4 and c.calls["all_gather"] == 0 and c.calls["barrier"] == 0, c.calls
    # local states restored, still shared inside the group, accumulation continues
    ss = mc.__getitem__("_SS", copy_state=False)
    pr = mc.__getitem__("_Prec", copy_state=False)
    assert ss.tp is pr.tp and not ss._is_synced and ss._to_sync and pr._to_sync
    assert ss.tp.tolist() == [(rank + 1) + 1, 2 * (rank + 1) + 1, 3 * (rank + 1) + 1]
    mc.update(torch.tensor([1, 1, 1]))
    assert mc.compute()["_SS"].tolist() == [7, 10, 13]


@pytest.mark.parametrize("body", [n for n in list(globals()) if n.startswith("_body_")])
def test_ddp(body):
    _run(body)
