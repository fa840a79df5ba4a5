"""Host-side semantics of Metric / MetricCollection on CPU (no kernels involved).
Mirrors what the reference pins in tests/unittests/bases/test_metric.py and test_collections.py."""
import pickle
from unittest.mock import Mock

import pytest
import torch

from helpers import DummyCounts, DummyDiff, DummyList, DummySum
from torchmetrics_b200.collections import MetricCollection
from torchmetrics_b200.metric import CompositionalMetric, Metric
from torchmetrics_b200.utilities.exceptions import TorchMetricsUserError


def test_kwargs_validation():
    for kw in ["compute_on_cpu", "dist_sync_on_step", "sync_on_compute", "compute_with_cache"]:
        with pytest.raises(ValueError, match=f"Expected keyword argument `{kw}` to be an? `bool` but got 3"):
            DummySum(**{kw: 3})
    with pytest.raises(ValueError, match="Expected keyword argument `dist_sync_fn` to be an callable function"):
        DummySum(dist_sync_fn=[2, 3])
    with pytest.raises(ValueError, match="Unexpected keyword arguments: `bar`, `foo`"):
        DummySum(foo=1, bar=2)


def test_add_state():
    m = DummySum()
    m.add_state("a", torch.tensor(0), "sum")
    assert m._reductions["a"](torch.tensor([1, 1])) == 2
    m.add_state("b", torch.tensor(0), "mean")
    assert m._reductions["b"](torch.tensor([1.0, 2.0])) == 1.5
    m.add_state("c", torch.tensor(0), "cat")
    assert m._reductions["c"]([torch.tensor([1]), torch.tensor([1])]).shape == (2,)
    m.add_state("d", torch.tensor(0), "max")
    m.add_state("e", torch.tensor(0), "min")
    with pytest.raises(ValueError, match="`dist_reduce_fx` must be callable or one of .*"):
        m.add_state("f", torch.tensor(0), "xyz")
    with pytest.raises(ValueError, match="`dist_reduce_fx` must be callable or one of .*"):
        m.add_state("g", torch.tensor(0), 42)
    with pytest.raises(ValueError, match="state variable must be a tensor or any empty list .*"):
        m.add_state("h", [torch.tensor(0)], "sum")
    with pytest.raises(ValueError, match="state variable must be a tensor or any empty list .*"):
        m.add_state("i", 42, "sum")
    m.add_state("j", torch.tensor(0), lambda x: x)
    assert not m._persistent["j"]
    m.add_state("k", torch.tensor(0), "sum", persistent=True)
    assert "k" in m.state_dict() and "a" not in m.state_dict()


def test_reset_update_compute_cache():
    m = DummySum()
    assert m._computed is None
    m.update(1)
    m.update(2)
    assert m.update_count == 2 and m.update_called
    assert m.compute() == 3
    assert m._computed == 3
    m.update(1)
    assert m._computed is None
    assert m.compute() == 4
    m.reset()
    assert m.x == 0 and m._computed is None and m.update_count == 0
    lm = DummyList()
    lm.update(torch.tensor(1))
    lm.reset()
    assert lm.x == []
    with pytest.warns(UserWarning, match="was called before the ``update`` method"):
        DummySum().compute()
    nocache = DummySum(compute_with_cache=False)
    nocache.update(1)
    nocache.compute()
    assert nocache._computed is None


@pytest.mark.parametrize("cls", [DummySum, DummyList])
def test_forward_modes(cls):
    m = cls()
    if cls is DummySum:
        assert m(5) == 5 and m._forward_cache == 5
        assert m(8) == 8 and m._forward_cache == 8
        assert m.compute() == 13 and m.update_count == 2
    else:
        assert m(torch.tensor([1])) == [torch.tensor([1])]
        m(torch.tensor([2]))
        assert len(m.x) == 2 and m.update_count == 2


def test_forward_reduce_all_reductions():
    m = DummyCounts(3)
    m(torch.tensor([1, 2, 3]))
    m(torch.tensor([3, 0, 1]))
    assert m.tp.tolist() == [4, 2, 4] and m.fp.tolist() == [8, 4, 8]
    assert m.hi.tolist() == [3, 2, 3] and m.lo.tolist() == [1, 0, 1]
    assert torch.allclose(m.avg, torch.tensor([2.0, 1.0, 2.0]))
    assert m.update_count == 2


def test_constants_are_immutable():
    m = DummySum()
    for name in ["is_differentiable", "higher_is_better", "full_state_update"]:
        with pytest.raises(RuntimeError, match=f"Can't change const `{name}`."):
            setattr(m, name, True)


def test_pickle_clone_hash_statedict():
    m = DummySum()
    m.update(1)
    m2 = pickle.loads(pickle.dumps(m))
    assert m2.compute() == 1
    m2.update(5)
    assert m2.compute() == 6 and m.compute() == 1
    c = m.clone()
    c.update(2)
    assert c.compute() == 3 and m.compute() == 1
    assert hash(m) != hash(c) and hash(DummyList()) != hash(DummyList())
    assert m.state_dict() == {}
    m.persistent(True)
    assert m.state_dict()["x"] == 1
    fresh = DummySum()
    fresh.persistent(True)
    fresh.load_state_dict(m.state_dict())
    assert fresh.x == 1


def test_dtype_and_device_moves():
    m = DummySum()
    assert m.x.dtype == torch.float32
    m.double()
    assert m.x.dtype == torch.float32  # casts only through set_dtype
    m.set_dtype(torch.float64)
    assert m.x.dtype == torch.float64 and m._defaults["x"].dtype == torch.float64 and m.dtype == torch.float64
    m.reset()
    assert m.x.dtype == torch.float64


def test_sync_state_machine_with_mock_backend():
    sync_fn = Mock(side_effect=lambda t, group=None: [t, t])
    m = DummyCounts(2, dist_sync_fn=sync_fn, distributed_available_fn=lambda: True)
    m.update(torch.tensor([1, 2]))
    assert m.compute() == 2 * (3 + 6)
    assert sync_fn.call_count == 5  # once per state tensor
    assert m.tp.tolist() == [1, 2] and not m._is_synced  # unsynced after compute
    m.sync(dist_sync_fn=sync_fn)
    assert m._is_synced and m.tp.tolist() == [2, 4]
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been synced."):
        m.sync(dist_sync_fn=sync_fn)
    with pytest.raises(TorchMetricsUserError, match="The Metric shouldn't be synced when performing ``forward``"):
        m(torch.tensor([1, 1]))
    m.unsync()
    assert m.tp.tolist() == [1, 2]
    with pytest.raises(TorchMetricsUserError, match="The Metric has already been un-synced."):
        m.unsync()
    with m.sync_context(dist_sync_fn=sync_fn):
        assert m.tp.tolist() == [2, 4]
    assert m.tp.tolist() == [1, 2]
    off = DummyCounts(2, dist_sync_fn=sync_fn, distributed_available_fn=lambda: True, sync_on_compute=False)
    off.update(torch.tensor([1, 2]))
    n = sync_fn.call_count
    off.compute()
    assert sync_fn.call_count == n


def test_filter_kwargs_and_composition():
    a, b = DummySum(), DummyDiff()
    assert a._filter_kwargs(x=1, y=2) == {"x": 1}
    comp = a + b
    assert isinstance(comp, CompositionalMetric)
    comp.update(x=3, y=1)
    assert comp.compute() == 2
    assert (a * 2).compute() == 6
    assert abs(b).compute() == 1
    comp.reset()
    assert a.x == 0 and b.x == 0


def test_metric_is_abstract():
    with pytest.raises(TypeError):
        Metric()  # type: ignore[abstract]


# ---------------------------------------------------------------- collections
class _A(DummyCounts):
    def compute(self):
        return self.tp.sum()


class _B(DummyCounts):
    def compute(self):
        return self.fp.sum()


class _C(DummySum):
    def compute(self):
        return self.x.sum()


def test_collection_compute_groups_share_state():
    mc = MetricCollection([_A(2), _B(2), _C()])
    assert len(mc.compute_groups) == 3
    mc.update(torch.tensor([1, 2]))
    assert mc.compute_groups == {0: ["_A", "_B"], 1: ["_C"]}
    a = mc.__getitem__("_A", copy_state=False)
    b = mc.__getitem__("_B", copy_state=False)
    assert a.tp is b.tp  # shared by reference
    mc.update(torch.tensor([1, 1]))
    assert b.tp.tolist() == [2, 3]
    assert mc.__getitem__("_B", copy_state=False).update_count == 2  # refreshed on access, as in the reference
    res = mc.compute()
    assert res["_A"] == 5 and res["_B"] == 10 and res["_C"] == 5
    # access hands out copies, and the next update re-links
    assert mc["_B"].tp is not mc["_A"].tp
    mc.update(torch.tensor([0, 1]))
    assert mc.__getitem__("_B", copy_state=False).tp.tolist() == [2, 4]
    mc.reset()
    assert mc.compute_groups == {0: ["_A", "_B"], 1: ["_C"]}
    ref = MetricCollection([_A(2), _B(2), _C()], compute_groups=False)
    for x in ([1, 2], [1, 1], [0, 1]):
        ref.update(torch.tensor(x))
        mc.update(torch.tensor(x))
    assert {k: int(v) for k, v in ref.compute().items()} == {k: int(v) for k, v in mc.compute().items()}


def test_collection_forward_prefix_postfix_and_errors():
    mc = MetricCollection({"a": _A(2), "c": _C()}, prefix="p_", postfix="_s")
    out = mc(torch.tensor([1, 2]))
    assert set(out) == {"p_a_s", "p_c_s"} and out["p_a_s"] == 3
    assert list(mc.keys()) == ["p_a_s", "p_c_s"] and list(mc.keys(keep_base=True)) == ["a", "c"]
    with pytest.raises(ValueError, match="Encountered two metrics both named _A"):
        MetricCollection([_A(), _A()])
    with pytest.raises(ValueError, match="Unknown input to MetricCollection."):
        MetricCollection(5)  # type: ignore[arg-type]
    with pytest.raises(ValueError, match="Expected input `prefix` to be a string"):
        MetricCollection([_A()], prefix=3)  # type: ignore[arg-type]
    with pytest.raises(ValueError, match="does not match a metric in the collection"):
        MetricCollection([_A(), _B()], compute_groups=[["_A", "nope"]])
    user = MetricCollection([_A(2), _B(2)], compute_groups=[["_A", "_B"]])
    user.update(torch.tensor([1, 2]))
    assert user.compute()["_B"] == 6
    nested = MetricCollection([MetricCollection([_A(2)], prefix="x_"), MetricCollection([_B(2)], postfix="_y")])
    nested.update(torch.tensor([1, 2]))
    assert set(nested.compute()) == {"x__A", "_B_y"}
    cl = mc.clone(prefix="q_")
    assert list(cl.keys()) == ["q_a_s", "q_c_s"]
