"""Dummy metrics with pure-torch updates, for testing the host-side state machine on CPU."""
import torch

from torchmetrics_b200.metric import Metric


class DummySum(Metric):
    full_state_update = False

    def __init__(self, **kw):
        super().__init__(**kw)
        self.add_state("x", torch.tensor(0.0), dist_reduce_fx="sum")

    def update(self, x):
        self.x = self.x + torch.as_tensor(x, dtype=torch.float32)

    def compute(self):
        return self.x


class DummyDiff(DummySum):
    def update(self, y):
        self.x = self.x - torch.as_tensor(y, dtype=torch.float32)


class DummyList(Metric):
    full_state_update = True

    def __init__(self, **kw):
        super().__init__(**kw)
        self.add_state("x", [], dist_reduce_fx="cat")

    def update(self, x=None):
        if x is not None:
            self.x.append(torch.as_tensor(x))

    def compute(self):
        return self.x


class DummyCounts(Metric):
    """int64 count states with every built-in reduction, as the classification metrics use them."""

    full_state_update = False

    def __init__(self, n=4, **kw):
        super().__init__(**kw)
        self.add_state("tp", torch.zeros(n, dtype=torch.long), dist_reduce_fx="sum")
        self.add_state("fp", torch.zeros(n, dtype=torch.long), dist_reduce_fx="sum")
        self.add_state("hi", torch.full((n,), -1.0), dist_reduce_fx="max")
        self.add_state("lo", torch.full((n,), 1e9), dist_reduce_fx="min")
        self.add_state("avg", torch.zeros(n), dist_reduce_fx="mean")

    def update(self, x):
        x = torch.as_tensor(x)
        self.tp += x.long()
        self.fp += 2 * x.long()
        self.hi = torch.maximum(self.hi, x.float())
        self.lo = torch.minimum(self.lo, x.float())
        self.avg = self.avg + x.float()

    def compute(self):
        return self.tp.sum() + self.fp.sum()
