"""Exact (sort-based) binary curve: descending sort of the scores, cumulative positives at the last index of every
run of equal scores (reference: functional/classification/precision_recall_curve.py:28-80)."""
from typing import Optional, Tuple

import torch
from torch import Tensor

from torchmetrics_b200 import _native


def exact_clf_curve(preds: Tensor, target: Tensor, pos_label: int = 1) -> Tuple[Tensor, Tensor, Tensor]:
    """(fps, tps, thresholds) with exact int64 counts; thresholds are the distinct scores, descending."""
    _native.require_cuda(preds, target)
    return _native.exact_curve(preds, target, pos_label)


def _fusable(state) -> bool:
    return (isinstance(state, (tuple, list)) and len(state) == 2 and isinstance(state[0], Tensor)
            and state[0].is_cuda and state[1].is_cuda and state[0].is_floating_point() and state[0].numel() > 0)


def fused_binary_auroc(state, pos_label: int = 1) -> Optional[Tensor]:
    """Exact AUROC of (preds, target) reduced on the device, or None when the caller should take the curve path
    (degenerate targets, where the reference warns and returns zeros)."""
    if not _fusable(state):
        return None
    preds, target = state
    area2, _, npos, _ = _native.exact_auroc_ap(preds, target, pos_label)
    nneg = preds.numel() - npos
    if npos <= 0 or nneg <= 0:
        return None
    return torch.tensor(area2 / (2.0 * npos * nneg), dtype=torch.float32, device=preds.device)


def fused_binary_average_precision(state, pos_label: int = 1) -> Optional[Tensor]:
    """Exact average precision reduced on the device, or None when there is no positive (reference: nan path)."""
    if not _fusable(state):
        return None
    preds, target = state
    _, apsum, npos, _ = _native.exact_auroc_ap(preds, target, pos_label)
    if npos <= 0:
        return None
    return torch.tensor(apsum / npos, dtype=torch.float32, device=preds.device)


# ---------------------------------------------------------------------------------------------------------------
# Distributed exact AUROC / average precision WITHOUT replication.
#
# The reference (and the default mode here) all_gathers every rank's scores and targets and then sorts the whole set
# on every rank (metric.py:427-457, then precision_recall_curve.py:28-80): W times the sort work and (W-1)/W of all
# samples received per rank.  Here the samples are range-partitioned by score instead:
#   1. every rank contributes a sample of its scores; the weighted quantiles of the gathered sample are W-1 splitters;
#   2. an all-to-all moves each (score, positive?) pair - 5 bytes - to the rank owning its score range, equal scores
#      always to the same rank, rank 0 taking the highest range;
#   3. each rank sorts only its range with K4b, told how many positives / negatives rank above it (one tiny
#      all_gather of per-range counts), so its two partial sums are its terms of the GLOBAL sums;
#   4. one all_reduce of four numbers finishes AUROC and AP.
# Every rank sends and receives about n/W samples once and sorts n/W samples.
# ---------------------------------------------------------------------------------------------------------------
_SPLITTER_SAMPLES = 4096


def partitioned_exact_sums(preds: Tensor, positive: Tensor, group=None) -> Tuple[float, float, float, float]:
    """(area2, ap_sum, num_positive, num_samples) of the union of all ranks' (preds, positive) - the quantities
    `_native.exact_auroc_ap` returns for the concatenated input - computed collectively.  `preds` float32 (n_local,)
    CUDA, `positive` (n_local,) uint8/bool; ranks may hold different (also zero) numbers of samples."""
    import torch.distributed as dist

    group = group if group is not None else dist.group.WORLD
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = preds.device
    preds = preds.detach().reshape(-1).to(torch.float32)
    positive = positive.reshape(-1).to(torch.uint8)
    n_local = preds.numel()

    # 1. splitters from a weighted sample (a rank's sample stands for n_local / S of its scores)
    S = _SPLITTER_SAMPLES
    mine = torch.full((S + 1,), float("nan"), device=dev)
    if n_local:
        idx = torch.randint(0, n_local, (S,), device=dev)
        mine[:S] = preds[idx]
    mine[S] = float(n_local)
    everyone = torch.empty(world * (S + 1), device=dev)
    dist.all_gather_into_tensor(everyone, mine, group=group)
    everyone = everyone.view(world, S + 1)
    counts = everyone[:, S]
    weights = (counts / S).unsqueeze(1).expand(world, S).reshape(-1)
    values = everyone[:, :S].reshape(-1)
    keep = ~values.isnan()
    values, weights = values[keep], weights[keep]
    if values.numel():
        order = torch.argsort(values)
        values, cum = values[order], torch.cumsum(weights[order], 0)
        marks = cum[-1] * torch.arange(1, world, device=dev) / world
        splitters = values[torch.searchsorted(cum, marks).clamp(max=values.numel() - 1)]
    else:
        splitters = torch.zeros(world - 1, device=dev)

    # 2. destination = score range; range 0 (rank 0) holds the highest scores, NaN ranks above everything.
    # Samples are regrouped by destination in two coalesced passes (tmb_range_partition); per destination every
    # rank publishes how many samples and how many positives it sends, so ONE all_gather tells every rank its
    # receive sizes and the counts ranked above its range.
    send_p, send_t, totals = _native.range_partition(preds, positive, splitters, 1)
    table = torch.empty(world, 2, world, dtype=torch.int64, device=dev)
    dist.all_gather_into_tensor(table.view(-1), totals.view(-1), group=group)
    table = table.tolist()  # [sender][0 = samples, 1 = positives][destination]; the split sizes must be host ints
    send_sizes = [int(v) for v in table[rank][0]]
    recv_sizes = [int(table[r][0][rank]) for r in range(world)]
    range_n = [sum(int(table[r][0][d]) for r in range(world)) for d in range(world)]
    range_p = [sum(int(table[r][1][d]) for r in range(world)) for d in range(world)]
    n_mine = sum(recv_sizes)
    recv_p = torch.empty(n_mine, dtype=torch.float32, device=dev)
    recv_t = torch.empty(n_mine, dtype=torch.uint8, device=dev)
    dist.all_to_all_single(recv_p, send_p, recv_sizes, send_sizes, group=group)
    dist.all_to_all_single(recv_t, send_t, recv_sizes, send_sizes, group=group)

    # 3. this range's terms of the global sums: K4b told how many positives / negatives rank above the range
    tp_off = sum(range_p[:rank])
    fp_off = sum(range_n[:rank]) - tp_off
    if n_mine:
        sums = _native.exact_auroc_ap(recv_p, recv_t, 1, tp_off, fp_off, as_tensor=True)[:2].clone()
    else:
        sums = torch.zeros(2, dtype=torch.float64, device=dev)

    # 4. global sums
    dist.all_reduce(sums, group=group)
    area2, apsum = sums.tolist()
    return area2, apsum, float(sum(range_p)), float(sum(range_n))
