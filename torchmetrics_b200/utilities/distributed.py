"""Cross-rank state exchange (reference: utilities/distributed.py:91-147, metric.py:427-457).

Two paths:
  * `gather_all_tensors` keeps the reference's per-tensor protocol (list of world_size tensors, uneven shapes
    padded and trimmed) for `dist_sync_fn` users and for states with custom / `None` reductions; it drops the
    reference's barrier.
  * `fused_state_sync` is the default path of `Metric.sync`: every sum / max / min / mean tensor state of the
    metrics handed in is packed into ONE flat buffer per (op, dtype) and reduced with ONE all_reduce (NCCL over
    NVLink on the GPU box, gloo in CPU tests); every cat state takes one shared size exchange plus one padded
    all_gather. Integer addition is exact and order-free, so the sum states are bit-identical to the reference's
    gather + stack + sum.
"""
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
import torch.nn.functional as F
from torch import Tensor


def distributed_available() -> bool:
    return dist.is_available() and dist.is_initialized()


def _simple_gather_all_tensors(result: Tensor, group: Any, world_size: int) -> List[Tensor]:
    gathered = [torch.zeros_like(result) for _ in range(world_size)]
    dist.all_gather(gathered, result, group)
    return gathered


def gather_all_tensors(result: Tensor, group: Optional[Any] = None) -> List[Tensor]:
    """Every rank receives the list of all ranks' tensors; shapes may differ per rank in any dimension."""
    if group is None:
        group = dist.group.WORLD
    result = result.contiguous()
    world_size = dist.get_world_size(group)
    if result.ndim == 0:
        return _simple_gather_all_tensors(result, group, world_size)
    local_size = torch.tensor(result.shape, device=result.device)
    sizes = [torch.zeros_like(local_size) for _ in range(world_size)]
    dist.all_gather(sizes, local_size, group=group)
    sizes_cpu = torch.stack(sizes).cpu()
    max_size = sizes_cpu.max(dim=0).values
    if bool((sizes_cpu == max_size).all()):
        return _simple_gather_all_tensors(result, group, world_size)
    pad = []
    for need in reversed((max_size - sizes_cpu.new_tensor(result.shape)).tolist()):
        pad += [0, int(need)]
    padded = F.pad(result, pad)
    gathered = [torch.zeros_like(padded) for _ in range(world_size)]
    dist.all_gather(gathered, padded, group)
    out = []
    for r, t in enumerate(gathered):
        out.append(t[tuple(slice(int(d)) for d in sizes_cpu[r].tolist())])
    return out


# -------------------------------------------------------------------------------------------------
# fused default sync
# -------------------------------------------------------------------------------------------------
_REDUCE_OPS = {"sum": dist.ReduceOp.SUM, "mean": dist.ReduceOp.SUM, "max": dist.ReduceOp.MAX, "min": dist.ReduceOp.MIN}

_DTYPE_CODES = [torch.float32, torch.float64, torch.float16, torch.bfloat16, torch.int64, torch.int32, torch.int16,
                torch.int8, torch.uint8, torch.bool]


_NARROW_INTS = (torch.bool, torch.uint8, torch.int8, torch.int16, torch.int32)
_HALF_FLOATS = (torch.float16, torch.bfloat16)


def _wire_dtype(op: str, dtype: torch.dtype) -> torch.dtype:
    """dtype a state travels and is reduced in.  The reference stacks the gathered copies and calls torch.sum /
    torch.mean on them (metric.py:449-456): sums of bool and narrow integer states come back as int64, and
    half-precision means are accumulated in float32 before they are rounded back."""
    if op == "sum" and dtype in _NARROW_INTS:
        return torch.int64
    if op == "mean" and dtype in _HALF_FLOATS:
        return torch.float32
    if dtype == torch.bool:
        return torch.uint8  # max / min of flags
    return dtype


# -------------------------------------------------------------------------------------------------
# K5: one-shot all-reduce of KB-sized arenas over NVLink peer memory (tmb_oneshot_allreduce)
# -------------------------------------------------------------------------------------------------
import os as _os

_ONESHOT: Dict[Any, Any] = {}
_ONESHOT_OFF = _os.environ.get("TM_B200_ONESHOT", "1") == "0"
_ONESHOT_OPS = {dist.ReduceOp.SUM: 0, dist.ReduceOp.MAX: 1, dist.ReduceOp.MIN: 2}
_ONESHOT_DTYPES = (torch.int64, torch.int32, torch.float32, torch.float64)


class _OneShot:
    """Per process group: this rank's symmetric buffer, the device table of every rank's mapping of it, the epoch."""

    def __init__(self, group: Any, device: torch.device) -> None:
        import torch.distributed._symmetric_memory as symm_mem

        from torchmetrics_b200 import _native

        self.lib = _native.load()
        nbytes = self.lib.tmb_oneshot_buffer_bytes()
        self.buf = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        hdl = symm_mem.rendezvous(self.buf, group.group_name)
        self.rank, self.world = hdl.rank, hdl.world_size
        self.peers = torch.tensor(list(hdl.buffer_ptrs), dtype=torch.int64, device=device)
        torch.cuda.synchronize(device)
        hdl.barrier()  # every rank's flags are zero before anybody's first call
        torch.cuda.synchronize(device)
        self.hdl = hdl
        self.epoch = 0
        self._status_ptr, self._dev = None, None

    def all_reduce(self, src: Tensor, dst: Tensor, rop: Any) -> None:
        """dst = reduce over ranks of src (same shape / dtype, contiguous; may be the same tensor)."""
        from torchmetrics_b200 import _native

        self.epoch += 1
        dev = src.device
        if self._status_ptr is None or self._dev != dev:
            self._dev, self._status_ptr = dev, _native.scratch(dev).status.data_ptr()
            self._peers_ptr, self._fn = self.peers.data_ptr(), self.lib.tmb_oneshot_allreduce
        stream = torch.cuda.current_stream(dev).cuda_stream
        if torch.cuda.current_device() == dev.index:
            rc = self._fn(self._peers_ptr, self.world, self.rank, src.data_ptr(), dst.data_ptr(), src.numel(),
                          _native._DTYPES[src.dtype], _ONESHOT_OPS[rop], self.epoch, self._status_ptr, stream)
        else:
            with torch.cuda.device(dev):
                rc = self._fn(self._peers_ptr, self.world, self.rank, src.data_ptr(), dst.data_ptr(), src.numel(),
                              _native._DTYPES[src.dtype], _ONESHOT_OPS[rop], self.epoch, self._status_ptr, stream)
        if rc:
            _native._check(rc, "tmb_oneshot_allreduce")

    def all_reduce_packed(self, parts: Sequence[Tensor], dst: Tensor, rop: Any) -> None:
        """dst (flat) = reduce over ranks of the concatenation of `parts` (device tensors of dst's dtype)."""
        from torchmetrics_b200 import _native

        if len(parts) == 1:
            return self.all_reduce(parts[0], dst, rop)
        self.epoch += 1
        dev = dst.device
        if self._status_ptr is None or self._dev != dev:
            self._dev, self._status_ptr = dev, _native.scratch(dev).status.data_ptr()
            self._peers_ptr, self._fn = self.peers.data_ptr(), self.lib.tmb_oneshot_allreduce
        srcs, counts, k = _packed_args(parts)
        with torch.cuda.device(dev):
            rc = self.lib.tmb_oneshot_allreduce_packed(self._peers_ptr, self.world, self.rank, srcs, counts, k,
                                                       dst.data_ptr(), _native._DTYPES[dst.dtype], _ONESHOT_OPS[rop],
                                                       self.epoch, self._status_ptr,
                                                       torch.cuda.current_stream(dev).cuda_stream)
        if rc:
            _native._check(rc, "tmb_oneshot_allreduce_packed")


def _packed_args(parts: Sequence[Tensor]):
    import ctypes

    k = len(parts)
    return ((ctypes.c_void_p * k)(*[p.data_ptr() for p in parts]), (ctypes.c_int64 * k)(*[p.numel() for p in parts]), k)


def _oneshot_for(device: torch.device, dtype: torch.dtype, nbytes: int, rop: Any, group: Any,
                 world: int = 1) -> Optional[_OneShot]:
    """The one-shot context when this reduction qualifies, else None: CUDA arena of at most 64 KB and at most 128 KB
    read per rank over all peers (one CTA does the reads; beyond that NCCL is faster: 64 KB x 8 ranks 56 vs 38 us),
    NCCL group, supported dtype and op; TM_B200_ONESHOT=0 switches it off.  The decision depends only on things every
    rank agrees on."""
    if (device.type != "cuda" or dtype not in _ONESHOT_DTYPES or rop not in _ONESHOT_OPS or nbytes > 65536
            or nbytes * world > 131072 or nbytes == 0 or _ONESHOT_OFF):
        return None
    key = id(group)
    ctx = _ONESHOT.get(key)
    if ctx is None:
        try:
            if "nccl" not in str(dist.get_backend(group)) or dist.get_world_size(group) > 64:
                raise RuntimeError("not an NCCL group of at most 64 ranks")
            ctx = _OneShot(group, device)
        except Exception:  # no symmetric memory on this system / group: NCCL it is, on every rank alike
            ctx = False
        _ONESHOT[key] = ctx
    return ctx or None


def fused_reduce(tensors: Sequence[Tuple[str, Tensor]], group: Optional[Any] = None) -> List[Tensor]:
    """All-reduce several tensors with as few collectives as possible.

    `tensors` is a list of (op, tensor) with op in sum/mean/max/min. Tensors sharing (op class, wire dtype, device)
    are packed into one flat arena, reduced in one all_reduce and returned as views of that arena (fresh memory: the
    inputs are left untouched so `unsync` can restore them).  Result dtypes are the reference's: int64 for sums of
    bool / narrow integer states, the state's own dtype otherwise.
    """
    if group is None:
        group = dist.group.WORLD
    world = dist.get_world_size(group)
    out: List[Optional[Tensor]] = [None] * len(tensors)
    buckets: Dict[Tuple[Any, torch.dtype, torch.device], List[int]] = {}
    for i, (op, t) in enumerate(tensors):
        buckets.setdefault((_REDUCE_OPS[op], _wire_dtype(op, t.dtype), t.device), []).append(i)
    for (rop, wire, device), idxs in buckets.items():
        single = tensors[idxs[0]][1].detach() if len(idxs) == 1 else None
        nbytes = sum(tensors[i][1].numel() for i in idxs) * wire.itemsize
        oneshot = _oneshot_for(device, wire, nbytes, rop, group, world)
        parts = [tensors[i][1].detach() for i in idxs]
        if (oneshot is not None and len(parts) <= 16
                and all(p.dtype == wire and p.is_contiguous() for p in parts)):
            # KB-sized states: the kernel reads the state tensors themselves, packs them on the way into this rank's
            # slot and writes the merged arena (no torch.cat / clone pass, no NCCL call)
            arena = torch.empty(nbytes // wire.itemsize, dtype=wire, device=device)
            oneshot.all_reduce_packed(parts, arena, rop)
        else:
            if single is not None:
                arena = single.reshape(-1).to(wire, copy=True)
            else:
                arena = torch.cat([tensors[i][1].detach().reshape(-1).to(wire) for i in idxs])
            if oneshot is not None:
                oneshot.all_reduce(arena, arena, rop)  # one kernel over NVLink peer memory instead of NCCL
            else:
                dist.all_reduce(arena, op=rop, group=group)
        off = 0
        for i in idxs:
            op, t = tensors[i]
            piece = arena[off:off + t.numel()].view(t.shape)
            off += t.numel()
            if op == "mean":
                piece = (piece / world).to(t.dtype)
            elif op != "sum" and piece.dtype != t.dtype:
                piece = piece.to(t.dtype)
            out[i] = piece
    return out  # type: ignore[return-value]


def fused_cat_gather(tensors: Sequence[Optional[Tensor]], group: Optional[Any] = None) -> List[Optional[Tensor]]:
    """Concatenate every rank's dim-0 slices of several cat states.

    One all_gather of a small int64 descriptor covers ALL states (length along dim 0, trailing shape, dtype), so a
    rank whose list state is still empty (`None` here) can take part with a matching empty tensor; then one padded
    `all_gather_into_tensor` per state, trimmed to the true lengths. Returns None for a state empty on every rank.
    """
    if group is None:
        group = dist.group.WORLD
    world = dist.get_world_size(group)
    rank_device = None
    for t in tensors:
        if t is not None:
            rank_device = t.device
            break
    if rank_device is None:
        rank_device = _group_device(group)
    MAXD = 6
    desc = torch.zeros(len(tensors), MAXD + 3, dtype=torch.int64)
    for i, t in enumerate(tensors):
        if t is None:
            continue
        if t.ndim - 1 > MAXD:
            raise ValueError("cat states with more than 7 dimensions are not supported by the fused gather")
        desc[i, 0] = 1
        desc[i, 1] = t.shape[0]
        desc[i, 2] = _DTYPE_CODES.index(t.dtype)
        desc[i, 3] = t.ndim - 1
        for d, s in enumerate(t.shape[1:]):
            desc[i, 4 + d] = s
    desc = desc.to(rank_device)
    # output is the dim-0 concatenation of every rank's block (the layout gloo and NCCL both accept)
    all_desc = torch.empty(world * desc.shape[0], desc.shape[1], dtype=torch.int64, device=rank_device)
    dist.all_gather_into_tensor(all_desc, desc, group=group)
    all_desc = all_desc.cpu().view(world, desc.shape[0], desc.shape[1])
    out: List[Optional[Tensor]] = []
    for i, t in enumerate(tensors):
        have = all_desc[:, i, 0].bool()
        if not bool(have.any()):
            out.append(None)
            continue
        src = int(torch.nonzero(have)[0])
        dtype = _DTYPE_CODES[int(all_desc[src, i, 2])]
        trailing = tuple(int(v) for v in all_desc[src, i, 4:4 + int(all_desc[src, i, 3])])
        lens = [int(all_desc[r, i, 1]) if bool(have[r]) else 0 for r in range(world)]
        if t is None:
            t = torch.empty((0, *trailing), dtype=dtype, device=rank_device)
        t = t.contiguous()
        max_len = max(lens)
        if t.shape[0] == max_len:
            send = t
        else:
            send = torch.zeros((max_len, *trailing), dtype=dtype, device=t.device)
            send[: t.shape[0]] = t
        recv = torch.empty((world * max_len, *trailing), dtype=dtype, device=t.device)
        if max_len > 0:
            dist.all_gather_into_tensor(recv, send, group=group)
        if all(n == max_len for n in lens):
            out.append(recv)
        else:
            out.append(torch.cat([recv[r * max_len: r * max_len + lens[r]] for r in range(world)], dim=0))
    return out


def _group_device(group: Any) -> torch.device:
    try:
        backend = dist.get_backend(group)
    except Exception:
        backend = "gloo"
    if "nccl" in str(backend):
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")
