"""Benchmark of the metric-state hot path (contract: see the task statement; one JSON line on stdout from rank 0).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1], the configuration the headline metric is quoted on): `MulticlassConfusionMatrix(
num_classes=1000).update` on 2**24-sample int64 batches, one batch per step per GPU (weak scaling: every rank has its
own batches).  `value` = samples consumed by update() per second over all ranks with the inputs already in HBM;
`e2e` = the same through the public Metric API with inputs in pinned HOST memory, the host->device copies and a
device->host read of the state inside the timed region.  `extra` carries the second half of the headline metric
(binned AUROC+AP, T=200) and, for N>1, the sync latencies.

`--impl reference` times the CPU restatement of the reference's algorithm (oracle/oracle.c, OpenMP, all host
threads) on the same workload.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NUM_CLASSES = 1000
BATCH = 1 << 24
BYTES_PER_SAMPLE = 16  # int64 pred + int64 target
METRIC = "update samples/s (MulticlassConfusionMatrix C=1000, 2^24-sample int64 batches)"


def _measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi sampled every 200 ms during the timed region (profiling recipe's clocks line)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                 "-i", str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 9:
                self.rows.append(parts)

    def stop(self):
        if self.proc is not None:
            time.sleep(0.25)
            self.proc.terminate()
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx = max(mx, float(r[2]))
            except ValueError:
                continue
            for name, v in zip(names, r[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


CONFIG = {"workload": "configs[1]: MulticlassConfusionMatrix num_classes=1000, 2^24-sample int64 batches",
          "samples_per_step_per_gpu": BATCH, "validate_args": False}


def _import_reference():
    """The unmodified reference (pure Python): TM_REFERENCE_SRC, /root/reference/src, or the copy pip-installed into
    baseline/_ref (which travels to the GPU box), with the 4-symbol lightning_utilities stand-in the golden-vector
    generator uses.  Returns the module or None."""
    shim = os.path.join(ROOT, "tests", "golden", "_refshim")
    for cand in (os.environ.get("TM_REFERENCE_SRC"), "/root/reference/src", os.path.join(ROOT, "baseline", "_ref")):
        if cand and os.path.isdir(os.path.join(cand, "torchmetrics")):
            sys.path[:0] = [cand, shim]
            try:
                import torchmetrics

                return torchmetrics, cand
            except Exception:
                del sys.path[:2]
    return None, None


def _reference_steps(args):
    """(value, seconds per step, description) of the reference's own CPU implementation, or None."""
    ref, where = _import_reference()
    if ref is None:
        return None
    import torch

    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.Generator().manual_seed(1234)
    batches = [(torch.randint(0, NUM_CLASSES, (BATCH,), generator=g), torch.randint(0, NUM_CLASSES, (BATCH,), generator=g))
               for _ in range(2)]
    metric = ref.classification.MulticlassConfusionMatrix(num_classes=NUM_CLASSES, validate_args=False)
    for i in range(args.warmup):
        metric.update(*batches[i % 2])
    t0 = time.perf_counter()
    for i in range(args.steps):
        metric.update(*batches[i % 2])
    dt = time.perf_counter() - t0
    assert int(metric.confmat.sum()) == (args.warmup + args.steps) * BATCH
    return args.steps * BATCH / dt, dt / args.steps, torch.get_num_threads(), (
        f"{args.steps} full 2^24-sample batches through torchmetrics {ref.__version__} "
        f"MulticlassConfusionMatrix.update on CPU tensors ({where})")


def run_reference(args, rank, world):
    """CPU arm: the reference's own code on the box's host cores (kind "reference"); if it cannot be imported, the
    C + OpenMP restatement oracle/oracle.c (kind "port")."""
    if rank != 0:
        return
    res = _reference_steps(args)
    if res is not None:
        value, step_s, cores, sample = res
        kind = "reference"
    else:
        import numpy as np

        from oracle import oracle_c as OC

        # torchrun exports OMP_NUM_THREADS=1 to its children; the CPU arm is meant to use every host core
        OC.set_num_threads(os.cpu_count() or 1)
        rng = np.random.default_rng(1234)
        batches = [(rng.integers(0, NUM_CLASSES, BATCH), rng.integers(0, NUM_CLASSES, BATCH)) for _ in range(2)]
        state = np.zeros((NUM_CLASSES, NUM_CLASSES), dtype=np.int64)
        for i in range(args.warmup):
            OC.confmat_update(*batches[i % 2], NUM_CLASSES, out=state)
        t0 = time.perf_counter()
        for i in range(args.steps):
            OC.confmat_update(*batches[i % 2], NUM_CLASSES, out=state)
        dt = time.perf_counter() - t0
        assert int(state.sum()) == (args.warmup + args.steps) * BATCH
        value, step_s, cores, kind = args.steps * BATCH / dt, dt / args.steps, OC.num_threads(), "port"
        sample = f"{args.steps} full 2^24-sample batches (oracle/oracle.c, OpenMP)"
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": dict(CONFIG),
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "note": "CPU arm runs on rank 0 only, one batch per step",
    }
    print(json.dumps(line), flush=True)


def reference_on_gpu(dev):
    """BASELINE.md section 5's second comparator: the unmodified reference with its tensors on the B200 (ATen bincount,
    compare-and-sum loops, CUB sort) - the reference ships no kernels of its own.  Device-timed, inputs resident."""
    ref, where = _import_reference()
    if ref is None:
        return {"unavailable": "reference not importable"}
    import torch

    R = ref.classification

    def timed(fn, reps):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out = {"source": where, "note": "torchmetrics 1.4.0dev classes, validate_args=False, CUDA tensors, ms per update"}
    g = torch.Generator(device=dev).manual_seed(99)
    n = 1 << 24
    p = torch.randint(0, NUM_CLASSES, (n,), device=dev, generator=g)
    t = torch.randint(0, NUM_CLASSES, (n,), device=dev, generator=g)
    m = R.MulticlassConfusionMatrix(num_classes=NUM_CLASSES, validate_args=False).to(dev)
    out["confmat_C1000_2p24_ms"] = timed(lambda: m.update(p, t), 5)
    del m, p, t
    sc = torch.rand(n, device=dev, generator=g)
    lb = (torch.rand(n, device=dev, generator=g) < 0.5).long()
    m = R.BinaryAUROC(thresholds=200, validate_args=False).to(dev)
    out["binned_auroc_T200_2p24_ms"] = timed(lambda: m.update(sc, lb), 2)
    m = R.BinaryAUROC(thresholds=None, validate_args=False).to(dev)
    m.update(sc, lb)

    def exact():
        m._computed = None
        return m.compute()

    out["exact_auroc_2p24_compute_ms"] = timed(exact, 2)
    del m, sc, lb
    mp = torch.rand(1 << 20, 100, device=dev, generator=g)
    mt = (torch.rand(1 << 20, 100, device=dev, generator=g) < 0.5).long()
    m = R.MultilabelStatScores(num_labels=100, validate_args=False).to(dev)
    out["multilabel_stat_scores_L100_2p20rows_ms"] = timed(lambda: m.update(mp, mt), 2)
    return out


def cpu_baseline_sample(seconds=10.0):
    """Bounded CPU sample of the same workload on the box's host cores: the reference's own code when it can be
    imported (kind "reference"), else the C/OpenMP port."""
    ref, where = _import_reference()
    if ref is not None:
        import torch

        torch.set_num_threads(os.cpu_count() or 1)
        g = torch.Generator().manual_seed(1234)
        p = torch.randint(0, NUM_CLASSES, (BATCH,), generator=g)
        t = torch.randint(0, NUM_CLASSES, (BATCH,), generator=g)
        m = ref.classification.MulticlassConfusionMatrix(num_classes=NUM_CLASSES, validate_args=False)
        m.update(p, t)
        t0, reps = time.perf_counter(), 0
        while time.perf_counter() - t0 < seconds and reps < 200:
            m.update(p, t)
            reps += 1
        dt = time.perf_counter() - t0
        return {"value": reps * BATCH / dt, "unit": "samples/s", "cores": torch.get_num_threads(), "kind": "reference",
                "sample": f"{reps} repetitions of one 2^24-sample batch (~{seconds:.0f} s) through torchmetrics "
                          f"{ref.__version__} MulticlassConfusionMatrix.update on CPU tensors ({where})"}
    import numpy as np

    from oracle import oracle_c as OC

    OC.set_num_threads(os.cpu_count() or 1)
    rng = np.random.default_rng(1234)
    p, t = rng.integers(0, NUM_CLASSES, BATCH), rng.integers(0, NUM_CLASSES, BATCH)
    st = np.zeros((NUM_CLASSES, NUM_CLASSES), dtype=np.int64)
    OC.confmat_update(p, t, NUM_CLASSES, out=st)
    t0, reps = time.perf_counter(), 0
    while time.perf_counter() - t0 < seconds and reps < 200:
        OC.confmat_update(p, t, NUM_CLASSES, out=st)
        reps += 1
    dt = time.perf_counter() - t0
    return {"value": reps * BATCH / dt, "unit": "samples/s", "cores": OC.num_threads(), "kind": "port",
            "sample": f"{reps} repetitions of one 2^24-sample batch (~{seconds:.0f} s), oracle/oracle.c with OpenMP"}


def measure_sync(tm, dist, dev, local_rank, world, confmat_metric, binned_metric, max_over_ranks):
    """SURVEY 8(d) "Sync us": for each state size, `Metric.sync()` of this package next to (a) the reference protocol
    on the same NCCL group (one gather per state: `Metric._sync_dist(gather_all_tensors)`, metric.py:427-457) and
    (b) the bare collective on a tensor of the same bytes.  Median of 30, host-timed around a stream sync AND
    device-timed with CUDA events, max over ranks."""
    import torch

    from torchmetrics_b200.utilities.distributed import gather_all_tensors

    def timed(fn, undo=None, reps=30):
        host, devt = [], []
        for _ in range(reps):
            torch.cuda.synchronize()
            dist.barrier(device_ids=[local_rank])
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            host.append((time.perf_counter() - t0) * 1e6)
            devt.append(a.elapsed_time(b) * 1e3)
            if undo:
                undo()
        return {"host_us": max_over_ranks(statistics.median(host)), "device_us": max_over_ranks(statistics.median(devt))}

    def metric_rows(m):
        def legacy():
            m._cache = {name: getattr(m, name) for name in m._defaults}
            m._sync_dist(gather_all_tensors)
            m._is_synced = True

        return {"ours": timed(m.sync, m.unsync), "reference_protocol": timed(legacy, m.unsync)}

    out = {}
    # 8 MB confusion matrix, 6.4 KB binned curve state
    for key, m, nbytes in (("confmat_8MB", confmat_metric, NUM_CLASSES * NUM_CLASSES * 8),
                           ("binned_6.4KB", binned_metric, 200 * 4 * 8)):
        buf = torch.zeros(nbytes // 8, dtype=torch.int64, device=dev)
        out[key] = dict(metric_rows(m), bare_all_reduce=timed(lambda: dist.all_reduce(buf)), bytes=nbytes)
    # 4 x (100,) int64 = 3.2 KB: the multilabel collection's shared state
    L = 100
    ss = tm.MultilabelStatScores(num_labels=L, validate_args=False).to(dev)
    ss.update(torch.rand(64, L, device=dev), torch.randint(0, 2, (64, L), device=dev))
    buf = torch.zeros(4 * L, dtype=torch.int64, device=dev)
    out["stat_scores_3.2KB"] = dict(metric_rows(ss), bare_all_reduce=timed(lambda: dist.all_reduce(buf)), bytes=3200)
    # cat states: 2^23 x (4 + 1) B per rank (scores fp32 + positives uint8) of the exact curve
    n = 1 << 23
    ex = tm.BinaryAUROC(thresholds=None, validate_args=False).to(dev)
    ex.update(torch.rand(n, device=dev), torch.randint(0, 2, (n,), device=dev))
    per_rank = sum(t.numel() * t.element_size() for t in (ex.preds[0], ex.target[0]))
    send = torch.zeros(per_rank, dtype=torch.uint8, device=dev)
    recv = torch.empty(per_rank * world, dtype=torch.uint8, device=dev)
    out["cat_2p23_per_rank"] = dict(metric_rows(ex),
                                    bare_all_gather=timed(lambda: dist.all_gather_into_tensor(recv, send)),
                                    bytes_per_rank=per_rank)
    out["note"] = ("ours = Metric.sync() (packed all_reduce / padded all_gather); reference_protocol = one "
                   "gather_all_tensors + stack + reduce per state on the same group; bare = the collective alone")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--skip-extra", action="store_true", help="only the headline workload")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3)

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))

    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    import torchmetrics_b200 as tm
    from torchmetrics_b200 import _native as N

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    N.load()

    def barrier():
        if world > 1:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def start_clock(ev):
        # Device-timed throughput loops start behind a ~0.3 ms device-side delay: the host enqueues the first steps
        # while the GPU spins, so the interval between the two events holds the K steps back to back and not the
        # host's latency to its first launch after the barrier (50-90 us, i.e. 5 % of a 20-step run of 65 us steps).
        # Host costs are what the `e2e` arm measures.
        torch.cuda._sleep(600_000)
        ev.record()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ------------------------------------------------------------------ headline: confusion matrix updates
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    batches = [(torch.randint(0, NUM_CLASSES, (BATCH,), device=dev, generator=g),
                torch.randint(0, NUM_CLASSES, (BATCH,), device=dev, generator=g)) for _ in range(2)]
    metric = tm.MulticlassConfusionMatrix(NUM_CLASSES, validate_args=False).to(dev)
    for i in range(args.warmup):
        metric.update(*batches[i % 2])
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    start_clock(e0)
    for i in range(args.steps):
        metric.update(*batches[i % 2])
    e1.record()
    barrier()
    step_ms = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    # parity of the timed state itself: the whole synced 1000 x 1000 matrix against torch.bincount(t*C+p) of the
    # batches this rank fed, summed over the ranks (confusion_matrix.py:333-337), bit for bit
    want = torch.zeros(NUM_CLASSES * NUM_CLASSES, dtype=torch.int64, device=dev)
    for j, (bp, bt) in enumerate(batches):
        times = sum(1 for i in range(args.warmup) if i % 2 == j) + sum(1 for i in range(args.steps) if i % 2 == j)
        want += times * torch.bincount(bt * NUM_CLASSES + bp, minlength=NUM_CLASSES * NUM_CLASSES)
    if world > 1:
        dist.all_reduce(want)
    synced = metric.compute()
    assert torch.equal(synced.reshape(-1), want), "synced confusion matrix differs from torch.bincount"
    assert int(synced.sum()) == (args.warmup + args.steps) * BATCH * world
    del want, synced
    value = BATCH * world / (step_ms * 1e-3)

    # dominant kernel alone (roofline.achieved): same launches through the C ABI, CUDA events on the launch stream
    cm = torch.zeros(NUM_CLASSES, NUM_CLASSES, dtype=torch.int64, device=dev)
    for i in range(3):
        N.confmat_update_(cm, *batches[i % 2], NUM_CLASSES)
    torch.cuda.synchronize()
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start_clock(k0)
    for i in range(args.steps):
        N.confmat_update_(cm, *batches[i % 2], NUM_CLASSES)
    k1.record()
    torch.cuda.synchronize()
    kernel_ms = k0.elapsed_time(k1) / args.steps
    peak, peak_src = _measured_peak()
    achieved = BATCH * BYTES_PER_SAMPLE / (kernel_ms * 1e-3) / 1e9
    traffic = None
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
            pk = tj.get("per_kernel", {})
            # DRAM bytes of the two launches of one update, from the round-2 `ncu --set full` capture
            a = pk.get("confmat_partition_tma_kernel@r2_confmat", {}).get("dram_bytes_per_launch")
            b = pk.get("confmat_bucket_hist_kernel@r2_confmat", {}).get("dram_bytes_per_launch")
            traffic = a + b if a and b else tj.get("confmat_partition_plus_hist_bytes_per_update")
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": "confmat_partition_tma_kernel<int64,int64> + confmat_bucket_hist_kernel (2 launches per update)", "achieved": achieved, "peak": peak,
                "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": BATCH * BYTES_PER_SAMPLE, "kernel_ms": kernel_ms,
                "frac_of_nominal_8TBps": achieved / 8000.0}

    # ------------------------------------------------------------------ e2e: host buffers through the public API
    e2e_steps = max(5, min(args.steps, 20))
    host = [(p.cpu().pin_memory(), t.cpu().pin_memory()) for p, t in batches]
    dbuf = [(torch.empty_like(batches[0][0]), torch.empty_like(batches[0][1])) for _ in range(2)]
    out_host = torch.empty(NUM_CLASSES, NUM_CLASSES, dtype=torch.int64).pin_memory()
    copy_stream = torch.cuda.Stream(device=dev)
    m2 = tm.MulticlassConfusionMatrix(NUM_CLASSES, validate_args=False).to(dev)
    main_stream = torch.cuda.current_stream(dev)

    def e2e_run(nsteps):
        ready = [torch.cuda.Event() for _ in range(2)]
        freed = [torch.cuda.Event() for _ in range(2)]

        def stage(i):
            with torch.cuda.stream(copy_stream):
                if i >= 2:
                    copy_stream.wait_event(freed[i % 2])
                dbuf[i % 2][0].copy_(host[i % 2][0], non_blocking=True)
                dbuf[i % 2][1].copy_(host[i % 2][1], non_blocking=True)
                ready[i % 2].record(copy_stream)

        stage(0)
        for i in range(nsteps):
            if i + 1 < nsteps:
                stage(i + 1)  # next step's copy overlaps this step's kernel
            main_stream.wait_event(ready[i % 2])
            m2.update(*dbuf[i % 2])
            freed[i % 2].record(main_stream)
            out_host.copy_(m2.confmat, non_blocking=True)  # the step's result (the state) back on the host
        torch.cuda.synchronize()

    e2e_run(3)
    m2.reset()
    barrier()
    t0 = time.perf_counter()
    e2e_run(e2e_steps)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    assert int(out_host.sum()) == e2e_steps * BATCH
    e2e = {"value": e2e_steps * BATCH * world / e2e_s, "unit": "samples/s",
           "h2d_bytes_per_step": BATCH * BYTES_PER_SAMPLE, "d2h_bytes_per_step": NUM_CLASSES * NUM_CLASSES * 8,
           "steps": e2e_steps, "note": "pinned host tensors -> .copy_ on a side stream (double-buffered) -> "
                                       "Metric.update -> state copied back; PCIe-bound"}
    del host, dbuf, m2
    # the sampler covers all three timed regions above (device-resident steps, kernel-only loop, e2e)
    clocks = sampler.stop() if rank == 0 else None
    gpu_launches = 2 * args.steps  # partition + bucket-histogram kernel per update in the timed region

    # ------------------------------------------------------------------ extra: binned AUROC+AP (config 3) and sync
    extra = {}
    if not args.skip_extra:
        # the default user path (validate_args=True) as a second line.  The value checks ride on the
        # update kernels; what is left is one 8-byte status read (a host sync) per update.
        mv = tm.MulticlassConfusionMatrix(NUM_CLASSES).to(dev)
        for i in range(3):
            mv.update(*batches[i % len(batches)])
        barrier()
        v0, v1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start_clock(v0)
        for i in range(20):
            mv.update(*batches[i % len(batches)])
        v1.record()
        barrier()
        vms = max_over_ranks(v0.elapsed_time(v1)) / 20
        extra["confmat_validate_args_true"] = {
            "workload": "configs[1] with validate_args=True (the default): label validation fused into the update "
                        "kernels, one status read per update",
            "value": BATCH * world / (vms * 1e-3), "unit": "samples/s", "ms_per_update": vms,
        }
        del mv
        md = tm.MulticlassConfusionMatrix(NUM_CLASSES, validate_args="deferred").to(dev)
        for i in range(3):
            md.update(*batches[i % len(batches)])
        barrier()
        v0, v1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start_clock(v0)
        for i in range(20):
            md.update(*batches[i % len(batches)])
        v1.record()
        barrier()
        dms = max_over_ranks(v0.elapsed_time(v1)) / 20
        md.compute()
        extra["confmat_validate_args_deferred"] = {
            "workload": "configs[1] with validate_args='deferred': the same fused validation, the status words copied "
                        "to pinned host memory behind each kernel and looked at one call later (no per-update sync)",
            "value": BATCH * world / (dms * 1e-3), "unit": "samples/s", "ms_per_update": dms,
        }
        del md
        del batches, cm
        torch.cuda.empty_cache()
        n_total = 1 << 28
        n_local = n_total // world
        g2 = torch.Generator(device=dev).manual_seed(99 + rank)
        preds = torch.rand(n_local, device=dev, generator=g2)
        target = (torch.rand(n_local, device=dev, generator=g2) < 0.3 + 0.4 * preds).long()
        coll = tm.MetricCollection([tm.BinaryAUROC(thresholds=200, validate_args=False),
                                    tm.BinaryAveragePrecision(thresholds=200, validate_args=False)]).to(dev)
        for _ in range(3):
            coll.update(preds, target)
        barrier()
        b0, b1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        start_clock(b0)
        for _ in range(reps):
            coll.update(preds, target)
        b1.record()
        barrier()
        bms = max_over_ranks(b0.elapsed_time(b1)) / reps
        out = coll.compute()
        extra["binned_auroc_ap_T200"] = {
            "workload": "configs[2]: BinaryAUROC+BinaryAveragePrecision thresholds=200, 2^28 fp32 scores + int64 "
                        "targets sharded over the ranks (strong scaling), one compute group => one update",
            "value": n_total / (bms * 1e-3), "unit": "samples/s", "ms_per_update": bms,
            "achieved_GBps_per_gpu": n_local * 12 / (bms * 1e-3) / 1e9,
            "frac_of_measured_peak": n_local * 12 / (bms * 1e-3) / 1e9 / peak,
            "auroc": float(out["BinaryAUROC"]), "ap": float(out["BinaryAveragePrecision"]),
        }
        if world > 1:
            extra["sync_us"] = measure_sync(tm, dist, dev, local_rank, world, metric,
                                            coll.__getitem__("BinaryAUROC", copy_state=False), max_over_ranks)
        del preds, target, coll
        torch.cuda.empty_cache()

        # configs[4]: MetricCollection of multilabel StatScores / Precision / Recall / F1, L = 100, 2^23 rows per rank
        L, rows = 100, 1 << 23
        g3 = torch.Generator(device=dev).manual_seed(7 + rank)
        mp = torch.rand(rows, L, device=dev, generator=g3)
        mt = torch.randint(0, 2, (rows, L), device=dev, generator=g3)
        mcoll = tm.MetricCollection([tm.MultilabelStatScores(num_labels=L, validate_args=False),
                                     tm.MultilabelPrecision(num_labels=L, validate_args=False),
                                     tm.MultilabelRecall(num_labels=L, validate_args=False),
                                     tm.MultilabelF1Score(num_labels=L, validate_args=False)]).to(dev)
        for _ in range(2):
            mcoll.update(mp, mt)
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 5
        start_clock(c0)
        for _ in range(reps):
            mcoll.update(mp, mt)
        c1.record()
        barrier()
        cms = max_over_ranks(c0.elapsed_time(c1)) / reps
        mcoll.compute()  # warm-up (lazy kernel loading of the small torch reducers)
        ctimes = []
        for _ in range(5):
            mcoll.update(mp, mt)
            barrier()  # ranks aligned: the number below is compute + sync, not waiting for a slower rank's update
            t0 = time.perf_counter()
            res = mcoll.compute()
            torch.cuda.synchronize()
            ctimes.append((time.perf_counter() - t0) * 1e3)
        compute_ms = max_over_ranks(statistics.median(ctimes))
        extra["multilabel_collection_L100"] = {
            "workload": "configs[4]: MetricCollection[MultilabelStatScores, Precision, Recall, F1] num_labels=100, "
                        "2^23-row batches per rank (weak), one compute group => one K2 pass per update",
            "value": rows * world / (cms * 1e-3), "unit": "samples/s", "ms_per_update": cms,
            "achieved_GBps_per_gpu": rows * L * 12 / (cms * 1e-3) / 1e9,
            "frac_of_measured_peak": rows * L * 12 / (cms * 1e-3) / 1e9 / peak,
            "compute_ms_incl_sync": compute_ms, "compute_groups": len(mcoll.compute_groups),
            "f1_macro": float(res["MultilabelF1Score"]),
        }
        del mp, mt, mcoll
        torch.cuda.empty_cache()

        # configs[3]: BinaryAUROC(thresholds=None), 2^26 scores in total; cat states gathered, sort + scan on every rank
        n_exact = (1 << 26) // world
        g4 = torch.Generator(device=dev).manual_seed(11 + rank)
        ep = torch.rand(n_exact, device=dev, generator=g4)
        et = (torch.rand(n_exact, device=dev, generator=g4) < 0.3 + 0.4 * ep).long()
        em = tm.BinaryAUROC(thresholds=None, validate_args=False).to(dev)
        em.update(ep, et)  # full-size warm-up: the caching allocator has to grow by ~2 GB the first time
        em.compute()
        em.reset()
        barrier()
        t0 = time.perf_counter()
        em.update(ep, et)
        torch.cuda.synchronize()
        upd_ms = (time.perf_counter() - t0) * 1e3
        barrier()
        t0 = time.perf_counter()
        val = float(em.compute())
        torch.cuda.synchronize()
        cmp_ms = max_over_ranks((time.perf_counter() - t0) * 1e3)
        extra["exact_auroc_2p26"] = {
            "workload": "configs[3]: BinaryAUROC(thresholds=None), 2^26 fp32 scores in total, cat-state all_gather "
                        "then radix sort + scan (K4) on every rank",
            "update_ms": upd_ms, "compute_ms_incl_gather": cmp_ms, "value": (1 << 26) / (cmp_ms * 1e-3),
            "unit": "samples/s (compute)", "auroc": val,
        }
        if world > 1:
            # the same scores without replication: range partition + local sort (distributed_exact="partition")
            pm = tm.BinaryAUROC(thresholds=None, validate_args=False, distributed_exact="partition").to(dev)
            pm.update(ep, et)
            pm.compute()
            ptimes = []
            for _ in range(3):
                pm._computed = None
                barrier()
                t0 = time.perf_counter()
                pval = float(pm.compute())
                torch.cuda.synchronize()
                ptimes.append((time.perf_counter() - t0) * 1e3)
            pms = max_over_ranks(statistics.median(ptimes))
            assert abs(pval - val) < 1e-6, (pval, val)
            extra["exact_auroc_2p26_partitioned"] = {
                "workload": "configs[3] with distributed_exact='partition': splitters from a sample, all-to-all range "
                            "partition (5 B/sample), local sort + reduce (K4b with offsets), scalar all_reduce",
                "compute_ms": pms, "value": (1 << 26) / (pms * 1e-3), "unit": "samples/s (compute)", "auroc": pval,
            }
            del pm
        del ep, et, em

    # ------------------------------------------------------------------ the reference's own torch ops on this GPU
    if not args.skip_extra and world == 1:
        try:
            extra["reference_on_same_gpu"] = reference_on_gpu(dev)
        except Exception as err:  # a comparator, never a reason to lose the bench line
            extra["reference_on_same_gpu"] = {"unavailable": f"{type(err).__name__}: {err}"[:200]}

    # ------------------------------------------------------------------ CPU baseline (rank 0, N=1 only)
    cpu_baseline = None
    if rank == 0 and world == 1:
        cpu_baseline = cpu_baseline_sample()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "int64", "data": "synthetic",
            "config": dict(CONFIG),
            "l2_policy": "inputs larger than L2: 268 MB read per step, two alternating batches (L2 126 MB)",
            "api": "torchmetrics_b200.MulticlassConfusionMatrix.update",
            "timing": "CUDA events on the launch stream around the K steps, max over ranks; the clock starts behind a "
                      "0.3 ms device-side delay so that the host's latency to the first launch is not counted",
            "clocks": clocks, "e2e": e2e, "gpu_launches": gpu_launches, "roofline": roofline,
            "cpu_baseline": cpu_baseline, "extra": extra,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
