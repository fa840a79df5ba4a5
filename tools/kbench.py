"""Kernel micro-benchmark (development tool): device-resident inputs, CUDA-event timing, algorithmic GB/s.
Usage: python tools/kbench.py [confmat|binned|stat|all]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from torchmetrics_b200 import _native as N  # noqa: E402

DEV = torch.device("cuda:0")
PEAK = 6551.7
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass


def timeit(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for a, b in evs:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in evs)
    return ts[len(ts) // 2], ts[0]


def report(name, nbytes, med, best, nsamples):
    print(f"{name:58s} med {med*1e3:9.1f} us  best {best*1e3:9.1f} us  {nbytes/med/1e6:8.1f} GB/s "
          f"({nbytes/med/1e6/PEAK*100:5.1f}% of {PEAK:.0f})  {nsamples/med/1e6:9.2f} Gsamples/s", flush=True)


def bench_confmat():
    for C, n in [(1000, 1 << 24), (1000, 1 << 26), (1000, 1 << 22), (1000, 1 << 20), (150, 1 << 24), (100, 1 << 24), (10, 1 << 24), (5, 1 << 24)]:
        g = torch.Generator(device=DEV).manual_seed(1)
        p = torch.randint(0, C, (n,), device=DEV, generator=g)
        t = torch.randint(0, C, (n,), device=DEV, generator=g)
        cm = torch.zeros(C, C, dtype=torch.int64, device=DEV)
        algos = [1, 3, 6] if C * C * 4 > 200 * 1024 else [1, 2, 6]
        for algo in algos:
            med, best = timeit(lambda: N.confmat_update_(cm, p, t, C, algo=algo))
            report(f"confmat C={C} n=2^{n.bit_length()-1} i64 uniform algo={algo}", n * 16, med, best, n)
        if C == 1000 and n == 1 << 24:
            p2 = torch.where(torch.rand(n, device=DEV) < 0.8, t, p)
            for algo in [1, 6]:
                med, best = timeit(lambda: N.confmat_update_(cm, p2, t, C, algo=algo))
                report(f"confmat C={C} n=2^24 i64 diag-heavy algo={algo}", n * 16, med, best, n)
            p3, t3 = p.int(), t.int()
            for algo in [1, 6]:
                med, best = timeit(lambda: N.confmat_update_(cm, p3, t3, C, algo=algo))
                report(f"confmat C={C} n=2^24 i32 uniform algo={algo}", n * 8, med, best, n)
            t4 = torch.where(torch.rand(n, device=DEV) < 0.5, torch.zeros_like(t), t)
            p4 = torch.where(torch.rand(n, device=DEV) < 0.9, t4, p)
            for algo in [1, 6]:
                med, best = timeit(lambda: N.confmat_update_(cm, p4, t4, C, algo=algo), iters=5)
                report(f"confmat C={C} n=2^24 i64 half-in-one-cell algo={algo}", n * 16, med, best, n)
            # reference-on-cuda comparator (ATen): key materialisation + bincount
            med, best = timeit(lambda: torch.bincount(t * C + p, minlength=C * C), iters=5)
            report("torch: bincount(t*C+p) [reference's ops on cuda]", n * 16, med, best, n)
        del p, t


def bench_binned():
    thr = torch.linspace(0, 1, 200, device=DEV)
    for n in [1 << 24, 1 << 26, 1 << 28]:
        g = torch.Generator(device=DEV).manual_seed(1)
        preds = torch.rand(n, device=DEV, generator=g)
        target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
        st = torch.zeros(200, 2, 2, dtype=torch.int64, device=DEV)
        for mode, name in [(N.PROB_AS_IS, "as-is"), (N.PROB_AUTO, "auto"), (N.PROB_SIGMOID, "sigmoid")]:
            med, best = timeit(lambda: N.binned_curve_update_(st, preds, target, thr, prob_mode=mode), iters=10)
            report(f"binned T=200 n=2^{n.bit_length()-1} f32+i64 {name}", n * 12, med, best, n)
        if n == 1 << 26:
            t8 = target.to(torch.uint8)
            med, best = timeit(lambda: N.binned_curve_update_(st, preds, t8, thr, prob_mode=N.PROB_AS_IS), iters=10)
            report("binned T=200 n=2^26 f32+u8 as-is", n * 5, med, best, n)
            for T in [50, 427, 1000]:
                thr2 = torch.linspace(0, 1, T, device=DEV)
                st2 = torch.zeros(T, 2, 2, dtype=torch.int64, device=DEV)
                med, best = timeit(lambda: N.binned_curve_update_(st2, preds, target, thr2, prob_mode=N.PROB_AS_IS),
                                   iters=10)
                report(f"binned T={T} n=2^26 f32+i64 as-is", n * 12, med, best, n)
        del preds, target


def bench_stat():
    for n, L in [(1 << 20, 100), (1 << 23, 100), (1 << 26, 1)]:
        g = torch.Generator(device=DEV).manual_seed(1)
        p = torch.rand(n, L, device=DEV, generator=g)
        t = torch.randint(0, 2, (n, L), device=DEV, generator=g)
        out = torch.zeros(4, L, dtype=torch.int64, device=DEV)
        for mode, name in [(N.PROB_AS_IS, "as-is"), (N.PROB_AUTO, "auto")]:
            med, best = timeit(lambda: N.stat_scores_update_(out, p, t, L, 0.5, None, mode), iters=10)
            report(f"stat_scores L={L} n=2^{n.bit_length()-1} f32+i64 {name}", n * L * 12, med, best, n)
        del p, t


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    print(torch.cuda.get_device_name(0), "SMs", N.load().tmb_sm_count(), "peak", PEAK)
    if what in ("confmat", "all"):
        bench_confmat()
    if what in ("binned", "all"):
        bench_binned()
    if what in ("stat", "all"):
        bench_stat()


def bench_exact():
    for n in [1 << 20, 1 << 24, 1 << 26]:
        g = torch.Generator(device=DEV).manual_seed(1)
        preds = torch.rand(n, device=DEV, generator=g)
        target = (torch.rand(n, device=DEV, generator=g) < 0.3 + 0.4 * preds).long()
        med, best = timeit(lambda: N.exact_curve(preds, target, 1), iters=5)
        report(f"exact_curve n=2^{n.bit_length()-1} f32+i64 (sort+scan, incl. D sync)", n * 12, med, best, n)
        med, best = timeit(lambda: torch.sort(preds, descending=True), iters=5)
        report(f"torch.sort(preds) alone [CUB radix sort, keys+indices]", n * 12, med, best, n)
        import torchmetrics_b200.functional as F
        med, best = timeit(lambda: F.binary_auroc(preds, target, thresholds=None, validate_args=False), iters=3)
        report(f"binary_auroc(thresholds=None) n=2^{n.bit_length()-1} end to end", n * 12, med, best, n)
        del preds, target


if __name__ == "__main__" and (len(sys.argv) > 1 and sys.argv[1] == "exact"):
    bench_exact()


def bench_binned_multi():
    thr = torch.linspace(0, 1, 200, device=DEV)
    for rows, cols in [(1 << 18, 100), (1 << 20, 100), (1 << 20, 10), (1 << 23, 10), (1 << 22, 4), (1 << 24, 4)]:
        g = torch.Generator(device=DEV).manual_seed(1)
        preds = torch.rand(rows, cols, device=DEV, generator=g)
        tl = torch.randint(0, 2, (rows, cols), device=DEV, generator=g)
        st = torch.zeros(200, cols, 2, 2, dtype=torch.int64, device=DEV)
        med, best = timeit(lambda: N.binned_curve_update_(st, preds, tl, thr, num_columns=cols, prob_mode=N.PROB_AS_IS),
                           iters=10)
        report(f"binned multilabel T=200 rows=2^{rows.bit_length()-1} L={cols} f32+i64", rows * cols * 12, med, best,
               rows * cols)
        tc = torch.randint(0, cols, (rows,), device=DEV, generator=g)
        med, best = timeit(lambda: N.binned_curve_update_(st, preds, tc, thr, num_columns=cols,
                                                          target_is_class_index=True, prob_mode=N.PROB_AS_IS), iters=10)
        report(f"binned multiclass T=200 rows=2^{rows.bit_length()-1} C={cols} f32+i64", rows * (cols * 4 + 8), med,
               best, rows * cols)
        del preds, tl, tc


if __name__ == "__main__" and (len(sys.argv) > 1 and sys.argv[1] == "binned_multi"):
    bench_binned_multi()


def bench_class_stats():
    for C, n in [(1000, 1 << 24), (1000, 1 << 26), (10, 1 << 24), (100, 1 << 24), (20000, 1 << 24)]:
        g = torch.Generator(device=DEV).manual_seed(1)
        t = torch.randint(0, C, (n,), device=DEV, generator=g)
        p = torch.where(torch.rand(n, device=DEV, generator=g) < 0.7, t, torch.randint(0, C, (n,), device=DEV, generator=g))
        out = torch.zeros(4, C, dtype=torch.int64, device=DEV)
        med, best = timeit(lambda: N.class_stats_update_(out, p, t, C))
        report(f"class_stats C={C} n=2^{n.bit_length()-1} i64 (70% correct)", n * 16, med, best, n)
        del p, t


if __name__ == "__main__" and (len(sys.argv) > 1 and sys.argv[1] == "class_stats"):
    bench_class_stats()


def bench_class_stats_argmax():
    for C, n, dt in [(1000, 8192, torch.float32), (1000, 1 << 17, torch.float32), (1000, 1 << 17, torch.bfloat16),
                     (100, 1 << 20, torch.float32), (10, 1 << 22, torch.float32)]:
        g = torch.Generator(device=DEV).manual_seed(1)
        s = torch.randn(n, C, device=DEV, generator=g).to(dt)
        t = torch.randint(0, C, (n,), device=DEV, generator=g)
        out = torch.zeros(4, C, dtype=torch.int64, device=DEV)
        med, best = timeit(lambda: N.class_stats_argmax_update_(out, s, t, C))
        report(f"class_stats_argmax C={C} n={n} {str(dt)[6:]}", n * (C * s.element_size() + 8), med, best, n)
        med, best = timeit(lambda: s.argmax(dim=1))
        report("   torch.argmax(dim=1) alone", n * C * s.element_size(), med, best, n)
        del s, t


if __name__ == "__main__" and (len(sys.argv) > 1 and sys.argv[1] == "class_stats_argmax"):
    bench_class_stats_argmax()
