#!/usr/bin/env python
"""linear_tc (tcgen05 3xTF32) vs the stock fp32 GEMM on the S3DIS tower's point-level shapes: time and
max relative error against float64.  Development aid."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 133399
g = torch.Generator(device="cuda").manual_seed(0)
shapes = [(M, 643, 256, False), (M, 256, 643, True), (M, 131, 512, False), (M, 512, 131, True), (M, 67, 128, False),
          (M, 35, 128, False), (M, 387, 13, False), (M, 64, 134, True), (23821, 518, 512, False), (23821, 96, 192, False),
          (300, 32, 32, False), (129, 8, 40, False)]
for (m, K, N, tr) in shapes:
    x = torch.randn(m, K, device="cuda", generator=g)
    w = torch.randn(K, N, device="cuda", generator=g).t() if tr else torch.randn(N, K, device="cuda", generator=g)
    b = torch.randn(N, device="cuda", generator=g)
    add = torch.randn(m, N, device="cuda", generator=g)
    y = linear_tc(x, w, b, add)
    torch.cuda.synchronize()
    rows = slice(0, min(m, 2048))
    ref = x[rows].double() @ w.double().t() + b.double() + add[rows].double()
    tail = slice(max(0, m - 300), m)
    ref_t = x[tail].double() @ w.double().t() + b.double() + add[tail].double()
    err = float(((y[rows].double() - ref).abs().max() / ref.abs().max()).item())
    err_t = float(((y[tail].double() - ref_t).abs().max() / ref_t.abs().max()).item())
    ys = torch.addmm(add, x, w.t()) + b
    err_s = float(((ys[rows].double() - ref).abs().max() / ref.abs().max()).item())

    def timeit(f, n=10):
        for _ in range(3):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            f()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n * 1e3

    t_tc = timeit(lambda: linear_tc(x, w, b))
    t_st = timeit(lambda: torch.nn.functional.linear(x, w, b))
    fl = 2.0 * m * K * N
    print(f"M={m:6d} K={K:4d} N={N:4d} tr={int(tr)}  tc {t_tc:8.1f} us ({fl / t_tc / 1e6:7.1f} TF/s)  stock {t_st:8.1f} us  "
          f"err tc {err:.2e} (tail {err_t:.2e})  stock {err_s:.2e}", flush=True)
