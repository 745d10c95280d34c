#!/usr/bin/env python
"""linear_tc_wgrad (tcgen05 3xTF32, MN-major operands) vs the stock paths on the tower's shapes."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import column_sum, linear_tc_wgrad, tn_tall  # noqa: E402

M0 = int(sys.argv[1]) if len(sys.argv) > 1 else 133399
g_ = torch.Generator(device="cuda").manual_seed(0)
shapes = [(300, 32, 32), (M0, 128, 32), (M0, 256, 643), (M0, 512, 131), (M0, 128, 67), (M0, 128, 35), (M0, 64, 134),
          (23821, 512, 518), (23821, 192, 96), (5000, 13, 387), (M0, 256, 640)]
for (m, N, K) in shapes:
    g = torch.randn(m, N, device="cuda", generator=g_)
    x = torch.randn(m, K, device="cuda", generator=g_)
    dw, db = linear_tc_wgrad(g, x, True)
    torch.cuda.synchronize()
    ref = g.double().t() @ x.double()
    refb = g.double().sum(0)
    err = float((dw.double() - ref).abs().max() / ref.abs().max())
    errb = float((db.double() - refb).abs().max() / refb.abs().max())
    st = tn_tall(g, x, 2048)
    err_s = float((st.double() - ref).abs().max() / ref.abs().max())

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

    t_tc = timeit(lambda: linear_tc_wgrad(g, x, True))
    t_st = timeit(lambda: (tn_tall(g, x, 2048), column_sum(g)))
    t_mm = timeit(lambda: g.t() @ x)
    fl = 2.0 * m * K * N
    print(f"M={m:6d} N={N:4d} K={K:4d}  tc {t_tc:8.1f} us ({fl / t_tc / 1e6:7.1f} TF/s)  tn_tall+colsum {t_st:8.1f} us  mm {t_mm:8.1f} us  "
          f"err dW {err:.2e} db {errb:.2e}  stock {err_s:.2e}", flush=True)
