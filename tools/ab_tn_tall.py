#!/usr/bin/env python
"""A/B: weight-gradient GEMM g^T x over M = 133399 rows, stock vs fused_ops.tn_tall slabs; column_sum vs sum(0)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import column_sum, tn_tall  # noqa: E402


def t(f, n=10):
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


M = 133399
for o, i in ((256, 643), (512, 131), (13, 387), (128, 67), (8, 134), (64, 134)):
    g, x = torch.randn(M, o, device="cuda"), torch.randn(M, i, device="cuda")
    print(f"dW [{o},{M}]x[{M},{i}]: stock {t(lambda: g.t() @ x):7.1f} us   tn_tall 2048 {t(lambda: tn_tall(g, x, 2048)):7.1f} us"
          f"   4096 {t(lambda: tn_tall(g, x, 4096)):7.1f} us   1024 {t(lambda: tn_tall(g, x, 1024)):7.1f} us")
for r, c in ((133399, 128), (133399, 256), (133399, 512), (23821, 192), (592, 768), (592, 288), (133399, 13), (1600000, 64), (133399, 35)):
    x = torch.randn(r, c, device="cuda")
    a, b = column_sum(x), x.double().sum(0)
    err = float((a.double() - b).abs().max() / b.abs().max())
    us = t(lambda: column_sum(x))
    print(f"sum [{r},{c}]: stock {t(lambda: x.sum(0)):7.1f} us   column_sum {us:7.1f} us = {r*c*4/us/1e3:6.0f} GB/s   rel err {err:.1e}")
