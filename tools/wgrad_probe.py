#!/usr/bin/env python
"""pcs_linear_tc_wgrad on one shape, plain and with every UMMA issued three times (with_bias >> 8 = 1; wrong results): the
difference / 2 is the tensor-pipe time of the launch.  usage: wgrad_probe.py M N K"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200._lib import lib, check  # noqa: E402
from pointcloudsegmentation_b200.fused_ops import _stream  # noqa: E402

M, N, K = (int(a) for a in sys.argv[1:4])
g = torch.randn(M, N, device="cuda")
x = torch.randn(M, K, device="cuda")
S = int(lib.pcs_linear_tc_wgrad_partials(M, N, K, 0))
part = torch.empty((S, N, K), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def run(wb):
    check(lib.pcs_linear_tc_wgrad(g.data_ptr(), M, N, N, x.data_ptr(), K, K, wb, part.data_ptr(), _stream()), "wgrad")


def t(wb, n=10):
    for _ in range(3):
        run(wb)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        flush.zero_()
        a.record()
        run(wb)
        b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[n // 2] * 1e3


t0, t3 = t(0), t(1 << 8)
print(f"M={M} N={N} K={K} splits={S}: plain {t0:8.1f} us ({2 * M * N * K / t0 / 1e6:6.1f} TFLOP/s fp32-equivalent), 3x UMMA {t3:8.1f} us -> tensor time ~{(t3 - t0) / 2:8.1f} us")
