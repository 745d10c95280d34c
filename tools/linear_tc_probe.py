#!/usr/bin/env python
"""Times pcs_linear_tc on given shapes with parts of the kernel switched off (the _dbg bits of csrc/linear_tc.cu:
4 = no output stores / addend, 8 = no A loads, 64 = no epilogue chunks, 128 = sums but no STG, 256 = STG into 1024 rows,
512 = 16-byte register stores where the TMA flavour would run, 1024 = one TMA staging tile per warp), to see which role bounds a layer.
usage: linear_tc_probe.py M K N [ldx]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200 import fused_ops as F  # noqa: E402

M, K, N = (int(a) for a in sys.argv[1:4])
ldx = int(sys.argv[4]) if len(sys.argv) > 4 else K
torch.manual_seed(0)
xb = torch.randn(M, ldx, device="cuda")
x = xb[:, :K]
w = torch.randn(N, K, device="cuda") * 0.05
add = torch.randn(M, N, device="cuda")
y = torch.empty(M, N, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def t(fn, n=10):
    for _ in range(3):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[n // 2] * 1e3


for name, kw in (("plain", {}), ("plain, 16-byte stores instead of TMA", dict(_dbg=512)), ("plain, one TMA staging tile", dict(_dbg=1024)),
                 ("bias+relu", dict(relu=True, bias=torch.zeros(N, device="cuda"))), ("no proxy fence in the producers (wrong results)", dict(_dbg=16)), ("addend", dict(addend=add)), ("addend+relu+bias", dict(addend=add, relu=True, bias=torch.zeros(N, device="cuda"))),
                 ("no stores", dict(_dbg=4)), ("addend, no A loads", dict(addend=add, _dbg=8)), ("no A loads", dict(_dbg=8)),
                 ("sums but no STG", dict(_dbg=128)), ("STG to 1024 rows (L2)", dict(_dbg=256)), ("no epilogue", dict(_dbg=64)), ("no A loads, no epilogue", dict(_dbg=72))):
    us = t(lambda: F.linear_tc(x, w, out=y, **kw))
    by = 4 * (M * K + M * N * (2 if "addend" in kw else 1))
    print(f"M={M} K={K} N={N} ldx={ldx} {name:26s} {us:8.1f} us  {by / us / 1e3:7.1f} GB/s  {2 * M * K * N / us / 1e6:6.1f} TFLOP/s")
