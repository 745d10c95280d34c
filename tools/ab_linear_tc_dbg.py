#!/usr/bin/env python
"""Which resource bounds k_linear_tc?  Times the kernel with parts switched off (results are then wrong)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(0)
for (m, K, N) in [(133399, 640, 256), (133399, 128, 512)]:
    x = torch.randn(m, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g)
    y = torch.empty(m, N, device="cuda")
    for dbg, name in [(0, "full"), (1, "no B copy"), (4, "no stores"), (8, "no x loads"), (13, "only MMAs"), (64, "no epilogue")]:
        for _ in range(2):
            linear_tc(x, w, None, out=y, _dbg=dbg)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            linear_tc(x, w, None, out=y, _dbg=dbg)
        e1.record()
        torch.cuda.synchronize()
        print(f"M={m} K={K} N={N} {name:14s} {e0.elapsed_time(e1) / 5 * 1e3:8.1f} us", flush=True)
