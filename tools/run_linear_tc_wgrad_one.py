#!/usr/bin/env python
"""Cost of one MN-major TF32 UMMA in linear_tc_wgrad: time with every UMMA issued 1x and 3x (development aid)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200 import fused_ops  # noqa: E402
from pointcloudsegmentation_b200._lib import check, lib  # noqa: E402

for (M, N, K) in [(133399, 256, 640), (133399, 512, 128)]:
    g = torch.randn(M, N, device="cuda"); x = torch.randn(M, K, device="cuda")
    S = int(lib.pcs_linear_tc_wgrad_partials(M, N, K, 0))
    part = torch.empty((S, N, K), device="cuda")
    for flag, name in [(0, "1x MMAs"), (1 << 8, "3x MMAs")]:
        f = lambda: check(lib.pcs_linear_tc_wgrad(g.data_ptr(), M, N, N, x.data_ptr(), K, K, flag, part.data_ptr(), fused_ops._stream()), "w")  # noqa: E731
        for _ in range(2):
            f()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            f()
        e1.record()
        torch.cuda.synchronize()
        print(f"M={M} N={N} K={K} S={S} {name}: {e0.elapsed_time(e1) / 5 * 1e3:.1f} us", flush=True)
