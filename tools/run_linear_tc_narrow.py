#!/usr/bin/env python
"""A few linear_tc launches for an ncu source-level capture (epilogue-bound shapes).  usage: run_linear_tc_narrow.py [M K N]..."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc  # noqa: E402

a = [int(v) for v in sys.argv[1:]]
shapes = [tuple(a[i:i + 3]) for i in range(0, len(a), 3)] or [(516607, 32, 96), (516607, 128, 512)]
g = torch.Generator(device="cuda").manual_seed(0)
for (m, K, N) in shapes:
    x = torch.randn(m, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g)
    y = torch.empty(m, N, device="cuda")
    linear_tc(x, w, None, out=y)
    torch.cuda.synchronize()
