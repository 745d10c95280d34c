#!/usr/bin/env python
"""One pcs_linear_tc_wgrad launch per shape (for an ncu source-level capture).  usage: run_wgrad_one.py [M N K]..."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc_wgrad  # noqa: E402

a = [int(v) for v in sys.argv[1:]]
shapes = [tuple(a[i:i + 3]) for i in range(0, len(a), 3)] or [(516607, 256, 644), (516607, 128, 32)]
for (M, N, K) in shapes:
    g = torch.randn(M, N, device="cuda")
    x = torch.randn(M, K, device="cuda")
    linear_tc_wgrad(g, x, False)
    torch.cuda.synchronize()
