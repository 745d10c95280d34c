#!/usr/bin/env python
"""Two narrow linear_tc launches for an ncu source-level capture: [516607,32] -> 96 (16-byte epilogue) and -> 99
(4-byte epilogue)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(0)
for (m, K, N) in [(516607, 32, 96), (516607, 32, 99)]:
    x = torch.randn(m, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g)
    y = torch.empty(m, N, device="cuda")
    linear_tc(x, w, None, out=y)
    torch.cuda.synchronize()
