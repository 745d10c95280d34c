#!/usr/bin/env python
"""One linear_tc launch per shape (for ncu)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200.fused_ops import linear_tc  # noqa: E402

g = torch.Generator(device="cuda").manual_seed(0)
for (m, K, N) in [(133399, 640, 256), (133399, 643, 256), (133399, 35, 128)]:
    x = torch.randn(m, K, device="cuda", generator=g)
    w = torch.randn(N, K, device="cuda", generator=g)
    y = torch.empty(m, N, device="cuda")
    for _ in range(2):
        linear_tc(x, w, None, out=y)
    torch.cuda.synchronize()
