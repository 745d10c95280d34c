#!/usr/bin/env python
"""A/B timing of dense_mlp3 forward variants (development aid)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from bench_ops import Timer  # noqa: E402
from pointcloudsegmentation_b200 import _lib, fused_ops  # noqa: E402

T = Timer(10)
rows = 1578346
for A, B, C, O in [(8, 8, 16, 32), (16, 16, 16, 48), (16, 16, 32, 64)]:
    g = torch.Generator(device="cuda").manual_seed(0)
    r = lambda *s: torch.randn(*s, device="cuda", generator=g)  # noqa: E731
    u, w2, w3, wo = r(rows, A + B + C + O), r(B, A) / 3, r(C, A + B) / 4, r(O, A + B + C) / 6
    macs = rows * (B * A + C * (B + A) + O * (A + B + C))
    out = {}
    ys = []
    for mode in (0, 1):
        _lib.lib.pcs_set_option(b"mlp_tensor_cores", mode)
        with torch.no_grad():
            ms = T(lambda: fused_ops.dense_mlp3(u, w2, w3, wo))
            ys.append(fused_ops.dense_mlp3(u, w2, w3, wo))
        out[mode] = ms
    ref = torch.relu(u[:, :A].double())
    h2 = torch.relu(u[:, A:A + B].double() + ref @ w2.double().t())
    h3 = torch.relu(u[:, A + B:A + B + C].double() + torch.cat([h2, ref], 1) @ w3.double().t())
    yr = u[:, A + B + C:].double() + torch.cat([h3, h2, ref], 1) @ wo.double().t()
    err = [float((y.double() - yr).abs().max() / yr.abs().max()) for y in ys]
    print(f"cfg ({A},{B},{C},{O}): " + "  ".join(f"mode{m} {1e3 * v:.1f} us {2 * macs / v / 1e9:.1f} TFLOP/s" for m, v in out.items()) + f"  rel err vs f64: mode0 {err[0]:.2e} mode1 {err[1]:.2e}", flush=True)
