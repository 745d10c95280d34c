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
        if hasattr(_lib._clib, "pcs_debug_mlp_const_weights"):
            _lib._clib.pcs_debug_mlp_const_weights(mode)
        with torch.no_grad():
            ms = T(lambda: fused_ops.dense_mlp3(u, w2, w3, wo))
            ys.append(fused_ops.dense_mlp3(u, w2, w3, wo))
        out[mode] = ms
    err = float((ys[0] - ys[1]).abs().max())
    print(f"cfg ({A},{B},{C},{O}): " + "  ".join(f"mode{m} {1e3 * v:.1f} us {2 * macs / v / 1e9:.1f} TFLOP/s" for m, v in out.items()) + f"  maxdiff {err:.2e}", flush=True)
