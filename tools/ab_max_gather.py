#!/usr/bin/env python
"""neighbor_max_feat_gather on step-like shapes: the TMA-staged stream kernel (seg_stream=1) vs the row walker (0)."""
import sys
from pathlib import Path

import numpy as np
import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pointcloudsegmentation_b200 import ops  # noqa: E402
from pointcloudsegmentation_b200._lib import lib  # noqa: E402

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def t(fn, n=8):
    for _ in range(2):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for a, b in ev:
        flush.zero_()
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    return sorted(a.elapsed_time(b) for a, b in ev)[n // 2] * 1e3


rng = np.random.default_rng(0)
for pn, deg, C in ((94916, 33, 64), (94916, 25.7, 48), (94916, 19, 48), (94916, 14, 48), (516607, 22, 32), (516607, 10, 32)):
    lens = np.clip(rng.poisson(deg, pn), 1, None).astype(np.int32)
    begs = np.concatenate([[0], np.cumsum(lens)[:-1]]).astype(np.int32)
    cs = int(lens.sum())
    x = torch.randn(cs, C, device="cuda")
    L, Bg = torch.from_numpy(lens).cuda(), torch.from_numpy(begs).cuda()
    by = 4 * C * cs + 8 * C * pn
    out = []
    for mode in (1, 0):
        lib.pcs_set_option(b"seg_stream", mode)
        us = t(lambda: ops.neighbor_max_feat_gather(x, L, Bg))
        out.append(f"seg_stream={mode}: {us:7.1f} us {by / us / 1e3:6.0f} GB/s")
    lib.pcs_set_option(b"seg_stream", 1)
    print(f"pn={pn} deg={deg} C={C} csum={cs}: " + " | ".join(out))
